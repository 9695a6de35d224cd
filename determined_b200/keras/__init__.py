"""``determined.keras`` (TFKerasTrial) is out of the north-star scope ("list
only"): TensorFlow is not part of the B200 stack and the reference's Keras path is Horovod-only.
Importing any name from here explains that instead of failing obscurely."""


def __getattr__(name: str):
    raise ImportError(f"determined_b200.keras.{name}: TensorFlow/Keras trials are not supported by the B200 build "
                      "(PyTorchTrial, DeepSpeedTrial and the Core API are); see DESIGN.md")
