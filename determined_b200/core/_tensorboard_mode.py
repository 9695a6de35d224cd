import enum


class TensorboardMode(enum.Enum):
    """AUTO: the harness syncs tfevents after each metrics report / checkpoint; MANUAL: the caller
    calls ``core_context.train.upload_tensorboard_files()`` (reference ``core/_tensorboard_mode.py``)."""

    AUTO = "AUTO"
    MANUAL = "MANUAL"
