"""model_hub: ready-made trials over third-party model libraries (reference ``model_hub/model_hub``).
``huggingface`` is provided; the reference's ``mmdetection`` adapter needs the mmcv/mmdet stack,
which is not available here."""
from determined_b200.model_hub import utils  # noqa: F401
