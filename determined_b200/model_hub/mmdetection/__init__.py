"""MMDetection adapter (reference ``model_hub/model_hub/mmdetection``)."""
from determined_b200.model_hub.mmdetection._callbacks import LrUpdaterCallback, build_lr_schedule  # noqa: F401
from determined_b200.model_hub.mmdetection._data import DatasetWithIndex, GroupSampler, build_dataloader  # noqa: F401
from determined_b200.model_hub.mmdetection._data_backends import FakeBackend, GCSBackend, S3Backend, sub_backend  # noqa: F401
from determined_b200.model_hub.mmdetection._trial import MMDetTrial, flatten_overrides  # noqa: F401
from determined_b200.model_hub.mmdetection.utils import (  # noqa: F401
    build_fp16_loss_scaler,
    get_config_pretrained_url_mapping,
    get_pretrained_ckpt_path,
)
