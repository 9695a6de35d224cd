"""``BatchMetricWriter`` under the reference's module path (``tensorboard/metric_writers/callback.py``)."""
from determined_b200.tensorboard.metric_writers import BatchMetricWriter  # noqa: F401
