from determined_b200.transformers._hf_callback import DetCallback, get_metric_type  # noqa: F401
