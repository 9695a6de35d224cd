"""HF trials (reference ``model_hub/model_hub/huggingface/__init__.py``)."""
from determined_b200.model_hub.huggingface._config_parser import (ConfigKwargs, DatasetKwargs, LRSchedulerKwargs, ModelKwargs, OptimizerKwargs,
                                                                  TokenizerKwargs, default_parse_config_tokenizer_model_kwargs,
                                                                  default_parse_optimizer_lr_scheduler_kwargs, parse_dict_to_dataclasses)
from determined_b200.model_hub.huggingface._trial import (BaseTransformerTrial, build_default_lr_scheduler, build_default_optimizer, build_using_auto,
                                                          group_parameters_for_optimizer)
from determined_b200.model_hub.huggingface._utils import ColumnDataset, default_load_dataset, remove_unused_columns  # noqa: E402
