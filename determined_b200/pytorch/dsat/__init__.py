"""Engine-config autotuning ("DSAT"; reference ``harness/determined/pytorch/dsat/__init__.py``)."""
from determined_b200.pytorch.dsat import defaults
from determined_b200.pytorch.dsat._utils import (
    dsat_reporting_context, get_base_parser, get_batch_config_from_mbs_gas_and_slots, get_dict_from_yaml_or_json_path,
    get_ds_config_from_hparams, get_full_parser, get_hf_args_with_overwrites, get_hf_ds_config_path_from_args,
    get_random_zero_optim_config, get_zero_stage_search_space, report_json_results, report_profile, smaller_is_better, update_hf_args)
from determined_b200.pytorch.dsat._dsat_search_method import (
    ASHADSATSearchData, ASHADSATSearchMethod, BaseDSATSearchMethod, BinarySearchDSATSearchMethod, DSATModelProfileInfoTrial,
    DSATSearchData, DSATTrial, DSATTrialTracker, RandomDSATSearchMethod, TestDSATSearchMethod)
from determined_b200.pytorch.dsat._run_dsat import get_custom_dsat_exp_conf_from_args, get_search_method_class
from determined_b200.pytorch.dsat.__main__ import get_search_runner_config_from_args
