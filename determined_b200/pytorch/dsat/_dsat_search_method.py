"""Search methods of the engine-config autotuner ("DSAT").

Parity: reference ``harness/determined/pytorch/dsat/_dsat_search_method.py`` — ``DSATTrial`` / ``DSATModelProfileInfoTrial`` /
``DSATTrialTracker`` ``:24-520``, ``BaseDSATSearchMethod`` ``:523-735`` and the ``random`` / ``binary`` / ``asha`` / ``_test``
methods ``:738-1432``.  The class, method and attribute names below are the extension surface users subclass (and the
reference's simulation suite, ``tests/experiment/pytorch/test_deepspeed_autotuning.py``, drives — it runs against this module
through ``tests/test_reference_user_code.py``); the implementation is organised differently:

* a trial is a node of a *lineage* (a chain of trials that vary only the micro-batch size of one random engine config); the
  micro-batch search is one window-narrowing rule (``_narrow``) shared by all three methods (``_LineageSearch``) instead of
  three copies: success raises the floor, failure lowers the ceiling, an empty window ends the lineage;
* memory model and batch-size ceiling per ZeRO stage come from the one-step model-profile trial, exactly as the stages of
  ``determined_b200.parallel.zero`` partition state (``profiles/r2_zero_memory_n2.json`` measures the engine against it).
"""
import abc
import argparse
import collections
import copy
import dataclasses
import json
import logging
import pathlib
import pickle
import random
import uuid
from typing import Any, Deque, Dict, Iterable, Iterator, List, Optional, Set, Tuple, Union

import numpy as np

from determined_b200 import searcher, util
from determined_b200.pytorch.dsat import _utils, defaults

logger = logging.getLogger("determined.pytorch")

MBS_KEY = "train_micro_batch_size_per_gpu"
Metric = Union[float, Dict[str, Any]]


def _ds_config_for(hparams: Dict[str, Any], model_dir: str) -> Dict[str, Any]:
    """The engine config a trial with these hyperparameters will run: the JSON file it names (in the model directory, or
    next to it when it was shipped with ``--include``) overlaid with the overwrite section."""
    try:
        return _utils.get_ds_config_from_hparams(hparams, model_dir)
    except FileNotFoundError:
        return _utils.get_ds_config_from_hparams(hparams)


@dataclasses.dataclass
class DSATSearchData:
    """Micro-batch-size window ``[lo, hi]`` still open for a lineage."""

    lo: int
    hi: int


class ASHADSATSearchData(DSATSearchData):
    def __init__(self, curr_rung: int, *args: Any, **kwargs: Any) -> None:
        super().__init__(*args, **kwargs)
        self.curr_rung = curr_rung


class DSATTrial:
    """One probe run: hyperparameters (with the engine-config overwrites), result, and its place in a lineage."""

    def __init__(self, hparams: Dict[str, Any], model_dir: str, slots_per_trial: int, length: int,
                 request_id: Optional[uuid.UUID] = None, parent: Optional["DSATTrial"] = None,
                 search_data: Optional[DSATSearchData] = None, searcher_metric_name: Optional[str] = None) -> None:
        self.hparams = hparams
        self.model_dir = model_dir
        self.slots_per_trial = slots_per_trial
        self.length = length
        self.request_id = request_id or uuid.uuid4()
        self.parent = parent
        self.search_data = search_data
        self.searcher_metric_name = searcher_metric_name
        self.metric: Metric = {}
        self.error = False
        self.running = False
        self.children: Set["DSATTrial"] = set()
        if parent is not None:
            parent.children.add(self)
        self.lineage_root: "DSATTrial" = self if parent is None else parent.lineage_root
        self.ds_config = _ds_config_for(hparams, model_dir)

    # -- result ----------------------------------------------------------------------------------------------------------
    @property
    def completed(self) -> bool:
        return bool(self.error or self.metric)

    @property
    def searcher_metric_val(self) -> Optional[float]:
        if self.searcher_metric_name is None:
            return None
        if isinstance(self.metric, float):
            return self.metric
        v = self.metric.get(self.searcher_metric_name)
        return None if v is None else float(v)

    # -- lineage ---------------------------------------------------------------------------------------------------------
    @property
    def lineage_set(self) -> Set["DSATTrial"]:
        """Every trial of the tree this trial belongs to."""
        seen: Set["DSATTrial"] = set()
        todo = [self.lineage_root]
        while todo:
            t = todo.pop()
            if t not in seen:
                seen.add(t)
                todo.extend(t.children)
        return seen

    @property
    def num_completed_trials_in_lineage(self) -> int:
        return sum(1 for t in self.lineage_set if t.completed)

    @property
    def error_in_direct_history(self) -> bool:
        """Did this trial or one of its ancestors fail (so the window's ceiling is a proven one)?"""
        t: Optional["DSATTrial"] = self
        while t is not None:
            if t.error:
                return True
            t = t.parent
        return False

    @property
    def mbs_in_lineage(self) -> Set[int]:
        return {t.mbs for t in self.lineage_set}

    # -- what it runs ----------------------------------------------------------------------------------------------------
    @property
    def stage(self) -> int:
        return int((self.ds_config.get("zero_optimization") or {}).get("stage", 0))

    @property
    def fp16(self) -> bool:
        return bool((self.ds_config.get("fp16") or {}).get("enabled"))

    @property
    def mbs(self) -> int:
        v = self.ds_config.get(MBS_KEY)
        assert isinstance(v, int), f"the engine config of a DSATTrial needs an integer `{MBS_KEY}`, got {v!r}"
        return v

    @property
    def create_and_val_ops(self) -> List[searcher.Operation]:
        return [searcher.Create(request_id=self.request_id, hparams=self.hparams, checkpoint=None),
                searcher.ValidateAfter(request_id=self.request_id, length=self.length)]

    def __repr__(self) -> str:
        state = "error" if self.error else "done" if self.completed else "running" if self.running else "new"
        return f"<{type(self).__name__} {str(self.request_id)[:8]} stage={self.stage} mbs={self.ds_config.get(MBS_KEY)} {state}>"


class DSATModelProfileInfoTrial(DSATTrial):
    """The one-step run that reports parameter counts, activation memory and device memory."""


class DSATTrialTracker:
    """All trials of one autotuning job, the queue of those not started yet, and the global limits."""

    def __init__(self, args: argparse.Namespace, exp_config: Dict[str, Any]) -> None:
        self.exp_config = exp_config
        self.max_trials: int = args.max_trials
        self.max_concurrent_trials: int = args.max_concurrent_trials
        self.max_slots: Optional[int] = args.max_slots
        self.model_dir: str = args.model_dir
        self.searcher_metric: str = args.metric
        self.start_profile_step: int = args.start_profile_step
        self.end_profile_step: int = args.end_profile_step
        self.zero_stages: Set[int] = set(args.zero_stages)
        self.slots_per_trial: int = exp_config["resources"]["slots_per_trial"]
        self.hparams: Dict[str, Any] = exp_config["hyperparameters"]
        self.smaller_is_better = _utils.smaller_is_better(self.searcher_metric)

        self.model_profile_info_trial: Optional[DSATTrial] = None
        self.num_trials_since_best_result = 0
        self.successful_stages: Set[int] = set()
        self.queue: Deque[DSATTrial] = collections.deque()
        self._trials: Dict[uuid.UUID, DSATTrial] = {}
        self._stage_mem: Optional[Dict[int, int]] = None
        self._stage_max_mbs: Optional[Dict[int, int]] = None

    # -- container -------------------------------------------------------------------------------------------------------
    def __len__(self) -> int:
        return len(self._trials)

    def __getitem__(self, request_id: uuid.UUID) -> DSATTrial:
        return self._trials[request_id]

    def __iter__(self) -> Iterator[Tuple[uuid.UUID, DSATTrial]]:
        return iter(self._trials.items())

    def __contains__(self, item: Union[uuid.UUID, DSATTrial]) -> bool:
        if isinstance(item, uuid.UUID):
            return item in self._trials
        if isinstance(item, DSATTrial):
            return self._trials.get(item.request_id) is item
        raise ValueError(f"expected a uuid.UUID or a DSATTrial, got {type(item).__name__}")

    # -- making trials ---------------------------------------------------------------------------------------------------
    def enforce_consistent_batch_config(self, hparams: Dict[str, Any]) -> None:
        """``train_batch_size = mbs x gas x slots`` must hold in what the engine reads: rewrite the overwrites in place."""
        batch = _utils.get_batch_config_from_mbs_gas_and_slots(_ds_config_for(hparams, self.model_dir), slots=self.slots_per_trial)
        hparams[defaults.OVERWRITE_KEY] = util.merge_dicts(hparams.get(defaults.OVERWRITE_KEY, {}), batch)

    def create_trial(self, hparams: Dict[str, Any], search_data: Optional[Any] = None, parent_trial: Optional[DSATTrial] = None) -> DSATTrial:
        self.enforce_consistent_batch_config(hparams)
        # the engine's probe reports (and exits) when its step counter reaches end_profile_step; one step of slack
        return DSATTrial(hparams=hparams, model_dir=self.model_dir, slots_per_trial=self.slots_per_trial, length=self.end_profile_step + 1,
                         parent=parent_trial, search_data=search_data, searcher_metric_name=self.searcher_metric)

    def create_model_profile_info_trial(self) -> DSATModelProfileInfoTrial:
        hp = copy.deepcopy(self.hparams)
        hp[defaults.OVERWRITE_KEY] = util.merge_dicts(hp.get(defaults.OVERWRITE_KEY, {}), defaults.MODEL_INFO_PROFILE_DS_CONFIG)
        self.enforce_consistent_batch_config(hp)
        t = DSATModelProfileInfoTrial(hparams=hp, model_dir=self.model_dir, slots_per_trial=self.slots_per_trial, length=1)
        self.model_profile_info_trial = t
        return t

    def queue_and_register_trial(self, trial: DSATTrial) -> None:
        twin = next((t for t in self.completed_trials if t.hparams == trial.hparams), None)
        if twin is not None:
            logger.warning("queueing a trial with the hyperparameters of the finished trial %s", twin.request_id)
        self._trials[trial.request_id] = trial
        self.queue.append(trial)

    # -- results ---------------------------------------------------------------------------------------------------------
    def update_trial_metric(self, trial: DSATTrial, metric: Metric) -> None:
        trial.metric = metric
        trial.running = False
        if isinstance(trial, DSATModelProfileInfoTrial):
            return                       # carries no searcher metric: not part of the early-stopping count
        self.successful_stages.add(trial.stage)
        self.num_trials_since_best_result = 0 if self.best_trial is trial else self.num_trials_since_best_result + 1

    def report_trial_early_exit(self, trial: DSATTrial) -> None:
        if self.best_trial is not None:      # the count only starts once there is a result to beat
            self.num_trials_since_best_result += 1
        trial.error = True
        trial.running = False

    def _best_of(self, trials: Iterable[DSATTrial]) -> Optional[DSATTrial]:
        scored = [t for t in trials if not isinstance(t, DSATModelProfileInfoTrial) and isinstance(t.metric, dict) and self.searcher_metric in t.metric]
        if not scored:
            return None
        pick = min if self.smaller_is_better else max
        return pick(scored, key=lambda t: float(t.metric[self.searcher_metric]))      # type: ignore[index]

    @property
    def best_trials_by_stage(self) -> Dict[int, Optional[DSATTrial]]:
        return {s: self._best_of(t for _, t in self if t.stage == s) for s in range(4)}

    @property
    def best_trial(self) -> Optional[DSATTrial]:
        return self._best_of(t for t in self.best_trials_by_stage.values() if t is not None)

    # -- counts and limits -----------------------------------------------------------------------------------------------
    @property
    def running_trials(self) -> List[DSATTrial]:
        return [t for _, t in self if t.running]

    @property
    def completed_trials(self) -> List[DSATTrial]:
        return [t for _, t in self if t.completed]

    @property
    def num_running_trials(self) -> int:
        return len(self.running_trials)

    @property
    def num_completed_trials(self) -> int:
        return len(self.completed_trials)

    @property
    def max_trials_queued(self) -> bool:
        return len(self.queue) >= self.max_trials

    @property
    def max_trials_are_running_or_closed(self) -> bool:
        return self.num_running_trials + self.num_completed_trials >= self.max_trials

    @property
    def should_be_failure(self) -> bool:
        """The job failed if the profile run failed, or if no tuned trial ever succeeded."""
        if self.model_profile_info_trial is not None and self.model_profile_info_trial.error:
            return True
        return all(t.error for _, t in self if t.completed and not isinstance(t, DSATModelProfileInfoTrial))

    @property
    def can_run_more_trials(self) -> bool:
        if not self.queue or self.max_trials_are_running_or_closed or self.num_running_trials >= self.max_concurrent_trials:
            return False
        if self.max_slots is not None:
            return (self.max_slots - self.num_running_trials * self.slots_per_trial) // self.slots_per_trial > 0
        return True

    # -- what the profile run measured -------------------------------------------------------------------------------------
    def _profiled(self, key: str) -> int:
        t = self.model_profile_info_trial
        assert t is not None, f"the model-profile trial must run before `{key}` is known"
        assert isinstance(t.metric, dict) and key in t.metric, f"the model-profile trial reported no `{key}`"
        assert isinstance(t.metric[key], int), f"`{key}` of the model-profile trial must be an int"
        return int(t.metric[key])

    @property
    def gpu_mem(self) -> int:
        return self._profiled("gpu_mem")

    @property
    def num_params(self) -> int:
        return self._profiled("num_params")

    @property
    def trainable_num_params(self) -> int:
        return self._profiled("trainable_num_params")

    @property
    def activation_mem_per_gpu(self) -> int:
        return self._profiled("activation_mem_per_gpu")

    @property
    def mem_per_gpu_per_stage(self) -> Dict[int, int]:
        """Bytes per GPU at micro-batch 1, per ZeRO stage: weights, fp32 gradients and Adam state (plus fp32 master
        weights under 16-bit training), each divided by the data-parallel size from the stage that partitions it."""
        if self._stage_mem is None:
            assert self.model_profile_info_trial is not None
            low = self.model_profile_info_trial.fp16
            n, nt, w = self.num_params, self.trainable_num_params, self.slots_per_trial
            weights = n * (2 if low else 4)
            grads = nt * 4
            optim = nt * ((4 if low else 0) + 4 + 4)
            static = {0: weights + grads + optim, 1: weights + grads + optim // w, 2: weights + (grads + optim) // w, 3: (weights + grads + optim) // w}
            self._stage_mem = {s: m + self.activation_mem_per_gpu for s, m in static.items()}
        return self._stage_mem

    @property
    def approx_max_mbs_per_stage(self) -> Dict[int, int]:
        """How many more micro-batch-1 activations fit next to the static state (at least 1)."""
        if self._stage_max_mbs is None:
            self._stage_max_mbs = {s: max((self.gpu_mem - m) // self.activation_mem_per_gpu, 1) for s, m in self.mem_per_gpu_per_stage.items()}
        return self._stage_max_mbs


class BaseDSATSearchMethod(searcher.SearchMethod):
    """Common event handling: the profile run first, every probe closed after its validation, the global limits
    (``max_trials``, ``max_concurrent_trials``, ``max_slots``) enforced when trials are taken off the queue.  A concrete
    method says which trials follow a success (``get_trials_after_validation_completed``) and a failure
    (``get_trials_after_early_exit``); it may also reorder the queue (``choose_next_trial_from_queue``) and stop early
    (``early_stopping_triggered``)."""

    def __init__(self, args: argparse.Namespace, exp_config: Dict[str, Any]) -> None:
        self.args = args
        self.exp_config = exp_config
        self.trial_tracker = DSATTrialTracker(args=args, exp_config=exp_config)
        self.rng = np.random.default_rng(seed=args.random_seed)
        random.seed(args.random_seed)

    # -- for subclasses ----------------------------------------------------------------------------------------------------
    @abc.abstractmethod
    def get_trials_after_validation_completed(self, searcher_state: searcher.SearcherState, last_trial: DSATTrial, metric: Metric) -> Iterable[DSATTrial]:
        """Trials to append to the queue, in order."""

    @abc.abstractmethod
    def get_trials_after_early_exit(self, searcher_state: searcher.SearcherState, last_trial: DSATTrial,
                                    exited_reason: searcher.ExitedReason) -> Iterable[DSATTrial]:
        """Trials to append to the queue, in order."""

    def choose_next_trial_from_queue(self) -> DSATTrial:
        return self.trial_tracker.queue.popleft()

    def early_stopping_triggered(self) -> bool:
        return False

    # -- SearchMethod ------------------------------------------------------------------------------------------------------
    def initial_operations(self, searcher_state: searcher.SearcherState) -> List[searcher.Operation]:
        tr = self.trial_tracker
        probe = tr.create_model_profile_info_trial()
        tr.queue_and_register_trial(probe)
        tr.queue.popleft()                  # registered, and started right away
        return probe.create_and_val_ops

    def on_trial_created(self, searcher_state: searcher.SearcherState, request_id: uuid.UUID) -> List[searcher.Operation]:
        return []

    def _enqueue(self, trials: Iterable[DSATTrial]) -> None:
        for t in trials:
            self.trial_tracker.queue_and_register_trial(t)

    def on_validation_completed(self, searcher_state: searcher.SearcherState, request_id: uuid.UUID, metric: Metric,
                                train_length: int) -> List[searcher.Operation]:
        tr = self.trial_tracker
        last = tr[request_id]
        tr.update_trial_metric(trial=last, metric=metric)
        if isinstance(last, DSATModelProfileInfoTrial):
            logger.info("device memory %d B; per-stage memory at micro-batch 1: %s; micro-batch ceilings: %s",
                        tr.gpu_mem, tr.mem_per_gpu_per_stage, tr.approx_max_mbs_per_stage)
        if not tr.max_trials_queued and not self.should_shutdown():
            self._enqueue(self.get_trials_after_validation_completed(searcher_state=searcher_state, last_trial=last, metric=metric))
        return [searcher.Close(request_id)]

    def on_trial_exited_early(self, searcher_state: searcher.SearcherState, request_id: uuid.UUID,
                              exited_reason: searcher.ExitedReason) -> List[searcher.Operation]:
        tr = self.trial_tracker
        last = tr[request_id]
        tr.report_trial_early_exit(last)
        ops: List[searcher.Operation] = []
        if exited_reason != searcher.ExitedReason.ERRORED:
            # a crash (typically out of memory) is information; an invalid hyperparameter or a user's cancel ends the job
            logger.info("shutting down: trial %s exited early (%s)", request_id, exited_reason)
            ops.append(searcher.Shutdown(failure=tr.should_be_failure))
        if not tr.max_trials_queued and not self.should_shutdown():
            self._enqueue(self.get_trials_after_early_exit(searcher_state=searcher_state, last_trial=last, exited_reason=exited_reason))
        return ops

    def on_trial_closed(self, searcher_state: searcher.SearcherState, request_id: uuid.UUID) -> List[searcher.Operation]:
        tr = self.trial_tracker
        if self.should_shutdown():
            if tr.best_trial is not None and self.args.run_full_experiment:
                self._submit_best()
            return [searcher.Shutdown(failure=tr.should_be_failure)]
        ops: List[searcher.Operation] = []
        while tr.can_run_more_trials:
            nxt = self.choose_next_trial_from_queue()
            nxt.running = True
            ops += nxt.create_and_val_ops
        return ops

    def _submit_best(self) -> None:
        """``--run-full-experiment``: the user's experiment again, with the best engine config found and the probing switched off."""
        from determined_b200.experimental import client

        best = self.trial_tracker.best_trial
        assert best is not None
        conf = util.merge_dicts(_utils.get_dict_from_yaml_or_json_path(self.args.config_path), {"hyperparameters": best.hparams})
        conf["hyperparameters"][defaults.OVERWRITE_KEY].pop("autotuning", None)
        conf["hyperparameters"].pop(defaults.USE_DSAT_MODE_KEY, None)
        client.create_experiment(conf, self.args.model_dir, self.args.include)

    def progress(self, searcher_state: searcher.SearcherState) -> float:
        return len(searcher_state.trials_closed) / self.trial_tracker.max_trials

    def should_shutdown(self) -> bool:
        tr = self.trial_tracker
        if tr.model_profile_info_trial is not None and tr.model_profile_info_trial.error:
            logger.info("shutting down: the model-profile trial failed; the configuration must at least run with micro-batch size 1")
            return True
        if self.early_stopping_triggered():
            logger.info("shutting down: early stopping")
            return True
        if tr.num_completed_trials >= tr.max_trials:
            logger.info("shutting down: all trials done")
            return True
        return False

    # -- persistence -------------------------------------------------------------------------------------------------------
    def save_method_state(self, path: pathlib.Path) -> None:
        with (path / "trial_tracker.pkl").open("wb") as f:
            pickle.dump(self.trial_tracker, f)
        with (path / "py_random_state.pkl").open("wb") as f:
            pickle.dump(random.getstate(), f)
        with (path / "np_rng.pkl").open("wb") as f:
            pickle.dump(self.rng, f)
        best = self.trial_tracker.best_trial
        if best is not None:
            cfg = copy.deepcopy(best.ds_config)
            cfg.pop("autotuning", None)
            (path / "best_ds_config.json").write_text(json.dumps(cfg))
            (path / "best_ds_metrics.json").write_text(json.dumps(best.metric))

    def load_method_state(self, path: pathlib.Path) -> None:
        with (path / "trial_tracker.pkl").open("rb") as f:
            self.trial_tracker = pickle.load(f)
        with (path / "py_random_state.pkl").open("rb") as f:
            random.setstate(pickle.load(f))
        with (path / "np_rng.pkl").open("rb") as f:
            self.rng = pickle.load(f)


# ---------------------------------------------------------------------------------------------------------------------
# lineage searches: one random engine config per lineage, the micro-batch size searched inside a narrowing window
# ---------------------------------------------------------------------------------------------------------------------
def _narrow(trial: DSATTrial) -> Optional[DSATSearchData]:
    """The window a successor of ``trial`` searches: a success raises the floor above its micro-batch size, a failure
    lowers the ceiling below it.  (May come out empty, ``lo > hi``: what that means is the method's decision.)"""
    if trial.search_data is None:
        return None
    w = copy.deepcopy(trial.search_data)
    if trial.error or trial.searcher_metric_val is None:
        w.hi = trial.mbs - 1
    else:
        w.lo = trial.mbs + 1
    return w


class _LineageSearch(BaseDSATSearchMethod):
    search_range_factor: float = 1.0

    # -- new lineages --------------------------------------------------------------------------------------------------------
    def _population(self) -> int:
        """How many lineages to start with: as many as may run at once."""
        n = self.args.max_concurrent_trials
        if self.args.max_slots:
            n = min(n, self.args.max_slots // self.trial_tracker.slots_per_trial)
        return n

    def _stage_pool(self) -> List[int]:
        return sorted(self.trial_tracker.zero_stages)

    def _hparams_for_stage(self, zero_stage: int) -> Dict[str, Any]:
        hp = copy.deepcopy(self.trial_tracker.hparams)
        hp[defaults.OVERWRITE_KEY] = util.merge_dicts(hp.get(defaults.OVERWRITE_KEY, {}),
                                                      {"zero_optimization": _utils.get_random_zero_optim_config(zero_stage)})
        return hp

    def _first_window(self, zero_stage: int) -> DSATSearchData:
        """1 .. the model's estimate of the largest micro-batch that fits at this stage (times ``search_range_factor``)."""
        return DSATSearchData(lo=1, hi=max(int(self.search_range_factor * self.trial_tracker.approx_max_mbs_per_stage[zero_stage]), 1))

    def _pick_mbs(self, window: DSATSearchData) -> int:
        return (window.lo + window.hi) // 2

    def get_random_hparams_and_search_data(self, zero_stage: int) -> Tuple[Dict[str, Any], DSATSearchData]:
        hp = self._hparams_for_stage(zero_stage)
        window = self._first_window(zero_stage)
        hp[defaults.OVERWRITE_KEY][MBS_KEY] = self._pick_mbs(window)
        return hp, window

    def get_random_trial(self) -> DSATTrial:
        hp, window = self.get_random_hparams_and_search_data(random.choice(self._stage_pool()))
        return self.trial_tracker.create_trial(hparams=hp, search_data=window)

    def get_trial_list_after_model_profile_info_run(self) -> List[DSATTrial]:
        return [self.get_random_trial() for _ in range(self._population())]

    # -- continuing a lineage ------------------------------------------------------------------------------------------------
    def _successor(self, trial: DSATTrial, window: DSATSearchData) -> DSATTrial:
        hp = copy.deepcopy(trial.hparams)
        hp[defaults.OVERWRITE_KEY][MBS_KEY] = self._pick_mbs(window)
        return self.trial_tracker.create_trial(hparams=hp, search_data=window, parent_trial=trial)

    def _continue_or_restart(self, last_trial: DSATTrial) -> List[DSATTrial]:
        """Plain bisection: the next trial of the lineage, or a fresh random lineage when its window is exhausted."""
        w = _narrow(last_trial)
        if w is None or w.lo > w.hi:
            return [self.get_random_trial()]
        return [self._successor(last_trial, w)]

    def get_trials_after_validation_completed(self, searcher_state: searcher.SearcherState, last_trial: DSATTrial,
                                              metric: Optional[Metric] = None) -> List[DSATTrial]:
        if isinstance(last_trial, DSATModelProfileInfoTrial):
            return self.get_trial_list_after_model_profile_info_run()
        return self.get_trial_list_after_successful_run(last_trial)

    def get_trial_list_after_successful_run(self, last_trial: DSATTrial) -> List[DSATTrial]:
        return self._continue_or_restart(last_trial)

    def get_trials_after_early_exit(self, searcher_state: searcher.SearcherState, last_trial: DSATTrial,
                                    exited_reason: searcher.ExitedReason) -> List[DSATTrial]:
        return self._continue_or_restart(last_trial)


class BinarySearchDSATSearchMethod(_LineageSearch):
    """Random engine configs, as many as may run concurrently; the micro-batch size of each found by bisection between 1
    and the memory model's ceiling (``--search-range-factor`` scales the ceiling).  A lineage whose window closes is
    replaced by a new random one until ``max_trials`` is spent."""

    def __init__(self, *args: Any, **kwargs: Any) -> None:
        super().__init__(*args, **kwargs)
        self.search_range_factor = self.args.search_range_factor


class RandomDSATSearchMethod(_LineageSearch):
    """Random engine configs with a randomised bisection (the next micro-batch size is drawn binomially around the
    window's middle, which covers more distinct sizes across lineages than always taking the midpoint) and aggressive
    pruning: a lineage is dropped after ``--trials-per-random-config`` probes, when it failed at its floor, when a stage
    1 / 2 config already works and this one is stage 3, or when same-stage configs already run batches this lineage has
    proven it cannot reach.  ``--early-stopping N`` ends the job after N probes without a new best."""

    def __init__(self, *args: Any, **kwargs: Any) -> None:
        super().__init__(*args, **kwargs)
        self.trials_per_random_config: int = self.args.trials_per_random_config
        self.early_stopping: Optional[int] = self.args.early_stopping

    def early_stopping_triggered(self) -> bool:
        return self.early_stopping is not None and self.trial_tracker.num_trials_since_best_result >= self.early_stopping

    def _pick_mbs(self, window: DSATSearchData) -> int:
        return self.get_random_mbs_from_search_data(window)

    def get_random_mbs_from_search_data(self, search_data: DSATSearchData) -> int:
        return int(search_data.lo + self.rng.binomial(search_data.hi - search_data.lo, 0.5))

    def _stage_pool(self) -> List[int]:
        tr = self.trial_tracker
        proven_cheap = {1, 2} & tr.successful_stages & tr.zero_stages      # stage 3 only pays when 1 and 2 do not fit
        return sorted(proven_cheap or tr.zero_stages)

    def _first_window(self, zero_stage: int) -> DSATSearchData:
        tr = self.trial_tracker
        best = tr.best_trials_by_stage[zero_stage]
        if best is None or best.search_data is None:
            return super()._first_window(zero_stage)
        # start above the best result of this stage; the ceiling is whatever is largest of: the best trial's own ceiling,
        # the largest batch any config of the stage has run, twice the new floor
        lo = best.mbs + 1
        largest_run = max(t.mbs for t in tr.completed_trials
                          if t.stage == best.stage and isinstance(t.metric, dict) and t.metric.get(tr.searcher_metric) is not None)
        w = copy.deepcopy(best.search_data)
        w.lo, w.hi = lo, max(largest_run, best.search_data.hi, 2 * lo)
        return w

    def should_stop_lineage(self, trial: DSATTrial) -> bool:
        w = trial.search_data
        assert w is not None
        if trial.error and trial.mbs <= w.lo:
            return True                                                             # not even the floor fits
        if trial.num_completed_trials_in_lineage >= self.trials_per_random_config:
            return True                                                             # budget of this config spent
        if trial.stage == 3 and {1, 2} & self.trial_tracker.successful_stages:
            return True                                                             # a cheaper stage works
        if trial.error_in_direct_history and any(o.mbs >= w.hi for _, o in self.trial_tracker
                                                 if o.stage == trial.stage and o.searcher_metric_val is not None):
            return True                                                             # others already run more than this one's proven ceiling
        return False

    def get_trial_list_after_successful_run(self, last_trial: DSATTrial) -> List[DSATTrial]:
        if last_trial.search_data is None or self.should_stop_lineage(last_trial):
            return [self.get_random_trial()]
        w = _narrow(last_trial)
        assert w is not None
        if w.lo > w.hi:
            w.hi *= 2               # the estimated ceiling was too low: it is only an estimate until something fails
        return [self._successor(last_trial, w)]

    def get_trials_after_early_exit(self, searcher_state: searcher.SearcherState, last_trial: DSATTrial,
                                    exited_reason: searcher.ExitedReason) -> List[DSATTrial]:
        if self.should_stop_lineage(last_trial):
            return [self.get_random_trial()]
        w = _narrow(last_trial)
        return [] if w is None else [self._successor(last_trial, w)]

    def choose_next_trial_from_queue(self) -> DSATTrial:
        """Queued trials whose lineage has meanwhile become pointless are swapped for fresh random ones."""
        q = self.trial_tracker.queue
        nxt = q.popleft()
        while self.should_stop_lineage(nxt):
            self.trial_tracker.queue_and_register_trial(self.get_random_trial())
            nxt = q.popleft()
        return nxt


class ASHADSATSearchMethod(_LineageSearch):
    """Asynchronous successive halving over random engine configs, where the resource is the number of bisection probes a
    config has been given: rung r allows ``min_binary_search_trials * divisor ** (asha_early_stopping + r)`` probes; of the
    lineages that finished a rung the best ``1 / divisor`` are promoted and continue their bisection in the next rung."""

    def __init__(self, *args: Any, **kwargs: Any) -> None:
        super().__init__(*args, **kwargs)
        self.divisor: int = self.args.divisor
        self.max_rungs: int = self.args.max_rungs
        self.min_binary_search_trials: int = self.args.min_binary_search_trials
        self.asha_early_stopping: int = self.args.asha_early_stopping
        self.search_range_factor = self.args.search_range_factor

    # -- windows carry the rung ----------------------------------------------------------------------------------------------
    def _first_window(self, zero_stage: int) -> ASHADSATSearchData:
        w = super()._first_window(zero_stage)
        return ASHADSATSearchData(lo=w.lo, hi=w.hi, curr_rung=0)

    def _population(self) -> int:
        return min(super()._population(), self.trial_tracker.max_trials)

    def max_trials_for_rung_idx(self, rung_idx: int) -> int:
        return int(self.min_binary_search_trials * self.divisor ** (self.asha_early_stopping + rung_idx))

    # -- lineages ------------------------------------------------------------------------------------------------------------
    @staticmethod
    def _rung(trial: DSATTrial) -> int:
        assert isinstance(trial.search_data, ASHADSATSearchData)
        return trial.search_data.curr_rung

    def get_latest_trial_in_lineage(self, trial: DSATTrial) -> DSATTrial:
        while trial.children:
            assert len(trial.children) == 1, "an ASHA lineage is a chain"
            trial = next(iter(trial.children))
        return trial

    def get_all_latest_trials_in_lineages(self) -> List[DSATTrial]:
        return [t for _, t in self.trial_tracker if not isinstance(t, DSATModelProfileInfoTrial) and not t.children
                and isinstance(t.search_data, ASHADSATSearchData)]

    def get_next_trial_in_lineage(self, trial: DSATTrial) -> Optional[DSATTrial]:
        latest = self.get_latest_trial_in_lineage(trial)
        w = _narrow(latest)
        if w is None or w.hi < w.lo:
            return None
        return self._successor(latest, w)

    def get_best_trial_in_lineage(self, trial: DSATTrial, max_rung_idx: Optional[int] = None) -> Optional[DSATTrial]:
        """Best result of the lineage, counting only probes run in rungs up to ``max_rung_idx``."""
        scored = [t for t in trial.lineage_set if t.searcher_metric_val is not None and (max_rung_idx is None or self._rung(t) <= max_rung_idx)]
        if not scored:
            return None
        pick = min if self.trial_tracker.smaller_is_better else max
        return pick(scored, key=lambda t: t.searcher_metric_val)

    def lineage_completed_rung(self, trial: DSATTrial, rung_idx: int) -> bool:
        latest = self.get_latest_trial_in_lineage(trial)
        w = latest.search_data
        assert isinstance(w, ASHADSATSearchData)
        if w.curr_rung > rung_idx:
            return True                                                    # already promoted past it
        if trial.num_completed_trials_in_lineage >= self.max_trials_for_rung_idx(rung_idx):
            return True                                                    # used up the rung's probes
        # a bisection that ended early (nothing left to try) has completed its current rung too
        return w.curr_rung == rung_idx and (w.hi == w.lo or (latest.error and latest.mbs == w.lo))

    @property
    def rungs(self) -> Dict[int, List[DSATTrial]]:
        """rung -> the latest trial of every lineage that has completed it."""
        out: Dict[int, List[DSATTrial]] = collections.defaultdict(list)
        for tip in self.get_all_latest_trials_in_lineages():
            r = 0
            while self.lineage_completed_rung(tip, r):
                out[r].append(tip)
                r += 1
        return out

    # -- promotion -----------------------------------------------------------------------------------------------------------
    def get_top_lineages_in_rung(self, rung_idx: int) -> List[DSATTrial]:
        """Best trial (within the rung) of each of the top ``1 / divisor`` lineages that completed the rung, best first."""
        done = self.rungs[rung_idx]
        k = len(done) // self.divisor
        if k == 0:
            return []
        bests = [b for b in (self.get_best_trial_in_lineage(tip, max_rung_idx=rung_idx) for tip in done) if b is not None]
        bests.sort(key=lambda t: t.searcher_metric_val, reverse=not self.trial_tracker.smaller_is_better)     # type: ignore[arg-type, return-value]
        return bests[:k]

    def get_next_promotable_lineage_in_rung(self, rung_idx: int) -> Optional[DSATTrial]:
        for best in self.get_top_lineages_in_rung(rung_idx):
            tip = self.get_latest_trial_in_lineage(best)
            if self._rung(tip) <= rung_idx:
                return tip
        return None

    def get_next_promotable_lineage(self) -> Optional[DSATTrial]:
        for rung_idx in reversed(range(self.max_rungs - 1)):               # nothing is promoted out of the top rung
            tip = self.get_next_promotable_lineage_in_rung(rung_idx)
            if tip is not None:
                return tip
        return None

    def get_next_trial(self, last_trial: DSATTrial) -> DSATTrial:
        """In this order: the lineage's own next probe while its rung has budget; a promoted lineage's next probe; a new
        random lineage."""
        nxt: Optional[DSATTrial] = None
        if not self.lineage_completed_rung(last_trial, self._rung(last_trial)):
            nxt = self.get_next_trial_in_lineage(last_trial)
        if nxt is None:
            tip = self.get_next_promotable_lineage()
            if tip is not None:
                nxt = self.get_next_trial_in_lineage(tip)
                if nxt is not None:
                    assert isinstance(nxt.search_data, ASHADSATSearchData)
                    nxt.search_data.curr_rung += 1
        return nxt if nxt is not None else self.get_random_trial()

    def get_trial_list_after_successful_run(self, last_trial: DSATTrial) -> List[DSATTrial]:
        return [self.get_next_trial(last_trial)]

    def get_trials_after_early_exit(self, searcher_state: searcher.SearcherState, last_trial: DSATTrial,
                                    exited_reason: searcher.ExitedReason) -> List[DSATTrial]:
        return [self.get_next_trial(last_trial)]

    def choose_next_trial_from_queue(self) -> DSATTrial:
        """Highest rung first (promotions matter most); among those the longest lineage, so that it reaches its next
        promotion soonest."""
        q = self.trial_tracker.queue
        top = max(self._rung(t) for t in q)
        nxt = max((t for t in q if self._rung(t) == top), key=lambda t: t.num_completed_trials_in_lineage)
        q.remove(nxt)
        return nxt


class TestDSATSearchMethod(BaseDSATSearchMethod):
    """For tests: after the profile run, ``max_trials - 1`` trials with micro-batch sizes 2, 3, ... and random stages."""

    __test__ = False       # not a pytest class

    def get_trials_after_validation_completed(self, searcher_state: searcher.SearcherState, last_trial: DSATTrial,
                                              metric: Optional[Metric] = None) -> List[DSATTrial]:
        if not isinstance(last_trial, DSATModelProfileInfoTrial):
            return []
        base = last_trial.hparams
        for key in ("model_info", "model_info_path"):          # the switches that made it a profile run
            base[defaults.OVERWRITE_KEY]["autotuning"].pop(key, None)
        out = []
        for mbs in range(2, self.trial_tracker.max_trials + 1):
            hp = copy.deepcopy(base)
            hp[defaults.OVERWRITE_KEY][MBS_KEY] = mbs
            hp[defaults.OVERWRITE_KEY]["zero_optimization"] = {"stage": random.choice(list(self.args.zero_stages))}
            out.append(self.trial_tracker.create_trial(hparams=hp, search_data=None, parent_trial=None))
        return out

    def get_trials_after_early_exit(self, searcher_state: searcher.SearcherState, last_trial: DSATTrial,
                                    exited_reason: searcher.ExitedReason) -> List[DSATTrial]:
        return []
