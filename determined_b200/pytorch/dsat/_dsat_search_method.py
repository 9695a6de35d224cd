"""Engine-config autotuning as custom search methods.

Parity: reference ``harness/determined/pytorch/dsat/_dsat_search_method.py``: ``DSATTrial``
``:24-160``, ``DSATTrialTracker`` (lineages, per-ZeRO-stage memory model ``:383-414``, approximate
max micro batch size ``:416-430``) ``:169-517``, ``BaseDSATSearchMethod`` (model-profile trial
first, then ``get_trials_after_model_profile_info_run`` / ``choose_next_trial_from_queue``, OOM
handling, early stopping) ``:520-736``, ``RandomDSATSearchMethod`` ``:746-962``,
``BinarySearchDSATSearchMethod`` ``:965-1085``, ``ASHADSATSearchMethod`` ``:1094-1383``.

The searched object is the engine config consumed by :mod:`determined_b200.parallel.zero` (or by
DeepSpeed): ``zero_optimization.stage`` and friends plus ``train_micro_batch_size_per_gpu``.
"""
import copy
import dataclasses
import logging
import pathlib
import pickle
import random
import uuid
from typing import Any, Deque, Dict, List, Optional, Tuple
import collections

from determined_b200 import searcher, util
from determined_b200.pytorch.dsat import _utils, defaults

logger = logging.getLogger("determined.pytorch.dsat")


class DSATTrial:
    def __init__(self, hparams: Dict[str, Any], model_dir: str, slots_per_trial: int, length: int,
                 request_id: Optional[uuid.UUID] = None, parent: Optional["DSATTrial"] = None,
                 search_data: Any = None, searcher_metric_name: Optional[str] = None) -> None:
        self.hparams, self.model_dir, self.slots_per_trial, self.length = hparams, model_dir, slots_per_trial, length
        self.request_id = request_id or uuid.uuid4()
        self.parent, self.search_data, self.searcher_metric_name = parent, search_data, searcher_metric_name
        self.children: List["DSATTrial"] = []
        if parent is not None:
            parent.children.append(self)
        self.metric: Any = {}
        self.error = False
        self.running = False
        self.ds_config = _utils.get_ds_config_from_hparams(hparams, model_dir)
        self._closed = False

    @property
    def completed(self) -> bool:
        return bool(self.error or self.metric)

    @property
    def lineage_root(self) -> "DSATTrial":
        t = self
        while t.parent is not None:
            t = t.parent
        return t

    @property
    def lineage_set(self) -> List["DSATTrial"]:
        out, stack = [], [self.lineage_root]
        while stack:
            t = stack.pop()
            out.append(t)
            stack.extend(t.children)
        return out

    @property
    def num_completed_trials_in_lineage(self) -> int:
        return sum(t.completed for t in self.lineage_set)

    @property
    def mbs(self) -> int:
        return int(self.ds_config["train_micro_batch_size_per_gpu"])

    @property
    def stage(self) -> int:
        return int((self.ds_config.get("zero_optimization") or {}).get("stage", 0))

    @property
    def searcher_metric_val(self) -> Optional[float]:
        if isinstance(self.metric, dict) and self.searcher_metric_name:
            return self.metric.get(self.searcher_metric_name)
        return None

    @property
    def create_and_val_ops(self) -> List[searcher.Operation]:
        return [searcher.Create(self.request_id, self.hparams, None), searcher.ValidateAfter(self.request_id, self.length)]


class DSATModelProfileInfoTrial(DSATTrial):
    """First trial of every search: runs with micro batch 1 to measure parameter counts and
    activation memory, which seed the memory model."""


class DSATTrialTracker:
    """All trials of one autotuning search + what was learnt from the model-profile trial."""

    def __init__(self, args: Any, exp_config: Dict[str, Any]) -> None:
        self.exp_config = exp_config
        self.max_trials, self.max_concurrent_trials = args.max_trials, args.max_concurrent_trials
        self.max_slots = args.max_slots
        self.model_dir = args.model_dir
        self.searcher_metric = args.metric
        self.start_profile_step, self.end_profile_step = args.start_profile_step, args.end_profile_step
        self.zero_stages = set(args.zero_stages)
        self.smaller_is_better = _utils.smaller_is_better(self.searcher_metric)
        self.slots_per_trial = int(exp_config["resources"]["slots_per_trial"])
        self.hparams = exp_config["hyperparameters"]
        self.all_trials_dict: Dict[uuid.UUID, DSATTrial] = {}
        self.queue: Deque[DSATTrial] = collections.deque()
        self.model_profile_info_trial: Optional[DSATTrial] = None
        self.num_trials_since_best_result = 0
        self.successful_stages: set = set()
        self._best_trial: Optional[DSATTrial] = None
        self._mem_per_gpu_per_stage: Optional[Dict[int, int]] = None

    def __len__(self) -> int:
        return len(self.all_trials_dict)

    def __getitem__(self, rid: uuid.UUID) -> DSATTrial:
        return self.all_trials_dict[rid]

    def __iter__(self) -> Any:
        return iter(self.all_trials_dict.items())

    def __contains__(self, rid: uuid.UUID) -> bool:
        return rid in self.all_trials_dict

    def create_trial(self, hparams: Dict[str, Any], search_data: Any = None, parent_trial: Optional[DSATTrial] = None) -> DSATTrial:
        """Build a trial whose overwrite block carries the autotuning markers + batch config."""
        hp = copy.deepcopy(hparams)
        ow = hp.setdefault(defaults.OVERWRITE_KEY, {})
        ow.setdefault("autotuning", {}).update({"enabled": True, "start_profile_step": self.start_profile_step,
                                                "end_profile_step": self.end_profile_step, "metric": self.searcher_metric})
        hp[defaults.USE_DSAT_MODE_KEY] = True
        base = _utils.get_ds_config_from_hparams(hp, self.model_dir)
        ow.update(_utils.get_batch_config_from_mbs_gas_and_slots(base, self.slots_per_trial))
        t = DSATTrial(hp, self.model_dir, self.slots_per_trial, self.end_profile_step, parent=parent_trial, search_data=search_data,
                      searcher_metric_name=self.searcher_metric)
        return t

    def create_model_profile_info_trial(self) -> DSATModelProfileInfoTrial:
        hp = copy.deepcopy(self.hparams)
        hp[defaults.OVERWRITE_KEY] = util.merge_dicts(hp.get(defaults.OVERWRITE_KEY, {}), defaults.MODEL_INFO_PROFILE_DS_CONFIG)
        hp[defaults.OVERWRITE_KEY].update({"train_batch_size": self.slots_per_trial, "gradient_accumulation_steps": 1})
        hp[defaults.USE_DSAT_MODE_KEY] = True
        t = DSATModelProfileInfoTrial(hp, self.model_dir, self.slots_per_trial, 1, searcher_metric_name=None)
        self.model_profile_info_trial = t
        return t

    def queue_and_register_trial(self, trial: DSATTrial) -> None:
        self.all_trials_dict[trial.request_id] = trial
        self.queue.append(trial)

    def update_trial_metric(self, trial: DSATTrial, metric: Any) -> None:
        trial.metric = metric
        trial.running = False
        if isinstance(trial, DSATModelProfileInfoTrial):
            self._mem_per_gpu_per_stage = None
            return
        self.successful_stages.add(trial.stage)
        val = trial.searcher_metric_val
        if val is not None and (self._best_trial is None or self._better(val, self._best_trial.searcher_metric_val)):
            self._best_trial = trial
            self.num_trials_since_best_result = 0
        else:
            self.num_trials_since_best_result += 1

    def _better(self, a: float, b: Optional[float]) -> bool:
        return b is None or (a < b if self.smaller_is_better else a > b)

    def report_trial_early_exit(self, trial: DSATTrial) -> None:
        trial.error, trial.running = True, False
        self.num_trials_since_best_result += 1

    @property
    def best_trial(self) -> Optional[DSATTrial]:
        return self._best_trial

    @property
    def num_completed_trials(self) -> int:
        return sum(t.completed for t in self.all_trials_dict.values())

    @property
    def num_running_trials(self) -> int:
        return sum(t.running for t in self.all_trials_dict.values())

    @property
    def max_trials_queued(self) -> bool:
        return len(self) >= self.max_trials

    @property
    def max_trials_are_running_or_closed(self) -> bool:
        return self.num_completed_trials + self.num_running_trials >= self.max_trials

    @property
    def can_run_more_trials(self) -> bool:
        if not self.queue or self.max_trials_are_running_or_closed:
            return False
        if self.num_running_trials >= self.max_concurrent_trials:
            return False
        if self.max_slots is not None and (self.num_running_trials + 1) * self.slots_per_trial > self.max_slots:
            return False
        return True

    @property
    def should_be_failure(self) -> bool:
        if self.model_profile_info_trial is not None and self.model_profile_info_trial.error:
            return True
        done = [t for t in self.all_trials_dict.values() if t.completed and not isinstance(t, DSATModelProfileInfoTrial)]
        return bool(done) and all(t.error for t in done) and self.max_trials_are_running_or_closed

    # -- memory model (reference :383-430) ---------------------------------------------------------
    def _info(self, key: str, default: int = 0) -> int:
        t = self.model_profile_info_trial
        assert t is not None and isinstance(t.metric, dict), "The model profile info Trial must be run before calling this method."
        return int(t.metric.get(key, default))

    @property
    def gpu_mem(self) -> int:
        return self._info("gpu_mem", 180 * 2**30)

    @property
    def num_params(self) -> int:
        return self._info("num_params")

    @property
    def trainable_num_params(self) -> int:
        return self._info("trainable_num_params", self.num_params)

    @property
    def activation_mem_per_gpu(self) -> int:
        return max(self._info("activation_mem_per_gpu", 1), 1)

    @property
    def mem_per_gpu_per_stage(self) -> Dict[int, int]:
        if self._mem_per_gpu_per_stage is None:
            t = self.model_profile_info_trial
            assert t is not None
            low = bool((t.ds_config.get("fp16") or {}).get("enabled") or (t.ds_config.get("bf16") or {}).get("enabled"))
            params = self.num_params * (2 if low else 4)
            grads = self.trainable_num_params * 4
            opt = self.trainable_num_params * ((4 if low else 0) + 8)    # master + Adam m, v
            n = self.slots_per_trial
            base = {0: params + grads + opt, 1: params + grads + opt // n, 2: params + (grads + opt) // n, 3: (params + grads + opt) // n}
            self._mem_per_gpu_per_stage = {s: m + self.activation_mem_per_gpu for s, m in base.items()}
        return self._mem_per_gpu_per_stage

    @property
    def approx_max_mbs_per_stage(self) -> Dict[int, int]:
        return {s: max((self.gpu_mem - m) // self.activation_mem_per_gpu, 1) for s, m in self.mem_per_gpu_per_stage.items()}


class BaseDSATSearchMethod(searcher.SearchMethod):
    def __init__(self, args: Any, exp_config: Dict[str, Any]) -> None:
        self.args, self.exp_config = args, exp_config
        self.trial_tracker = DSATTrialTracker(args, exp_config)
        self.rng = random.Random(args.random_seed)
        self._tracker_file = "trial_tracker.pkl"

    # -- to implement ------------------------------------------------------------------------------
    def get_trials_after_model_profile_info_run(self) -> List[DSATTrial]:
        raise NotImplementedError

    def choose_next_trial_from_queue(self) -> DSATTrial:
        return self.trial_tracker.queue.popleft()

    def get_new_searcher_ops_list(self, last_trial: DSATTrial) -> List[DSATTrial]:
        """Given the trial that just finished (or failed), return follow-up trials."""
        raise NotImplementedError

    def should_stop(self) -> bool:
        return False

    # -- SearchMethod ------------------------------------------------------------------------------
    def initial_operations(self, searcher_state: searcher.SearcherState) -> List[searcher.Operation]:
        t = self.trial_tracker.create_model_profile_info_trial()
        self.trial_tracker.queue_and_register_trial(t)
        self.trial_tracker.queue.popleft()
        t.running = True
        return t.create_and_val_ops

    def on_trial_created(self, searcher_state: searcher.SearcherState, request_id: uuid.UUID) -> List[searcher.Operation]:
        return []

    def _drain_queue(self) -> List[searcher.Operation]:
        ops: List[searcher.Operation] = []
        tr = self.trial_tracker
        while tr.can_run_more_trials:
            nxt = self.choose_next_trial_from_queue()
            nxt.running = True
            ops += nxt.create_and_val_ops
        return ops

    def on_validation_completed(self, searcher_state: searcher.SearcherState, request_id: uuid.UUID, metric: Any, train_length: int) -> List[searcher.Operation]:
        tr = self.trial_tracker
        last = tr[request_id]
        if isinstance(metric, dict) and metric.get("OOM"):
            tr.report_trial_early_exit(last)
        else:
            tr.update_trial_metric(last, metric if isinstance(metric, dict) else {tr.searcher_metric: metric})
        ops: List[searcher.Operation] = [searcher.Close(request_id)]
        if isinstance(last, DSATModelProfileInfoTrial) and not last.error:
            new = self.get_trials_after_model_profile_info_run()
        elif tr.should_be_failure or self.should_stop():
            new = []
        else:
            new = self.get_new_searcher_ops_list(last)
        for t in new:
            if not tr.max_trials_queued:
                tr.queue_and_register_trial(t)
        return ops + self._drain_queue()

    def on_trial_closed(self, searcher_state: searcher.SearcherState, request_id: uuid.UUID) -> List[searcher.Operation]:
        tr = self.trial_tracker
        if tr.should_be_failure:
            return [searcher.Shutdown(failure=True)]
        nothing_left = not tr.queue and tr.num_running_trials == 0
        if nothing_left or (tr.max_trials_are_running_or_closed and tr.num_running_trials == 0) or (self.should_stop() and tr.num_running_trials == 0):
            if tr.best_trial is not None:
                logger.info("autotuning finished: best %s = %s with config overwrite %s", tr.searcher_metric,
                            tr.best_trial.searcher_metric_val, tr.best_trial.hparams.get(defaults.OVERWRITE_KEY))
            return [searcher.Shutdown()]
        return []

    def on_trial_exited_early(self, searcher_state: searcher.SearcherState, request_id: uuid.UUID, exited_reason: searcher.ExitedReason) -> List[searcher.Operation]:
        tr = self.trial_tracker
        last = tr[request_id]
        tr.report_trial_early_exit(last)
        if isinstance(last, DSATModelProfileInfoTrial):
            return [searcher.Shutdown(failure=True)]
        if exited_reason != searcher.ExitedReason.ERRORED:
            return [searcher.Shutdown(failure=True)]
        for t in self.get_new_searcher_ops_list(last):
            if not tr.max_trials_queued:
                tr.queue_and_register_trial(t)
        return self._drain_queue()

    def progress(self, searcher_state: searcher.SearcherState) -> float:
        return min(self.trial_tracker.num_completed_trials / max(self.trial_tracker.max_trials, 1), 1.0)

    def save_method_state(self, path: pathlib.Path) -> None:
        with (path / self._tracker_file).open("wb") as f:
            pickle.dump((self.trial_tracker, self.rng.getstate()), f)

    def load_method_state(self, path: pathlib.Path) -> None:
        with (path / self._tracker_file).open("rb") as f:
            self.trial_tracker, st = pickle.load(f)
        self.rng.setstate(st)

    # -- helpers -----------------------------------------------------------------------------------
    def _hparams_with(self, stage: int, mbs: int, zero_cfg: Optional[Dict[str, Any]] = None) -> Dict[str, Any]:
        hp = copy.deepcopy(self.trial_tracker.hparams)
        ow = hp.setdefault(defaults.OVERWRITE_KEY, {})
        z = dict(zero_cfg or {})
        z["stage"] = stage
        ow["zero_optimization"] = z
        ow["train_micro_batch_size_per_gpu"] = int(mbs)
        return hp


@dataclasses.dataclass
class DSATSearchData:
    lo: int
    hi: int


class RandomDSATSearchMethod(BaseDSATSearchMethod):
    """Random ZeRO configs; inside each config's lineage the micro batch size is refined by a
    bounded binary search (``trials_per_random_config`` trials), with optional early stopping."""

    def __init__(self, args: Any, exp_config: Dict[str, Any]) -> None:
        super().__init__(args, exp_config)
        self.trials_per_random_config = int(getattr(args, "trials_per_random_config", defaults.AUTOTUNING_ARG_DEFAULTS["trials-per-random-config"]))
        self.early_stopping = getattr(args, "early_stopping", None)

    def _random_trial(self) -> DSATTrial:
        tr = self.trial_tracker
        stage = self.rng.choice(sorted(tr.zero_stages))
        zero = _utils.get_random_zero_optim_config(stage, self.rng)
        hi = max(int(tr.approx_max_mbs_per_stage[stage]), 1)
        mbs = self.rng.randint(1, hi)
        return tr.create_trial(self._hparams_with(stage, mbs, zero), search_data=DSATSearchData(lo=1, hi=2 * hi - 1))

    def get_trials_after_model_profile_info_run(self) -> List[DSATTrial]:
        n = min(self.trial_tracker.max_concurrent_trials, self.trial_tracker.max_trials - 1)
        return [self._random_trial() for _ in range(max(n, 1))]

    def should_stop(self) -> bool:
        return self.early_stopping is not None and self.trial_tracker.num_trials_since_best_result >= self.early_stopping

    def get_new_searcher_ops_list(self, last: DSATTrial) -> List[DSATTrial]:
        tr = self.trial_tracker
        sd: DSATSearchData = last.search_data
        if last.num_completed_trials_in_lineage >= self.trials_per_random_config or sd is None:
            return [self._random_trial()]
        lo, hi = sd.lo, sd.hi
        if last.error:
            hi = last.mbs - 1
        else:
            lo = last.mbs + 1
        if lo > hi:
            return [self._random_trial()]
        mbs = (lo + hi) // 2
        z = dict(last.ds_config.get("zero_optimization") or {})
        return [tr.create_trial(self._hparams_with(last.stage, mbs, z), search_data=DSATSearchData(lo, hi), parent_trial=last)]


class BinarySearchDSATSearchMethod(BaseDSATSearchMethod):
    """Per ZeRO stage, binary-search the largest micro batch size that fits (best throughput is
    usually near the memory limit); default ZeRO knobs otherwise."""

    def __init__(self, args: Any, exp_config: Dict[str, Any]) -> None:
        super().__init__(args, exp_config)
        self.search_range_factor = float(getattr(args, "search_range_factor", 1.0))

    def get_trials_after_model_profile_info_run(self) -> List[DSATTrial]:
        tr = self.trial_tracker
        out = []
        for stage in sorted(tr.zero_stages):
            hi = max(int(self.search_range_factor * tr.approx_max_mbs_per_stage[stage]), 1)
            mbs = (1 + hi) // 2
            out.append(tr.create_trial(self._hparams_with(stage, mbs), search_data=DSATSearchData(1, hi)))
        return out

    def get_new_searcher_ops_list(self, last: DSATTrial) -> List[DSATTrial]:
        sd: DSATSearchData = last.search_data
        if sd is None:
            return []
        lo, hi = sd.lo, sd.hi
        if last.error:
            hi = last.mbs - 1
        else:
            lo = last.mbs + 1
        if lo > hi:
            return []
        return [self.trial_tracker.create_trial(self._hparams_with(last.stage, (lo + hi) // 2, dict(last.ds_config.get("zero_optimization") or {})),
                                                search_data=DSATSearchData(lo, hi), parent_trial=last)]


@dataclasses.dataclass
class ASHADSATSearchData(DSATSearchData):
    curr_rung: int = 0


class ASHADSATSearchMethod(BaseDSATSearchMethod):
    """Successive halving over random ZeRO configs where the "resource" is the number of binary
    search trials a lineage may spend on finding its best micro batch size: rung r grants
    ``min_binary_search_trials * divisor**r`` trials; only the top 1/divisor lineages of a rung are
    promoted."""

    def __init__(self, args: Any, exp_config: Dict[str, Any]) -> None:
        super().__init__(args, exp_config)
        self.divisor = int(getattr(args, "divisor", 2))
        self.max_rungs = int(getattr(args, "max_rungs", 5))
        self.min_bs_trials = int(getattr(args, "min_binary_search_trials", 3))
        self.asha_early_stopping = int(getattr(args, "asha_early_stopping", 0))
        self.search_range_factor = float(getattr(args, "search_range_factor", 1.0))
        self.rungs: Dict[int, List[DSATTrial]] = {r: [] for r in range(self.max_rungs)}   # finished lineage roots per rung
        self.promoted: set = set()

    def max_trials_for_rung(self, r: int) -> int:
        return self.min_bs_trials * self.divisor ** r

    def should_stop(self) -> bool:
        return self.asha_early_stopping > 0 and self.trial_tracker.num_trials_since_best_result >= self.asha_early_stopping

    def _new_lineage(self) -> DSATTrial:
        tr = self.trial_tracker
        stage = self.rng.choice(sorted(tr.zero_stages))
        hi = max(int(self.search_range_factor * tr.approx_max_mbs_per_stage[stage]), 1)
        return tr.create_trial(self._hparams_with(stage, (1 + hi) // 2, _utils.get_random_zero_optim_config(stage, self.rng)),
                               search_data=ASHADSATSearchData(1, hi, 0))

    def get_trials_after_model_profile_info_run(self) -> List[DSATTrial]:
        n = min(self.trial_tracker.max_concurrent_trials, self.trial_tracker.max_trials - 1)
        return [self._new_lineage() for _ in range(max(n, 1))]

    def _best_in_lineage(self, t: DSATTrial) -> Optional[float]:
        vals = [x.searcher_metric_val for x in t.lineage_set if x.searcher_metric_val is not None]
        if not vals:
            return None
        return min(vals) if self.trial_tracker.smaller_is_better else max(vals)

    def _promotable(self, r: int) -> Optional[DSATTrial]:
        done = [t for t in self.rungs[r] if self._best_in_lineage(t) is not None]
        done.sort(key=lambda t: self._best_in_lineage(t), reverse=not self.trial_tracker.smaller_is_better)   # type: ignore
        for t in done[: len(self.rungs[r]) // self.divisor]:
            if (r, t.request_id) not in self.promoted:
                return t
        return None

    def get_new_searcher_ops_list(self, last: DSATTrial) -> List[DSATTrial]:
        tr = self.trial_tracker
        sd: ASHADSATSearchData = last.search_data
        budget = self.max_trials_for_rung(sd.curr_rung)
        lo, hi = sd.lo, sd.hi
        if last.error:
            hi = last.mbs - 1
        else:
            lo = last.mbs + 1
        if last.num_completed_trials_in_lineage < budget and lo <= hi:
            return [tr.create_trial(self._hparams_with(last.stage, (lo + hi) // 2, dict(last.ds_config.get("zero_optimization") or {})),
                                    search_data=ASHADSATSearchData(lo, hi, sd.curr_rung), parent_trial=last)]
        # lineage finished its rung
        self.rungs[sd.curr_rung].append(last.lineage_root)
        for r in range(self.max_rungs - 2, -1, -1):
            cand = self._promotable(r)
            if cand is not None:
                self.promoted.add((r, cand.request_id))
                best = max((x for x in cand.lineage_set if x.completed and not x.error), key=lambda x: x.mbs, default=cand)
                lo2, hi2 = best.mbs + 1, max(best.search_data.hi, best.mbs + 1) * 2
                return [tr.create_trial(self._hparams_with(best.stage, (lo2 + hi2) // 2, dict(best.ds_config.get("zero_optimization") or {})),
                                        search_data=ASHADSATSearchData(lo2, hi2, r + 1), parent_trial=best)]
        return [self._new_lineage()]


class TestDSATSearchMethod(BaseDSATSearchMethod):
    """Deterministic tiny search used by end-to-end tests (``_test``): one trial per stage."""

    def get_trials_after_model_profile_info_run(self) -> List[DSATTrial]:
        return [self.trial_tracker.create_trial(self._hparams_with(s, 1)) for s in sorted(self.trial_tracker.zero_stages)]

    def get_new_searcher_ops_list(self, last: DSATTrial) -> List[DSATTrial]:
        return []
