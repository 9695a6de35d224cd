"""Fused optimizers over flat buffers (native kernels in ``csrc/optim.cu``).

``fuse_optimizer_(opt, layout)`` converts a ``torch.optim.{SGD,Adam,AdamW,Adadelta}`` *in place*
(object identity, ``param_groups`` and hyper-parameter dicts are preserved, so LR schedulers and
user references keep working) into one whose ``step()`` is a single kernel launch per param
group.  ``state_dict()`` stays interchangeable with the stock torch optimizer so checkpoints
follow the reference's ``optimizers_state_dict`` contract (``_pytorch_trial.py:1292-1329``).
"""
import logging
from typing import Any, Callable, Dict, List, Optional

import torch

from determined_b200.ops import _lib
from determined_b200.parallel.flat import FlatLayout

logger = logging.getLogger("determined.ops")

_SUPPORTED = (torch.optim.SGD, torch.optim.Adam, torch.optim.AdamW, torch.optim.Adadelta)


def _kind(opt: torch.optim.Optimizer) -> Optional[str]:
    # exact classes only: subclasses may override step()
    t = type(opt)
    if t is torch.optim.SGD:
        return "sgd"
    if t is torch.optim.AdamW:
        return "adamw"
    if t is torch.optim.Adam:
        return "adam"
    if t is torch.optim.Adadelta:
        return "adadelta"
    return None


def can_fuse(opt: torch.optim.Optimizer) -> bool:
    k = _kind(opt)
    if k is None:
        return False
    for g in opt.param_groups:
        if g.get("amsgrad") or g.get("differentiable"):
            return False
        if k in ("adam", "adamw") and isinstance(g.get("lr"), torch.Tensor):
            return False
        for p in g["params"]:
            if p.requires_grad and (p.dtype != torch.float32 or not p.is_cuda or p.is_sparse):
                return False
    return True


class _GroupState:
    def __init__(self, device: torch.device) -> None:
        self.lr = torch.zeros((), dtype=torch.float32, device=device)
        self.lr_host: Optional[float] = None
        self.step = torch.zeros((), dtype=torch.int32, device=device)
        self.bufs: Dict[str, torch.Tensor] = {}


class _FusedMixin:
    """Mixed in front of the torch optimizer class by :func:`fuse_optimizer_`."""

    _detb200_kind: str
    _detb200_layout: FlatLayout
    _detb200_groups: List[_GroupState]

    # -- hyper-parameters that live on the device (so a captured graph sees updates) -------------
    def sync_hparams(self) -> None:
        """Push host-side lr changes (LR schedulers mutate ``param_groups[i]['lr']``) to the
        device scalars.  Must run outside CUDA-graph capture."""
        for g, gs in zip(self.param_groups, self._detb200_groups):  # type: ignore[attr-defined]
            lr = float(g["lr"])
            if gs.lr_host != lr:
                gs.lr.fill_(lr)
                gs.lr_host = lr

    def _state_buf(self, gi: int, name: str) -> torch.Tensor:
        """Flat state tensor ``name`` (one per optimizer, spanning all groups); on first use for a
        group, per-parameter views are published through the regular torch ``state`` mapping."""
        lay = self._detb200_layout
        flats = self.__dict__.setdefault("_detb200_flats", {})
        if name not in flats:
            flats[name] = torch.zeros(lay.total, dtype=torch.float32, device=lay.device)
        gs = self._detb200_groups[gi]
        if name not in gs.bufs:
            gs.bufs[name] = flats[name]
            for s in lay.slots:
                if s.group == gi:
                    self.state[s.param][name] = lay.state_view(flats[name], s)  # type: ignore[attr-defined]
        return gs.bufs[name]

    @torch.no_grad()
    def step(self, closure: Optional[Callable] = None, *, grad_scale: Optional[torch.Tensor] = None,
             skip: Optional[torch.Tensor] = None, static_scale: float = 1.0, zero_grad: bool = False) -> Any:
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = _lib.load()
        capturing = torch.cuda.is_current_stream_capturing()
        if not capturing:
            self.sync_hparams()
        st = _lib.stream_ptr()
        lay = self._detb200_layout
        kind = self._detb200_kind
        gsp, skp = _lib.ptr(grad_scale), _lib.ptr(skip)
        for gi, (g, gs) in enumerate(zip(self.param_groups, self._detb200_groups)):  # type: ignore[attr-defined]
            seg = lay.segments[gi]
            if seg.numel == 0:
                continue
            p = lay.flat_param.data_ptr() + 4 * seg.offset
            gr = lay.flat_grad.data_ptr() + 4 * seg.offset
            mx = int(bool(g.get("maximize", False)))
            wd = float(g.get("weight_decay", 0.0))
            if kind in ("adam", "adamw"):
                m = self._state_buf(gi, "exp_avg").data_ptr() + 4 * seg.offset
                v = self._state_buf(gi, "exp_avg_sq").data_ptr() + 4 * seg.offset
                _lib.check(lib.detb200_optim_tick(gs.step.data_ptr(), skp, st), "optim_tick")
                b1, b2 = g["betas"]
                _lib.check(lib.detb200_adam_flat(p, gr, m, v, None, seg.numel, gs.lr.data_ptr(), gs.step.data_ptr(),
                                                 gsp, skp, float(b1), float(b2), float(g["eps"]), wd,
                                                 int(kind == "adamw"), float(static_scale), int(zero_grad), mx, st),
                           "adam_flat")
            elif kind == "sgd":
                mom = float(g.get("momentum", 0.0))
                buf = None
                if mom != 0.0:
                    buf = self._state_buf(gi, "momentum_buffer").data_ptr() + 4 * seg.offset
                _lib.check(lib.detb200_optim_tick(gs.step.data_ptr(), skp, st), "optim_tick")
                _lib.check(lib.detb200_sgd_flat(p, gr, buf, seg.numel, gs.lr.data_ptr(), gs.step.data_ptr(), gsp, skp,
                                                mom, float(g.get("dampening", 0.0)), wd, int(bool(g.get("nesterov", False))),
                                                float(static_scale), int(zero_grad), mx, st), "sgd_flat")
            elif kind == "adadelta":
                sq = self._state_buf(gi, "square_avg").data_ptr() + 4 * seg.offset
                acc = self._state_buf(gi, "acc_delta").data_ptr() + 4 * seg.offset
                _lib.check(lib.detb200_optim_tick(gs.step.data_ptr(), skp, st), "optim_tick")
                _lib.check(lib.detb200_adadelta_flat(p, gr, sq, acc, seg.numel, gs.lr.data_ptr(), gsp, skp,
                                                     float(g["rho"]), float(g["eps"]), wd, float(static_scale),
                                                     int(zero_grad), mx, st), "adadelta_flat")
        return loss

    def zero_grad(self, set_to_none: bool = True) -> None:
        # gradients are persistent views into the flat communication buffer: one memset
        self._detb200_layout.flat_grad.zero_()

    # -- checkpoint compatibility with stock torch optimizers ------------------------------------
    def state_dict(self) -> Dict[str, Any]:
        """Same structure as the stock torch optimizer's (so either can load the other's): state
        tensors are detached copies, and a per-parameter ``step`` is synthesised from the group's
        device-side counter."""
        steps = [int(gs.step.item()) for gs in self._detb200_groups]
        sd = super().state_dict()  # type: ignore[misc]
        group_of = {}
        for gi, g in enumerate(sd["param_groups"]):
            for pid in g["params"]:
                group_of[pid] = gi
        new_state: Dict[Any, Dict[str, Any]] = {}
        for pid, st in sd["state"].items():
            entry = {k: (v.detach().clone() if isinstance(v, torch.Tensor) else v) for k, v in st.items()}
            if self._detb200_kind != "sgd" and entry:
                entry["step"] = torch.tensor(float(steps[group_of.get(pid, 0)]))
            if entry:
                new_state[pid] = entry
        sd["state"] = new_state
        sd["detb200_fused"] = {"kind": self._detb200_kind, "steps": steps}
        return sd

    def load_state_dict(self, state_dict: Dict[str, Any]) -> None:
        sd = dict(state_dict)
        fused_meta = sd.pop("detb200_fused", None)
        groups = sd["param_groups"]
        if len(groups) != len(self.param_groups):  # type: ignore[attr-defined]
            raise ValueError("loaded state dict has a different number of parameter groups")
        saved_state = sd.get("state", {})
        # hyper-parameters
        for g, sg in zip(self.param_groups, groups):  # type: ignore[attr-defined]
            if len(g["params"]) != len(sg["params"]):
                raise ValueError("loaded state dict contains a parameter group that doesn't match the size of optimizer's group")
            for k, v in sg.items():
                if k != "params":
                    g[k] = v
        names = {"adam": ("exp_avg", "exp_avg_sq"), "adamw": ("exp_avg", "exp_avg_sq"),
                 "sgd": ("momentum_buffer",), "adadelta": ("square_avg", "acc_delta")}[self._detb200_kind]
        steps = [0] * len(self._detb200_groups)
        for gi, (g, sg) in enumerate(zip(self.param_groups, groups)):  # type: ignore[attr-defined]
            for p, pid in zip(g["params"], sg["params"]):
                st = saved_state.get(pid)
                if st is None:
                    continue
                slot = self._detb200_layout.slot_of(p)
                if slot is None:
                    continue
                for name in names:
                    if st.get(name) is not None:
                        self._state_buf(gi, name)
                        self.state[p][name].copy_(st[name])  # type: ignore[attr-defined]
                if "step" in st:
                    steps[gi] = max(steps[gi], int(float(st["step"])))
        if fused_meta and "steps" in fused_meta and len(fused_meta["steps"]) == len(steps):
            steps = [max(a, int(b)) for a, b in zip(steps, fused_meta["steps"])]
        if self._detb200_kind == "sgd" and not any(steps):
            # torch SGD stores no step; a loaded momentum buffer means "not the first step"
            steps = [1 if any("momentum_buffer" in saved_state.get(pid, {}) for pid in sg["params"]) else 0
                     for sg in groups]
        for gs, n in zip(self._detb200_groups, steps):
            gs.step.fill_(n)
            gs.lr_host = None
        self.sync_hparams()


def fuse_optimizer_(opt: torch.optim.Optimizer, layout: FlatLayout) -> bool:
    """Convert ``opt`` in place.  Returns False (and leaves it untouched) if unsupported."""
    if not can_fuse(opt):
        return False
    kind = _kind(opt)
    assert kind is not None
    if opt.state:
        # carry over existing state (optimizer was stepped before wrapping): rare, keep generic
        logger.info("optimizer already has state; copying it into flat buffers")
    old_state = {p: dict(s) for p, s in opt.state.items()}
    base = type(opt)
    opt.__class__ = type("Fused" + base.__name__, (_FusedMixin, base), {})
    opt._detb200_kind = kind            # type: ignore[attr-defined]
    opt._detb200_layout = layout        # type: ignore[attr-defined]
    opt._detb200_groups = [_GroupState(layout.device) for _ in opt.param_groups]  # type: ignore[attr-defined]
    opt.state.clear()
    if old_state:
        names = {"adam": ("exp_avg", "exp_avg_sq"), "adamw": ("exp_avg", "exp_avg_sq"),
                 "sgd": ("momentum_buffer",), "adadelta": ("square_avg", "acc_delta")}[kind]
        for s in layout.slots:
            st = old_state.get(s.param)
            if not st:
                continue
            for name in names:
                if st.get(name) is not None:
                    opt._state_buf(s.group, name)  # type: ignore[attr-defined]
                    opt.state[s.param][name].copy_(st[name])
            if "step" in st:
                opt._detb200_groups[s.group].step.fill_(int(float(st["step"])))  # type: ignore[attr-defined]
            elif kind == "sgd" and st.get("momentum_buffer") is not None:
                opt._detb200_groups[s.group].step.fill_(1)  # type: ignore[attr-defined]
    opt.sync_hparams()                  # type: ignore[attr-defined]
    return True


# ---------------------------------------------------------------------------------------------
# gradient norm / clipping helpers
# ---------------------------------------------------------------------------------------------
class NormScratch:
    """Per-device scratch + result scalars for :func:`sumsq`."""

    def __init__(self, device: torch.device) -> None:
        lib = _lib.load()
        self.scratch = torch.zeros(int(lib.detb200_sumsq_scratch_bytes()), dtype=torch.uint8, device=device)
        self.sumsq = torch.zeros(1, dtype=torch.float32, device=device)
        self.coef = torch.ones(1, dtype=torch.float32, device=device)
        self.norm = torch.zeros(1, dtype=torch.float32, device=device)
        self.nonfinite = torch.zeros(1, dtype=torch.int32, device=device)


def sumsq(x: torch.Tensor, scratch: NormScratch, accumulate: bool = False, check_finite: bool = False) -> torch.Tensor:
    """``scratch.sumsq (+)= sum(x**2)`` for a flat fp32 tensor; returns ``scratch.sumsq``."""
    assert x.dtype == torch.float32 and x.is_cuda and x.is_contiguous()
    lib = _lib.load()
    _lib.check(lib.detb200_sumsq(x.data_ptr(), x.numel(), scratch.scratch.data_ptr(), scratch.sumsq.data_ptr(),
                                 scratch.nonfinite.data_ptr() if check_finite else None, int(accumulate),
                                 _lib.stream_ptr()), "sumsq")
    return scratch.sumsq


def clip_coef(terms: torch.Tensor, max_norm: float, scratch: NormScratch, extra_scale: float = 1.0) -> torch.Tensor:
    """``scratch.coef = extra_scale*min(1, max_norm/(sqrt(sum(terms))+1e-6))``; also fills ``scratch.norm``."""
    lib = _lib.load()
    _lib.check(lib.detb200_clip_coef(terms.data_ptr(), terms.numel(), float(max_norm), float(extra_scale),
                                     scratch.coef.data_ptr(), scratch.norm.data_ptr(), _lib.stream_ptr()), "clip_coef")
    return scratch.coef
