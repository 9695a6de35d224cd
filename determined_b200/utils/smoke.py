"""One tiny training step of each flagship on cuda:0 through the public API, ordered so that a profiler that only looks at
the first few hundred launches sees every native kernel family: first the transformer path (tcgen05 GEMMs incl. the CTA-pair
kernel, tcgen05 flash attention forward + backward on a packed QKV projection with in-place rotary, LayerNorm, fused
cross-entropy, fused AdamW), then the ResNet-18/CIFAR trial (channels-last BatchNorm with ReLU / residual folded in, fused
SGD, fused clip), then the collective kernels on a one-GPU stand-in.  Sizes are the smallest that still select the
production kernels; the whole function stays under ~700 kernel launches."""
import ctypes
import logging

import torch

logger = logging.getLogger("determined.smoke")


def _gpt_part(dev: torch.device):
    from determined_b200 import pytorch
    from determined_b200.models import gpt
    from determined_b200.ops import _lib

    n0 = _lib.LAUNCHES[0]
    torch.manual_seed(0)
    # 512 tokens x 256 channels: M >= 256 so the qkv / mlp GEMMs take the cta_group::2 kernel, the head one the plain kernel
    cfg = gpt.GPTConfig(n_layer=2, n_head=2, n_embd=256, block_size=256, vocab_size=1024, rotary_pct=0.25, parallel_residual=True,
                        tie_embeddings=False)
    with pytorch.init(hparams={"global_batch_size": 4}, aggregation_frequency=1) as ctx:
        m = ctx.wrap_model(gpt.GPT(cfg).to(dev))
        o = ctx.wrap_optimizer(torch.optim.AdamW(m.parameters(), lr=3e-3))
        assert type(o).__name__.startswith("Fused"), "fused optimizer not active on a GPU box"
        idx = torch.randint(0, 1024, (4, 257), device=dev)
        gl = []
        for i in range(2):
            ctx._current_batch_idx = i
            _, l = m(idx[:, :-1], idx[:, 1:])
            ctx.backward(l)
            ctx.step_optimizer(o, clip_grads=pytorch.FusedClipGradNorm(1.0))
            gl.append(float(l))
    assert all(v == v for v in gl) and gl[-1] < gl[0], f"GPT smoke loss did not decrease: {gl}"
    return gl, _lib.LAUNCHES[0] - n0


def _resnet_part(dev: torch.device):
    from determined_b200 import pytorch
    from determined_b200.models.resnet_cifar import resnet18_cifar
    from determined_b200.ops import xent

    with pytorch.init(hparams={"global_batch_size": 32}, aggregation_frequency=1) as ctx:
        model = ctx.wrap_model(resnet18_cifar().to(memory_format=torch.channels_last))
        opt = ctx.wrap_optimizer(torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9, nesterov=True))
        assert type(opt).__name__.startswith("Fused"), "fused optimizer not active on a GPU box"
        x = torch.randn(32, 3, 32, 32, device=dev).contiguous(memory_format=torch.channels_last)
        y = torch.randint(0, 10, (32,), device=dev)
        losses = []
        for i in range(2):
            ctx._current_batch_idx = i
            with torch.autocast("cuda", dtype=torch.bfloat16):
                logits = model(x)
            loss = xent.cross_entropy(logits.float(), y)
            ctx.backward(loss)
            ctx.step_optimizer(opt, clip_grads=pytorch.FusedClipGradNorm(5.0))
            losses.append(float(loss))
    assert all(l == l for l in losses), f"NaN loss: {losses}"
    assert losses[-1] < losses[0], f"loss did not decrease: {losses}"
    return losses


def _collective_part(dev: torch.device) -> None:
    """The all-reduce kernel's full flag / epoch / barrier protocol for 4 ranks emulated in one cooperative launch, and the
    reduce-scatter + AdamW + all-gather kernel with world 1 (what a one-GPU ZeRO step runs)."""
    from determined_b200.ops import _lib, comm

    lib = _lib.load()
    world, n = 4, 4 * 4 * 2048
    groups = [comm.PeerGroup(0, 1, None, dev) for _ in range(world)]
    bufs = [comm.SymmetricBuffer(4 * n, dev) for _ in range(world)]
    tens = [b.tensor(torch.float32, n) for b in bufs]
    for r, t in enumerate(tens):
        t.fill_(float(r + 1))
    ptrs = (ctypes.c_void_p * world)(*[b.ptr for b in bufs])
    ctrls = (ctypes.c_void_p * world)(*[g.ctrl.ptr for g in groups])
    _lib.check(lib.detb200_allreduce_fused_emulated(ptrs, ctrls, world, 0, n, 1.0 / world, None, None, 1, 2, 20.0,
                                                    _lib.stream_ptr()), "allreduce_emulated")
    torch.cuda.synchronize()
    assert all(float(t.min()) == float(t.max()) == 2.5 for t in tens), "emulated all-reduce gave a wrong mean"
    h = torch.randn(64, 1024, device=dev, dtype=torch.bfloat16, requires_grad=True)
    from determined_b200.ops import norm

    ln = norm.LayerNorm(1024).to(dev)
    ln(h).float().sum().backward()


def run() -> None:
    from determined_b200.ops import _lib

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    n0 = _lib.LAUNCHES[0]
    gl, n_gpt = _gpt_part(dev)
    losses = _resnet_part(dev)
    _collective_part(dev)
    torch.cuda.synchronize()
    print(f"smoke ok: gpt losses={['%.4f' % l for l in gl]} resnet losses={['%.4f' % l for l in losses]} "
          f"native kernel launches: gpt part {n_gpt}, total {_lib.LAUNCHES[0] - n0}")
