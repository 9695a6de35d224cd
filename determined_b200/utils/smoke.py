"""One tiny forward+backward+fused-step of the flagship (ResNet-18/CIFAR trial) on cuda:0, through
the public API, plus one call of every other native kernel family so a broken build shows up here."""
import logging

import torch

logger = logging.getLogger("determined.smoke")


def run() -> None:
    from determined_b200 import pytorch
    from determined_b200.models.resnet_cifar import resnet18_cifar
    from determined_b200.ops import norm, xent

    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    with pytorch.init(hparams={"global_batch_size": 32}, aggregation_frequency=1) as ctx:
        model = ctx.wrap_model(resnet18_cifar().to(memory_format=torch.channels_last))
        opt = ctx.wrap_optimizer(torch.optim.SGD(model.parameters(), lr=0.05, momentum=0.9, nesterov=True))
        assert type(opt).__name__.startswith("Fused"), "fused optimizer not active on a GPU box"
        x = torch.randn(32, 3, 32, 32, device=dev).contiguous(memory_format=torch.channels_last)
        y = torch.randint(0, 10, (32,), device=dev)
        losses = []
        for i in range(3):
            ctx._current_batch_idx = i
            with torch.autocast("cuda", dtype=torch.bfloat16):
                logits = model(x)
            loss = xent.cross_entropy(logits.float(), y)
            ctx.backward(loss)
            ctx.step_optimizer(opt, clip_grads=pytorch.FusedClipGradNorm(5.0))
            losses.append(float(loss))
        assert all(l == l for l in losses), f"NaN loss: {losses}"
        assert losses[-1] < losses[0], f"loss did not decrease: {losses}"
    h = torch.randn(64, 1024, device=dev, dtype=torch.bfloat16, requires_grad=True)
    ln = norm.LayerNorm(1024).to(dev)
    ln(h).float().sum().backward()
    # the transformer path: tcgen05 GEMMs (incl. the CTA-pair kernel and fused MLP epilogues), tcgen05 flash attention
    # forward + backward on a packed QKV projection with in-place rotary, LayerNorm, fused cross-entropy
    from determined_b200.models import gpt
    from determined_b200.ops import _lib

    n0 = _lib.LAUNCHES[0]
    torch.manual_seed(0)
    m = gpt.GPT(gpt.GPTConfig(n_layer=2, n_head=2, n_embd=256, block_size=256, vocab_size=1024, rotary_pct=0.25, parallel_residual=True,
                              tie_embeddings=False)).to(dev)
    o = torch.optim.AdamW(m.parameters(), lr=3e-3)
    idx = torch.randint(0, 1024, (4, 257), device=dev)
    gl = []
    for _ in range(3):
        _, l = m(idx[:, :-1], idx[:, 1:])
        o.zero_grad()
        l.backward()
        o.step()
        gl.append(float(l))
    assert all(v == v for v in gl) and gl[-1] < gl[0], f"GPT smoke loss did not decrease: {gl}"
    torch.cuda.synchronize()
    print(f"smoke ok: resnet losses={['%.4f' % l for l in losses]} gpt losses={['%.4f' % l for l in gl]} "
          f"native kernel launches in the GPT part: {_lib.LAUNCHES[0] - n0}")
