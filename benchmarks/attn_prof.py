"""Phase timing of one softmax / compute warp of the attention kernels (DETB200_ATTN_PROF=1 python benchmarks/attn_prof.py)."""
import math
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from determined_b200.ops import attention as A  # noqa: E402

q = torch.randn(8, 16, 2048, 128, device="cuda").to(torch.bfloat16)
for _ in range(2):
    A.attention_forward(q, q, q, False, 1 / math.sqrt(128), "bhsd")
qkv = torch.randn(8, 2048, 3, 16, 128, device="cuda").to(torch.bfloat16)
o, lse = A.attention_forward(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], False, 1 / math.sqrt(128), "bshd")
dqkv = torch.empty_like(qkv)
for _ in range(2):
    A.attention_backward_packed(qkv, o, lse, torch.randn_like(o), dqkv, False, 1 / math.sqrt(128))
torch.cuda.synchronize()
