import sys, math, torch
sys.path.insert(0, "/root/repo")
from determined_b200.ops import attention as A
q = torch.randn(8, 16, 2048, 128, device="cuda").to(torch.bfloat16)
for _ in range(2):
    A.attention_forward(q, q, q, False, 1 / math.sqrt(128), "bhsd")
torch.cuda.synchronize()
