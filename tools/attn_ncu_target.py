"""ncu target: one tcgen05 attention launch at cfg4's T-level shape (B=16, L=3004, fp16) through the C-ABI test hook."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from matcha_b200._lib import Plan  # noqa: E402

b, l = 16, int(os.environ.get("ATTN_L", "3004"))
g = torch.Generator().manual_seed(0)
qkv = torch.randn(b, l, 384, generator=g)
qkv[:, :, :128] *= 0.3
plan = Plan("fp16", 160, use_graph=False)
y = plan.test_attention(qkv.cuda(), torch.ones(b, l).cuda())
torch.cuda.synchronize()
print("ok", float(y.abs().sum()))
