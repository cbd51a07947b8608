"""Single-launch timing of the attention kernel (MB200_PROBE=reps) over (B, L) shapes; MB200_ATTN=mma for the old kernel."""
import os
import sys

os.environ.setdefault("MB200_PROBE", "20")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from matcha_b200._lib import Plan  # noqa: E402

shapes = [(16, 3004), (16, 1502), (64, 452), (64, 226), (32, 904), (32, 452), (1, 452), (16, 128), (16, 256), (16, 1024)]
for b, l in shapes:
    g = torch.Generator().manual_seed(0)
    qkv = torch.randn(b, l, 384, generator=g)
    qkv[:, :, :128] *= 0.3
    plan = Plan("fp16", 160, use_graph=False)
    plan.test_attention(qkv.cuda(), torch.ones(b, l).cuda())
    items = b * 2 * ((l + 255) // 256)
    print(f"   items {items} ({items / 148:.2f} per CTA), key tiles {(l + 127) // 128}", file=sys.stderr)
