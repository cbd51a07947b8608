"""Per-phase cycle counters of the tcgen05 attention kernel (MB200_ATTN_DBG=1), CTA 0, one launch at B=16, L given.  Needs a MB200_PROBE_BUILD=1 build of the library.
softmax warp 2 (query tile 0): [mask vote, wait S, streaming pass, two-pass redo, wait PV, store+arrive, -, -, -, item epilogue]
issuer: [item set-up, polling loop]"""
import os
import sys

os.environ["MB200_ATTN_DBG"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from matcha_b200._lib import Plan  # noqa: E402

for l in (3004, 1502, 452):
    b = 16
    g = torch.Generator().manual_seed(0)
    qkv = torch.randn(b, l, 384, generator=g)
    qkv[:, :, :128] *= 0.3
    plan = Plan("fp16", 160, use_graph=False)
    plan.test_attention(qkv.cuda(), torch.ones(b, l).cuda())
    c = plan.debug_probe(2)
    items = b * 2 * ((l + 255) // 256)
    n_items = (items + 147) // 148
    tiles = n_items * ((l + 127) // 128)
    names = ["mask_vote", "wait_S", "stream", "redo", "wait_PV", "store+arrive", "-", "-", "-", "epilogue"]
    print(f"L={l}: items/CTA {n_items}, key tiles {tiles}")
    print("  softmax per tile:", {n: round(v / tiles) for n, v in zip(names, c[0][:10])}, "total", round(sum(c[0][:10]) / tiles))
    print("  issuer per tile:", {n: round(v / tiles) for n, v in zip(["setup", "loop"], c[1][:2])})
