"""Developer probe of the fused feed-forward kernel: MB200_FF_DBG=<n> python tools/ff_probe.py  (n-th ff op built; cfg2 sizes:
n = 0, 5 are the T-level blocks, 1..4 the T/2-level ones).  Prints per-role cycle counters of a few CTAs."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from matcha_b200 import synthetic  # noqa: E402
from matcha_b200._lib import Plan  # noqa: E402

os.environ.setdefault("MB200_FF_DBG", "0")
plan = Plan("fp16", 160, use_graph=False)
plan.load_estimator(synthetic.estimator_state_dict(160, seed=1234))
B, T_PAD = 64, 452
g = torch.Generator().manual_seed(0)
mu, z = torch.randn(B, 80, T_PAD, generator=g).cuda(), torch.randn(B, 80, T_PAD, generator=g).cuda()
mask = torch.ones(B, 1, T_PAD).cuda()
for _ in range(3):
    v = plan.estimator_forward(z, mask, mu, torch.tensor(0.3))
torch.cuda.synchronize()
rows = plan.debug_probe(148)
names = ["prod total", "prod wait x_empty", "prod wait w_empty", "mma total", "mma wait x_full", "mma wait t1_empty", "mma wait w_full",
         "mma wait mid_full", "mma wait t2_empty", "epi total", "epi wait t1_full", "epi wait mid_empty", "epi wait t2_full", "tiles"]
for cta in (0, 1, 77, 147):
    print(f"[cta {cta}] " + ", ".join(f"{n} {rows[cta][i]}" for i, n in enumerate(names)))
