"""Developer probe of the fused res_conv/GroupNorm/LayerNorm/QKV kernel: MB200_RQ_DBG=<n> python tools/rq_probe.py
(n-th op built at cfg2 sizes: 0, 5 = T level).  Prints the epilogue warp's timestamps (cycles since its start) for tile 0."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from matcha_b200 import synthetic  # noqa: E402
from matcha_b200._lib import Plan  # noqa: E402

os.environ.setdefault("MB200_RQ_DBG", "1")
plan = Plan("fp16", 160, use_graph=False)
plan.load_estimator(synthetic.estimator_state_dict(160, seed=1234))
B, T_PAD = 64, 452
g = torch.Generator().manual_seed(0)
mu, z = torch.randn(B, 80, T_PAD, generator=g).cuda(), torch.randn(B, 80, T_PAD, generator=g).cuda()
mask = torch.ones(B, 1, T_PAD).cuda()
for _ in range(3):
    plan.estimator_forward(z, mask, mu, torch.tensor(0.3))
torch.cuda.synchronize()
rows = plan.debug_probe(148)
names = ["requests issued", "M0 done (t0_full)", "E0 math+res store done", "x tile written", "M1 done (t1_full)", "E1 done", "all tiles"]
for cta in (0, 1, 77):
    print(f"[cta {cta}] " + ", ".join(f"{n} {rows[cta][i]}" for i, n in enumerate(names)))
