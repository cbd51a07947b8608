"""Developer probe (MB200_PROBE_BUILD=1 build): per-role and per-epilogue-phase cycle counters of the n-th conv op of the
estimator at cfg2's shape.  usage: python tools/est_op_probe.py <op index> [...]; each index runs in its own process."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def one(idx):
    os.environ["MB200_DBG"] = "0"
    os.environ["MB200_DBG_OP"] = str(idx)
    import torch
    from matcha_b200._lib import Plan
    from matcha_b200 import synthetic as weights
    b, t = 64, 452
    plan = Plan("fp16", 160, use_graph=False)
    plan.load_estimator(weights.estimator_state_dict(160, seed=1234))
    g = torch.Generator().manual_seed(0)
    mu = torch.randn(b, 80, t, generator=g).cuda()
    x = torch.randn(b, 80, t, generator=g).cuda()
    mask = torch.ones(b, 1, t).cuda()
    for _ in range(3):
        plan.estimator_forward(x, mask, mu, torch.full((b,), 0.3).cuda(), None)
    torch.cuda.synchronize()
    c = plan.debug_probe(148)
    names = ["producer", "mma", "epi-w2", "epi-w6", "kernel"]
    for cta in (0, 147):
        r = c[cta]
        print(f"  [cta {cta:3d}] " + " ".join(f"{n} {r[2 * k]}/{r[2 * k + 1]}" for k, n in enumerate(names)))
        print("            epilogue w2: header %d, tmem %d, staging %d, arithmetic %d, stores %d" % tuple(r[10:15]))


if __name__ == "__main__":
    if len(sys.argv) == 3 and sys.argv[1] == "--one":
        one(int(sys.argv[2]))
    else:
        for idx in sys.argv[1:]:
            subprocess.run([sys.executable, __file__, "--one", idx], check=False)
