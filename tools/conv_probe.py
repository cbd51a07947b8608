"""Developer probe: times single conv launches of the persistent kernel (vocoder-shaped) with pipeline stages knocked
out (needs a library built with MB200_PROBE_BUILD=1).  Each configuration runs in its own process because the knobs are
read from the environment when the plan builds the launch.

    MB200_PROBE_BUILD=1 python -m matcha_b200.build && python tools/conv_probe.py
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

CASES = {  # name: (B, C, T, K, dil)
    "s2k11": (8, 128, 28928, 11, 1),
    "s2k3": (8, 128, 28928, 3, 1),
    "s3k11": (4, 64, 115712, 11, 1),
    "s4k11": (2, 32, 462848, 11, 1),
    "s4k3": (2, 32, 462848, 3, 1),
    "s1k11": (16, 256, 3616, 11, 1),
    "e_c3": (64, 256, 452, 3, 1),          # estimator Block1D conv at full resolution
    "e_ff1": (64, 256, 452, 1, 1, 1024),   # feed-forward up projection
    "e_ff2": (64, 1024, 452, 1, 1, 256),
    "e_c3_mid": (64, 256, 113, 3, 1),
    "e_c3_b256": (256, 256, 452, 3, 1),     # same layer, 4x the tiles: steady-state cost per tile
    "e_c3_b512": (512, 256, 452, 3, 1),
    "e_c3_half": (64, 256, 226, 3, 1),      # T/2 level: 128 tiles, one per CTA
}


def one(name):
    import torch
    from matcha_b200._lib import Plan
    b, c, t, k, dil = CASES[name][:5]
    cout = CASES[name][5] if len(CASES[name]) > 5 else c
    plan = Plan("fp16", 160, use_graph=False)
    g = torch.Generator().manual_seed(0)
    x = torch.randn(b, c, t, generator=g)
    w = torch.randn(cout, c, k, generator=g) / (c * k) ** 0.5
    bias = torch.randn(cout, generator=g)
    os.environ["MB200_PROBE"] = os.environ.get("MB200_PROBE", "20")
    only = os.environ.get("CONV_PROBE_ONLY")   # "epi,dbg": run a single configuration (for ncu)
    for epi, dbg in [tuple(only.split(","))] if only else [("0", "0"), ("2", "0"), ("2", "1"), ("2", "2"), ("2", "4"), ("2", "8"), ("2", "12"), ("2", "3"), ("2", "7"), ("2", "15")]:
        os.environ["MB200_PROBE_EPI"] = epi
        os.environ["MB200_DBG"] = dbg
        sys.stderr.write(f"{name} epi={epi} dbg={dbg:>2} ")
        sys.stderr.flush()
        plan.test_conv1d(x, w, bias, dilation=dil)


if __name__ == "__main__":
    for name in (sys.argv[1:] or list(CASES)):
        one(name)
