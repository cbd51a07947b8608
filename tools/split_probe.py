"""Experiment: one plan at B = 64 against two plans at B = 32 on two streams (tail filling between independent chains).
`python tools/split_probe.py`"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import CONFIGS  # noqa: E402
from matcha_b200 import synthetic  # noqa: E402
from matcha_b200._lib import Plan  # noqa: E402

est, gen = synthetic.estimator_state_dict(160, seed=1234), synthetic.generator_state_dict(1234)
B, T_VALID, N_STEPS = CONFIGS["cfg2"]["batch"], CONFIGS["cfg2"]["t_valid"], CONFIGS["cfg2"]["n_steps"]
MEL_MEAN, MEL_STD = CONFIGS["cfg2"]["stats"]
T_PAD = (T_VALID + 3) // 4 * 4
g = torch.Generator().manual_seed(0)
mu, z = torch.randn(B, 80, T_PAD, generator=g).cuda(), torch.randn(B, 80, T_PAD, generator=g).cuda()
mask = torch.ones(B, 1, T_PAD).cuda()


def mk():
    p = Plan("fp16", 160)
    p.load_estimator(est)
    p.load_generator(gen)
    return p


def timed(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    e.record()
    torch.cuda.synchronize()
    return a.elapsed_time(e) / n


res = {}
p0 = mk()
res["one_plan_B64_ms"] = timed(lambda: p0.synthesise(mu, mask, z, N_STEPS, T_VALID, MEL_MEAN, MEL_STD))
for parts in (2, 4):
    plans = [mk() for _ in range(parts)]
    streams = [torch.cuda.Stream() for _ in range(parts)]
    h = B // parts
    sl = [(mu[i * h:(i + 1) * h].contiguous(), mask[i * h:(i + 1) * h].contiguous(), z[i * h:(i + 1) * h].contiguous()) for i in range(parts)]

    def run():
        cur = torch.cuda.current_stream()
        for p, s, (a, b, c) in zip(plans, streams, sl):
            s.wait_stream(cur)
            with torch.cuda.stream(s):
                p.synthesise(a, b, c, N_STEPS, T_VALID, MEL_MEAN, MEL_STD)
        for s in streams:
            cur.wait_stream(s)

    res[f"{parts}_plans_B{h}_ms"] = timed(run)
    del plans
print(json.dumps(res))
