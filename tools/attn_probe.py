"""A/B of the attention kernels inside the real decode (MB200_ATTN=mma | umma): event-timed attention region per Euler
decode at cfg2 (B=64, T=452) and cfg4 (B=16, T=3004), from the plan's built-in profiler.  Writes gpurun_out/attn_probe.json."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from matcha_b200 import synthetic as weights  # noqa: E402
from matcha_b200._lib import Plan  # noqa: E402

out = {}
sd = weights.estimator_state_dict(160, seed=1234)
for (name, b, t, steps) in [("cfg2", 64, 452, 10), ("cfg4", 16, 3004, 10)]:
    g = torch.Generator().manual_seed(1)
    mu, z = torch.randn(b, 80, t, generator=g).cuda(), torch.randn(b, 80, t, generator=g).cuda()
    mask = weights.prefix_mask([t - 1] * b, t).cuda()
    res = {}
    for kern in ("mma", "umma"):
        os.environ["MB200_ATTN"] = kern
        plan = Plan("fp16", 160)
        plan.load_estimator(sd)
        res[kern + "_dec"] = plan.cfm_decode(mu, mask, z, steps)
        plan.set_profiling(True)
        plan.cfm_decode(mu, mask, z, steps)
        torch.cuda.synchronize()
        prof = plan.get_profile()
        plan.set_profiling(False)
        n, ms, fl = prof[(0, 1)]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            plan.cfm_decode(mu, mask, z, steps)
        e1.record()
        torch.cuda.synchronize()
        out[f"{name}_{kern}"] = {"attention_launches": n, "attention_ms": ms, "attention_tflops": fl / ms / 1e9,
                                 "decode_ms": e0.elapsed_time(e1) / 3}
        del plan
    d = (res["mma_dec"].double() - res["umma_dec"].double()).norm() / res["mma_dec"].double().norm()
    out[f"{name}_rel_l2_between_kernels"] = float(d)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/attn_probe.json", "w"), indent=1)
print(json.dumps(out, indent=1))
