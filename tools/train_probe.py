"""Times CFM.compute_loss forward+backward (csrc/train.cu) at the BASELINE cfg5 shape: B=32, T=512, per-sample t.

    python tools/train_probe.py            (needs a B200)
"""
import json
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from matcha_b200._lib import Plan
from matcha_b200 import synthetic as weights

B, T = 32, 512
dev = torch.device("cuda")
g = torch.Generator().manual_seed(2)
lengths = [int(v) for v in (torch.randint(64, 129, (B,), generator=g) * 4)]
x1, mu, z = (torch.randn(B, 80, T, generator=g).to(dev) for _ in range(3))
tt = torch.rand(B, 1, 1, generator=g).to(dev)
mask = weights.prefix_mask(lengths, T).to(dev)
plan = Plan("fp32", 160, use_graph=False)
named = [(k[len("estimator."):], v.to(dev).contiguous()) for k, v in weights.estimator_state_dict(160, seed=9).items()]
res = {}
for backward in (False, True):
    for _ in range(2):
        plan.compute_loss(named, x1, mask, mu, z, tt, None, 1e-4, backward=backward, want_dmu=backward)
    torch.cuda.synchronize()
    n0 = plan.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        plan.compute_loss(named, x1, mask, mu, z, tt, None, 1e-4, backward=backward, want_dmu=backward)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    res["fwd+bwd" if backward else "fwd"] = {"ms": ms, "launches": (plan.launch_count() - n0) // 5}
valid = sum(lengths)
flop_fwd = B * T * (10_985_472 + 1536 * T)
res["frames_per_s_fwd_bwd"] = valid / (res["fwd+bwd"]["ms"] * 1e-3)
res["tflops_fwd_bwd"] = 3 * flop_fwd / (res["fwd+bwd"]["ms"] * 1e-3) / 1e12
res["config"] = {"B": B, "T": T, "valid_frames": valid, "dtype": "f32 data, GEMMs " + {"tf32x3": "3xTF32 tensor-core (fp32-accurate)", "tf32": "plain TF32 tensor-core", "ffma": "FFMA"}.get(os.environ.get("MB200_TRAIN_GEMM", "tf32x3"), "?"), "gemm": os.environ.get("MB200_TRAIN_GEMM", "tf32x3"), "workspace_gb": plan.workspace_bytes() / 1e9}
print(json.dumps(res))
