"""Developer check: fused vs two-launch ResBlock steps, where and how much they differ."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from matcha_b200._lib import Plan
from oracle import weights
prec = sys.argv[1] if len(sys.argv) > 1 else "bf16"
sd = weights.generator_state_dict(7, weight_norm=False)
for (B, L) in [(3, 1), (2, 13), (1, 40)]:
    mel = torch.randn(B, 80, L, generator=torch.Generator().manual_seed(11)) * 1.5 - 4.0
    out = {}
    for fuse in ("1", "0"):
        os.environ["MB200_FUSE_PAIRS"] = fuse
        plan = Plan(prec, 160, use_graph=False)
        plan.load_generator(sd)
        out[fuse] = plan.hifigan_forward(mel).cpu()
        out[fuse + "b"] = plan.hifigan_forward(mel).cpu()
    d = (out["1"] - out["0"]).abs()
    print(prec, B, L, "max diff", float(d.max()), "n diff", int((d > 0).sum()), "of", d.numel(), "self-repeat", float((out["1"] - out["1b"]).abs().max()),
          float((out["0"] - out["0b"]).abs().max()), "first idx", (d > 0).nonzero()[:3].tolist())
