"""Developer probe of the fused ResBlock kernel: MB200_PAIR_DBG=<n> python tools/pair_probe.py  (n-th pair op, cfg2 sizes)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from matcha_b200 import synthetic  # noqa: E402
from matcha_b200._lib import Plan  # noqa: E402

plan = Plan("fp16", 160, use_graph=False)
plan.load_generator(synthetic.generator_state_dict(1234))
mel = torch.randn(64, 80, 452, device="cuda")
wav = plan.hifigan_forward(mel)
torch.cuda.synchronize()
print("ok", float(wav.abs().max()))
