"""One eager (no CUDA graph) pass of the cfg2 hot path: the command ncu wraps.  `python tools/ncu_target.py [precision]`"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from bench import B, MEL_MEAN, MEL_STD, N_STEPS, T_VALID, make_inputs  # noqa: E402
from matcha_b200 import synthetic  # noqa: E402
from matcha_b200._lib import Plan  # noqa: E402

prec = sys.argv[1] if len(sys.argv) > 1 else "fp16"
plan = Plan(prec, 160, use_graph=False)
plan.load_estimator(synthetic.estimator_state_dict(160, seed=1234))
plan.load_generator(synthetic.generator_state_dict(1234))
mu, mask, z = (t.cuda() for t in make_inputs(B))
dec, mel, wav = plan.synthesise(mu, mask, z, N_STEPS, T_VALID, MEL_MEAN, MEL_STD)
torch.cuda.synchronize()
print("ok", float(wav.abs().max()), plan.launch_count())
