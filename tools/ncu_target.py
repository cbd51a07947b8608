"""One eager (no CUDA graph) pass of a bench workload's hot path: the command ncu wraps.
    python tools/ncu_target.py [precision] [config]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from matcha_b200 import synthetic  # noqa: E402
from matcha_b200._lib import Plan  # noqa: E402

prec = sys.argv[1] if len(sys.argv) > 1 else "fp16"
cfg = sys.argv[2] if len(sys.argv) > 2 else "cfg2"
c = bench.CONFIGS[cfg]
plan = Plan(prec, c["c_in"], use_graph=False)
plan.load_estimator(synthetic.estimator_state_dict(c["c_in"], seed=1234))
plan.load_generator(synthetic.generator_state_dict(1234))
batches, frames, info = bench.workload(cfg, 0, 1)
mean, std = c["stats"]
for bt in batches[:1]:
    dev = lambda v: None if v is None else v.cuda()  # noqa: E731
    dec, mel, wav = plan.synthesise(dev(bt["mu"]), dev(bt["mask"]), dev(bt["z"]), c["n_steps"], bt["length"], mean, std, spks=dev(bt["spks"]))
torch.cuda.synchronize()
print("ok", float(wav.abs().max()), plan.launch_count())
