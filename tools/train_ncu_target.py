"""One compute_loss forward + backward at the cfg5 shape: the command ncu wraps for the training launch list."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from matcha_b200 import synthetic as weights  # noqa: E402
from matcha_b200._lib import Plan  # noqa: E402

B, T = 32, 512
dev = torch.device("cuda")
g = torch.Generator().manual_seed(2)
lengths = [int(v) for v in (torch.randint(64, 129, (B,), generator=g) * 4)]
x1, mu, z = (torch.randn(B, 80, T, generator=g).to(dev) for _ in range(3))
tt = torch.rand(B, 1, 1, generator=g).to(dev)
mask = weights.prefix_mask(lengths, T).to(dev)
plan = Plan("fp32", 160, use_graph=False)
named = [(k[len("estimator."):], v.to(dev).contiguous()) for k, v in weights.estimator_state_dict(160, seed=9).items()]
out = plan.compute_loss(named, x1, mask, mu, z, tt, None, 1e-4, backward=True, want_dmu=True)
torch.cuda.synchronize()
print("ok", float(out["loss"]))
