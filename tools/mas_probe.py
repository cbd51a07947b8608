"""Timing of the CUDA monotonic alignment search against the reference's Cython/OpenMP code (oracle/_ref) including the
device -> host -> device round trip the reference wrapper makes.  `python tools/mas_probe.py`"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from matcha_b200 import monotonic_align  # noqa: E402
from oracle.build_ref import load_ref_mas  # noqa: E402

core = load_ref_mas()
out = []
for name, b, tx, ty in (("training batch (B=32, 170 tokens, 512 frames)", 32, 170, 512), ("long (B=16, 667 tokens, 2001 frames)", 16, 667, 2001)):
    g = torch.Generator().manual_seed(1)
    value = (torch.randn(b, tx, ty, generator=g) * 40 - 120).cuda()
    mask = torch.ones(b, tx, ty, device="cuda")
    for _ in range(3):
        path = monotonic_align.maximum_path(value, mask)
    torch.cuda.synchronize()
    a, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        path = monotonic_align.maximum_path(value, mask)
    e.record()
    torch.cuda.synchronize()
    row = {"case": name, "kernel_ms": round(a.elapsed_time(e) / 10, 4)}
    if core is not None:
        def ref():
            v = (value * mask).cpu().numpy().astype(np.float32)
            p = np.zeros_like(v).astype(np.int32)
            m = mask.cpu().numpy()
            core.maximum_path_c(p, v, m.sum(1)[:, 0].astype(np.int32), m.sum(2)[:, 0].astype(np.int32))
            return torch.from_numpy(p).to(device="cuda", dtype=torch.float32)
        ref()
        t0 = time.perf_counter()
        for _ in range(3):
            rp = ref()
        torch.cuda.synchronize()
        row["reference_cython_roundtrip_ms"] = round((time.perf_counter() - t0) / 3 * 1e3, 3)
        row["identical"] = bool(torch.equal(rp, path))
        row["host_cores"] = os.cpu_count()
    out.append(row)
print(json.dumps(out))
