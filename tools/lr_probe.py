"""Timing of the length-regulation kernels (mb200_length_regulate) against the torch formulation of the reference
(generate_path + dense matmul) on the GPU, at BASELINE cfg2 / cfg4 shapes.  `python tools/lr_probe.py`"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from matcha_b200 import _lib  # noqa: E402
from matcha_b200 import matcha_tts as mt  # noqa: E402


def timed(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def reference_formulation(mu_x, w, x_mask, y_len, t_pad):
    y_mask = mt.sequence_mask(y_len, t_pad).unsqueeze(1).to(x_mask.dtype)
    attn_mask = x_mask.unsqueeze(1).unsqueeze(-1) * y_mask.unsqueeze(2)
    attn = mt.generate_path(w, attn_mask.squeeze(1))
    return torch.matmul(attn.transpose(1, 2), mu_x.transpose(1, 2)).transpose(1, 2), attn


out = []
for name, b, lx in (("cfg2", 64, 301), ("cfg4", 16, 2001)):
    w = torch.full((b, lx), 1.5, device="cuda")
    x_mask = torch.ones(b, lx, device="cuda")
    y_len = torch.clamp_min(w.sum(1), 1).long()
    t_pad = mt.fix_len_compatibility(int(y_len.max()))
    mu_x = torch.randn(b, 80, lx, device="cuda")
    ms_k = timed(lambda: _lib.length_regulate(mu_x, w, x_mask, y_len, t_pad))
    ms_t = timed(lambda: reference_formulation(mu_x, w, x_mask, y_len, t_pad))
    bytes_moved = 4 * (b * lx * t_pad + 2 * b * 80 * t_pad + b * 80 * lx)
    out.append({"case": name, "B": b, "Lx": lx, "T_pad": t_pad, "kernel_ms": round(ms_k, 4), "torch_formulation_ms": round(ms_t, 4),
                "algorithmic_MB": round(bytes_moved / 1e6, 1), "GB_per_s": round(bytes_moved / ms_k / 1e6, 1)})
print(json.dumps(out))
