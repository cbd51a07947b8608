"""TEST INFRASTRUCTURE ONLY -- the oracle (oracle/matcha_oracle.py) run on CUDA tensors at BASELINE.json's full sizes.

SURVEY.md section 8(c) step 4: "oracle runs fp32 on CPU and, on the GPU box, fp32 CUDA with
torch.backends.cuda.matmul.allow_tf32 = False; cudnn.allow_tf32 = False".  The oracle is device-agnostic torch, so the
same functions that are pinned to the reference-generated fixtures on the CPU (tests/test_oracle.py) give the full-size
answer in about a second on the GPU.  Used by tests/test_full_size_gpu.py (parity at cfg2 / cfg3 / cfg4 sizes) and by
bench.py's `library_baseline` leg (stock torch-CUDA eager as "the library bar", SURVEY.md section 8(d) last row; reported,
never the product).  Nothing under matcha_b200/ imports this file.
"""
import contextlib
import time

import torch

from matcha_b200 import synthetic as weights
from oracle import matcha_oracle as mo

LJ = (-5.536622, 2.116101)      # configs/data/ljspeech.yaml:19-20
VCTK = (-6.630575, 2.482914)    # configs/data/vctk.yaml:13-14

# BASELINE.json configs[1..3]; cfg3's B is one GPU's share of the 256 at 4 / 8 GPUs
CONFIGS = {
    "cfg2": dict(b=64, t_valid=451, n_steps=10, c_in=160, stats=LJ),
    "cfg3": dict(b=64, t_valid=451, n_steps=4, c_in=224, stats=VCTK),
    "cfg3_b32": dict(b=32, t_valid=451, n_steps=4, c_in=224, stats=VCTK),
    "cfg4": dict(b=16, t_valid=3001, n_steps=10, c_in=160, stats=LJ),
}


@contextlib.contextmanager
def tf32(enabled):
    """fp32 everywhere (enabled=False: the oracle's arithmetic) or stock TF32 tensor-core eager (the faster library bar)."""
    old = (torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32)
    torch.backends.cuda.matmul.allow_tf32 = torch.backends.cudnn.allow_tf32 = enabled
    try:
        yield
    finally:
        torch.backends.cuda.matmul.allow_tf32, torch.backends.cudnn.allow_tf32 = old


def make_case(name, ragged=True, seed=1234):
    """Seeded inputs of one configuration (CPU tensors) + the state_dicts.  ragged: the last utterances are shorter, so
    padded frames (attended to, inside the GroupNorm statistics) are exercised at full size."""
    c = CONFIGS[name]
    g = torch.Generator().manual_seed(seed)
    b, t_valid = c["b"], c["t_valid"]
    t_pad = (t_valid + 3) // 4 * 4
    n_rag = b // 4 if ragged else 0
    lens = [t_valid] * (b - n_rag) + [int(v) for v in torch.randint(t_valid // 8, t_valid, (n_rag,), generator=g)]
    case = dict(c)
    case.update(name=name, t_pad=t_pad, lens=lens,
                mu=torch.randn(b, 80, t_pad, generator=g), z=torch.randn(b, 80, t_pad, generator=g) * 0.667,
                spks=torch.randn(b, 64, generator=g) if c["c_in"] == 224 else None,
                mask=weights.prefix_mask(lens, t_pad),
                est_sd=weights.estimator_state_dict(c["c_in"], seed=seed), gen_sd=weights.generator_state_dict(seed))
    return case


def oracle_cuda(case, device="cuda", allow_tf32=False):
    """decode -> crop to the longest valid length -> denormalise -> vocoder -> clamp, all by the oracle on `device`.
    Returns (dec (B,80,T_pad), mel (B,80,t_valid), wav (B,1,256 t_valid)) on the device."""
    dev = torch.device(device)
    mv = lambda t: None if t is None else t.to(dev)  # noqa: E731
    if case.get("_dev") != str(dev):     # weights go up once per case, not once per call
        case["_est"] = {k: v.to(dev) for k, v in case["est_sd"].items()}
        case["_gen"] = {k: v.to(dev) for k, v in case["gen_sd"].items()}
        case["_in"] = tuple(mv(case[k]) for k in ("z", "mu", "mask", "spks"))
        case["_dev"] = str(dev)
    est_sd, gen_sd = case["_est"], case["_gen"]
    z, mu, mask, spks = case["_in"]
    mean, std = case["stats"]
    with torch.inference_mode(), tf32(allow_tf32):
        dec = mo.solve_euler(est_sd, z, mu, mask, case["n_steps"], spks)
        mel = dec[:, :, :case["t_valid"]] * std + mean
        wav = mo.hifigan_forward(gen_sd, mel).clamp(-1, 1)
    return dec, mel, wav


def time_library(case, allow_tf32, steps=3, warmup=1, device="cuda"):
    """Stock torch-CUDA eager (cuDNN / cuBLAS / ATen kernels) on the same workload: ms per step, CUDA-event timed."""
    for _ in range(warmup):
        oracle_cuda(case, device, allow_tf32)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(steps):
        oracle_cuda(case, device, allow_tf32)
    e1.record()
    torch.cuda.synchronize()
    return max(e0.elapsed_time(e1), (time.perf_counter() - t0) * 1e3) / steps


def rel_l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())
