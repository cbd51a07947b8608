"""Developer bring-up script for the GPU box: runs each kernel group in its own process (a trapped kernel poisons the
CUDA context) and prints error norms against torch / the oracle.  Not part of the product or the test-suite.

    python tools/gpu_check.py            # all groups
    python tools/gpu_check.py conv:fp16  # one group
"""
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def rel(a, b):
    return float((a - b).norm() / (b.norm() + 1e-30))


def q(x, prec):
    import torch
    if prec == "fp32":
        return x
    return x.to(torch.bfloat16 if prec == "bf16" else torch.float16).float()


def group_conv(prec, simt):
    import torch
    import torch.nn.functional as F
    from matcha_b200._lib import Plan
    plan = Plan(prec, 160, force_simt=simt, use_graph=False)
    g = torch.Generator().manual_seed(0)
    cases = [  # (B, Cin, Cout, T, K, dil, stride, transposed)
        (2, 256, 256, 140, 3, 1, 1, False), (1, 64, 64, 300, 7, 3, 1, False), (2, 32, 32, 513, 11, 5, 1, False),
        (2, 256, 256, 24, 3, 1, 2, False), (2, 256, 256, 70, 4, 1, 2, True), (1, 512, 256, 9, 16, 1, 8, True),
        (1, 128, 64, 200, 4, 1, 2, True), (1, 64, 32, 130, 4, 1, 2, True), (3, 160, 256, 24, 3, 1, 1, False),
        (2, 256, 1024, 130, 1, 1, 1, False), (2, 1024, 256, 130, 1, 1, 1, False), (2, 256, 384, 4, 1, 1, 1, False),
        (2, 128, 256, 129, 1, 1, 1, False), (2, 80, 512, 37, 7, 1, 1, False), (2, 256, 80, 140, 1, 1, 1, False),
        (1, 128, 128, 1000, 11, 5, 1, False), (1, 32, 32, 2500, 11, 5, 1, False), (2, 64, 64, 2200, 7, 3, 1, False),
        (1, 128, 128, 2100, 3, 1, 1, False), (1, 128, 128, 2304, 11, 1, 1, False), (1, 64, 32, 1100, 4, 1, 2, True),
    ]
    for (b, cin, cout, t, k, dil, stride, tr) in cases:
        x = torch.randn(b, cin, t, generator=g)
        w = torch.randn(cin, cout, k, generator=g) if tr else torch.randn(cout, cin, k, generator=g)
        w = w / (cin * k) ** 0.5
        bias = torch.randn(cout, generator=g)
        xq, wq = q(x, prec).double(), q(w, prec).double()
        if tr:
            ref = F.conv_transpose1d(xq, wq, bias.double(), stride=stride, padding=(k - stride) // 2)
        else:
            ref = F.conv1d(xq, wq, bias.double(), stride=stride, dilation=dil, padding=dil * (k - 1) // 2)
        try:
            y = plan.test_conv1d(x, w, bias, dilation=dil, stride=stride, transposed=tr).cpu().double()
            print(f"conv {prec} simt={simt} {(b, cin, cout, t, k, dil, stride, tr)} rel={rel(y, ref):.3e} "
                  f"maxabs={float((y - ref).abs().max()):.3e}", flush=True)
        except Exception as e:  # noqa
            print(f"conv {prec} simt={simt} {(b, cin, cout, t, k, dil, stride, tr)} FAILED: {e}", flush=True)
            raise


def group_attn(prec):
    import torch
    from matcha_b200._lib import Plan
    plan = Plan(prec, 160, use_graph=False)
    g = torch.Generator().manual_seed(1)
    for (b, t, lens) in [(2, 24, [24, 9]), (2, 140, [140, 77]), (1, 452, [400]), (1, 1, [1])]:
        qkv = torch.randn(b, t, 384, generator=g)
        qkv[:, :, :128] *= 0.3
        mask = (torch.arange(t)[None] < torch.tensor(lens)[:, None]).float()
        x = q(qkv, prec).double()
        qq, kk, vv = (x[:, :, i * 128:(i + 1) * 128].reshape(b, t, 2, 64).transpose(1, 2) for i in range(3))
        logits = qq @ kk.transpose(-1, -2) + mask[:, None, None, :].double()
        ref = (torch.softmax(logits, -1) @ vv).transpose(1, 2).reshape(b, t, 128)
        y = plan.test_attention(qkv, mask).cpu().double()
        print(f"attn {prec} {(b, t)} rel={rel(y, ref):.3e} maxabs={float((y - ref).abs().max()):.3e}", flush=True)


def group_est(prec, simt):
    import torch
    from matcha_b200._lib import Plan
    from oracle import matcha_oracle as mo, weights
    for name in ["est_lj_ragged", "est_lj_min", "est_vctk_ragged", "est_lj_t140"]:
        fx = torch.load(os.path.join(ROOT, "tests", "golden", name + ".pt"), weights_only=False)
        sd = weights.estimator_state_dict(fx["c_in"], seed=fx["seed"])
        plan = Plan(prec, fx["c_in"], force_simt=simt, use_graph=True)
        plan.load_estimator(sd)
        v = plan.estimator_forward(fx["z"], fx["mask"], fx["mu"], fx["t_scalar"], fx["spks"]).cpu()
        vb = plan.estimator_forward(fx["z"], fx["mask"], fx["mu"], fx["t_batched"], fx["spks"]).cpu()
        dec = plan.cfm_decode(fx["mu"], fx["mask"], fx["z"], fx["n_steps"], fx["spks"]).cpu()
        dec2 = plan.cfm_decode(fx["mu"], fx["mask"], fx["z"], fx["n_steps"], fx["spks"]).cpu()  # graph replay
        print(f"est {prec} simt={simt} {name}: v rel={rel(v, fx['v_scalar']):.3e} max={float((v - fx['v_scalar']).abs().max()):.3e} | "
              f"vb rel={rel(vb, fx['v_batched']):.3e} max={float((vb - fx['v_batched']).abs().max()):.3e} | "
              f"dec rel={rel(dec, fx['decoded']):.3e} max={float((dec - fx['decoded']).abs().max()):.3e} | "
              f"replay max={float((dec2 - dec).abs().max()):.3e}", flush=True)


def group_gen(prec, simt):
    import torch
    from matcha_b200._lib import Plan
    from oracle import weights
    for name in ["gen_plain", "gen_weightnorm"]:
        fx = torch.load(os.path.join(ROOT, "tests", "golden", name + ".pt"), weights_only=False)
        sd = weights.generator_state_dict(fx["seed"], weight_norm=fx["weight_norm"])
        plan = Plan(prec, 160, force_simt=simt, use_graph=True)
        plan.load_generator(sd)
        wav = plan.hifigan_forward(fx["mel"]).cpu()
        wav2 = plan.hifigan_forward(fx["mel"]).cpu()
        print(f"gen {prec} simt={simt} {name}: rel={rel(wav, fx['wav']):.3e} max={float((wav - fx['wav']).abs().max()):.3e} "
              f"replay max={float((wav2 - wav).abs().max()):.3e} refmax={float(fx['wav'].abs().max()):.3f}", flush=True)


GROUPS = {
    "conv:fp32": lambda: group_conv("fp32", False),
    "conv:bf16simt": lambda: group_conv("bf16", True),
    "conv:bf16": lambda: group_conv("bf16", False),
    "conv:fp16": lambda: group_conv("fp16", False),
    "attn:fp32": lambda: group_attn("fp32"),
    "attn:fp16": lambda: group_attn("fp16"),
    "est:fp32": lambda: group_est("fp32", False),
    "est:fp16simt": lambda: group_est("fp16", True),
    "est:fp16": lambda: group_est("fp16", False),
    "est:bf16": lambda: group_est("bf16", False),
    "gen:fp32": lambda: group_gen("fp32", False),
    "gen:fp16simt": lambda: group_gen("fp16", True),
    "gen:fp16": lambda: group_gen("fp16", False),
    "gen:bf16": lambda: group_gen("bf16", False),
}

if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] in GROUPS:
        GROUPS[sys.argv[1]]()
        sys.exit(0)
    names = sys.argv[1:] or list(GROUPS)
    for n in names:
        t0 = time.time()
        try:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), n], capture_output=True, text=True, timeout=300)
            out = r.stdout + ("\n[stderr]\n" + r.stderr[-3000:] if r.returncode else "")
            print(f"=== {n}: exit {r.returncode} ({time.time() - t0:.1f}s)\n{out}", flush=True)
        except subprocess.TimeoutExpired:
            print(f"=== {n}: TIMEOUT", flush=True)
