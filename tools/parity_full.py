"""Full-size parity record: the CUDA path in fp32 / fp16 / bf16 against the oracle on CUDA (fp32, TF32 off) at cfg2 / cfg3 /
cfg4, plus run-to-run determinism and the library bar (stock torch-CUDA eager).  Writes gpurun_out/parity_full.json.

    python tools/parity_full.py [cfg2 cfg3 cfg3_b32 cfg4] [--no-fp32] [--no-lib]
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from matcha_b200._lib import Plan  # noqa: E402
from oracle import full_size as fs  # noqa: E402


def main():
    names = [a for a in sys.argv[1:] if not a.startswith("--")] or ["cfg2", "cfg3", "cfg3_b32", "cfg4"]
    precs = ["fp16", "bf16"] + ([] if "--no-fp32" in sys.argv else ["fp32"])
    out = {}
    for name in names:
        case = fs.make_case(name)
        o_dec, o_mel, o_wav = fs.oracle_cuda(case)
        rec = {"b": case["b"], "t_pad": case["t_pad"], "n_steps": case["n_steps"], "c_in": case["c_in"]}
        if "--no-lib" not in sys.argv:
            frames = case["b"] * case["t_valid"]
            for tf in (False, True):
                ms = fs.time_library(case, tf, steps=2, warmup=1)
                rec[f"library_tf32_{'on' if tf else 'off'}"] = {"ms_per_step": ms, "frames_per_s": frames / ms * 1e3}
            t_dec, _, t_wav = fs.oracle_cuda(case, allow_tf32=True)
            rec["library_tf32_on"]["rel_l2_vs_fp32"] = {"dec": fs.rel_l2(t_dec, o_dec), "wav": fs.rel_l2(t_wav, o_wav)}
            del t_dec, t_wav
        mean, std = case["stats"]
        for prec in precs:
            plan = Plan(prec, case["c_in"])
            plan.load_estimator(case["est_sd"])
            plan.load_generator(case["gen_sd"])
            args = (case["mu"], case["mask"], case["z"], case["n_steps"], case["t_valid"], mean, std)
            dec, mel, wav = plan.synthesise(*args, spks=case["spks"])
            dec2, mel2, wav2 = plan.synthesise(*args, spks=case["spks"])
            wav_v = plan.hifigan_forward(o_mel)   # the vocoder alone, from the oracle's exact mel
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(3):
                plan.synthesise(*args, spks=case["spks"])
            e1.record()
            torch.cuda.synchronize()
            rec[prec] = {
                "dec_rel_l2": fs.rel_l2(dec, o_dec), "dec_max_abs": float((dec - o_dec).abs().max()),
                "wav_rel_l2": fs.rel_l2(wav, o_wav), "wav_max_abs": float((wav - o_wav).abs().max()),
                "vocoder_only_wav_rel_l2": fs.rel_l2(wav_v.clamp(-1, 1), o_wav),
                "run_to_run_equal": bool(torch.equal(dec, dec2) and torch.equal(wav, wav2)),
                "run_to_run_dec_rel_l2": fs.rel_l2(dec2, dec), "ms_per_step": e0.elapsed_time(e1) / 3,
                "wav_abs_max": float(o_wav.abs().max()), "dec_abs_max": float(o_dec.abs().max()),
            }
            print(name, prec, json.dumps(rec[prec]), flush=True)
            del plan, dec, mel, wav, dec2, mel2, wav2, wav_v
            torch.cuda.empty_cache()
        out[name] = rec
        del o_dec, o_mel, o_wav, case
        torch.cuda.empty_cache()
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/parity_full.json", "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
