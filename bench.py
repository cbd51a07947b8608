#!/usr/bin/env python
"""Benchmark of the Matcha-TTS synthesis hot path (CFM 10-step Euler decode + HiFi-GAN v1) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--precision fp16|bf16|fp32] [--impl reference]

Workload = BASELINE.json configs[1] ("cfg2"): LJSpeech hyper-parameters, random-init weights
(seed 1234), batch 64 utterances of 150 phonemes -> 451 valid mel frames each (T_pad 452), 10 Euler steps, vocoder on
the 451-frame crop.  One "step" = one pass of the hot path over one such batch.  Prints ONE JSON line (rank 0).

 * value     : valid mel frames/s, inputs resident in HBM, CUDA-event timed, max over ranks.
 * e2e       : same through the host-buffer C-ABI call (pinned host inputs H2D, mel+wav D2H inside the timed region).
 * roofline  : conv-GEMM kernel family (tcgen05), algorithmic FLOPs / event-timed duration vs the measured bf16 peak.
 * cpu_baseline : the CPU oracle (oracle/matcha_oracle.py, torch CPU ops, all host threads) on a bounded sample.
 * --impl reference : the reference arm = the same CPU oracle timed as the headline of its own JSON line.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

B, N_PHONEMES, T_VALID, T_PAD, N_STEPS = 64, 150, 451, 452, 10
MEL_MEAN, MEL_STD = -5.536622, 2.116101          # configs/data/ljspeech.yaml:19-20
D_EST, F_VOC = 10_985_472, 614_105_088            # SURVEY.md Appendix B (FLOPs per frame per step / per frame)
CPU_SAMPLE_B = 2                                  # utterances per CPU-oracle step (bounded sample of the workload)


def flops_per_valid_frame(t_pad=T_PAD, n_steps=N_STEPS):
    """F = n_steps * (D + 1536 * T_pad) + 614,105,088 (estimator runs over T_pad frames)."""
    return n_steps * (D_EST + 1536 * t_pad) * (t_pad / T_VALID) + F_VOC


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return {"tflops": p.get("bf16_tflops_sustained", 1400.0), "tflops_burst": p.get("bf16_tflops"),
                "hbm_gbs": p.get("hbm_gbs"), "source": "MEASURED_PEAKS.json (bf16 sustained)"}
    return {"tflops": 1590.0, "tflops_burst": 1590.0, "hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """Samples SM clocks / throttle reasons with nvidia-smi while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0])); mx.append(float(parts[1])); power.append(float(parts[2]))
            except ValueError:
                continue
            for n, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        busy = [s for s, p in zip(sm, power) if p > 300] or sm
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def make_inputs(batch, seed=1234):
    import torch
    from matcha_b200 import synthetic as weights
    g = torch.Generator().manual_seed(seed)
    mu = torch.randn(batch, 80, T_PAD, generator=g)
    z = torch.randn(batch, 80, T_PAD, generator=g) * 0.667     # temperature 0.667 (cli.py:240)
    mask = weights.prefix_mask([T_VALID] * batch, T_PAD)
    return mu, mask, z


def cpu_oracle_step(est_sd, gen_sd, mu, mask, z):
    import torch
    from oracle import matcha_oracle as mo
    with torch.inference_mode():
        dec = mo.solve_euler(est_sd, z, mu, mask, N_STEPS)
        mel = dec[:, :, :T_VALID] * MEL_STD + MEL_MEAN
        return mo.hifigan_forward(gen_sd, mel).clamp(-1, 1)


def run_reference(args):
    """Reference arm: the CPU restatement of the reference path (the reference itself is PyTorch-CPU code that cannot
    travel to the GPU box; the oracle is pinned against it by tests/test_oracle.py) on all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from matcha_b200 import synthetic as weights
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    est_sd, gen_sd = weights.estimator_state_dict(160, seed=1234), weights.generator_state_dict(1234)
    mu, mask, z = make_inputs(CPU_SAMPLE_B)
    for _ in range(args.warmup):
        cpu_oracle_step(est_sd, gen_sd, mu, mask, z)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_oracle_step(est_sd, gen_sd, mu, mask, z)
    dt = (time.perf_counter() - t0) / args.steps
    value = CPU_SAMPLE_B * T_VALID / dt
    sample = f"{CPU_SAMPLE_B} utterances x {T_VALID} frames per step (of the 64-utterance batch), fp32, {cores} threads"
    print(json.dumps({
        "impl": "reference", "metric": "mel_frames_per_s", "value": value, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "rtf": dt / (CPU_SAMPLE_B * T_VALID * 256 / 22050.0),
        "config": {"workload": "cfg2: LJSpeech Matcha-TTS hot path, 150 phonemes -> 451 frames (T_pad 452), 10 Euler "
                               "steps + HiFi-GAN v1 (bounded sample of the batch-64 workload)", "sample_batch": CPU_SAMPLE_B},
        "cpu_baseline": {"value": value, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def run_ours(args):
    import torch
    import torch.distributed as dist
    from matcha_b200._lib import Plan
    from matcha_b200 import synthetic as weights  # random-init weights; oracle/ is only touched by the cpu_baseline leg

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (B200); there is no CPU fallback for the product path")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # NCCL logs to stdout by default; stdout carries the JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)

    plan = Plan(args.precision, 160, use_graph=True)
    plan.load_estimator(weights.estimator_state_dict(160, seed=1234))
    plan.load_generator(weights.generator_state_dict(1234))
    mu_h, mask_h, z_h = (t.contiguous().pin_memory() for t in make_inputs(B, seed=1234 + rank))
    mu, mask, z = mu_h.to(dev), mask_h.to(dev), z_h.to(dev)
    mel_h = torch.empty(B, 80, T_VALID).pin_memory()
    wav_h = torch.empty(B, 1, 256 * T_VALID).pin_memory()
    gather_mel = [torch.empty(B, 80, T_VALID, device=dev) for _ in range(world)] if world > 1 and rank == 0 else None
    gather_wav = [torch.empty(B, 1, 256 * T_VALID, device=dev) for _ in range(world)] if world > 1 and rank == 0 else None

    def step():
        dec, mel, wav = plan.synthesise(mu, mask, z, N_STEPS, T_VALID, MEL_MEAN, MEL_STD)
        if world > 1:  # the path's only exchange: final gather of mels / waveforms to rank 0 over NVLink
            dist.gather(mel, gather_mel, dst=0)
            dist.gather(wav, gather_wav, dst=0)
        return wav

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    sync_all()
    launches0 = plan.launch_count()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    sync_all()
    ms = e0.elapsed_time(e1) / args.steps
    launches = (plan.launch_count() - launches0) // args.steps
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)

    # ---- end to end through the host-buffer C-ABI call (pinned inputs H2D + mel/wav D2H inside the timed region)
    for _ in range(2):
        plan.synthesise_host(mu_h, mask_h, z_h, N_STEPS, T_VALID, MEL_MEAN, MEL_STD, mel_out=mel_h, wav_out=wav_h)
    sync_all()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        plan.synthesise_host(mu_h, mask_h, z_h, N_STEPS, T_VALID, MEL_MEAN, MEL_STD, mel_out=mel_h, wav_out=wav_h)
    ms_e2e = (time.perf_counter() - t0) * 1e3 / args.steps   # the call synchronises before returning
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([ms_e2e], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_e2e = float(t)
    h2d = (mu_h.numel() + z_h.numel() + mask_h.numel()) * 4
    d2h = (mel_h.numel() + wav_h.numel()) * 4

    # ---- per-kernel-family timing (CUDA events around every launch, on the launching stream), rank 0
    roofline, breakdown = None, None
    if rank == 0:
        plan.set_profiling(True)
        for _ in range(2):
            plan.synthesise(mu, mask, z, N_STEPS, T_VALID, MEL_MEAN, MEL_STD)
        torch.cuda.synchronize()
        prof = plan.get_profile()
        plan.set_profiling(False)
        peaks = measured_peaks()
        names = {0: "estimator", 1: "vocoder_stage1(C256,+conv_pre)", 2: "vocoder_stage2(C128)", 3: "vocoder_stage3(C64)",
                 4: "vocoder_stage4(C32)", 5: "other"}
        kinds = {0: "conv_gemm", 1: "attention", 2: "elementwise", 3: "memset"}
        breakdown, g_ms, g_fl, g_n, tot_ms = {}, 0.0, 0.0, 0, 0.0
        for (r, k), (n, pms, fl) in sorted(prof.items()):
            n, pms, fl = n / 2, pms / 2, fl / 2
            breakdown[f"{names[r]}/{kinds[k]}"] = {"launches": n, "ms": round(pms, 4),
                                                   "tflops": round(fl / pms / 1e9, 1) if pms and fl else None}
            tot_ms += pms
            if k == 0:
                g_ms, g_fl, g_n = g_ms + pms, g_fl + fl, g_n + n
        achieved = g_fl / g_ms / 1e9 if g_ms else 0.0
        traffic, traffic_src = None, None
        try:   # DRAM bytes per launch of the same kernel family from the committed ncu pass (bench.py cannot run under ncu)
            with open(os.path.join(ROOT, "profiles", "r01_dram_traffic.json")) as f:
                tj = json.load(f)
            traffic, traffic_src = tj["dram_bytes_per_launch"], tj["source"]
        except (OSError, KeyError, ValueError):
            pass
        roofline = {"bound": "tensor", "achieved": achieved,
                    "kernel": "tcgen05 conv-GEMM family: conv_gemm_umma2_kernel + resblock_pair_kernel + ff_fused_kernel + resqkv_fused_kernel (all instantiations)",
                    "peak": peaks["tflops"], "unit": "TFLOP/s", "frac": achieved / peaks["tflops"], "traffic": traffic,
                    "traffic_unit": "bytes/launch (dram__bytes_read.sum + dram__bytes_write.sum)", "traffic_source": traffic_src,
                    "peak_source": peaks["source"], "launches_per_step": g_n, "avg_launch_us": g_ms * 1e3 / max(g_n, 1),
                    "share_of_step_time": g_ms / tot_ms if tot_ms else None,
                    "flops_per_step": g_fl}

    # ---- CPU baseline: the oracle on this box's host cores, bounded sample (rank 0, N = 1 only)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        est_sd, gen_sd = weights.estimator_state_dict(160, seed=1234), weights.generator_state_dict(1234)
        cmu, cmask, cz = make_inputs(CPU_SAMPLE_B)
        cpu_oracle_step(est_sd, gen_sd, cmu, cmask, cz)
        times = []
        t_end = time.perf_counter() + 20.0
        while len(times) < 5 and (time.perf_counter() < t_end or not times):
            t0 = time.perf_counter()
            cpu_oracle_step(est_sd, gen_sd, cmu, cmask, cz)
            times.append(time.perf_counter() - t0)
        cpu = {"value": CPU_SAMPLE_B * T_VALID / min(times), "unit": "frames/s", "cores": cores, "kind": "port",
               "sample": f"{CPU_SAMPLE_B} utterances x {T_VALID} frames (of the 64), 10 steps + vocoder, fp32 torch-CPU "
                         f"oracle, best of {len(times)}"}

    if rank == 0:
        frames = world * B * T_VALID
        value = frames / (ms * 1e-3)
        total_flops = B * (T_PAD * N_STEPS * (D_EST + 1536 * T_PAD) + T_VALID * F_VOC)
        peaks = measured_peaks()
        out = {
            "metric": "mel_frames_per_s", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": {"fp16": "f16", "bf16": "bf16", "fp32": "f32"}[args.precision],
            "data": "synthetic",
            "rtf": (ms * 1e-3) / (frames * 256 / 22050.0),
            "config": {"workload": "cfg2: LJSpeech Matcha-TTS hot path (CFM 10-step Euler decode + HiFi-GAN v1), batch 64 "
                                   "per GPU, 150 phonemes -> 451 valid frames (T_pad 452), random-init weights seed 1234",
                       "batch_per_gpu": B, "t_pad": T_PAD, "n_steps": N_STEPS, "precision": args.precision,
                       "l2": "working set (>3 GB of activations per step) exceeds the 126 MB L2; no explicit flush",
                       "parallelism": f"dp{world}" + (" + final NCCL gather of mel/wav to rank 0" if world > 1 else "")},
            "e2e": {"value": frames / (ms_e2e * 1e-3), "unit": "frames/s", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches) * args.steps,
            "gpu_launches_per_step": int(launches),
            "whole_step": {"algorithmic_tflop": total_flops / 1e12, "tflops": total_flops / (ms * 1e-3) / 1e12,
                           "frac_of_peak": total_flops / (ms * 1e-3) / 1e12 / peaks["tflops"]},
            "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks, "breakdown": breakdown,
            "workspace_gb": plan.workspace_bytes() / 1e9,
        }
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--precision", default="fp16", choices=["fp16", "bf16", "fp32"])
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        if args.steps > 6:
            args.steps = 6          # bounded: each CPU step is seconds long
        args.warmup = min(args.warmup, 1) or 1
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
