#!/usr/bin/env python
"""Benchmark of the Matcha-TTS synthesis hot path (CFM Euler decode + HiFi-GAN v1) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config cfg2|cfg2b|cfg3|cfg4|cfg5] [--precision fp16|bf16|fp32]
                    [--impl reference]

Workloads = BASELINE.json configs; random-init weights and synthetic inputs, seed 1234:
  cfg2  (default, the headline) LJSpeech, 64 utterances per GPU of 150 phonemes -> 451 valid mel frames (T_pad 452),
        10 Euler steps + vocoder on the 451-frame crop.  Weak scaling (64 per GPU).
  cfg2b the length-bucketed variant: a pool of 1024 utterances with log-normal lengths, sorted by length, cut into 16
        batches of 64 (T_pad per batch = the longest, rounded up to the graph bucket), batches dealt to ranks by longest-
        processing-time; one step = the whole pool.  Strong scaling.
  cfg3  VCTK (C_in 224, speaker vectors), 256 utterances in all sharded 256/N per GPU, 4 Euler steps.  Strong scaling.
  cfg4  long form: 16 utterances per GPU of 1000 phonemes -> 3001 frames (T_pad 3004), 10 steps.  Weak scaling.
  cfg5  training: CFM.compute_loss forward + backward, 32 utterances per GPU of 512 frames (ragged), gradient all-reduce.
One "step" = one pass of the path over the workload.  Prints ONE JSON line (rank 0).

 * value     : valid mel frames/s, inputs resident in HBM, CUDA-event timed, max over ranks.
 * e2e       : same through the host-buffer C-ABI call (pinned host inputs H2D, mel+wav D2H inside the timed region).
 * roofline  : conv-GEMM kernel family (tcgen05), algorithmic FLOPs / event-timed duration vs the measured bf16 peak.
 * cpu_baseline : the CPU oracle (oracle/matcha_oracle.py, torch CPU ops, all host threads) on a bounded sample.
 * library_baseline : the same oracle on this GPU in stock torch-CUDA eager (cuDNN / cuBLAS), TF32 off and on: "the
                library bar" of SURVEY.md section 8(d) - reported beside the product, never the product.
 * --impl reference : the reference arm = the same CPU oracle timed as the headline of its own JSON line.
"""
import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

LJ = (-5.536622, 2.116101)       # configs/data/ljspeech.yaml:19-20
VCTK = (-6.630575, 2.482914)     # configs/data/vctk.yaml:13-14
D_EST = {160: 10_985_472, 224: 11_116_544}   # estimator dense FLOPs per frame per step
F_VOC = 614_105_088                          # HiFi-GAN v1 FLOPs per mel frame
T_BUCKET = 32                                # cfg2b: T_pad is rounded up to a multiple of this (one CUDA graph per bucket)

CONFIGS = {
    "cfg2": dict(c_in=160, n_steps=10, stats=LJ, scaling="weak", batch=64, t_valid=451,
                 text="cfg2: LJSpeech Matcha-TTS hot path (CFM 10-step Euler decode + HiFi-GAN v1), batch 64 per GPU, 150 "
                      "phonemes -> 451 valid frames (T_pad 452), random-init weights seed 1234"),
    "cfg2b": dict(c_in=160, n_steps=10, stats=LJ, scaling="strong", pool=1024, batch=64,
                  text="cfg2b: LJSpeech, pool of 1024 utterances, phonemes ~ clip(lognormal(ln 100, 0.45), 20, 300) -> 3 frames "
                       "per phoneme + 1, globally length-sorted, 16 batches of 64 (T_pad = longest rounded up to 32), batches "
                       "dealt to ranks by longest-processing-time, 10 Euler steps + HiFi-GAN v1"),
    "cfg3": dict(c_in=224, n_steps=4, stats=VCTK, scaling="strong", total=256, t_valid=451,
                 text="cfg3: VCTK Matcha-TTS (C_in 224, speaker vectors), 256 utterances sharded over the GPUs, 150 phonemes -> "
                      "451 frames (T_pad 452), 4 Euler steps + HiFi-GAN v1"),
    "cfg4": dict(c_in=160, n_steps=10, stats=LJ, scaling="weak", batch=16, t_valid=3001,
                 text="cfg4: LJSpeech long form, batch 16 per GPU, 1000 phonemes -> 3001 valid frames (T_pad 3004), 10 Euler "
                      "steps + HiFi-GAN v1"),
    "cfg5": dict(c_in=160, scaling="weak", batch=32, t_pad=512,
                 text="cfg5: CFM.compute_loss forward + backward of the estimator, batch 32 per GPU, T 512, lengths uniform in "
                      "{256..512 step 4}, per-sample t, gradient all-reduce over the GPUs"),
}


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


def ceil_to(n, m):
    return (n + m - 1) // m * m


def batch_flops(c_in, n_steps, b, t_pad, length):
    """the estimator runs over T_pad frames per step, the vocoder over the cropped length."""
    return b * (t_pad * n_steps * (D_EST[c_in] + 1536 * t_pad) + length * F_VOC)


def make_batch(c_in, lens, t_pad, seed):
    """One batch of synthetic hot-path inputs (CPU tensors): mu, z ~ N(0,1) (z at temperature 0.667, cli.py:240), prefix
    masks, speaker vectors for VCTK.  `length` = the longest valid length = what the vocoder sees (matcha_tts.py:136)."""
    import torch
    from matcha_b200 import synthetic as weights
    g = torch.Generator().manual_seed(seed)
    b = len(lens)
    return dict(mu=torch.randn(b, 80, t_pad, generator=g), z=torch.randn(b, 80, t_pad, generator=g) * 0.667,
                mask=weights.prefix_mask(list(lens), t_pad), spks=torch.randn(b, 64, generator=g) if c_in == 224 else None,
                lens=list(lens), t_pad=t_pad, length=max(lens))


def pool_lengths(n=1024, seed=1234):
    """cfg2b: phonemes N_i = clip(round(lognormal(ln 100, 0.45)), 20, 300).  With the survey's
    recipe (2N+1 tokens, one frame per token, length_scale 1.5) an utterance of N phonemes has 3N+1 frames (150 -> 451)."""
    import torch
    g = torch.Generator().manual_seed(seed)
    n_ph = torch.exp(math.log(100.0) + 0.45 * torch.randn(n, generator=g)).round().clamp(20, 300).long()
    return [int(3 * v + 1) for v in n_ph]


def workload(cfg_name, rank, world):
    """This rank's batches of one step of the configuration, and the global frame count of a step."""
    c = CONFIGS[cfg_name]
    if cfg_name == "cfg2b":
        from matcha_b200 import dist as mdist
        lens = sorted(pool_lengths(c["pool"]), reverse=True)
        groups = [lens[i:i + c["batch"]] for i in range(0, len(lens), c["batch"])]
        t_pads = [ceil_to(g[0], T_BUCKET) for g in groups]
        cost = [batch_flops(c["c_in"], c["n_steps"], len(g), tp, g[0]) for g, tp in zip(groups, t_pads)]
        # batches are the units dealt to ranks (longest-processing-time on their FLOPs); bucketing is decided globally,
        # so which GPU runs a batch does not change its result
        mine = mdist.partition([int(x / 1e9) for x in cost], world)[rank]
        batches = [make_batch(c["c_in"], groups[i], t_pads[i], 1234 + i) for i in sorted(mine)]
        # the final gather's buffer shape is the same on every rank and known from the global partition: no size exchange
        b_max = max(sum(len(groups[i]) for i in part) for part in mdist.partition([int(x / 1e9) for x in cost], world))
        return batches, sum(lens), dict(n_batches=len(groups), t_pads=sorted(set(t_pads)), gather_rows=b_max, gather_len=lens[0],
                                        pad_waste=1.0 - sum(lens) / sum(len(g) * tp for g, tp in zip(groups, t_pads)))
    t_valid = c["t_valid"]
    t_pad = ceil_to(t_valid, 4)
    if c["scaling"] == "strong":      # cfg3: the 256 utterances are dealt to the ranks
        if c["total"] % world:
            raise SystemExit(f"{cfg_name}: {c['total']} utterances do not divide over {world} GPUs")
        b = c["total"] // world
        frames = c["total"] * t_valid
    else:
        b = c["batch"]
        frames = world * b * t_valid
    return [make_batch(c["c_in"], [t_valid] * b, t_pad, 1234 + rank)], frames, dict(batch_per_gpu=b, t_pad=t_pad)


def cpu_sample(cfg_name):
    """Bounded sample of the workload for the CPU legs: a couple of utterances of the configuration's shape."""
    c = CONFIGS[cfg_name]
    if cfg_name == "cfg2b":
        lens = sorted(pool_lengths(c["pool"]), reverse=True)
        lens = [lens[len(lens) // 2], lens[len(lens) // 2 + 1]]          # two median-length utterances
        t_pad = ceil_to(lens[0], T_BUCKET)
    else:
        lens = [c["t_valid"]] * (1 if c["t_valid"] > 1000 else 2)
        t_pad = ceil_to(c["t_valid"], 4)
    return make_batch(c["c_in"], lens, t_pad, 1234), lens


def cpu_oracle_step(cfg_name, est_sd, gen_sd, batch):
    import torch
    from oracle import matcha_oracle as mo
    c = CONFIGS[cfg_name]
    mean, std = c["stats"]
    with torch.inference_mode():
        dec = mo.solve_euler(est_sd, batch["z"], batch["mu"], batch["mask"], c["n_steps"], batch["spks"])
        mel = dec[:, :, :batch["length"]] * std + mean
        return mo.hifigan_forward(gen_sd, mel).clamp(-1, 1)


def run_reference(args):
    """Reference arm: the CPU restatement of the reference path (the reference itself is PyTorch-CPU code that cannot
    travel to the GPU box; the oracle is pinned against it by tests/test_oracle.py) on all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    from matcha_b200 import synthetic as weights
    if args.config == "cfg5":
        print(json.dumps({"impl": "reference", "unavailable": "cfg5 (training) has no CPU reference arm; run --config cfg2"}))
        return
    c = CONFIGS[args.config]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    est_sd, gen_sd = weights.estimator_state_dict(c["c_in"], seed=1234), weights.generator_state_dict(1234)
    batch, lens = cpu_sample(args.config)
    for _ in range(args.warmup):
        cpu_oracle_step(args.config, est_sd, gen_sd, batch)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_oracle_step(args.config, est_sd, gen_sd, batch)
    dt = (time.perf_counter() - t0) / args.steps
    frames = sum(lens)
    value = frames / dt
    sample = f"{len(lens)} utterance(s) x {lens[0]} frames per step (a bounded sample of the workload), fp32, {cores} threads"
    print(json.dumps({
        "impl": "reference", "metric": "mel_frames_per_s", "value": value, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": c["scaling"], "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "rtf": dt / (frames * 256 / 22050.0),
        "config": {"workload": c["text"] + " (bounded sample)", "sample_batch": len(lens)},
        "cpu_baseline": {"value": value, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def run_ours(args):
    import torch
    import torch.distributed as dist
    from matcha_b200._lib import Plan
    from matcha_b200 import synthetic as weights  # random-init weights; oracle/ is only touched by the baseline legs

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (B200); there is no CPU fallback for the product path")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # NCCL logs to stdout by default; stdout carries the JSON line
        import datetime
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(seconds=120))
    dev = torch.device("cuda", local)
    c = CONFIGS[args.config]
    mean, std = c["stats"]
    n_steps = c["n_steps"]

    plan = Plan(args.precision, c["c_in"], use_graph=True)
    plan.load_estimator(weights.estimator_state_dict(c["c_in"], seed=1234))
    plan.load_generator(weights.generator_state_dict(1234))
    batches, frames, shape_info = workload(args.config, rank, world)
    for bt in batches:
        for k in ("mu", "mask", "z", "spks"):
            if bt[k] is not None:
                bt[k + "_h"] = bt[k].contiguous().pin_memory()
                bt[k] = bt[k + "_h"].to(dev)
            else:
                bt[k + "_h"] = None
        # two sets of pinned output buffers: the end-to-end loop keeps two batches in flight
        bt["mel_h"] = [torch.empty(len(bt["lens"]), 80, bt["length"]).pin_memory() for _ in range(2)]
        bt["wav_h"] = [torch.empty(len(bt["lens"]), 1, 256 * bt["length"]).pin_memory() for _ in range(2)]

    # The path's only exchange: the final gather of mels / waveforms to rank 0 over NVLink.  It is issued asynchronously
    # (NCCL's own stream) and overlaps the NEXT step's decode; every rank's contribution has the same shape per batch
    # slot in cfg2 / cfg3 / cfg4, and cfg2b gathers its (ragged) batches one gather_padded per step.
    pending = []

    def gather_async(mels, wavs):
        if world == 1:
            return
        if args.config == "cfg2b":   # ragged shards: one padded buffer per rank, (rows, length) fixed by the global partition
            buf = torch.zeros(shape_info["gather_rows"], 1, 256 * shape_info["gather_len"], device=dev)
            row = 0
            for w in wavs:
                buf[row:row + w.shape[0], :, :w.shape[-1]] = w
                row += w.shape[0]
            out = [torch.empty_like(buf) for _ in range(world)] if rank == 0 else None
            pending.append(dist.gather(buf, out, dst=0, async_op=True))
            return
        mel, wav = torch.cat(mels), torch.cat(wavs)
        gm = [torch.empty_like(mel) for _ in range(world)] if rank == 0 else None
        gw = [torch.empty_like(wav) for _ in range(world)] if rank == 0 else None
        pending.append(dist.gather(mel, gm, dst=0, async_op=True))
        pending.append(dist.gather(wav, gw, dst=0, async_op=True))

    def step():
        mels, wavs = [], []
        for bt in batches:
            dec, mel, wav = plan.synthesise(bt["mu"], bt["mask"], bt["z"], n_steps, bt["length"], mean, std, spks=bt["spks"])
            mels.append(mel)
            wavs.append(wav)
        while len(pending) > 2:       # at most one step's gathers in flight behind the compute stream
            pending.pop(0).wait()
        gather_async(mels, wavs)

    def drain():
        while pending:
            pending.pop(0).wait()

    def sync_all():
        drain()
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
    drain()                           # the last step's gather is inside the timed region
    e1.record()
    sync_all()
    ms = e0.elapsed_time(e1) / args.steps
    launches = (plan.launch_count() - launches0) // args.steps
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)

    # ---- end to end through the host-buffer C-ABI calls (pinned inputs H2D + mel/wav D2H of EVERY step inside the timed
    # region).  A serving loop double-buffers: batch i+1 is submitted before batch i is waited for, so its upload and batch
    # i's read-back overlap the other's compute (mb200_synthesise_host_submit / _wait, two slots).
    jobs = [bt for _ in range(args.steps) for bt in batches]

    def run_host(job_list):
        for i, bt in enumerate(job_list):
            if i >= 2:
                plan.synthesise_host_wait(i % 2)
            plan.synthesise_host_submit(bt["mu_h"], bt["mask_h"], bt["z_h"], n_steps, bt["length"], mean, std,
                                        bt["mel_h"][i % 2], bt["wav_h"][i % 2], spks=bt["spks_h"], slot=i % 2)
        for i in range(max(0, len(job_list) - 2), len(job_list)):
            plan.synthesise_host_wait(i % 2)

    run_host(batches * 2)
    sync_all()
    t0 = time.perf_counter()
    run_host(jobs)
    ms_e2e = (time.perf_counter() - t0) * 1e3 / args.steps   # every output is in host memory when run_host returns
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([ms_e2e], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_e2e = float(t)
    h2d = sum(bt[k].numel() * 4 for bt in batches for k in ("mu_h", "z_h", "mask_h", "spks_h") if bt[k] is not None)
    d2h = sum((bt["mel_h"][0].numel() + bt["wav_h"][0].numel()) * 4 for bt in batches)

    # ---- per-kernel-family timing (CUDA events around every launch, on the launching stream), rank 0
    roofline, breakdown = None, None
    if rank == 0:
        plan.set_profiling(True)
        for _ in range(2):
            for bt in batches:
                plan.synthesise(bt["mu"], bt["mask"], bt["z"], n_steps, bt["length"], mean, std, spks=bt["spks"])
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
        if args.config == "cfg2":
            try:   # DRAM bytes per launch of the same kernel family from the committed ncu pass (bench.py cannot run under ncu)
                tpath = os.path.join(ROOT, "profiles", "r02_dram_traffic.json")
                if not os.path.exists(tpath):
                    tpath = os.path.join(ROOT, "profiles", "r01_dram_traffic.json")
                with open(tpath) as f:
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

    # ---- baselines on this box (rank 0, N = 1 only): the CPU oracle on the host cores, and the oracle on this GPU in
    # stock torch-CUDA eager ("the library bar")
    cpu, library = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        est_sd, gen_sd = weights.estimator_state_dict(c["c_in"], seed=1234), weights.generator_state_dict(1234)
        sample, lens = cpu_sample(args.config)
        cpu_oracle_step(args.config, est_sd, gen_sd, sample)
        times = []
        t_end = time.perf_counter() + 20.0
        while len(times) < 5 and (time.perf_counter() < t_end or not times):
            t0 = time.perf_counter()
            cpu_oracle_step(args.config, est_sd, gen_sd, sample)
            times.append(time.perf_counter() - t0)
        cpu = {"value": sum(lens) / min(times), "unit": "frames/s", "cores": cores, "kind": "port",
               "sample": f"{len(lens)} utterance(s) x {lens[0]} frames of the workload, {n_steps} steps + vocoder, fp32 torch-CPU "
                         f"oracle, best of {len(times)}"}
    if rank == 0 and world == 1 and not args.no_library_baseline:
        from oracle import full_size as fs
        del plan                       # the library run needs the activations' memory (fp32 eager keeps whole tensors)
        torch.cuda.empty_cache()
        library = {"what": "oracle/matcha_oracle.py on this GPU in stock torch-CUDA eager (cuDNN / cuBLAS / ATen), fp32 I/O, "
                           "the same batches; reported, never the product"}
        est_d = {k: v.to(dev) for k, v in weights.estimator_state_dict(c["c_in"], seed=1234).items()}
        gen_d = {k: v.to(dev) for k, v in weights.generator_state_dict(1234).items()}
        my_frames = sum(sum(bt["lens"]) for bt in batches)
        for tf in (False, True):
            def lib_step():
                with fs.tf32(tf):
                    for bt in batches:
                        cpu_oracle_step(args.config, est_d, gen_d, bt)
            lib_step()
            torch.cuda.synchronize()
            l0, l1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            l0.record()
            for _ in range(2):
                lib_step()
            l1.record()
            torch.cuda.synchronize()
            lms = max(l0.elapsed_time(l1), (time.perf_counter() - t0) * 1e3) / 2
            library["tf32_on" if tf else "tf32_off"] = {"ms_per_step": lms, "frames_per_s": my_frames / lms * 1e3}
        plan = None

    if rank == 0:
        value = frames / (ms * 1e-3)
        total_flops = sum(batch_flops(c["c_in"], n_steps, len(bt["lens"]), bt["t_pad"], bt["length"]) for bt in batches)
        peaks = measured_peaks()
        cfg = {"workload": c["text"], "n_steps": n_steps, "precision": args.precision,
               "l2": "working set (GBs of activations per step) exceeds the 126 MB L2; no explicit flush",
               "parallelism": f"dp{world}" + (" + final NCCL gather of mel/wav to rank 0, asynchronous, overlapping the next step" if world > 1 else "")}
        cfg.update(shape_info)
        out = {
            "metric": "mel_frames_per_s", "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": c["scaling"],
            "vs_baseline": None, "dtype": {"fp16": "f16", "bf16": "bf16", "fp32": "f32"}[args.precision],
            "data": "synthetic", "rtf": (ms * 1e-3) / (frames * 256 / 22050.0), "config": cfg,
            "e2e": {"value": frames / (ms_e2e * 1e-3), "unit": "frames/s", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches) * args.steps, "gpu_launches_per_step": int(launches),
            "whole_step": {"algorithmic_tflop_rank0": total_flops / 1e12, "tflops_rank0": total_flops / (ms * 1e-3) / 1e12,
                           "frac_of_peak": total_flops / (ms * 1e-3) / 1e12 / peaks["tflops"]},
            "roofline": roofline, "cpu_baseline": cpu, "library_baseline": library, "clocks": clocks, "breakdown": breakdown,
        }
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def run_train(args):
    """cfg5: CFM.compute_loss forward + backward (csrc/train.cu) on 32 utterances per GPU, then the training path's one
    real exchange, the gradient all-reduce (matcha_b200.dist.allreduce_gradients, NCCL).  Metric: valid training frames/s."""
    import types

    import torch
    import torch.distributed as dist

    import matcha_b200 as mb
    from matcha_b200 import dist as mdist
    from matcha_b200 import synthetic as weights

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (B200); there is no CPU fallback for the product path")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        import datetime
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(seconds=120))
    dev = torch.device("cuda", local)
    c = CONFIGS["cfg5"]
    b, t_pad = c["batch"], c["t_pad"]
    dec = dict(channels=[256, 256], dropout=0.05, attention_head_dim=64, n_blocks=1, num_mid_blocks=2, num_heads=2,
               act_fn="snakebeta")
    cfm = mb.CFM(160, 80, types.SimpleNamespace(solver="euler", sigma_min=1e-4), dec, precision="fp32")
    cfm.load_state_dict(weights.estimator_state_dict(160, seed=1234), strict=True)
    cfm = cfm.to(dev).train()      # training mode: the transformer blocks' Dropout(0.05) is inside the timed step
    params = list(cfm.estimator.parameters())
    g = torch.Generator().manual_seed(1234 + rank)
    lengths = [int(v) for v in (torch.randint(64, 129, (b,), generator=g) * 4)]
    x1_h, mu_h = (torch.randn(b, 80, t_pad, generator=g).pin_memory() for _ in range(2))
    mask_h = weights.prefix_mask(lengths, t_pad).pin_memory()
    x1, mu, mask = x1_h.to(dev), mu_h.to(dev), mask_h.to(dev)
    valid = torch.tensor(float(sum(lengths)), device=dev)

    def step(host=False):
        if host:   # end to end: this step's batch comes up from pinned host memory, the loss goes back
            a, m, k = x1_h.to(dev, non_blocking=True), mu_h.to(dev, non_blocking=True), mask_h.to(dev, non_blocking=True)
        else:
            a, m, k = x1, mu, mask
        for p in params:
            p.grad = None
        loss, _ = cfm.compute_loss(a, k, m)
        loss.backward()
        mdist.allreduce_gradients(params, valid_frames=valid)
        return float(loss) if host else loss

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    sync_all()
    plan = cfm.estimator.train_plan()
    n0 = plan.launch_count()
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
    launches = (plan.launch_count() - n0) // args.steps
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step(host=True)
    ms_e2e = (time.perf_counter() - t0) * 1e3 / args.steps
    clocks = sampler.stop() if rank == 0 else None
    tot = torch.tensor([ms, ms_e2e, float(sum(lengths))], device=dev)
    if world > 1:
        mx = tot[:2].clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        fr = tot[2:].clone()
        dist.all_reduce(fr)
        ms, ms_e2e, frames = float(mx[0]), float(mx[1]), float(fr[0])
    else:
        frames = float(tot[2])
    if rank == 0:
        peaks = measured_peaks()
        flops = 3.0 * b * t_pad * (D_EST[160] + 1536 * t_pad)     # forward + (dgrad + wgrad) ~ 3x the forward GEMM work
        achieved = flops / (ms * 1e-3) / 1e12
        print(json.dumps({
            "metric": "training_mel_frames_per_s", "value": frames / (ms * 1e-3), "unit": "frames/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32 (3xTF32 mma.sync GEMMs, fp32 data)", "data": "synthetic",
            "config": {"workload": c["text"], "batch_per_gpu": b, "t_pad": t_pad,
                       "parallelism": f"dp{world}" + (" + NCCL gradient all-reduce (32 MB buckets) after the backward" if world > 1 else "")},
            "e2e": {"value": frames / (ms_e2e * 1e-3), "unit": "frames/s", "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": (x1_h.numel() + mu_h.numel() + mask_h.numel()) * 4, "d2h_bytes_per_step": 4},
            "gpu_launches": int(launches) * args.steps, "gpu_launches_per_step": int(launches),
            "roofline": {"bound": "tensor", "achieved": achieved, "peak": peaks["tflops"], "unit": "TFLOP/s",
                         "frac": achieved / peaks["tflops"], "traffic": None, "peak_source": peaks["source"],
                         "kernel": "whole training step (tgemm_kernel<3> family is 86 % of it)", "flops_per_step": flops},
            "cpu_baseline": None, "clocks": clocks,
        }))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default="cfg2", choices=sorted(CONFIGS))
    ap.add_argument("--precision", default="fp16", choices=["fp16", "bf16", "fp32"])
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-library-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        if args.steps > 6:
            args.steps = 6          # bounded: each CPU step is seconds long
        args.warmup = min(args.warmup, 1) or 1
        run_reference(args)
    elif args.config == "cfg5":
        run_train(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
