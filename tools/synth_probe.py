"""End-to-end MatchaTTS.synthesise (encoder kernels -> durations -> length regulation -> CFM decode) + vocoder latency, and
the text encoder alone (CUDA kernels vs the torch formulation on the same GPU).  Writes gpurun_out/synth_probe.json."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import matcha_b200 as mb  # noqa: E402
from matcha_b200 import matcha_tts as mt  # noqa: E402
from matcha_b200 import synthetic  # noqa: E402


def timed(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3


out = {}
model = mb.MatchaTTS(**mt.default_config(n_spks=1))
model.encoder.load_state_dict(synthetic.text_encoder_state_dict(n_spks=1, seed=1234), strict=True)
model.decoder.load_state_dict(synthetic.estimator_state_dict(160, seed=1234), strict=True)
# deterministic durations as in SURVEY.md section 8(d): logw = 0 -> one frame per token, length_scale 1.5
with torch.no_grad():
    model.encoder.proj_w.proj.weight.zero_()
    model.encoder.proj_w.proj.bias.zero_()
model = model.cuda().eval()
gen = mb.Generator(mb.AttrDict(mb.HIFIGAN_V1), precision="fp16")
gen.load_state_dict(synthetic.generator_state_dict(1234, weight_norm=True), strict=True)
gen = gen.cuda().eval()
g = torch.Generator().manual_seed(0)
for b, n_ph in [(1, 150), (64, 150), (1, 1000)]:
    lx = 2 * n_ph + 1
    x = torch.randint(1, 178, (b, lx), generator=g).cuda()
    lens = torch.full((b,), lx).cuda()

    def full():
        o = model.synthesise(x, lens, n_timesteps=10, temperature=0.667, length_scale=1.5)
        return gen(o["mel"]).clamp(-1, 1)

    def enc_cuda():
        with torch.inference_mode():
            return model.encoder(x, lens)

    def enc_torch():
        with torch.no_grad():
            return model.encoder(x, lens)      # grad mode "enabled" inside no_grad is False -> still CUDA; force torch below

    wav = full()
    ms_full = timed(full, n=5 if b > 1 else 20)
    ms_enc = timed(enc_cuda)
    model.encoder.train()                       # training mode = the torch formulation (dropout is active but timing is the same)
    with torch.no_grad():
        ms_torch = timed(lambda: model.encoder(x, lens))
    model.encoder.eval()
    frames = int(wav.shape[-1] // 256)
    out[f"B{b}_N{n_ph}"] = {"synthesise_plus_vocoder_ms": ms_full, "frames_per_utt": frames, "rtf": ms_full * 1e-3 / (b * frames * 256 / 22050),
                            "text_encoder_cuda_ms": ms_enc, "text_encoder_torch_cuda_ms": ms_torch}
    print(f"B{b}_N{n_ph}", out[f"B{b}_N{n_ph}"], flush=True)
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/synth_probe.json", "w"), indent=1)
