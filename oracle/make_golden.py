"""TEST INFRASTRUCTURE ONLY -- generate tests/golden/*.pt by running the UNMODIFIED reference modules
(imported from /root/reference through oracle/ref_shims.py) on seeded synthetic weights and inputs.

Run in the build container only:  python -m oracle.make_golden
Weights are not stored: they are regenerated from the seed by oracle/weights.py (and loaded into the
reference modules with strict=True, which also pins the state_dict key layout of SURVEY.md Appendix A).
Each fixture stores the small inputs and the reference outputs.
"""
import os
import types

import torch

from oracle import ref_shims, weights

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
DEC = dict(channels=[256, 256], dropout=0.05, attention_head_dim=64, n_blocks=1, num_mid_blocks=2,
           num_heads=2, act_fn="snakebeta")  # configs/model/decoder/default.yaml:1-7


def build_cfm(c_in, seed):
    CFM = ref_shims.load_reference()[0]
    n_spks, spk = (1, 64) if c_in == 160 else (109, c_in - 160)
    cfm = CFM(160, 80, types.SimpleNamespace(solver="euler", sigma_min=1e-4), DEC, n_spks=n_spks,
              spk_emb_dim=spk).eval()
    cfm.load_state_dict(weights.estimator_state_dict(c_in, seed=seed), strict=True)
    return cfm


def build_generator(seed, weight_norm):
    _, Generator, AttrDict, v1, _ = ref_shims.load_reference()
    gen = Generator(AttrDict(v1)).eval()
    sd = weights.generator_state_dict(seed, weight_norm=weight_norm)
    if not weight_norm:
        gen.remove_weight_norm()
    gen.load_state_dict(sd, strict=True)
    return gen


def estimator_case(name, c_in, lengths, t_pad, seed, n_steps, t_scalar):
    g = torch.Generator().manual_seed(seed + 1)
    b = len(lengths)
    cfm = build_cfm(c_in, seed)
    mu = torch.randn(b, 80, t_pad, generator=g)
    z = torch.randn(b, 80, t_pad, generator=g) * 0.667
    mask = weights.prefix_mask(lengths, t_pad)
    spks = torch.randn(b, c_in - 160, generator=g) if c_in > 160 else None
    tb = torch.rand(b, generator=g)
    with torch.inference_mode():
        v_scalar = cfm.estimator(z, mask, mu, torch.tensor(t_scalar), spks)
        v_batched = cfm.estimator(z, mask, mu, tb, spks)
        dec = cfm.solve_euler(z, torch.linspace(0, 1, n_steps + 1), mu, mask, spks, None)
    # compute_loss with its two random draws pinned (flow_matching.py:108,110)
    x1 = torch.randn(b, 80, t_pad, generator=g)
    t_draw = torch.rand(b, 1, 1, generator=g)
    z_draw = torch.randn(b, 80, t_pad, generator=g)
    orig_rand, orig_randn_like = torch.rand, torch.randn_like
    torch.rand = lambda *a, **k: t_draw.clone()
    torch.randn_like = lambda *a, **k: z_draw.clone()
    try:
        with torch.no_grad():
            loss, y = cfm.compute_loss(x1, mask, mu, spks)
    finally:
        torch.rand, torch.randn_like = orig_rand, orig_randn_like
    fx = dict(kind="estimator", c_in=c_in, seed=seed, lengths=list(lengths), t_pad=t_pad, n_steps=n_steps,
              t_scalar=t_scalar, mu=mu, z=z, mask=mask, spks=spks, t_batched=tb, v_scalar=v_scalar.clone(),
              v_batched=v_batched.clone(), decoded=dec.clone(), x1=x1, t_draw=t_draw, z_draw=z_draw,
              loss=loss.clone(), y=y.clone())
    torch.save(fx, os.path.join(OUT, name + ".pt"))
    print(name, "v", float(v_scalar.abs().max()), "dec", float(dec.abs().max()), "loss", float(loss))


GRAD_FULL = ("final_proj.weight", "final_block.block.1.weight", "mid_blocks.0.1.0.ff.net.0.alpha",
             "mid_blocks.0.1.0.ff.net.0.beta", "time_mlp.linear_1.bias", "down_blocks.0.1.0.norm1.weight",
             "up_blocks.0.0.mlp.1.bias", "down_blocks.0.2.conv.bias", "up_blocks.0.2.conv.bias")


def grad_case(name, c_in, lengths, t_pad, seed):
    """BASECFM.compute_loss under autograd in the UNMODIFIED reference module (flow_matching.py:87-118): the loss, every
    parameter gradient's L2 norm and sum, a few small gradients in full, and d loss / d mu (the encoder side)."""
    g = torch.Generator().manual_seed(seed + 1)
    b = len(lengths)
    cfm = build_cfm(c_in, seed)             # .eval(): the transformer blocks' Dropout(0.05) (decoder.py:207) is off
    x1 = torch.randn(b, 80, t_pad, generator=g)
    mu = torch.randn(b, 80, t_pad, generator=g).requires_grad_(True)
    mask = weights.prefix_mask(lengths, t_pad)
    spks = torch.randn(b, c_in - 160, generator=g).requires_grad_(True) if c_in > 160 else None
    t_draw = torch.rand(b, 1, 1, generator=g)
    z_draw = torch.randn(b, 80, t_pad, generator=g)
    orig_rand, orig_randn_like = torch.rand, torch.randn_like
    torch.rand = lambda *a, **k: t_draw.clone()
    torch.randn_like = lambda *a, **k: z_draw.clone()
    try:
        loss, y = cfm.compute_loss(x1, mask, mu, spks)
        loss.backward()
    finally:
        torch.rand, torch.randn_like = orig_rand, orig_randn_like
    grads = {k[len("estimator."):]: p.grad.detach().clone() for k, p in cfm.named_parameters()}
    fx = dict(kind="grads", c_in=c_in, seed=seed, lengths=list(lengths), t_pad=t_pad, x1=x1, mu=mu.detach().clone(), mask=mask,
              spks=None if spks is None else spks.detach().clone(), t_draw=t_draw, z_draw=z_draw, loss=loss.detach().clone(),
              grad_norm={k: float(v.double().norm()) for k, v in grads.items()},
              grad_sum={k: float(v.double().sum()) for k, v in grads.items()},
              grad_full={k: grads[k] for k in GRAD_FULL}, dmu=mu.grad.detach().clone(),
              dspks=None if spks is None else spks.grad.detach().clone())
    torch.save(fx, os.path.join(OUT, name + ".pt"))
    print(name, "loss", float(loss), "params", len(grads), "|dmu|", float(mu.grad.norm()))


def denoiser_case(name, seed, lengths, strength):
    """The UNMODIFIED reference Denoiser (hifigan/denoiser.py) over the reference Generator: bias_spec and the denoised
    audio for a ragged set of seeded waveforms (each its own call: the reference takes (L,) or equal-length (B, L))."""
    ref_shims.install()
    from matcha.hifigan.denoiser import Denoiser
    gen = build_generator(seed, weight_norm=False)
    den = Denoiser(gen, mode="zeros")
    g = torch.Generator().manual_seed(seed + 3)
    audios = [torch.tanh(torch.randn(n, generator=g) * 0.3) for n in lengths]
    outs = [den(a, strength=strength).clone() for a in audios]
    batch = torch.stack([torch.tanh(torch.randn(3000, generator=g) * 0.3) for _ in range(3)])
    fx = dict(kind="denoiser", seed=seed, strength=strength, bias_spec=den.bias_spec.clone(), audios=audios, outs=outs,
              batch=batch, batch_out=den(batch, strength=strength).clone())
    torch.save(fx, os.path.join(OUT, name + ".pt"))
    print(name, "bias max", float(den.bias_spec.max()), "out", [tuple(o.shape) for o in outs], tuple(fx["batch_out"].shape))


def generator_case(name, b, t, seed, weight_norm):
    g = torch.Generator().manual_seed(seed + 2)
    gen = build_generator(seed, weight_norm)
    mel = torch.randn(b, 80, t, generator=g) * 2.1 - 5.5  # LJSpeech-like mel statistics
    with torch.inference_mode():
        wav = gen(mel)
    fx = dict(kind="generator", seed=seed, weight_norm=weight_norm, mel=mel, wav=wav.clone())
    torch.save(fx, os.path.join(OUT, name + ".pt"))
    print(name, "wav", float(wav.abs().max()), float(wav.abs().mean()))


def encoder_case(name, n_spks, lengths, seed, length_scale):
    """Upstream of the hot path: the reference TextEncoder + the duration / alignment steps of
    MatchaTTS.synthesise (matcha_tts.py:115-131), run with the reference's own functions."""
    ref_shims.install()
    from matcha.models.components.text_encoder import TextEncoder
    from matcha.utils.model import fix_len_compatibility, generate_path, sequence_mask
    ns = types.SimpleNamespace
    params = ns(n_feats=80, n_channels=192, filter_channels=768, filter_channels_dp=256, n_heads=2, n_layers=6,
                kernel_size=3, p_dropout=0.1, spk_emb_dim=64, n_spks=n_spks, prenet=True)
    dp = ns(filter_channels_dp=256, kernel_size=3, p_dropout=0.1)
    enc = TextEncoder("RoPE Encoder", params, dp, 178, n_spks, 64).eval()
    enc.load_state_dict(weights.text_encoder_state_dict(n_spks=n_spks, seed=seed), strict=True)
    g = torch.Generator().manual_seed(seed)
    b, lx = len(lengths), max(lengths)
    x = torch.randint(1, 178, (b, lx), generator=g)
    x_lengths = torch.tensor(lengths)
    for i, n in enumerate(lengths):
        x[i, n:] = 0
    spks = torch.randn(b, 64, generator=g) if n_spks > 1 else None
    with torch.inference_mode():
        mu_x, logw, x_mask = enc(x, x_lengths, spks)
        w_ceil = torch.ceil(torch.exp(logw) * x_mask) * length_scale
        y_lengths = torch.clamp_min(torch.sum(w_ceil, [1, 2]), 1).long()
        y_max = y_lengths.max()
        t_pad = fix_len_compatibility(y_max)
        y_mask = sequence_mask(y_lengths, t_pad).unsqueeze(1).to(x_mask.dtype)
        attn_mask = x_mask.unsqueeze(-1) * y_mask.unsqueeze(2)
        attn = generate_path(w_ceil.squeeze(1), attn_mask.squeeze(1)).unsqueeze(1)
        mu_y = torch.matmul(attn.squeeze(1).transpose(1, 2), mu_x.transpose(1, 2)).transpose(1, 2)
    fx = dict(n_spks=n_spks, seed=seed, length_scale=length_scale, x=x, x_lengths=x_lengths, spks=spks,
              mu_x=mu_x.clone(), logw=logw.clone(), x_mask=x_mask.clone(), y_lengths=y_lengths.clone(), t_pad=int(t_pad),
              attn=attn.clone().to(torch.uint8), mu_y=mu_y.clone())
    torch.save(fx, os.path.join(OUT, name + ".pt"))
    print(name, "Lx", lx, "y_lengths", y_lengths.tolist(), "t_pad", int(t_pad))


def mas_case(name, x_lengths, y_lengths, seed):
    """Monotonic alignment search through the reference's own Cython code (compiled by oracle/build_ref.py), called the
    way matcha/utils/monotonic_align/__init__.py:7-22 calls it."""
    import numpy as np
    from oracle.build_ref import build, load_ref_mas
    build()
    core = load_ref_mas()
    rng = np.random.default_rng(seed)
    b, tx, ty = len(x_lengths), max(x_lengths), max(y_lengths)
    value = (rng.standard_normal((b, tx, ty)) * 40.0 - 120.0).astype(np.float32)      # log-prior-like magnitudes
    mask = np.zeros((b, tx, ty), dtype=np.float32)
    for i, (nx, ny) in enumerate(zip(x_lengths, y_lengths)):
        mask[i, :nx, :ny] = 1.0
    v = (value * mask).astype(np.float32)
    path = np.zeros_like(v).astype(np.int32)
    t_x_max = mask.sum(1)[:, 0].astype(np.int32)
    t_y_max = mask.sum(2)[:, 0].astype(np.int32)
    core.maximum_path_c(path, v, t_x_max, t_y_max)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), value=value, x_lengths=np.array(x_lengths), y_lengths=np.array(y_lengths),
                        path=path.astype(np.uint8))
    print(name, "Tx", tx, "Ty", ty, "ones per utterance", path.sum((1, 2)).tolist())


def forward_case(name, n_spks, x_lengths, y_lengths, seed):
    """The reference MatchaTTS.forward (matcha_tts.py:150-245) - encoder, log-prior, its own Cython MAS (oracle/_ref),
    duration / prior / flow-matching losses - run unmodified, with the draws of compute_loss recorded."""
    from oracle.build_ref import build
    build()
    from matcha_b200 import matcha_tts as mt       # only for the shipped hyper-parameters (default_config)
    _, _, _, _, MatchaTTS = ref_shims.load_reference()
    cfg = mt.default_config(n_spks=n_spks)
    model = MatchaTTS(n_vocab=178, n_spks=n_spks, spk_emb_dim=64, n_feats=80, encoder=cfg["encoder"],
                      decoder=dict(cfg["decoder"]), cfm=cfg["cfm"], data_statistics=cfg["data_statistics"], out_size=None).eval()
    c_in = 160 if n_spks == 1 else 224
    enc_sd = weights.text_encoder_state_dict(n_spks=n_spks, seed=seed)
    est_sd = weights.estimator_state_dict(c_in, seed=seed + 1)
    sd = {"encoder." + k: v for k, v in enc_sd.items()}
    sd.update({"decoder." + k: v for k, v in est_sd.items()})
    sd["mel_mean"], sd["mel_std"] = torch.tensor(-5.536622), torch.tensor(2.116101)
    g = torch.Generator().manual_seed(seed)
    if n_spks > 1:
        sd["spk_emb.weight"] = torch.randn(n_spks, 64, generator=g)
    model.load_state_dict(sd, strict=True)
    b, lx, ty = len(x_lengths), max(x_lengths), max(y_lengths)
    x = torch.randint(1, 178, (b, lx), generator=g)
    for i, n in enumerate(x_lengths):
        x[i, n:] = 0
    y = torch.randn(b, 80, ty, generator=g)
    for i, n in enumerate(y_lengths):
        y[i, :, n:] = 0
    spks = torch.randint(0, n_spks, (b,), generator=g) if n_spks > 1 else None
    draws = {}
    real_rand, real_randn_like = torch.rand, torch.randn_like

    def rand(*a, **k):
        draws["t"] = real_rand(*a, **k)
        return draws["t"]

    def randn_like(*a, **k):
        draws["z"] = real_randn_like(*a, **k)
        return draws["z"]

    torch.manual_seed(seed)
    torch.rand, torch.randn_like = rand, randn_like
    try:
        dur_loss, prior_loss, diff_loss = model(x, torch.tensor(x_lengths), y, torch.tensor(y_lengths), spks=spks)
        (dur_loss + prior_loss + diff_loss).backward()
    finally:
        torch.rand, torch.randn_like = real_rand, real_randn_like
    named = dict(model.named_parameters())
    keep = ["encoder.proj_m.weight", "encoder.proj_w.proj.weight", "encoder.encoder.norm_layers_2.5.gamma",
            "encoder.prenet.proj.bias", "decoder.estimator.final_proj.weight", "decoder.estimator.mid_blocks.0.0.mlp.1.bias"]
    fx = dict(n_spks=n_spks, seed=seed, x=x, x_lengths=torch.tensor(x_lengths), y=y, y_lengths=torch.tensor(y_lengths), spks=spks,
              spk_emb=sd.get("spk_emb.weight"), t=draws["t"].clone(), z=draws["z"].clone(),
              dur_loss=dur_loss.detach().clone(), prior_loss=prior_loss.detach().clone(), diff_loss=diff_loss.detach().clone(),
              grads={k: named[k].grad.clone() for k in keep})
    torch.save(fx, os.path.join(OUT, name + ".pt"))
    print(name, "losses", float(dur_loss), float(prior_loss), float(diff_loss))


def main():
    import sys
    os.makedirs(OUT, exist_ok=True)
    torch.manual_seed(0)
    torch.set_num_threads(8)
    if "forward" in sys.argv[1:]:          # MatchaTTS.forward (training losses), the caller of 8(f) row 4 and 8(a) a15
        forward_case("fwd_lj_ragged", 1, (13, 7, 10), (40, 24, 33), seed=81)
        forward_case("fwd_vctk_ragged", 109, (9, 12), (28, 36), seed=82)
        return
    if "mas" in sys.argv[1:]:
        mas_case("mas_ragged", (23, 9, 1, 17), (61, 40, 5, 17), seed=71)
        mas_case("mas_long", (60, 41), (181, 120), seed=72)
        return
    if "denoiser" in sys.argv[1:]:
        denoiser_case("denoiser_ragged", 51, (513, 1024, 2303, 5000), strength=0.05)
        return
    if "grads" in sys.argv[1:]:            # only the compute_loss gradient fixtures (added after the forward ones)
        grad_case("grad_lj_ragged", 160, (24, 17, 9), 24, seed=41)
        grad_case("grad_vctk_ragged", 224, (36, 20), 36, seed=42)
        return
    estimator_case("est_lj_ragged", 160, (24, 17, 9), 24, seed=11, n_steps=3, t_scalar=0.3)
    estimator_case("est_lj_min", 160, (4, 1), 4, seed=12, n_steps=2, t_scalar=0.0)
    estimator_case("est_vctk_ragged", 224, (20, 13), 20, seed=13, n_steps=2, t_scalar=0.75)
    estimator_case("est_lj_t140", 160, (140, 77), 140, seed=14, n_steps=2, t_scalar=0.5)
    generator_case("gen_plain", 2, 9, seed=21, weight_norm=False)
    generator_case("gen_weightnorm", 1, 5, seed=22, weight_norm=True)
    encoder_case("enc_lj_ragged", 1, (23, 9, 16), seed=31, length_scale=1.0)
    encoder_case("enc_vctk_scaled", 109, (12, 20), seed=32, length_scale=1.5)


if __name__ == "__main__":
    main()
