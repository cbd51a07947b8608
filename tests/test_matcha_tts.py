"""`matcha_b200.MatchaTTS`: the synthesis-side drop-in (SURVEY §8(b) boundary #1).

CPU tests pin the upstream of the hot path (text encoder, durations, length regulation) against golden vectors made by
the UNMODIFIED reference TextEncoder / generate_path (oracle/make_golden.py: encoder_case).  The GPU test runs the
whole `synthesise` call and checks the decoder output against the CPU oracle fed the same mu_y.
"""
import os

import pytest
import torch

import matcha_b200 as mb
from matcha_b200 import matcha_tts as mt
from matcha_b200 import synthetic

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _fx(name):
    return torch.load(os.path.join(GOLDEN, name + ".pt"), weights_only=False)


def _encoder(fx):
    cfg = mt.default_config(n_spks=fx["n_spks"])
    enc = mt.TextEncoder(cfg["encoder"].encoder_type, cfg["encoder"].encoder_params,
                         cfg["encoder"].duration_predictor_params, 178, fx["n_spks"], 64).eval()
    enc.load_state_dict(synthetic.text_encoder_state_dict(n_spks=fx["n_spks"], seed=fx["seed"]), strict=True)
    return enc


@pytest.mark.parametrize("name", ["enc_lj_ragged", "enc_vctk_scaled"])
def test_text_encoder_matches_reference_golden(name):
    fx = _fx(name)
    enc = _encoder(fx)
    with torch.inference_mode():
        mu_x, logw, x_mask = enc(fx["x"], fx["x_lengths"], fx["spks"])
    assert torch.equal(x_mask, fx["x_mask"])
    assert float((mu_x - fx["mu_x"]).abs().max()) < 2e-4
    assert float((logw - fx["logw"]).abs().max()) < 2e-4


def test_length_regulation_has_no_cpu_path():
    """north_star: no CPU fallback in the product.  `length_regulate` is the CUDA kernel pair only (the GPU tests below pin
    it bit-exactly to the reference-generated fixtures; the torch restatement of the alignment matmul is the oracle's,
    pinned in tests/test_oracle.py)."""
    fx = _fx("enc_lj_ragged")
    w_ceil = torch.ceil(torch.exp(fx["logw"]) * fx["x_mask"]) * fx["length_scale"]
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        mt.length_regulate(fx["mu_x"], w_ceil, fx["x_mask"], fx["y_lengths"], fx["t_pad"])


def test_fix_len_and_generate_path_edges():
    assert [mt.fix_len_compatibility(n) for n in (1, 4, 5, 451, 3001)] == [4, 4, 8, 452, 3004]
    dur = torch.tensor([[2.0, 0.0, 3.0]])
    mask = torch.ones(1, 3, 8)
    path = mt.generate_path(dur, mask)
    assert path[0].tolist() == [[1, 1, 0, 0, 0, 0, 0, 0], [0] * 8, [0, 0, 1, 1, 1, 0, 0, 0]]


def test_module_surface_and_state_dict_layout():
    for n_spks in (1, 109):
        model = mb.MatchaTTS(**mt.default_config(n_spks=n_spks))
        keys = set(model.state_dict().keys())
        assert {"mel_mean", "mel_std", "encoder.emb.weight", "encoder.prenet.proj.weight",
                "encoder.encoder.attn_layers.5.conv_o.bias", "encoder.encoder.norm_layers_2.0.gamma",
                "encoder.proj_w.norm_2.beta", "encoder.proj_m.weight",
                "decoder.estimator.final_proj.weight"} <= keys
        assert ("spk_emb.weight" in keys) == (n_spks > 1)
        enc_keys = {k[len("encoder."):] for k in keys if k.startswith("encoder.")}
        assert enc_keys == set(synthetic.text_encoder_state_dict(n_spks=n_spks).keys())
        c_in = 160 if n_spks == 1 else 224
        assert model.state_dict()["decoder.estimator.time_mlp.linear_1.weight"].shape == (1024, c_in)
    with pytest.raises(RuntimeError):      # training forward needs CUDA tensors: the alignment search has no CPU fallback
        x = torch.randint(1, 178, (1, 5))
        model(x, torch.tensor([5]), torch.randn(1, 80, 8), torch.tensor([8]),
              spks=torch.tensor([3]) if n_spks > 1 else None)


@pytest.mark.gpu
@pytest.mark.parametrize("n_spks", [1, 109])
def test_synthesise_end_to_end_against_oracle(n_spks):
    from oracle import matcha_oracle as mo
    c_in = 160 if n_spks == 1 else 224
    model = mb.MatchaTTS(**mt.default_config(n_spks=n_spks))
    enc_sd = synthetic.text_encoder_state_dict(n_spks=n_spks, seed=5)
    est_sd = synthetic.estimator_state_dict(c_in, seed=6)
    sd = {"encoder." + k: v for k, v in enc_sd.items()}
    sd.update({"decoder." + k: v for k, v in est_sd.items()})
    sd["mel_mean"], sd["mel_std"] = torch.tensor(-5.5), torch.tensor(2.1)
    g = torch.Generator().manual_seed(3)
    if n_spks > 1:
        sd["spk_emb.weight"] = torch.randn(n_spks, 64, generator=g)
    model.load_state_dict(sd, strict=True)
    model = model.cuda().eval()
    model.decoder.estimator.precision = "fp32"
    x = torch.randint(1, 178, (3, 19), generator=g)
    x_lengths = torch.tensor([19, 7, 12])
    spk_ids = torch.tensor([4, 100, 0]) if n_spks > 1 else None
    out = model.synthesise(x.cuda(), x_lengths.cuda(), n_timesteps=3, temperature=0.0,
                           spks=None if spk_ids is None else spk_ids.cuda(), length_scale=1.0)
    y_len = out["mel_lengths"].cpu()
    y_max = int(y_len.max())
    assert out["decoder_outputs"].shape == (3, 80, y_max) and out["encoder_outputs"].shape == (3, 80, y_max)
    assert out["attn"].shape[0] == 3 and out["attn"].shape[1] == 1 and isinstance(out["rtf"], float)
    # oracle: same mu_y (padded to the estimator's length), zero noise (temperature 0), CPU fp32 Euler decode
    t_pad = mt.fix_len_compatibility(y_max)
    mu_y = torch.zeros(3, 80, t_pad)
    mu_y[:, :, :y_max] = out["encoder_outputs"].cpu()
    mask = mt.sequence_mask(y_len, t_pad).unsqueeze(1).float()
    spk_vec = None if spk_ids is None else sd["spk_emb.weight"][spk_ids]
    ref = mo.solve_euler(est_sd, torch.zeros_like(mu_y), mu_y, mask, 3, spks=spk_vec)[:, :, :y_max]
    assert float((out["decoder_outputs"].cpu() - ref).abs().max()) <= 1e-3
    assert torch.allclose(out["mel"].cpu(), out["decoder_outputs"].cpu() * 2.1 - 5.5, atol=1e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["enc_lj_ragged", "enc_vctk_scaled"])
def test_cuda_length_regulation_matches_reference_golden(name):
    """`mb200_length_regulate` (prefix sum + gather) against the unmodified reference's generate_path + matmul: bit-exact."""
    fx = _fx(name)
    w_ceil = torch.ceil(torch.exp(fx["logw"]) * fx["x_mask"]) * fx["length_scale"]
    mu_y, attn, y_mask = mt.length_regulate(fx["mu_x"].cuda(), w_ceil.cuda(), fx["x_mask"].cuda(), fx["y_lengths"].cuda(),
                                            fx["t_pad"])
    assert mu_y.is_cuda and attn.shape == (fx["mu_x"].shape[0], fx["mu_x"].shape[2], fx["t_pad"])
    assert torch.equal(attn.cpu().to(torch.uint8), fx["attn"].squeeze(1))
    assert torch.equal(mu_y.cpu(), fx["mu_y"])
    assert y_mask.shape == (fx["mu_x"].shape[0], 1, fx["t_pad"])
    assert torch.equal(y_mask.cpu().squeeze(1).sum(1).long(), fx["y_lengths"])


@pytest.mark.gpu
@pytest.mark.parametrize("b,lx,scale", [(4, 2001, 1.5), (64, 301, 1.5), (3, 57, 0.37), (2, 1, 1.0), (5, 130, 0.9)])
def test_cuda_length_regulation_matches_oracle(b, lx, scale):
    """Seeded cases up to BASELINE cfg4's size (Lx 2001 -> 3001 frames) and the edges: tokens with zero frames, an
    utterance whose durations are all zero (y_length clamps to 1 and its only frame has no owner), a non-dyadic
    length_scale (prefix sums inexact in fp32), Lx = 1, ragged token counts."""
    from oracle import matcha_oracle as mo
    g = torch.Generator().manual_seed(100 * b + lx)
    x_len = torch.randint(max(1, lx // 3), lx + 1, (b,), generator=g)
    x_len[0] = lx
    x_mask = (torch.arange(lx)[None, :] < x_len[:, None]).float()
    dur = torch.randint(0, 4, (b, lx), generator=g).float() if scale != 1.5 else torch.full((b, lx), 1.0)
    if b > 1:
        dur[1] = 0.0                                  # an utterance with no frames at all
    w = dur * x_mask * scale
    y_len = torch.clamp_min(w.sum(1), 1).long()
    t_pad = mt.fix_len_compatibility(int(y_len.max()))
    mu_x = torch.randn(b, 80, lx, generator=g)
    ref_mu, ref_attn, ref_ym = mo.length_regulate(mu_x, w, x_mask, y_len, t_pad)
    mu_y, attn, y_mask = mt.length_regulate(mu_x.cuda(), w.unsqueeze(1).cuda(), x_mask.unsqueeze(1).cuda(), y_len.cuda(), t_pad)
    assert torch.equal(attn.cpu(), ref_attn)
    assert torch.equal(mu_y.cpu(), ref_mu)
    assert torch.equal(y_mask.cpu().squeeze(1), ref_ym)
    # size-independent properties: every frame below y_length of an utterance with frames has exactly one owner,
    # owners are monotone in the frame index, padded frames have none
    own = attn.cpu().sum(1)
    for i in range(b):
        n = int(y_len[i])
        if float(w[i].sum()) >= 1:
            assert torch.equal(own[i, :n], torch.ones(n))
            tok = attn.cpu()[i, :, :n].argmax(0)
            assert bool((tok[1:] >= tok[:-1]).all())
        assert float(own[i, n:].abs().sum()) == 0.0


@pytest.mark.gpu
def test_cuda_length_regulation_unaligned_frame_count():
    """t_pad that is not a multiple of 4 (rows of the alignment lose their 16-byte alignment: scalar store path)."""
    from matcha_b200 import _lib
    from oracle import matcha_oracle as mo
    g = torch.Generator().manual_seed(9)
    b, lx = 3, 41
    x_mask = torch.ones(b, lx)
    w = torch.randint(0, 5, (b, lx), generator=g).float() * 0.77
    y_len = torch.clamp_min(w.sum(1), 1).long()
    mu_x = torch.randn(b, 80, lx, generator=g)
    for t_pad in (int(y_len.max()) + 1, int(y_len.max()) + 2, int(y_len.max()) + 3):
        ref_mu, ref_attn, ref_ym = mo.length_regulate(mu_x, w, x_mask, y_len, t_pad)
        mu_y, attn, y_mask = _lib.length_regulate(mu_x.cuda(), w.cuda(), x_mask.cuda(), y_len.cuda(), t_pad)
        assert torch.equal(attn.cpu(), ref_attn) and torch.equal(mu_y.cpu(), ref_mu) and torch.equal(y_mask.cpu(), ref_ym)
        mu_only, none, _ = _lib.length_regulate(mu_x.cuda(), w.cuda(), x_mask.cuda(), y_len.cuda(), t_pad, want_attn=False)
        assert none is None and torch.equal(mu_only.cpu(), ref_mu)


def _training_model(fx):
    n_spks = fx["n_spks"]
    model = mb.MatchaTTS(**mt.default_config(n_spks=n_spks))
    enc_sd = synthetic.text_encoder_state_dict(n_spks=n_spks, seed=fx["seed"])
    est_sd = synthetic.estimator_state_dict(160 if n_spks == 1 else 224, seed=fx["seed"] + 1)
    sd = {"encoder." + k: v for k, v in enc_sd.items()}
    sd.update({"decoder." + k: v for k, v in est_sd.items()})
    sd["mel_mean"], sd["mel_std"] = torch.tensor(-5.536622), torch.tensor(2.116101)
    if n_spks > 1:
        sd["spk_emb.weight"] = fx["spk_emb"]
    model.load_state_dict(sd, strict=True)
    return model.eval(), est_sd


@pytest.mark.parametrize("name", ["fwd_lj_ragged", "fwd_vctk_ragged"])
def test_training_losses_oracle_chain_matches_reference_forward(name):
    """Encoder (this package, CPU) -> log-prior -> oracle MAS -> losses with the oracle's compute_loss, against the
    unmodified reference `MatchaTTS.forward` (its own Cython MAS): oracle/make_golden.py forward_case."""
    import math
    import numpy as np
    from oracle import mas as omas
    from oracle import matcha_oracle as mo
    fx = _fx(name)
    model, est_sd = _training_model(fx)
    with torch.inference_mode():
        spks = model.spk_emb(fx["spks"]) if fx["n_spks"] > 1 else None
        mu_x, logw, x_mask = model.encoder(fx["x"], fx["x_lengths"], spks)
        y, y_mask = fx["y"], mt.sequence_mask(fx["y_lengths"], fx["y"].shape[-1]).unsqueeze(1).float()
        attn_mask = (x_mask.unsqueeze(-1) * y_mask.unsqueeze(2)).squeeze(1)
        factor = -0.5 * torch.ones_like(mu_x)
        log_prior = (torch.matmul(factor.transpose(1, 2), y ** 2) - torch.matmul(2.0 * (factor * mu_x).transpose(1, 2), y)
                     + torch.sum(factor * mu_x ** 2, 1).unsqueeze(-1) - 0.5 * math.log(2 * math.pi) * 80)
        attn = torch.from_numpy(omas.maximum_path(log_prior.numpy(), attn_mask.numpy()).astype(np.float32))
        logw_ = torch.log(1e-8 + attn.sum(-1).unsqueeze(1)) * x_mask
        dur = torch.sum((logw - logw_) ** 2) / torch.sum(fx["x_lengths"])
        mu_y = torch.matmul(attn.transpose(1, 2), mu_x.transpose(1, 2)).transpose(1, 2)
        prior = torch.sum(0.5 * ((y - mu_y) ** 2 + math.log(2 * math.pi)) * y_mask) / (torch.sum(y_mask) * 80)
        diff, _ = mo.compute_loss(est_sd, y, y_mask, mu_y, fx["t"], fx["z"], spks=spks)
    for got, key in ((dur, "dur_loss"), (prior, "prior_loss"), (diff, "diff_loss")):
        assert abs(float(got) - float(fx[key])) <= 2e-4 * abs(float(fx[key])), key


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["fwd_lj_ragged", "fwd_vctk_ragged"])
def test_training_forward_matches_reference_forward(name, monkeypatch):
    """`MatchaTTS.forward` on the GPU (CUDA monotonic alignment search + CUDA compute_loss forward/backward) against the
    unmodified reference forward: the three losses and a sample of encoder- and decoder-side gradients."""
    fx = _fx(name)
    model, _ = _training_model(fx)
    model = model.cuda()
    t, z = fx["t"].cuda(), fx["z"].cuda()
    monkeypatch.setattr(torch, "rand", lambda *a, **k: t)
    monkeypatch.setattr(torch, "randn_like", lambda *a, **k: z)
    spks = fx["spks"].cuda() if fx["n_spks"] > 1 else None
    dur, prior, diff = model(fx["x"].cuda(), fx["x_lengths"].cuda(), fx["y"].cuda(), fx["y_lengths"].cuda(), spks=spks)
    for got, key in ((dur, "dur_loss"), (prior, "prior_loss"), (diff, "diff_loss")):
        assert abs(float(got) - float(fx[key])) <= 1e-3 * abs(float(fx[key])), key      # fp32 path, tolerance 1e-3 relative
    (dur + prior + diff).backward()
    named = dict(model.named_parameters())
    rels = {}
    for k, ref in fx["grads"].items():
        g = named[k].grad
        assert g is not None, k
        rels[k] = float((g.cpu() - ref).norm() / ref.norm().clamp_min(1e-12))
    # decoder parameters and the encoder's output layers: 2e-3.  The prenet bias sits under six encoder layers and its
    # gradient is a sum over all positions with heavy cancellation (measured 0.6-1.4 % between torch's CUDA and CPU
    # autograd of the same encoder fed the CUDA-computed d(loss)/d(mu_y)): 3e-2.
    for k, rel in rels.items():
        assert rel <= (3e-2 if k == "encoder.prenet.proj.bias" else 2e-3), rels


@pytest.mark.gpu
def test_training_forward_segment_cut(monkeypatch):
    """`out_size`: a window as long as the batch reproduces the uncut losses; a shorter one trains on zero-padded random
    windows (finite losses, gradients reach encoder and decoder)."""
    fx = _fx("fwd_lj_ragged")
    model, _ = _training_model(fx)
    model = model.cuda()
    t, z = fx["t"].cuda(), fx["z"].cuda()
    monkeypatch.setattr(torch, "rand", lambda *a, **k: t)
    monkeypatch.setattr(torch, "randn_like", lambda *a, **k: z)
    args = (fx["x"].cuda(), fx["x_lengths"].cuda(), fx["y"].cuda(), fx["y_lengths"].cuda())
    full = model(*args, out_size=fx["y"].shape[-1])
    for got, key in zip(full, ("dur_loss", "prior_loss", "diff_loss")):
        assert abs(float(got) - float(fx[key])) <= 1e-3 * abs(float(fx[key])), key
    monkeypatch.undo()
    import random
    random.seed(0)
    dur, prior, diff = model(*args, out_size=16)
    assert all(torch.isfinite(v) for v in (dur, prior, diff))
    (dur + prior + diff).backward()
    named = dict(model.named_parameters())
    assert float(named["encoder.proj_m.weight"].grad.abs().sum()) > 0
    assert float(named["decoder.estimator.final_proj.weight"].grad.abs().sum()) > 0
