"""Pins oracle/matcha_oracle.py (the CPU restatement) against the reference:
 (1) committed golden vectors produced by the unmodified reference modules (oracle/make_golden.py);
 (2) live, in the build container, against the reference modules through oracle/ref_shims.py.
CPU only (-m "not gpu")."""
import glob
import os
import types

import pytest
import torch

from oracle import matcha_oracle as mo
from oracle import weights

TOL = 2e-5  # fp32 CPU vs fp32 CPU, different op order (explicit softmax vs SDPA)


def _load(golden_dir, name):
    return torch.load(os.path.join(golden_dir, name + ".pt"), weights_only=False)


EST_CASES = ["est_lj_ragged", "est_lj_min", "est_vctk_ragged", "est_lj_t140"]


@pytest.mark.parametrize("name", EST_CASES)
def test_estimator_matches_golden(golden_dir, name):
    fx = _load(golden_dir, name)
    sd = weights.estimator_state_dict(fx["c_in"], seed=fx["seed"])
    v = mo.estimator_forward(sd, fx["z"], fx["mask"], fx["mu"], torch.tensor(fx["t_scalar"]), fx["spks"])
    assert (v - fx["v_scalar"]).abs().max() < TOL
    vb = mo.estimator_forward(sd, fx["z"], fx["mask"], fx["mu"], fx["t_batched"], fx["spks"])
    assert (vb - fx["v_batched"]).abs().max() < TOL


@pytest.mark.parametrize("name", EST_CASES)
def test_euler_decode_matches_golden(golden_dir, name):
    fx = _load(golden_dir, name)
    sd = weights.estimator_state_dict(fx["c_in"], seed=fx["seed"])
    dec = mo.solve_euler(sd, fx["z"], fx["mu"], fx["mask"], fx["n_steps"], fx["spks"])
    assert (dec - fx["decoded"]).abs().max() < 5 * TOL


@pytest.mark.parametrize("name", EST_CASES[:3])
def test_compute_loss_matches_golden(golden_dir, name):
    fx = _load(golden_dir, name)
    sd = weights.estimator_state_dict(fx["c_in"], seed=fx["seed"])
    loss, y = mo.compute_loss(sd, fx["x1"], fx["mask"], fx["mu"], fx["t_draw"], fx["z_draw"], fx["spks"])
    assert torch.equal(y, fx["y"]) or (y - fx["y"]).abs().max() < 1e-6
    assert abs(float(loss) - float(fx["loss"])) < 1e-4 * abs(float(fx["loss"]))


@pytest.mark.parametrize("name", ["gen_plain", "gen_weightnorm"])
def test_generator_matches_golden(golden_dir, name):
    fx = _load(golden_dir, name)
    sd = weights.generator_state_dict(fx["seed"], weight_norm=fx["weight_norm"])
    wav = mo.hifigan_forward(sd, fx["mel"])
    assert wav.shape == fx["wav"].shape
    assert (wav - fx["wav"]).abs().max() < TOL


def test_euler_schedule_matches_linspace_loop():
    ts, dts = mo.euler_schedule(10)
    assert ts[0] == 0.0 and len(ts) == 10
    assert abs(ts[-1] + dts[-1] - 1.0) < 1e-6


def test_padding_changes_valid_frames():
    """padded frames take part in attention and GroupNorm, so the same
    utterance alone vs padded in a batch differs -- the oracle must reproduce that, not 'fix' it."""
    sd = weights.estimator_state_dict(160, seed=5)
    g = torch.Generator().manual_seed(0)
    mu, z = torch.randn(1, 80, 16, generator=g), torch.randn(1, 80, 16, generator=g)
    t = torch.tensor(0.5)
    alone = mo.estimator_forward(sd, z[:, :, :8], weights.prefix_mask([8], 8), mu[:, :, :8], t)
    padded = mo.estimator_forward(sd, z, weights.prefix_mask([8], 16), mu, t)
    assert (alone - padded[:, :, :8]).abs().max() > 1e-2
    assert padded[:, :, 8:].abs().max() == 0


# ----------------------------------------------------------------------------- live pin (container only)
@pytest.mark.reference
def test_live_reference_estimator_and_state_dict_layout():
    from oracle import ref_shims
    CFM = ref_shims.load_reference()[0]
    dec = dict(channels=[256, 256], dropout=0.05, attention_head_dim=64, n_blocks=1, num_mid_blocks=2,
               num_heads=2, act_fn="snakebeta")
    cfm = CFM(160, 80, types.SimpleNamespace(solver="euler"), dec, n_spks=109, spk_emb_dim=64).eval()
    sd = weights.estimator_state_dict(224, seed=77)
    assert set(sd) == set(cfm.state_dict())
    cfm.load_state_dict(sd, strict=True)
    g = torch.Generator().manual_seed(3)
    b, t_pad = 2, 12
    mu, z = torch.randn(b, 80, t_pad, generator=g), torch.randn(b, 80, t_pad, generator=g)
    spks = torch.randn(b, 64, generator=g)
    mask = weights.prefix_mask([12, 7], t_pad)
    with torch.inference_mode():
        ref = cfm.solve_euler(z, torch.linspace(0, 1, 5), mu, mask, spks, None)
    got = mo.solve_euler(sd, z, mu, mask, 4, spks)
    assert (ref - got).abs().max() < 5 * TOL


@pytest.mark.reference
def test_live_reference_generator_both_layouts():
    from oracle import ref_shims
    _, Generator, AttrDict, v1, _ = ref_shims.load_reference()
    gen = Generator(AttrDict(v1)).eval()
    sd = weights.generator_state_dict(31, weight_norm=True)
    assert set(sd) == set(gen.state_dict())
    gen.load_state_dict(sd, strict=True)
    mel = torch.randn(1, 80, 6) * 2 - 5
    with torch.inference_mode():
        ref = gen(mel)
    assert (mo.hifigan_forward(sd, mel) - ref).abs().max() < TOL
    gen.remove_weight_norm()
    sd2 = {k: v.clone() for k, v in gen.state_dict().items()}
    assert len(sd2) == 156
    assert (mo.hifigan_forward(sd2, mel) - ref).abs().max() < TOL


def test_golden_fixtures_are_small(golden_dir):
    total = sum(os.path.getsize(p) for p in glob.glob(os.path.join(golden_dir, "*.pt")))
    assert 0 < total < 4 << 20


@pytest.mark.parametrize("name", ["grad_lj_ragged", "grad_vctk_ragged"])
def test_compute_loss_gradients_match_reference_golden(golden_dir, name):
    """The oracle's compute_loss under autograd against gradients the unmodified reference module produced
    (oracle/make_golden.py grads): loss, every parameter gradient's norm and sum, a few gradients in full, d/d mu."""
    fx = torch.load(os.path.join(golden_dir, name + ".pt"), weights_only=False)
    sd = {k: v.clone().requires_grad_(True) for k, v in weights.estimator_state_dict(fx["c_in"], seed=fx["seed"]).items()}
    mu = fx["mu"].clone().requires_grad_(True)
    spks = None if fx["spks"] is None else fx["spks"].clone().requires_grad_(True)
    loss, _ = mo.compute_loss(sd, fx["x1"], fx["mask"], mu, fx["t_draw"], fx["z_draw"], spks)
    loss.backward()
    assert abs(float(loss.detach()) - float(fx["loss"])) <= 1e-5 * abs(float(fx["loss"]))
    assert len(fx["grad_norm"]) == 180
    for k, n in fx["grad_norm"].items():
        g = sd["estimator." + k].grad.double()
        assert abs(float(g.norm()) - n) <= 1e-4 * n + 1e-9, k
        assert abs(float(g.sum()) - fx["grad_sum"][k]) <= 1e-4 * n * g.numel() ** 0.5 + 1e-9, k
    for k, g in fx["grad_full"].items():
        got = sd["estimator." + k].grad
        assert float((got - g).norm() / g.norm()) <= 1e-4, k
    assert float((mu.grad - fx["dmu"]).norm() / fx["dmu"].norm()) <= 1e-4
    if spks is not None:
        assert float((spks.grad - fx["dspks"]).norm() / fx["dspks"].norm()) <= 1e-4


def test_denoiser_matches_reference_golden(golden_dir):
    """oracle denoiser (hifigan/denoiser.py restated) against the unmodified reference Denoiser over the reference
    Generator (oracle/make_golden.py denoiser): bias spectrum and denoised audio, ragged lengths and a batch."""
    fx = torch.load(os.path.join(golden_dir, "denoiser_ragged.pt"), weights_only=False)
    sd = weights.generator_state_dict(fx["seed"], weight_norm=False)
    bias = mo.denoiser_bias(sd)
    assert bias.shape == (1, 513, 1)
    assert float((bias - fx["bias_spec"]).abs().max()) <= 1e-4 * float(fx["bias_spec"].abs().max())
    for a, want in zip(fx["audios"], fx["outs"]):
        got = mo.denoise(a, fx["bias_spec"], fx["strength"])
        assert got.shape == want.shape and float((got - want).abs().max()) <= 1e-6
    got = mo.denoise(fx["batch"], fx["bias_spec"], fx["strength"])
    assert float((got - fx["batch_out"]).abs().max()) <= 1e-6


@pytest.mark.parametrize("name", ["enc_lj_ragged", "enc_vctk_scaled"])
def test_length_regulate_matches_reference_golden(golden_dir, name):
    """The oracle's alignment + matmul restatement against the unmodified reference's generate_path / matmul."""
    fx = torch.load(os.path.join(golden_dir, name + ".pt"), weights_only=False)
    w = (torch.ceil(torch.exp(fx["logw"]) * fx["x_mask"]) * fx["length_scale"]).squeeze(1)
    mu_y, attn, y_mask = mo.length_regulate(fx["mu_x"], w, fx["x_mask"].squeeze(1), fx["y_lengths"], fx["t_pad"])
    assert torch.equal(attn.to(torch.uint8), fx["attn"].squeeze(1))
    assert torch.equal(mu_y, fx["mu_y"])
    assert torch.equal(y_mask.sum(1).long(), fx["y_lengths"])
