"""HiFi-GAN bias denoiser (matcha/hifigan/denoiser.py) on the CUDA path against the golden
vectors of the unmodified reference Denoiser and against the CPU oracle.  fp32 path: max-abs <= 1e-4 on O(1) audio
(the radix-2 shared-memory FFT and torch's pocketfft / cuFFT differ by fp32 rounding only).  Needs a B200."""
import os

import pytest
import torch

import matcha_b200 as mb
from matcha_b200._lib import Plan
from oracle import matcha_oracle as mo
from oracle import weights

pytestmark = pytest.mark.gpu
TOL = 1e-4


def test_denoise_kernel_matches_reference_golden(golden_dir):
    fx = torch.load(os.path.join(golden_dir, "denoiser_ragged.pt"), weights_only=False)
    plan = Plan("fp32", 160, use_graph=False)
    bias = fx["bias_spec"].reshape(-1)
    for a, want in zip(fx["audios"], fx["outs"]):
        got = plan.denoise(a.reshape(1, -1), bias, fx["strength"]).cpu()
        assert got.shape == want.shape
        assert float((got - want).abs().max()) <= TOL, (a.numel(), float((got - want).abs().max()))
    got = plan.denoise(fx["batch"], bias, fx["strength"]).cpu()
    assert got.shape == fx["batch_out"].shape
    assert float((got - fx["batch_out"]).abs().max()) <= TOL
    # strength 0 is the identity away from the edges' window-envelope rounding (STFT -> iSTFT reconstructs the signal)
    a = fx["audios"][-1]
    ident = plan.denoise(a.reshape(1, -1), bias, 0.0).cpu()
    assert float((ident[0] - a[:ident.shape[1]]).abs().max()) <= TOL


def test_denoiser_module_is_a_drop_in(golden_dir):
    """Denoiser(vocoder) over the CUDA Generator: bias_spec matches the reference's (computed there with the reference
    Generator in fp32) and forward() matches the oracle at the cli's strength; 1-D audio returns (1, L') as upstream."""
    fx = torch.load(os.path.join(golden_dir, "denoiser_ragged.pt"), weights_only=False)
    gen = mb.Generator(mb.AttrDict(mb.HIFIGAN_V1), precision="fp32")
    gen.remove_weight_norm()
    gen.load_state_dict(weights.generator_state_dict(fx["seed"], weight_norm=False), strict=True)
    gen = gen.cuda().eval()
    den = mb.Denoiser(gen, mode="zeros")
    assert den.bias_spec.shape == (1, 513, 1)
    ref = fx["bias_spec"]
    assert float((den.bias_spec.cpu() - ref).abs().max()) <= 2e-3 * float(ref.abs().max())   # through the fp32 CUDA vocoder
    g = torch.Generator().manual_seed(4)
    wav = torch.tanh(torch.randn(2, 256 * 37, generator=g) * 0.5)
    out = den(wav.cuda(), strength=0.00025).cpu()
    want = mo.denoise(wav, den.bias_spec.cpu(), 0.00025)
    assert out.shape == want.shape == (2, 256 * 37)
    assert float((out - want).abs().max()) <= TOL
    one = den(wav[0].cuda(), strength=0.01).cpu()
    assert one.shape == (1, 256 * 37)
    assert float((one - mo.denoise(wav[0], den.bias_spec.cpu(), 0.01)).abs().max()) <= TOL
    with pytest.raises(NotImplementedError):
        mb.Denoiser(gen, filter_length=512)
    with pytest.raises(RuntimeError):
        den(torch.zeros(1, 400).cuda())


def test_denoise_full_size_properties_cfg2():
    """cfg2 vocoder output size (64 x 115 456 samples): linear in scale (gain depends on |S| only through the bias term, so
    with bias 0 the map is the identity), idempotent-monotone in strength (energy does not grow), finite."""
    plan = Plan("fp32", 160, use_graph=False)
    g = torch.Generator().manual_seed(9)
    wav = torch.tanh(torch.randn(64, 256 * 451, generator=g) * 0.4).cuda()
    bias = torch.rand(513, generator=g).cuda() * 20
    out0 = plan.denoise(wav, torch.zeros(513).cuda(), 1.0)
    assert float((out0 - wav).abs().max()) <= TOL
    e_prev = float(wav.pow(2).sum())
    for strength in (0.01, 0.1, 1.0):
        out = plan.denoise(wav, bias, strength)
        assert torch.isfinite(out).all()
        e = float(out.pow(2).sum())
        assert e <= e_prev * (1 + 1e-5)
        e_prev = e
