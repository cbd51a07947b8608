"""The text encoder + duration predictor as CUDA kernels (csrc/text_encoder.cu), through the C ABI
(`mb200_load_text_encoder` / `mb200_text_encoder_forward`).

Pinned to the golden vectors made by the UNMODIFIED reference TextEncoder (oracle/make_golden.py: encoder_case; tolerance
2e-4, the one the CPU test of the torch restatement uses), and - at BASELINE.json's token counts, where no fixture exists -
to that torch restatement run in fp32 on the CPU (itself pinned to the same fixtures by tests/test_matcha_tts.py).  The
durations the path derives from logw must come out IDENTICAL: they decide every utterance's frame count."""
import os

import pytest
import torch

from matcha_b200 import matcha_tts as mt
from matcha_b200 import synthetic

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = 2e-4


def _encoder(n_spks, seed):
    cfg = mt.default_config(n_spks=n_spks)
    enc = mt.TextEncoder(cfg["encoder"].encoder_type, cfg["encoder"].encoder_params,
                         cfg["encoder"].duration_predictor_params, 178, n_spks, 64).eval()
    enc.load_state_dict(synthetic.text_encoder_state_dict(n_spks=n_spks, seed=seed), strict=True)
    return enc


def _frames(logw, x_mask, scale):
    return torch.clamp_min(torch.sum(torch.ceil(torch.exp(logw) * x_mask) * scale, [1, 2]), 1).long()


@pytest.mark.parametrize("name", ["enc_lj_ragged", "enc_vctk_scaled"])
def test_cuda_text_encoder_matches_reference_golden(name):
    fx = torch.load(os.path.join(GOLDEN, name + ".pt"), weights_only=False)
    enc = _encoder(fx["n_spks"], fx["seed"]).cuda()
    with torch.inference_mode():
        mu_x, logw, x_mask = enc(fx["x"].cuda(), fx["x_lengths"].cuda(), None if fx["spks"] is None else fx["spks"].cuda())
    assert mu_x.is_cuda and getattr(enc, "_plan", None) is not None          # the kernels ran, not the torch formulation
    assert torch.equal(x_mask.cpu(), fx["x_mask"])
    assert float((mu_x.cpu() - fx["mu_x"]).abs().max()) < TOL
    assert float((logw.cpu() - fx["logw"]).abs().max()) < TOL
    assert torch.equal(_frames(logw.cpu(), x_mask.cpu(), fx["length_scale"]), fx["y_lengths"])


@pytest.mark.parametrize("n_spks,b,lx", [(1, 8, 301), (109, 4, 301), (1, 2, 2001), (1, 3, 1), (109, 2, 70)])
def test_cuda_text_encoder_matches_torch_formulation_at_size(n_spks, b, lx):
    """cfg2 / cfg3 token count (150 phonemes -> 301 tokens), cfg4's (1000 -> 2001), a single token, ragged lengths."""
    g = torch.Generator().manual_seed(lx + n_spks)
    enc = _encoder(n_spks, 77)
    x = torch.randint(0, 178, (b, lx), generator=g)
    lens = torch.tensor([lx] + [int(v) for v in torch.randint(1, lx + 1, (b - 1,), generator=g)])
    spks = torch.randn(b, 64, generator=g) if n_spks > 1 else None
    with torch.inference_mode():
        want = enc(x, lens, spks)                                            # torch, fp32, CPU
        got = enc.cuda()(x.cuda(), lens.cuda(), None if spks is None else spks.cuda())
        again = enc(x.cuda(), lens.cuda(), None if spks is None else spks.cuda())
    assert getattr(enc, "_plan", None) is not None
    assert torch.equal(got[2].cpu(), want[2])
    assert float((got[0].cpu() - want[0]).abs().max()) < TOL
    assert float((got[1].cpu() - want[1]).abs().max()) < TOL
    assert torch.equal(_frames(got[1].cpu(), got[2].cpu(), 1.0), _frames(want[1], want[2], 1.0))
    assert all(torch.equal(a, c) for a, c in zip(again, got))      # run to run: identical


def test_training_mode_keeps_the_torch_path():
    enc = _encoder(1, 5).cuda().train()
    x = torch.randint(0, 178, (2, 9)).cuda()
    mu_x, logw, _ = enc(x, torch.tensor([9, 4]).cuda())
    assert mu_x.requires_grad and logw.requires_grad and getattr(enc, "_plan", None) is None
