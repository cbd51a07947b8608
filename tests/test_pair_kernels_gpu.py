"""CTA pairs (tcgen05 cta_group::2, 2-CTA clusters) against the single-CTA kernels: conv_gemm_umma2 (N = 256 and N = 128
layers), the fused ResBlock step (HiFi-GAN stages 3-4) and the fused post-attention kernel.  A pair runs the SAME MMAs in the
same order on the same operands - it only fetches each weight tile half from each CTA's shared memory - so every output must
be EQUAL, not close.  Sizes include odd tile counts (the last pair then has a ghost tile) and single-tile launches (no pair).
The GroupNorm statistics summed by their consumer (MB200_GN_FINISH unset) against the ordered sum inside the conv are
checked the same way: both add the same per-32-row partial sums in index order."""
import pytest
import torch

from matcha_b200 import synthetic as weights

pytestmark = pytest.mark.gpu

PAIR_VARS = ("MB200_PAIR", "MB200_RES_PAIR", "MB200_FF_PAIR")
OPTIONAL_PAIR_VARS = ("MB200_RQ_PAIR", "MB200_PAIR_RESIDENT")   # built, measured neutral, off by default: covered here


def _plan(prec):
    from matcha_b200._lib import Plan
    return Plan(prec, 160, use_graph=False)


@pytest.mark.parametrize("prec", ["fp16", "bf16", "fp32"])
@pytest.mark.parametrize("b,t,lengths", [(1, 4, [3]), (3, 140, [140, 77, 9]), (5, 452, [451, 452, 300, 9, 200])])
def test_estimator_with_cta_pairs_equals_single_cta_kernels(monkeypatch, prec, b, t, lengths):
    sd = weights.estimator_state_dict(160, seed=5)
    g = torch.Generator().manual_seed(3)
    x, mu = torch.randn(b, 80, t, generator=g), torch.randn(b, 80, t, generator=g)
    mask = weights.prefix_mask(lengths, t)
    out = {}
    for pairs in ("on", "off"):
        for v in PAIR_VARS:
            if pairs == "off":
                monkeypatch.setenv(v, "0")
            else:
                monkeypatch.delenv(v, raising=False)
        for v in OPTIONAL_PAIR_VARS:
            monkeypatch.setenv(v, "1" if pairs == "on" else "0")
        plan = _plan(prec)
        plan.load_estimator(sd)
        out[pairs] = plan.estimator_forward(x, mask, mu, torch.tensor(0.4)).cpu()
    assert torch.isfinite(out["on"]).all()
    assert torch.equal(out["on"], out["off"])


@pytest.mark.parametrize("prec", ["fp16", "bf16", "fp32"])
@pytest.mark.parametrize("B,L", [(1, 1), (1, 9), (3, 37)])
def test_vocoder_with_cta_pairs_equals_single_cta_kernels(monkeypatch, prec, B, L):
    sd = weights.generator_state_dict(7, weight_norm=False)
    mel = torch.randn(B, 80, L, generator=torch.Generator().manual_seed(11)) * 1.5 - 4.0
    out = {}
    for pairs in ("on", "off"):
        for v in PAIR_VARS:
            if pairs == "off":
                monkeypatch.setenv(v, "0")
            else:
                monkeypatch.delenv(v, raising=False)
        for v in OPTIONAL_PAIR_VARS:
            monkeypatch.setenv(v, "1" if pairs == "on" else "0")
        plan = _plan(prec)
        plan.load_generator(sd)
        out[pairs] = plan.hifigan_forward(mel).cpu()
    assert out["on"].shape == (B, 1, 256 * L)
    assert torch.isfinite(out["on"]).all()
    assert torch.equal(out["on"], out["off"])


@pytest.mark.parametrize("prec", ["fp16", "fp32"])
def test_statistics_summed_by_the_consumer_equal_the_ordered_sum_in_the_conv(monkeypatch, prec):
    sd = weights.estimator_state_dict(160, seed=5)
    g = torch.Generator().manual_seed(4)
    b, t = 4, 300
    x, mu = torch.randn(b, 80, t, generator=g), torch.randn(b, 80, t, generator=g)
    mask = weights.prefix_mask([300, 123, 299, 31], t)
    out = {}
    for finish in ("consumer", "conv"):
        if finish == "conv":
            monkeypatch.setenv("MB200_GN_FINISH", "1")
        else:
            monkeypatch.delenv("MB200_GN_FINISH", raising=False)
        plan = _plan(prec)
        plan.load_estimator(sd)
        out[finish] = plan.estimator_forward(x, mask, mu, torch.tensor(0.7)).cpu()
    assert torch.equal(out["consumer"], out["conv"])
