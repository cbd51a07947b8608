"""BASECFM.compute_loss (flow_matching.py:87-118) forward + backward on the CUDA training path, against the CPU oracle
under torch autograd (same weights, same injected t and z).  The path is fp32 throughout, so the bar is the fp32 one:
loss and velocity max-abs <= 1e-3, every parameter gradient within 2e-3 relative L2 of autograd's.  Needs a B200."""
import os
import types

import pytest
import torch

import matcha_b200 as mb
from matcha_b200._lib import Plan
from oracle import matcha_oracle as mo
from oracle import weights

pytestmark = pytest.mark.gpu

DEC = dict(channels=[256, 256], dropout=0.05, attention_head_dim=64, n_blocks=1, num_mid_blocks=2, num_heads=2,
           act_fn="snakebeta")
CFM_PARAMS = types.SimpleNamespace(solver="euler", sigma_min=1e-4)
GRAD_REL_L2 = 2e-3


def rel_l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))


def _case(b, t, lengths, c_in, seed):
    g = torch.Generator().manual_seed(seed)
    x1 = torch.randn(b, 80, t, generator=g)
    mu = torch.randn(b, 80, t, generator=g)
    z = torch.randn(b, 80, t, generator=g)
    tt = torch.rand(b, 1, 1, generator=g)
    mask = weights.prefix_mask(lengths, t)
    spks = torch.randn(b, c_in - 160, generator=g) if c_in > 160 else None
    return x1, mask, mu, z, tt, spks


def _oracle(sd, x1, mask, mu, z, tt, spks):
    sd = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    mu = mu.clone().requires_grad_(True)
    spks = None if spks is None else spks.clone().requires_grad_(True)
    loss, y = mo.compute_loss(sd, x1, mask, mu, tt, z, spks)
    loss.backward()
    return float(loss.detach()), y.detach(), {k: v.grad for k, v in sd.items()}, mu.grad, None if spks is None else spks.grad


@pytest.mark.parametrize("b,t,lengths,c_in", [(2, 24, [24, 13], 160), (3, 52, [52, 40, 7], 224), (2, 136, [130, 136], 160)])
def test_compute_loss_forward_backward_matches_autograd_oracle(b, t, lengths, c_in):
    sd = weights.estimator_state_dict(c_in, seed=7)
    x1, mask, mu, z, tt, spks = _case(b, t, lengths, c_in, 11)
    loss_o, y_o, grads_o, dmu_o, dspk_o = _oracle(sd, x1, mask, mu, z, tt, spks)

    dev = torch.device("cuda")
    plan = Plan("fp32", c_in, use_graph=False)
    named = [(k[len("estimator."):], v.to(dev).contiguous()) for k, v in sd.items()]
    out = plan.compute_loss(named, x1, mask, mu, z, tt, spks, 1e-4, backward=True, want_dmu=True, want_dspks=c_in > 160)
    torch.cuda.synchronize()
    assert abs(float(out["loss"]) - loss_o) <= 1e-3 * max(1.0, abs(loss_o)), (float(out["loss"]), loss_o)
    assert float((out["y"].cpu() - y_o).abs().max()) <= 1e-5
    with torch.no_grad():
        v_o = mo.estimator_forward(sd, y_o, mask, mu, tt.reshape(-1), spks)
    assert float((out["v"].cpu() - v_o).abs().max()) <= 1e-3
    worst = ("", 0.0)
    for k, g in out["grads"].items():
        want = grads_o["estimator." + k]
        assert torch.isfinite(g).all(), k
        r = rel_l2(g.cpu(), want)
        if r > worst[1]:
            worst = (k, r)
    assert worst[1] <= GRAD_REL_L2, worst
    assert rel_l2(out["dmu"].cpu(), dmu_o) <= GRAD_REL_L2
    if c_in > 160:
        assert rel_l2(out["dspks"].cpu(), dspk_o) <= GRAD_REL_L2


def test_compute_loss_module_is_an_autograd_node():
    """Drop-in use: CFM.compute_loss(...)[0].backward() fills .grad of the estimator's parameters (and of mu, which the
    text encoder produced) with the CUDA-computed gradients; an SGD step then lowers the loss."""
    dev = torch.device("cuda")
    sd = weights.estimator_state_dict(160, seed=3)
    cfm = mb.CFM(160, 80, CFM_PARAMS, DEC, n_spks=1, spk_emb_dim=64).to(dev)
    cfm.load_state_dict(sd, strict=True)
    cfm.eval()      # no dropout: the oracle below runs without it (training-mode dropout is pinned further down)
    x1, mask, mu, z, tt, _ = _case(2, 40, [40, 29], 160, 5)
    x1, mask, z, tt = x1.to(dev), mask.to(dev), z.to(dev), tt.to(dev)
    mu = mu.to(dev).requires_grad_(True)
    loss, y = cfm.compute_loss(x1, mask, mu, t=tt, z=z)
    assert loss.requires_grad and not y.requires_grad
    loss.backward()
    loss_o, _, grads_o, dmu_o, _ = _oracle(sd, x1.cpu(), mask.cpu(), mu.detach().cpu(), z.cpu(), tt.cpu(), None)
    assert abs(float(loss) - loss_o) <= 1e-3 * max(1.0, abs(loss_o))
    for k, p in cfm.named_parameters():
        assert p.grad is not None, k
        assert rel_l2(p.grad.cpu(), grads_o[k]) <= GRAD_REL_L2, k
    assert rel_l2(mu.grad.cpu(), dmu_o) <= GRAD_REL_L2
    with torch.no_grad():
        for p in cfm.parameters():
            p -= 1e-3 * p.grad
    loss2, _ = cfm.compute_loss(x1, mask, mu.detach(), t=tt, z=z)
    assert float(loss2) < float(loss)


def test_compute_loss_property_cfg5_shape():
    """BASELINE cfg5 shape (B=32, T=512, ragged lengths, per-sample t): gradients are finite, padded frames carry no
    gradient into mu, and the loss is invariant to what sits in x1's valid region ordering of the batch (permutation
    of utterances permutes nothing in the summed loss)."""
    dev = torch.device("cuda")
    sd = weights.estimator_state_dict(160, seed=9)
    b, t = 32, 512
    g = torch.Generator().manual_seed(2)
    lengths = [int(v) for v in (torch.randint(64, 129, (b,), generator=g) * 4)]
    x1, mask, mu, z, tt, _ = _case(b, t, lengths, 160, 21)
    plan = Plan("fp32", 160, use_graph=False)
    named = [(k[len("estimator."):], v.to(dev).contiguous()) for k, v in sd.items()]
    out = plan.compute_loss(named, x1, mask, mu, z, tt, None, 1e-4, backward=True, want_dmu=True)
    perm = torch.randperm(b, generator=g)
    out2 = plan.compute_loss(named, x1[perm], mask[perm], mu[perm], z[perm], tt[perm], None, 1e-4, backward=True)
    torch.cuda.synchronize()
    assert torch.isfinite(out["loss"]).all()
    assert abs(float(out["loss"]) - float(out2["loss"])) <= 1e-4 * abs(float(out["loss"]))
    dmu = out["dmu"].cpu()
    assert torch.isfinite(dmu).all()
    assert float((dmu * (1 - mask)).abs().max()) == 0.0
    for k, gr in out["grads"].items():
        assert torch.isfinite(gr).all(), k
        assert rel_l2(out2["grads"][k], gr) <= 1e-3, k


@pytest.mark.parametrize("name", ["grad_lj_ragged", "grad_vctk_ragged"])
def test_compute_loss_gradients_match_reference_golden(golden_dir, name):
    """The CUDA backward against gradients produced by the UNMODIFIED reference CFM.compute_loss under autograd
    (tests/golden/grad_*.pt, generated by oracle/make_golden.py grads in the build container)."""
    fx = torch.load(os.path.join(golden_dir, name + ".pt"), weights_only=False)
    sd = weights.estimator_state_dict(fx["c_in"], seed=fx["seed"])
    dev = torch.device("cuda")
    plan = Plan("fp32", fx["c_in"], use_graph=False)
    named = [(k[len("estimator."):], v.to(dev).contiguous()) for k, v in sd.items()]
    out = plan.compute_loss(named, fx["x1"], fx["mask"], fx["mu"], fx["z_draw"], fx["t_draw"], fx["spks"], 1e-4, backward=True,
                            want_dmu=True, want_dspks=fx["spks"] is not None)
    assert abs(float(out["loss"]) - float(fx["loss"])) <= 1e-4 * abs(float(fx["loss"]))
    assert set(out["grads"]) == set(fx["grad_norm"])
    for k, n in fx["grad_norm"].items():
        g = out["grads"][k].double()
        assert abs(float(g.norm()) - n) <= GRAD_REL_L2 * n + 1e-9, (k, float(g.norm()), n)
    for k, g in fx["grad_full"].items():
        assert rel_l2(out["grads"][k].cpu(), g) <= GRAD_REL_L2, k
    assert rel_l2(out["dmu"].cpu(), fx["dmu"]) <= GRAD_REL_L2
    if fx["spks"] is not None:
        assert rel_l2(out["dspks"].cpu(), fx["dspks"]) <= GRAD_REL_L2


@pytest.mark.parametrize("b,t,lengths,c_in,p", [(2, 24, [24, 13], 160, 0.05), (3, 52, [52, 40, 7], 224, 0.3)])
def test_training_dropout_matches_autograd_oracle_with_the_same_masks(b, t, lengths, c_in, p):
    """Dropout of the transformer blocks in training mode (decoder.py:206 -> transformer.py:124, :200; rate 0.05 shipped).
    The CUDA path draws its keep masks from a counter-based hash of (seed, site, element); the oracle applies the SAME masks
    at the reference's two positions per block and differentiates with torch autograd: loss, velocity and every gradient
    must agree at the fp32 bar.  (Which elements are dropped is this package's stream, not torch's Philox - as with any two
    seeds of the reference; where, how often and how the survivors are scaled is the reference's.)"""
    sd = weights.estimator_state_dict(c_in, seed=7)
    x1, mask, mu, z, tt, spks = _case(b, t, lengths, c_in, 11)
    seed = 987654321
    sdo = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    mu_o = mu.clone().requires_grad_(True)
    loss_o, _ = mo.compute_loss(sdo, x1, mask, mu_o, tt, z, spks, drop=(p, seed))
    loss_o.backward()
    loss_eval, _ = mo.compute_loss(sd, x1, mask, mu, tt, z, spks)

    dev = torch.device("cuda")
    plan = Plan("fp32", c_in, use_graph=False)
    named = [(k[len("estimator."):], v.to(dev).contiguous()) for k, v in sd.items()]
    plan.set_train_dropout(p, seed)
    out = plan.compute_loss(named, x1, mask, mu, z, tt, spks, 1e-4, backward=True, want_dmu=True)
    again = plan.compute_loss(named, x1, mask, mu, z, tt, spks, 1e-4, backward=True, want_dmu=True)
    plan.set_train_dropout(0.0, 0)
    ev = plan.compute_loss(named, x1, mask, mu, z, tt, spks, 1e-4, backward=False)
    torch.cuda.synchronize()
    assert abs(float(out["loss"]) - float(loss_o)) <= 1e-3 * max(1.0, abs(float(loss_o)))
    assert abs(float(loss_o) - float(loss_eval)) > 1e-4                        # the masks did something
    assert abs(float(ev["loss"]) - float(loss_eval)) <= 1e-3 * max(1.0, abs(float(loss_eval)))   # p = 0 is eval mode again
    # same seed, same masks (the training kernels reduce with atomics, so two runs agree to rounding, not bit for bit)
    assert abs(float(out["loss"]) - float(again["loss"])) <= 1e-6 * abs(float(out["loss"]))
    worst = max(rel_l2(out["grads"][k[len("estimator."):]].cpu(), v.grad) for k, v in sdo.items())
    assert worst <= GRAD_REL_L2, worst
    assert rel_l2(out["dmu"].cpu(), mu_o.grad) <= GRAD_REL_L2


def test_module_training_mode_applies_dropout_and_eval_does_not():
    cfm = mb.CFM(160, 80, CFM_PARAMS, DEC, precision="fp32")
    cfm.load_state_dict(weights.estimator_state_dict(160, seed=7), strict=True)
    cfm = cfm.cuda()
    x1, mask, mu, z, tt, _ = (None if v is None else v.cuda() for v in _case(2, 24, [24, 13], 160, 11))
    assert cfm.estimator.dropout == 0.05
    losses = {}
    for mode in ("eval", "train_a", "train_a2", "train_b"):
        cfm.train(mode != "eval")
        cfm.estimator.dropout_seed = {"train_a": 1, "train_a2": 1, "train_b": 2}.get(mode)
        loss, _ = cfm.compute_loss(x1, mask, mu, t=tt, z=z)
        losses[mode] = float(loss)
    close = lambda a, b: abs(a - b) <= 1e-6 * abs(a)      # noqa: E731  (atomics: equal to rounding)
    assert close(losses["train_a"], losses["train_a2"])
    assert not close(losses["train_a"], losses["train_b"]) and not close(losses["train_b"], losses["eval"])
    cfm.estimator.dropout_seed = None        # seeds now come from torch's generator
    cfm.train()
    torch.manual_seed(5); l1 = float(cfm.compute_loss(x1, mask, mu, t=tt, z=z)[0])
    torch.manual_seed(5); l2 = float(cfm.compute_loss(x1, mask, mu, t=tt, z=z)[0])
    assert close(l1, l2)
