"""Drop-in replacements for the reference's CFM / Decoder modules, backed by the sm_100a kernels.

Mirrors (paths under the reference tree):
  * matcha/models/components/flow_matching.py:11-132  BASECFM / CFM  (forward, solve_euler, compute_loss)
  * matcha/models/components/decoder.py:201-443       Decoder        (constructor arguments, parameter tree, forward)
The modules below hold parameters under exactly the reference's state_dict keys so that
`load_state_dict(reference.state_dict(), strict=True)` works; none of them computes anything in PyTorch -- forward
calls go through the C ABI (matcha_b200/_lib.py) and fail loudly without the CUDA library or a GPU.
"""
import math

import torch
import torch.nn as nn

from . import _lib


class _Params(nn.Module):
    """A parameter container: never called, only named."""

    def forward(self, *a, **k):  # pragma: no cover
        raise RuntimeError("parameter container of the CUDA path; not callable")


def _conv(cin, cout, k, transposed=False):
    return (nn.ConvTranspose1d if transposed else nn.Conv1d)(cin, cout, k)


def _block1d(cin, cout):
    m = _Params()
    m.block = nn.ModuleList([_conv(cin, cout, 3), nn.GroupNorm(8, cout), nn.Identity()])  # conv, norm, (Mish)
    return m


def _resnet(cin, cout, temb):
    m = _Params()
    m.mlp = nn.ModuleList([nn.Identity(), nn.Linear(temb, cout)])  # (Mish), Linear -> key "mlp.1"
    m.block1 = _block1d(cin, cout)
    m.block2 = _block1d(cout, cout)
    m.res_conv = _conv(cin, cout, 1)
    return m


class _SnakeBeta(_Params):
    def __init__(self, cin, cout):
        super().__init__()
        self.alpha = nn.Parameter(torch.zeros(cout))
        self.beta = nn.Parameter(torch.zeros(cout))
        self.proj = nn.Linear(cin, cout)


def _transformer(dim, heads, head_dim):
    inner = heads * head_dim
    m = _Params()
    m.norm1 = nn.LayerNorm(dim)
    attn = _Params()
    attn.to_q = nn.Linear(dim, inner, bias=False)
    attn.to_k = nn.Linear(dim, inner, bias=False)
    attn.to_v = nn.Linear(dim, inner, bias=False)
    attn.to_out = nn.ModuleList([nn.Linear(inner, dim), nn.Identity()])
    m.attn1 = attn
    m.norm3 = nn.LayerNorm(dim)
    ff = _Params()
    ff.net = nn.ModuleList([_SnakeBeta(dim, 4 * dim), nn.Identity(), nn.Linear(4 * dim, dim)])
    m.ff = ff
    return m


def _wrap(name, module):
    m = _Params()
    setattr(m, name, module)
    return m


class Decoder(nn.Module):
    """U-Net velocity estimator.  Same constructor keywords as decoder.py:201-215; only the shipped configuration
    (channels (256,256), 1 transformer block per level, 2 mid blocks, 2 heads x 64, snakebeta) has kernels."""

    def __init__(self, in_channels, out_channels, channels=(256, 256), dropout=0.05, attention_head_dim=64,
                 n_blocks=1, num_mid_blocks=2, num_heads=4, act_fn="snake", down_block_type="transformer",
                 mid_block_type="transformer", up_block_type="transformer", precision="fp16"):
        super().__init__()
        channels = tuple(channels)
        for bt in (down_block_type, mid_block_type, up_block_type):
            if bt not in ("transformer", "conformer"):
                raise ValueError(f"Unknown block type {bt}")
        unsupported = []
        if channels != (256, 256):
            unsupported.append(f"channels={channels}")
        if (n_blocks, num_mid_blocks, num_heads, attention_head_dim) != (1, 2, 2, 64):
            unsupported.append("n_blocks/num_mid_blocks/num_heads/attention_head_dim != 1/2/2/64")
        if act_fn != "snakebeta":
            unsupported.append(f"act_fn={act_fn}")
        if (down_block_type, mid_block_type, up_block_type) != ("transformer",) * 3:
            unsupported.append("conformer blocks")
        if out_channels != 80 or in_channels not in (160, 224):
            unsupported.append(f"in/out channels {in_channels}/{out_channels}")
        if unsupported:
            raise NotImplementedError("matcha_b200 has kernels for the shipped decoder configuration only "
                                      "(configs/model/decoder/default.yaml); unsupported: " + ", ".join(unsupported))
        self.in_channels, self.out_channels, self.precision = in_channels, out_channels, precision
        self.dropout = float(dropout)     # transformer-block dropout, active in .train() on the CUDA training path
        temb = channels[0] * 4
        tm = _Params()
        tm.linear_1 = nn.Linear(in_channels, temb)
        tm.linear_2 = nn.Linear(temb, temb)
        self.time_mlp = tm
        c = channels[0]
        self.down_blocks = nn.ModuleList([
            nn.ModuleList([_resnet(in_channels, c, temb), nn.ModuleList([_transformer(c, num_heads, attention_head_dim)]),
                           _wrap("conv", _conv(c, c, 3))]),
            nn.ModuleList([_resnet(c, c, temb), nn.ModuleList([_transformer(c, num_heads, attention_head_dim)]),
                           _conv(c, c, 3)])])
        self.mid_blocks = nn.ModuleList([
            nn.ModuleList([_resnet(c, c, temb), nn.ModuleList([_transformer(c, num_heads, attention_head_dim)])])
            for _ in range(num_mid_blocks)])
        self.up_blocks = nn.ModuleList([
            nn.ModuleList([_resnet(2 * c, c, temb), nn.ModuleList([_transformer(c, num_heads, attention_head_dim)]),
                           _wrap("conv", _conv(c, c, 4, transposed=True))]),
            nn.ModuleList([_resnet(2 * c, c, temb), nn.ModuleList([_transformer(c, num_heads, attention_head_dim)]),
                           _conv(c, c, 3)])])
        self.final_block = _block1d(c, c)
        self.final_proj = _conv(c, out_channels, 1)
        self._init_like_reference()
        self._plans = {}
        self._packed_version = None
        self._register_load_state_dict_pre_hook(lambda *a, **k: self._invalidate())

    def _init_like_reference(self):
        # decoder.py:345-361: kaiming-normal Conv1d / Linear with zero bias, unit GroupNorm; ConvTranspose1d and
        # LayerNorm keep torch defaults.
        for m in self.modules():
            if isinstance(m, (nn.Conv1d, nn.Linear)):
                nn.init.kaiming_normal_(m.weight, nonlinearity="relu")
                if m.bias is not None:
                    nn.init.constant_(m.bias, 0)
            elif isinstance(m, nn.GroupNorm):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)

    def _invalidate(self):
        self._packed_version = None

    def _version(self):
        return tuple(p._version for p in self.parameters()) + (id(self),)

    def plan(self, precision=None):
        """The CUDA plan holding this module's packed weights (re-packed when parameters changed)."""
        precision = precision or self.precision
        version = self._version()
        if self._packed_version != version:
            self._plans = {}
            self._packed_version = version
        if precision not in self._plans:
            dev = next(self.parameters()).device
            if dev.type != "cuda":
                raise RuntimeError("matcha_b200.Decoder runs on CUDA only: move the module to a B200 (no CPU fallback)")
            p = _lib.Plan(precision, self.in_channels, device=dev)
            p.load_estimator({k: v for k, v in self.state_dict().items()}, prefix="")
            self._plans[precision] = p
        return self._plans[precision]

    def train_plan(self):
        """The plan the training path runs on.  compute_loss reads the module's parameters in place (no packed copy), so
        one plan serves every optimiser step."""
        if getattr(self, "_train_plan", None) is None:
            dev = next(self.parameters()).device
            if dev.type != "cuda":
                raise RuntimeError("matcha_b200.Decoder runs on CUDA only: move the module to a B200 (no CPU fallback)")
            self._train_plan = _lib.Plan("fp32", self.in_channels, device=dev)
        return self._train_plan

    @torch.no_grad()
    def forward(self, x, mask, mu, t, spks=None, cond=None):
        """decoder.py:363: x, mu (B,80,T); mask (B,1,T); t () or (B,); spks (B,64) or None -> (B,80,T)."""
        return self.plan().estimator_forward(x, mask, mu, t, spks)


class BASECFM(nn.Module):
    """flow_matching.py:11-118 with the Euler loop, estimator and velocity update on the CUDA path."""

    def __init__(self, n_feats, cfm_params, n_spks=1, spk_emb_dim=128):
        super().__init__()
        self.n_feats = n_feats
        self.n_spks = n_spks
        self.spk_emb_dim = spk_emb_dim
        self.solver = cfm_params.solver
        self.sigma_min = getattr(cfm_params, "sigma_min", 1e-4)
        self.estimator = None

    @torch.inference_mode()
    def forward(self, mu, mask, n_timesteps, temperature=1.0, spks=None, cond=None):
        z = torch.randn_like(mu) * temperature
        # the reference builds t_span = linspace(0, 1, n + 1) here (flow_matching.py:52); the CUDA path evaluates exactly
        # that grid inside its graph, so no tensor is built, moved or compared (no device -> host sync on this call)
        return self.estimator.plan().cfm_decode(mu, mask, z, int(n_timesteps), spks)

    def solve_euler(self, x, t_span, mu, mask, spks, cond):
        n = len(t_span) - 1
        # Is this the uniform grid every caller of the reference uses?  Decided on the host, with a tolerance (CUDA and
        # CPU linspace may differ by an ulp), against the schedule the kernels use.  A t_span that lives on the device costs
        # one small read-back here; BASECFM.forward never comes this way.
        span = t_span.detach().to("cpu", torch.float32)
        uniform = n >= 1 and bool(torch.allclose(span, torch.linspace(0, 1, n + 1), rtol=0.0, atol=2e-7))
        if uniform:  # whole loop in one CUDA graph
            return self.estimator.plan().cfm_decode(mu, mask, x, n, spks)
        # arbitrary grid: same recurrence, one estimator launch program per step (slow path, said once)
        if not getattr(self, "_warned_slow_grid", False):
            import warnings
            warnings.warn("matcha_b200: solve_euler got a non-uniform t_span; running one estimator call per step instead of "
                          "the captured Euler graph", stacklevel=2)
            self._warned_slow_grid = True
        t, dt = t_span[0], t_span[1] - t_span[0]
        for step in range(1, len(t_span)):
            x = x + dt * self.estimator(x, mask, mu, t, spks, cond)
            t = t + dt
            if step < len(t_span) - 1:
                dt = t_span[step + 1] - t
        return x

    def compute_loss(self, x1, mask, mu, spks=None, cond=None, t=None, z=None):
        """flow_matching.py:87-118.  Same draws as the reference (`torch.rand([b,1,1])` then `torch.randn_like(x1)`, unless
        t / z are injected for parity tests); y, u, the estimator forward with per-sample t, the loss AND the estimator's
        whole backward pass run in CUDA (csrc/train.cu).  The returned loss is an autograd node: `loss.backward()`
        delivers the CUDA-computed gradients to the estimator's parameters, to `mu` (the encoder side) and to `spks`."""
        b = mu.shape[0]
        if t is None:
            t = torch.rand([b, 1, 1], device=mu.device, dtype=mu.dtype)
        if z is None:
            z = torch.randn_like(x1)
        names = [k for k, _ in self.estimator.named_parameters()]
        params = [p for _, p in self.estimator.named_parameters()]
        loss, y = _ComputeLoss.apply(self.estimator, names, float(self.sigma_min), x1, mask, mu, spks, t, z, *params)
        return loss, y


class _ComputeLoss(torch.autograd.Function):
    """Bridges the CUDA forward+backward of compute_loss into autograd: gradients are produced by the kernels during
    forward() (one fused launch sequence) and handed out, scaled by the incoming gradient, in backward()."""

    @staticmethod
    def forward(ctx, estimator, names, sigma_min, x1, mask, mu, spks, t, z, *params):
        nig = ctx.needs_input_grad          # (estimator, names, sigma_min, x1, mask, mu, spks, t, z, *params)
        need = bool(nig[5] or nig[6] or any(nig[9:]))
        plan = estimator.train_plan()
        # decoder.py:206: the transformer blocks' dropout is active in training mode.  The seed comes from torch's generator,
        # so torch.manual_seed makes a training run reproducible; `estimator.dropout_seed` (tests) overrides it.
        if estimator.training and estimator.dropout > 0:
            seed = getattr(estimator, "dropout_seed", None)
            if seed is None:
                seed = int(torch.randint(0, 2 ** 62, (1,)).item())
            plan.set_train_dropout(estimator.dropout, seed)
        else:
            plan.set_train_dropout(0.0, 0)
        out = plan.compute_loss(zip(names, (p.detach() for p in params)), x1.detach(), mask, mu.detach(), z, t,
                                None if spks is None else spks.detach(), sigma_min, backward=need,
                                want_dmu=bool(nig[5]), want_dspks=bool(nig[6]))
        ctx.saved = (out["grads"], out["dmu"], out["dspks"], names) if need else None
        ctx.mark_non_differentiable(out["y"])
        return out["loss"], out["y"]

    @staticmethod
    def backward(ctx, g_loss, _g_y):
        if ctx.saved is None:
            raise RuntimeError("compute_loss was run without gradients enabled")
        grads, dmu, dspks, names = ctx.saved
        scale = lambda g: None if g is None else g * g_loss
        # inputs: estimator, names, sigma_min, x1, mask, mu, spks, t, z, *params
        return (None, None, None, None, None, scale(dmu), scale(dspks), None, None, *[scale(grads[k]) for k in names])


class CFM(BASECFM):
    def __init__(self, in_channels, out_channel, cfm_params, decoder_params, n_spks=1, spk_emb_dim=64,
                 precision="fp16"):
        super().__init__(n_feats=in_channels, cfm_params=cfm_params, n_spks=n_spks, spk_emb_dim=spk_emb_dim)
        in_channels = in_channels + (spk_emb_dim if n_spks > 1 else 0)
        self.estimator = Decoder(in_channels=in_channels, out_channels=out_channel, precision=precision,
                                 **dict(decoder_params))
