"""`MatchaTTS` drop-in for synthesis (SURVEY.md §8(b) boundary #1; reference matcha/models/matcha_tts.py:24-148).

What runs where:
  * the hot path -- `self.decoder(mu_y, y_mask, n_timesteps, temperature, spks)` (matcha_tts.py:135) -- is the CUDA path of
    this package (`matcha_b200.cfm.CFM` -> libmatcha_b200.so);
  * the text encoder and duration predictor are PyTorch written for this package (not yet CUDA
    kernels); length regulation is CUDA (`mb200_length_regulate`, §8(f) row 2).  The encoder keeps the reference's parameter names
    (`encoder.emb`, `encoder.prenet.*`, `encoder.encoder.attn_layers.N.conv_q ...`, `encoder.proj_m`, `encoder.proj_w.*`,
    `spk_emb`, `decoder.estimator.*`) so `load_state_dict(strict=True)` of a reference checkpoint's `state_dict` works;
  * length regulation does not form the dense (B, Lx, T) alignment matmul (matcha_tts.py:129-131): `mu_y` is a gather
    of `mu_x` by the token index of each frame, which is what the 0/1 alignment computes; `attn` is still returned for
    API compatibility.

Training: `forward` returns the reference's three losses with the monotonic alignment search (§8(f) row 4) and the
decoder loss + backward (§8(a) row a15) in CUDA; the optimiser / Lightning loop stays outside.
"""
from __future__ import annotations

import datetime as dt
import math
import random
from types import SimpleNamespace

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import monotonic_align
from .cfm import CFM


def _get(cfg, name, default=None):
    """Config access for dicts, namespaces and OmegaConf-like objects."""
    if isinstance(cfg, dict):
        return cfg.get(name, default)
    return getattr(cfg, name, default)


def sequence_mask(length, max_length=None):
    """(B,) lengths -> (B, max_length) bool prefix masks (reference utils/model.py:7-11)."""
    if max_length is None:
        max_length = int(length.max())
    return torch.arange(max_length, dtype=length.dtype, device=length.device)[None, :] < length[:, None]


def fix_len_compatibility(length, num_downsamplings_in_unet=2):
    """Round up to a multiple of 2**n (reference utils/model.py:14-20)."""
    f = 2 ** num_downsamplings_in_unet
    return int(math.ceil(float(length) / f) * f)


def generate_path(duration, mask):
    """0/1 monotonic alignment (B, Lx, T) from integer-valued durations (reference utils/model.py:29-41)."""
    b, t_x, t_y = mask.shape
    cum = torch.cumsum(duration, 1)
    path = sequence_mask(cum.reshape(b * t_x), t_y).to(mask.dtype).view(b, t_x, t_y)
    path = path - F.pad(path, (0, 0, 1, 0))[:, :-1]
    return path * mask


class ChannelNorm(nn.Module):
    """LayerNorm over the channel axis of (B, C, T), eps 1e-4, parameters `gamma`/`beta` (text_encoder.py:16-33)."""

    def __init__(self, channels, eps=1e-4):
        super().__init__()
        self.eps = eps
        self.gamma = nn.Parameter(torch.ones(channels))
        self.beta = nn.Parameter(torch.zeros(channels))

    def forward(self, x):
        return F.layer_norm(x.transpose(1, 2), (x.shape[1],), self.gamma, self.beta, self.eps).transpose(1, 2)


class _Prenet(nn.Module):  # text_encoder.py:36-67; ReLU -> Dropout(0.5) after every norm (:50, built at :356)
    def __init__(self, ch, k=5, n_layers=3, p_dropout=0.5):
        super().__init__()
        self.p_dropout = p_dropout
        self.conv_layers = nn.ModuleList(nn.Conv1d(ch, ch, k, padding=k // 2) for _ in range(n_layers))
        self.norm_layers = nn.ModuleList(ChannelNorm(ch) for _ in range(n_layers))
        self.proj = nn.Conv1d(ch, ch, 1)
        nn.init.zeros_(self.proj.weight)
        nn.init.zeros_(self.proj.bias)

    def forward(self, x, x_mask):
        h = x
        for conv, norm in zip(self.conv_layers, self.norm_layers):
            h = F.dropout(torch.relu(norm(conv(h * x_mask))), self.p_dropout, self.training)
        return (x + self.proj(h)) * x_mask


class _DurationPredictor(nn.Module):  # text_encoder.py:70-93; dropout after each norm (:77, :88, :92)
    def __init__(self, cin, filt, k, p_dropout=0.0):
        super().__init__()
        self.p_dropout = p_dropout
        self.conv_1 = nn.Conv1d(cin, filt, k, padding=k // 2)
        self.norm_1 = ChannelNorm(filt)
        self.conv_2 = nn.Conv1d(filt, filt, k, padding=k // 2)
        self.norm_2 = ChannelNorm(filt)
        self.proj = nn.Conv1d(filt, 1, 1)

    def forward(self, x, x_mask):
        x = F.dropout(self.norm_1(torch.relu(self.conv_1(x * x_mask))), self.p_dropout, self.training)
        x = F.dropout(self.norm_2(torch.relu(self.conv_2(x * x_mask))), self.p_dropout, self.training)
        return self.proj(x * x_mask) * x_mask


def _rope(x, d, base=10000.0):
    """Rotary embedding on the first `d` features of (B, H, T, Dh), positions along T (text_encoder.py:96-166)."""
    t = x.shape[2]
    theta = 1.0 / (base ** (torch.arange(0, d, 2, device=x.device).float() / d))
    ang = torch.arange(t, device=x.device).float()[:, None] * theta[None, :]
    ang = torch.cat([ang, ang], dim=1)                       # (T, d)
    cos, sin = ang.cos()[None, None], ang.sin()[None, None]
    xr, xp = x[..., :d], x[..., d:]
    half = d // 2
    rot = torch.cat([-xr[..., half:], xr[..., :half]], dim=-1)
    return torch.cat([xr * cos + rot * sin, xp], dim=-1)


class _SelfAttention(nn.Module):  # text_encoder.py:169-237 (no proximal bias in the shipped configs)
    def __init__(self, ch, n_heads, p_dropout=0.0):
        super().__init__()
        assert ch % n_heads == 0
        self.p_dropout = p_dropout     # on the attention probabilities (:207, :224)
        self.n_heads, self.k_channels = n_heads, ch // n_heads
        self.conv_q, self.conv_k, self.conv_v, self.conv_o = (nn.Conv1d(ch, ch, 1) for _ in range(4))
        for c in (self.conv_q, self.conv_k, self.conv_v):
            nn.init.xavier_uniform_(c.weight)
        self.rope_dims = int(self.k_channels * 0.5)

    def forward(self, x, attn_mask):
        b, c, t = x.shape
        split = lambda y: y.view(b, self.n_heads, self.k_channels, t).transpose(2, 3)   # (B, H, T, Dh)
        q, k, v = split(self.conv_q(x)), split(self.conv_k(x)), split(self.conv_v(x))
        q, k = _rope(q, self.rope_dims), _rope(k, self.rope_dims)
        scores = torch.matmul(q, k.transpose(-2, -1)) / math.sqrt(self.k_channels)
        scores = scores.masked_fill(attn_mask == 0, -1e4)    # replaces, does not add (text_encoder.py:222)
        o = torch.matmul(F.dropout(torch.softmax(scores, dim=-1), self.p_dropout, self.training), v)
        return self.conv_o(o.transpose(2, 3).reshape(b, c, t))


class _FFN(nn.Module):  # text_encoder.py:240-257; dropout after the ReLU (:243, :253)
    def __init__(self, ch, filt, k, p_dropout=0.0):
        super().__init__()
        self.p_dropout = p_dropout
        self.conv_1 = nn.Conv1d(ch, filt, k, padding=k // 2)
        self.conv_2 = nn.Conv1d(filt, ch, k, padding=k // 2)

    def forward(self, x, x_mask):
        h = F.dropout(torch.relu(self.conv_1(x * x_mask)), self.p_dropout, self.training)
        return self.conv_2(h * x_mask) * x_mask


class _Encoder(nn.Module):  # text_encoder.py:260-310; dropout on both sub-layer outputs (:266, :319-322)
    def __init__(self, ch, filt, n_heads, n_layers, k, p_dropout=0.0):
        super().__init__()
        self.p_dropout = p_dropout
        self.attn_layers = nn.ModuleList(_SelfAttention(ch, n_heads, p_dropout) for _ in range(n_layers))
        self.norm_layers_1 = nn.ModuleList(ChannelNorm(ch) for _ in range(n_layers))
        self.ffn_layers = nn.ModuleList(_FFN(ch, filt, k, p_dropout) for _ in range(n_layers))
        self.norm_layers_2 = nn.ModuleList(ChannelNorm(ch) for _ in range(n_layers))

    def forward(self, x, x_mask):
        attn_mask = x_mask.unsqueeze(2) * x_mask.unsqueeze(-1)
        for attn, n1, ffn, n2 in zip(self.attn_layers, self.norm_layers_1, self.ffn_layers, self.norm_layers_2):
            x = x * x_mask
            x = n1(x + F.dropout(attn(x, attn_mask), self.p_dropout, self.training))
            x = n2(x + F.dropout(ffn(x, x_mask), self.p_dropout, self.training))
        return x * x_mask


class TextEncoder(nn.Module):
    """Phoneme ids -> (mu_x (B,n_feats,Lx), logw (B,1,Lx), x_mask (B,1,Lx)); reference text_encoder.py:313-410."""

    def __init__(self, encoder_type, encoder_params, duration_predictor_params, n_vocab, n_spks=1, spk_emb_dim=128):
        super().__init__()
        p = encoder_params
        self.n_vocab, self.n_spks, self.spk_emb_dim = n_vocab, n_spks, spk_emb_dim
        self.n_feats, self.n_channels = _get(p, "n_feats"), _get(p, "n_channels")
        self.emb = nn.Embedding(n_vocab, self.n_channels)
        nn.init.normal_(self.emb.weight, 0.0, self.n_channels ** -0.5)
        self.prenet = _Prenet(self.n_channels) if _get(p, "prenet", True) else None
        ch = self.n_channels + (spk_emb_dim if n_spks > 1 else 0)
        self.encoder = _Encoder(ch, _get(p, "filter_channels"), _get(p, "n_heads"), _get(p, "n_layers"),
                                _get(p, "kernel_size"), float(_get(p, "p_dropout", 0.0) or 0.0))
        self.proj_m = nn.Conv1d(ch, self.n_feats, 1)
        d = duration_predictor_params
        self.proj_w = _DurationPredictor(ch, _get(d, "filter_channels_dp"), _get(d, "kernel_size"),
                                         float(_get(d, "p_dropout", 0.0) or 0.0))

    def _cuda_plan(self):
        """The fp32 plan holding this module's packed weights for the CUDA kernels (csrc/text_encoder.cu); re-packed when
        a parameter changed."""
        from . import _lib
        version = tuple(p._version for p in self.parameters())
        if getattr(self, "_plan_version", None) != version or getattr(self, "_plan", None) is None:
            dev = next(self.parameters()).device
            plan = _lib.Plan("fp32", 160, device=dev)
            plan.load_text_encoder(self.state_dict(), n_heads=self.encoder.attn_layers[0].n_heads)
            self._plan, self._plan_version = plan, version
        return self._plan

    def forward(self, x, x_lengths, spks=None):
        """Inference on a GPU (no gradients wanted, module in eval mode) runs the hand-written kernels (SURVEY.md section 8(f)
        row 3).  Training keeps the torch formulation below: it needs autograd and dropout, and north_star leaves the
        encoder's training to PyTorch."""
        if x.is_cuda and not self.training and not torch.is_grad_enabled() and self.prenet is not None:
            return self._cuda_plan().text_encoder_forward(x, x_lengths, spks)
        h = self.emb(x).transpose(1, 2) * math.sqrt(self.n_channels)
        x_mask = sequence_mask(x_lengths, h.shape[2]).unsqueeze(1).to(h.dtype)
        if self.prenet is not None:
            h = self.prenet(h, x_mask)
        if self.n_spks > 1:
            h = torch.cat([h, spks.unsqueeze(-1).expand(-1, -1, h.shape[-1])], dim=1)
        h = self.encoder(h, x_mask)
        return self.proj_m(h) * x_mask, self.proj_w(h.detach(), x_mask), x_mask


def length_regulate(mu_x, w_ceil, x_mask, y_lengths, t_pad):
    """mu_y (B, n_feats, t_pad), attn (B, Lx, t_pad) and y_mask (B, 1, t_pad) from per-token frame counts.

    Same result as `attn = generate_path(...)`, `mu_y = attn^T @ mu_x` (matcha_tts.py:126-131) without the dense matmul:
    frame j of utterance b copies the first token whose duration prefix sum exceeds j, and frames at or beyond
    y_lengths[b] are zero.  Runs the `mb200_length_regulate` kernels (csrc/length_regulate.cu, prefix sum + gather,
    bit-exact against the reference-generated fixtures).  CUDA tensors only: there is no CPU path in this package (the
    torch restatement that the CPU tests pin to the same fixtures lives in oracle/matcha_oracle.py).
    """
    from . import _lib
    mu_y, attn, y_mask = _lib.length_regulate(mu_x, w_ceil.squeeze(1), x_mask.squeeze(1), y_lengths, t_pad)
    return mu_y.to(mu_x.dtype), attn.to(x_mask.dtype), y_mask.unsqueeze(1).to(x_mask.dtype)


class MatchaTTS(nn.Module):
    """Synthesis-side drop-in for the reference `MatchaTTS` (constructor and `synthesise` contract: matcha_tts.py:24-148)."""

    def __init__(self, n_vocab, n_spks, spk_emb_dim, n_feats, encoder, decoder, cfm, data_statistics, out_size=None,
                 optimizer=None, scheduler=None, prior_loss=True):
        super().__init__()
        self.n_vocab, self.n_spks, self.spk_emb_dim, self.n_feats = n_vocab, n_spks, spk_emb_dim, n_feats
        self.out_size, self.prior_loss = out_size, prior_loss
        if n_spks > 1:
            self.spk_emb = nn.Embedding(n_spks, spk_emb_dim)
        enc_params = _get(encoder, "encoder_params")
        self.encoder = TextEncoder(_get(encoder, "encoder_type"), enc_params, _get(encoder, "duration_predictor_params"),
                                   n_vocab, n_spks, spk_emb_dim)
        nf = _get(enc_params, "n_feats")
        self.decoder = CFM(in_channels=2 * nf, out_channel=nf, cfm_params=cfm, decoder_params=decoder, n_spks=n_spks,
                           spk_emb_dim=spk_emb_dim)
        self.update_data_statistics(data_statistics)
        # Padded mel length = y_max rounded up to a multiple of this.  4 is the reference's (utils/model.py:14-20) and the
        # default: padded frames are attended to and enter the GroupNorm statistics, so the padded
        # length is part of the result.  A serving caller may set 32 or 64 to bound the number of distinct shapes (one
        # CUDA graph each, at most kMaxCachedPrograms kept): results then equal the reference's run with that padding.
        self.t_pad_multiple = 4

    def update_data_statistics(self, data_statistics):   # baselightningmodule.py:20-28
        if data_statistics is None:
            data_statistics = {"mel_mean": 0.0, "mel_std": 1.0}
        self.register_buffer("mel_mean", torch.tensor(float(_get(data_statistics, "mel_mean"))))
        self.register_buffer("mel_std", torch.tensor(float(_get(data_statistics, "mel_std"))))

    @torch.inference_mode()
    def synthesise(self, x, x_lengths, n_timesteps, temperature=1.0, spks=None, length_scale=1.0):
        t0 = dt.datetime.now()
        if self.n_spks > 1:
            spks = self.spk_emb(spks.long())
        mu_x, logw, x_mask = self.encoder(x, x_lengths, spks)
        w_ceil = torch.ceil(torch.exp(logw) * x_mask) * length_scale
        y_lengths = torch.clamp_min(torch.sum(w_ceil, [1, 2]), 1).long()
        y_max_length = int(y_lengths.max())
        t_pad = fix_len_compatibility(y_max_length)
        if self.t_pad_multiple > 4:
            t_pad = -(-t_pad // self.t_pad_multiple) * self.t_pad_multiple
        mu_y, attn, y_mask = length_regulate(mu_x, w_ceil, x_mask, y_lengths, t_pad)
        encoder_outputs = mu_y[:, :, :y_max_length]
        decoder_outputs = self.decoder(mu_y, y_mask, n_timesteps, temperature, spks)[:, :, :y_max_length]
        seconds = (dt.datetime.now() - t0).total_seconds()
        return {
            "encoder_outputs": encoder_outputs,
            "decoder_outputs": decoder_outputs,
            "attn": attn.unsqueeze(1)[:, :, :y_max_length],   # the reference slices the Lx axis here (matcha_tts.py:144)
            "mel": decoder_outputs * self.mel_std + self.mel_mean,
            "mel_lengths": y_lengths,
            "rtf": seconds * 22050 / (decoder_outputs.shape[-1] * 256),
        }

    def forward(self, x, x_lengths, y, y_lengths, spks=None, out_size=None, cond=None):
        """The three training losses `(dur_loss, prior_loss, diff_loss)` (reference matcha_tts.py:150-245).

        Encoder and the scalar losses are torch; the alignment between tokens and mel frames is found on the GPU
        (`monotonic_align.maximum_path`, csrc/mas.cu - the reference round-trips through the host here) and the decoder
        loss with its whole backward pass runs in CUDA (`CFM.compute_loss`, csrc/train.cu).  Gradients reach the encoder
        through `mu_y` (prior and flow-matching loss) and `logw` (duration loss), as upstream.
        """
        if self.n_spks > 1:
            spks = self.spk_emb(spks)
        mu_x, logw, x_mask = self.encoder(x, x_lengths, spks)
        y_mask = sequence_mask(y_lengths, y.shape[-1]).unsqueeze(1).to(x_mask)
        with torch.no_grad():
            pair_mask = (x_mask.unsqueeze(-1) * y_mask.unsqueeze(2)).squeeze(1)            # (B, Lx, T)
            attn = monotonic_align.maximum_path(_log_prior(mu_x, y, self.n_feats), pair_mask).detach()
        # duration head against the log of the aligned frame counts (utils/model.py:44-46)
        frames_per_token = attn.sum(-1).unsqueeze(1)
        dur_loss = torch.sum((logw - torch.log(1e-8 + frames_per_token) * x_mask) ** 2) / torch.sum(x_lengths)
        if out_size is not None:
            attn, y, y_mask = _random_segments(attn, y, y_lengths, out_size, y_mask.dtype)
        mu_y = torch.matmul(attn.transpose(1, 2), mu_x.transpose(1, 2)).transpose(1, 2)
        diff_loss, _ = self.decoder.compute_loss(x1=y, mask=y_mask, mu=mu_y, spks=spks, cond=cond)
        prior_loss = 0
        if self.prior_loss:
            nll = 0.5 * ((y - mu_y) ** 2 + math.log(2 * math.pi)) * y_mask
            prior_loss = nll.sum() / (y_mask.sum() * self.n_feats)
        return dur_loss, prior_loss, diff_loss


def _log_prior(mu_x, y, n_feats):
    """log N(y_j; mu_i, I) for every (token i, frame j): (B, Lx, T).  The reference (matcha_tts.py:185-190) forms the
    |y_j|^2 term with a (B, Lx, 80) x (B, 80, T) matmul against a constant; here it is a channel sum broadcast over tokens."""
    y_sq = (y * y).sum(1, keepdim=True)                      # (B, 1, T)
    mu_sq = (mu_x * mu_x).sum(1).unsqueeze(-1)               # (B, Lx, 1)
    cross = torch.matmul(mu_x.transpose(1, 2), y)            # (B, Lx, T)
    return -0.5 * (y_sq - 2.0 * cross + mu_sq) - 0.5 * math.log(2 * math.pi) * n_feats


def _random_segments(attn, y, y_lengths, out_size, mask_dtype):
    """One random window of `out_size` frames per utterance, zero-padded where the utterance is shorter (the Grad-TTS
    memory hack the reference keeps, matcha_tts.py:203-225).  Returns the cut alignment, mel and frame mask."""
    lengths = [int(v) for v in y_lengths.tolist()]
    kept = [min(n, out_size) for n in lengths]
    starts = [random.randrange(n - out_size) if n > out_size else 0 for n in lengths]
    attn_seg = attn.new_zeros(attn.shape[0], attn.shape[1], out_size)
    y_seg = y.new_zeros(y.shape[0], y.shape[1], out_size)
    for i, (s0, n) in enumerate(zip(starts, kept)):
        y_seg[i, :, :n] = y[i, :, s0:s0 + n]
        attn_seg[i, :, :n] = attn[i, :, s0:s0 + n]
    seg_mask = sequence_mask(torch.tensor(kept, device=y.device), out_size).unsqueeze(1).to(mask_dtype)
    return attn_seg, y_seg, seg_mask


def default_config(n_spks=1):
    """The shipped hyper-parameters (configs/model/matcha.yaml, encoder/default.yaml, decoder/default.yaml, cfm/default.yaml)."""
    ns = SimpleNamespace
    enc = ns(encoder_type="RoPE Encoder",
             encoder_params=ns(n_feats=80, n_channels=192, filter_channels=768, filter_channels_dp=256, n_heads=2,
                               n_layers=6, kernel_size=3, p_dropout=0.1, spk_emb_dim=64, n_spks=n_spks, prenet=True),
             duration_predictor_params=ns(filter_channels_dp=256, kernel_size=3, p_dropout=0.1))
    dec = dict(channels=(256, 256), dropout=0.05, attention_head_dim=64, n_blocks=1, num_mid_blocks=2, num_heads=2,
               act_fn="snakebeta")
    cfm = ns(name="CFM", solver="euler", sigma_min=1e-4)
    return dict(n_vocab=178, n_spks=n_spks, spk_emb_dim=64, n_feats=80, encoder=enc, decoder=dec, cfm=cfm,
                data_statistics=dict(mel_mean=-5.536622, mel_std=2.116101), out_size=None)
