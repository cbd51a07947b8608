"""HiFi-GAN v1 `Generator` drop-in (reference: matcha/hifigan/models.py:149-206, config matcha/hifigan/config.py:1-28).

Same constructor (`Generator(h)` with the attribute-dict `h`), same parameter names in both layouts (weight_g/weight_v
under weight-norm, plain `weight` after `remove_weight_norm()`), same `forward(mel (B,80,T)) -> (B,1,256*T)`.
The forward pass is the CUDA launch program of csrc/vocoder.cu; the modules here only own the parameters.
"""
import torch
import torch.nn as nn
from torch.nn.utils import remove_weight_norm, weight_norm

from . import _lib

LRELU_SLOPE = 0.1

V1 = dict(resblock="1", upsample_rates=[8, 8, 2, 2], upsample_kernel_sizes=[16, 16, 4, 4], upsample_initial_channel=512,
          resblock_kernel_sizes=[3, 7, 11], resblock_dilation_sizes=[[1, 3, 5], [1, 3, 5], [1, 3, 5]])


class AttrDict(dict):
    """dict with attribute access (matcha/hifigan/env.py:7-10)."""

    def __init__(self, *a, **k):
        super().__init__(*a, **k)
        self.__dict__ = self


def _init(m, mean=0.0, std=0.01):  # hifigan/xutils.py:25-28
    if isinstance(m, (nn.Conv1d, nn.ConvTranspose1d)):
        m.weight.data.normal_(mean, std)


class _ResBlockParams(nn.Module):
    def __init__(self, channels, kernel_size, dilations):
        super().__init__()
        self.convs1 = nn.ModuleList([weight_norm(nn.Conv1d(channels, channels, kernel_size, dilation=d,
                                                           padding=(kernel_size * d - d) // 2)) for d in dilations])
        self.convs2 = nn.ModuleList([weight_norm(nn.Conv1d(channels, channels, kernel_size,
                                                           padding=(kernel_size - 1) // 2)) for _ in dilations])
        self.convs1.apply(_init)
        self.convs2.apply(_init)

    def remove_weight_norm(self):
        for layer in list(self.convs1) + list(self.convs2):
            remove_weight_norm(layer)


class Generator(nn.Module):
    def __init__(self, h, precision="fp16"):
        super().__init__()
        self.h = h
        get = (lambda k: h[k]) if isinstance(h, dict) else (lambda k: getattr(h, k))
        cfg = {k: get(k) for k in V1}
        same = (str(cfg["resblock"]) == "1" and cfg["upsample_initial_channel"] == 512
                and all(list(cfg[k]) == V1[k] for k in ("upsample_rates", "upsample_kernel_sizes", "resblock_kernel_sizes"))
                and [list(d) for d in cfg["resblock_dilation_sizes"]] == V1["resblock_dilation_sizes"])
        if not same:
            raise NotImplementedError("matcha_b200 has kernels for the HiFi-GAN v1 generator only (hifigan/config.py v1)")
        self.precision = precision
        self.num_kernels = 3
        self.num_upsamples = 4
        self.conv_pre = weight_norm(nn.Conv1d(80, 512, 7, 1, padding=3))
        self.ups = nn.ModuleList()
        ch = 512
        for u, k in zip(V1["upsample_rates"], V1["upsample_kernel_sizes"]):
            self.ups.append(weight_norm(nn.ConvTranspose1d(ch, ch // 2, k, u, padding=(k - u) // 2)))
            ch //= 2
        self.resblocks = nn.ModuleList()
        for i in range(4):
            ch = 512 // (2 ** (i + 1))
            for k, d in zip(V1["resblock_kernel_sizes"], V1["resblock_dilation_sizes"]):
                self.resblocks.append(_ResBlockParams(ch, k, d))
        self.conv_post = weight_norm(nn.Conv1d(ch, 1, 7, 1, padding=3))
        self.ups.apply(_init)
        self.conv_post.apply(_init)
        self._plans = {}
        self._packed_version = None

    def remove_weight_norm(self):
        for layer in self.ups:
            remove_weight_norm(layer)
        for blk in self.resblocks:
            blk.remove_weight_norm()
        remove_weight_norm(self.conv_pre)
        remove_weight_norm(self.conv_post)
        self._packed_version = None

    def _version(self):
        return tuple((k, p._version) for k, p in self.named_parameters())

    def plan(self, precision=None):
        precision = precision or self.precision
        version = self._version()
        if self._packed_version != version:
            self._plans = {}
            self._packed_version = version
        if precision not in self._plans:
            dev = next(self.parameters()).device
            if dev.type != "cuda":
                raise RuntimeError("matcha_b200.Generator runs on CUDA only: move the module to a B200 (no CPU fallback)")
            p = _lib.Plan(precision, 160, device=dev)
            # parameters only: under weight-norm the derived `.weight` attribute is not in the state_dict
            p.load_generator(dict(self.state_dict()))
            self._plans[precision] = p
        return self._plans[precision]

    @torch.no_grad()
    def forward(self, x):
        return self.plan().hifigan_forward(x)


class Denoiser(nn.Module):
    """Drop-in for matcha/hifigan/denoiser.py:10-64 ("removes model bias from audio"): same constructor
    (`Denoiser(vocoder, filter_length=1024, n_overlap=4, win_length=1024, mode="zeros")`), same `bias_spec` buffer
    (1, 513, 1) and `forward(audio, strength=0.0005)`.  The bias audio comes from the vocoder passed in (the CUDA
    Generator above, or anything callable on a (1, 80, 88) mel); STFT, spectral subtraction and iSTFT are one CUDA
    kernel (csrc/denoiser.cu).  Only the shipped STFT geometry (1024 / hop 256 / 1024) has a kernel."""

    def __init__(self, vocoder, filter_length=1024, n_overlap=4, win_length=1024, mode="zeros"):
        super().__init__()
        if (filter_length, n_overlap, win_length) != (1024, 4, 1024):
            raise NotImplementedError("matcha_b200.Denoiser has a kernel for filter_length = win_length = 1024, n_overlap = 4 only")
        self.filter_length, self.hop_length, self.win_length = filter_length, filter_length // n_overlap, win_length
        p = next(vocoder.parameters())
        dtype, device = p.dtype, p.device
        if device.type != "cuda":
            raise RuntimeError("matcha_b200.Denoiser runs on CUDA only (no CPU fallback)")
        self.device = device
        if mode == "zeros":
            mel_input = torch.zeros((1, 80, 88), dtype=dtype, device=device)
        elif mode == "normal":
            mel_input = torch.randn((1, 80, 88), dtype=dtype, device=device)
        else:
            raise Exception(f"Mode {mode} if not supported")   # message as in denoiser.py:24
        self._plan = _lib.Plan("fp32", 160, device=device, use_graph=False)
        with torch.no_grad():
            bias_audio = vocoder(mel_input).float().squeeze(0)                     # (1, 22528)
            bias = self._plan.denoiser_bias(bias_audio.reshape(-1))
        self.register_buffer("bias_spec", bias.reshape(1, 513, 1))

    @torch.inference_mode()
    def forward(self, audio, strength=0.0005):
        a = audio.reshape(1, -1) if audio.dim() == 1 else audio    # (L,) comes back as (1, L'): bias_spec broadcasts a batch axis
        if a.dim() != 2:
            raise ValueError("audio must be (L,) or (B, L)")
        return self._plan.denoise(a, self.bias_spec.reshape(-1), strength)
