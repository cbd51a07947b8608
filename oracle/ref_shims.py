"""TEST INFRASTRUCTURE ONLY -- import the UNMODIFIED reference modules from /root/reference.

Only usable in the build container (the GPU box has no /root/reference).  Used by
`oracle/make_golden.py` to generate the committed fixtures under tests/golden/ and by
the CPU tests that pin `oracle/matcha_oracle.py` against the real reference code.

The reference imports several packages that are absent here (diffusers 0.25.0, lightning,
conformer, hydra, ...).  Everything except three diffusers pieces is import-time
decoration; those three are restated from the pinned diffusers==0.25.0 release
(requirements.txt:38) as SURVEY.md Appendix C describes:
  * Attention + AttnProcessor2_0  (used at matcha/models/components/transformer.py:196-204, :266-271)
  * LoRACompatibleLinear == nn.Linear without a LoRA layer (transformer.py:48, :126)
  * get_activation("silu") == nn.SiLU (matcha/models/components/decoder.py:92)
"""
import importlib
import os
import sys
import types

import torch
import torch.nn as nn
import torch.nn.functional as F

REFERENCE_ROOT = os.environ.get("MATCHA_REFERENCE_ROOT", "/root/reference")


def reference_available():
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "matcha"))


class _ShimAttention(nn.Module):
    """diffusers 0.25.0 Attention with the default AttnProcessor2_0, self-attention only."""

    def __init__(self, query_dim, heads=8, dim_head=64, dropout=0.0, bias=False,
                 cross_attention_dim=None, upcast_attention=False, **_):
        super().__init__()
        inner = heads * dim_head
        self.heads = heads
        self.to_q = nn.Linear(query_dim, inner, bias=bias)
        self.to_k = nn.Linear(query_dim, inner, bias=bias)
        self.to_v = nn.Linear(query_dim, inner, bias=bias)
        self.to_out = nn.ModuleList([nn.Linear(inner, query_dim), nn.Dropout(dropout)])

    def forward(self, hidden_states, encoder_hidden_states=None, attention_mask=None, **_):
        b, l, _c = hidden_states.shape
        h = self.heads
        if attention_mask is not None:
            # prepare_attention_mask: (B, L) -> repeat_interleave(heads) -> view (B, H, 1, L)
            attention_mask = attention_mask.repeat_interleave(h, dim=0).view(b, h, -1, attention_mask.shape[-1])
        q = self.to_q(hidden_states).view(b, -1, h, self.to_q.out_features // h).transpose(1, 2)
        k = self.to_k(hidden_states).view(b, -1, h, self.to_q.out_features // h).transpose(1, 2)
        v = self.to_v(hidden_states).view(b, -1, h, self.to_q.out_features // h).transpose(1, 2)
        o = F.scaled_dot_product_attention(q, k, v, attn_mask=attention_mask, dropout_p=0.0, is_causal=False)
        o = o.transpose(1, 2).reshape(b, -1, self.to_q.out_features)
        return self.to_out[1](self.to_out[0](o))


def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


class _Anything:
    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Anything()

    def __getattr__(self, _):
        return _Anything()


def _get_activation(name):
    return {"silu": nn.SiLU, "swish": nn.SiLU, "mish": nn.Mish, "gelu": nn.GELU, "relu": nn.ReLU}[name]()


class _LightningModule(nn.Module):
    def save_hyperparameters(self, *a, **k):
        pass

    def log(self, *a, **k):
        pass


def install():
    """Install the sys.modules stubs and put the reference on sys.path (idempotent)."""
    if not reference_available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    if "diffusers.models.attention_processor" in sys.modules and getattr(
            sys.modules["diffusers"], "_mb200_shim", False):
        return
    d = _mod("diffusers", _mb200_shim=True)
    _mod("diffusers.models")
    _mod("diffusers.models.activations", get_activation=_get_activation)
    _mod("diffusers.models.attention", GEGLU=_Anything, GELU=_Anything, AdaLayerNorm=_Anything,
         AdaLayerNormZero=_Anything, ApproximateGELU=_Anything)
    _mod("diffusers.models.attention_processor", Attention=_ShimAttention)
    _mod("diffusers.models.lora", LoRACompatibleLinear=nn.Linear)
    _mod("diffusers.utils")
    _mod("diffusers.utils.torch_utils", maybe_allow_in_graph=lambda cls: cls)
    _mod("conformer", ConformerBlock=type("ConformerBlock", (nn.Module,), {}))
    _mod("lightning", LightningModule=_LightningModule, Callback=object, Trainer=object,
         LightningDataModule=object)
    _mod("lightning.pytorch", Callback=object, LightningDataModule=object, LightningModule=_LightningModule,
         Trainer=object)
    _mod("lightning.pytorch.utilities", grad_norm=lambda *a, **k: {},
         rank_zero_only=lambda f: f)
    _mod("lightning.pytorch.loggers", Logger=object)
    _mod("lightning.pytorch.utilities.rank_zero", rank_zero_only=lambda f: f)
    for name in ("hydra", "hydra.core", "hydra.core.hydra_config", "hydra.utils", "omegaconf", "rich",
                 "rich.prompt", "rich.syntax", "rich.tree", "rootutils", "wget", "gdown", "matplotlib",
                 "matplotlib.pyplot", "matplotlib.pylab", "librosa", "librosa.filters", "librosa.util",
                 "phonemizer", "soundfile", "piper_phonemize", "unidecode", "inflect"):
        if name not in sys.modules:
            try:
                importlib.import_module(name)
            except Exception:  # absent: stub it
                _mod(name)
    for name, attrs in (("hydra.core.hydra_config", {"HydraConfig": _Anything}),
                        ("hydra.utils", {"instantiate": _Anything()}),
                        ("omegaconf", {"DictConfig": dict, "OmegaConf": _Anything(), "open_dict": _Anything()}),
                        ("rich.prompt", {"Prompt": _Anything}),
                        ("matplotlib", {"use": lambda *a, **k: None})):
        m = sys.modules[name]
        if isinstance(m, types.ModuleType) and not getattr(m, "__file__", None):
            m.__dict__.update(attrs)
    del d
    # Cython MAS (matcha/utils/monotonic_align/core.pyx): the reference's own code compiled into oracle/_ref by
    # oracle/build_ref.py when that has been run, else a stub (synthesis never calls it).
    from oracle.build_ref import load_ref_mas
    core = load_ref_mas()
    if core is not None:
        sys.modules["matcha.utils.monotonic_align.core"] = core
    else:
        _mod("matcha.utils.monotonic_align.core", maximum_path_c=_Anything())
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)


def load_reference():
    """Returns (CFM, Generator, AttrDict, hifigan_v1_cfg, MatchaTTS-or-None) from the reference tree."""
    install()
    from matcha.models.components.flow_matching import CFM
    from matcha.hifigan.models import Generator
    from matcha.hifigan.env import AttrDict
    from matcha.hifigan.config import v1
    try:
        from matcha.models.matcha_tts import MatchaTTS
    except Exception:  # pragma: no cover - depends on which extra stubs are needed
        MatchaTTS = None
    return CFM, Generator, AttrDict, v1, MatchaTTS
