"""CPU-side checks (-m "not gpu"): the C-ABI library builds/loads and exports everything include/matcha_b200.h declares,
host logic (Euler schedule, state_dict layouts, configuration errors), and the no-fallback rule."""
import ctypes
import os
import re
import types

import pytest
import torch

import matcha_b200 as mb
from matcha_b200 import _lib
from oracle import matcha_oracle as mo
from oracle import weights

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEC = dict(channels=[256, 256], dropout=0.05, attention_head_dim=64, n_blocks=1, num_mid_blocks=2, num_heads=2,
           act_fn="snakebeta")
CFM_PARAMS = types.SimpleNamespace(solver="euler", sigma_min=1e-4)


@pytest.fixture(scope="module")
def lib():
    from matcha_b200.build import build
    build(verbose=False)
    return _lib.load()


def test_library_exports_every_declared_symbol(lib):
    header = open(os.path.join(ROOT, "include", "matcha_b200.h")).read()
    declared = sorted(set(re.findall(r"\b(mb200_[a-z0-9_]+)\s*\(", header)))
    assert len(declared) >= 15
    assert sorted(_lib.EXPORTS) == declared
    for name in declared:
        assert hasattr(lib, name), name
    assert b"sm_100a" in lib.mb200_version()


def test_library_is_sm100a_tensor_core_code():
    """The shipped kernels must be real Blackwell code: tcgen05 MMA, TMEM loads and TMA in the SASS."""
    import shutil
    import subprocess
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    sass = subprocess.run(["cuobjdump", "-sass", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in sass or "SM100" in sass.upper()
    for mnemonic in ("UTCHMMA", "LDTM", "UTMALDG"):
        assert mnemonic in sass, mnemonic


@pytest.mark.parametrize("n", [1, 2, 3, 4, 7, 10, 16, 50, 51, 100, 333])
def test_euler_schedule_matches_oracle(lib, n):
    ts, dts = _lib.euler_schedule(n)
    ots, odts = mo.euler_schedule(n)
    assert ts == ots and dts == odts  # bit-exact fp32: same linspace, same accumulation order


def test_errors_are_reported_not_swallowed(lib):
    assert lib.mb200_euler_schedule(0, None, None) != 0
    assert b"bad arguments" in lib.mb200_last_error()
    cfg = _lib.Cfg(7, 160, 0, 0)
    h = ctypes.c_void_p()
    assert lib.mb200_plan_create(ctypes.byref(cfg), ctypes.byref(h)) != 0
    assert lib.mb200_last_error()


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the behaviour on a machine without a GPU")
def test_no_cpu_fallback():
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _lib.Plan("fp16", 160)
    cfm = mb.CFM(160, 80, CFM_PARAMS, DEC)
    x = torch.zeros(1, 80, 4)
    with pytest.raises(RuntimeError, match="CUDA only"):
        cfm.estimator(x, torch.ones(1, 1, 4), x, torch.tensor(0.5))
    gen = mb.Generator(mb.AttrDict(mb.HIFIGAN_V1))
    with pytest.raises(RuntimeError, match="CUDA only"):
        gen(torch.zeros(1, 80, 4))


@pytest.mark.parametrize("c_in,n_spks", [(160, 1), (224, 109)])
def test_cfm_state_dict_layout(c_in, n_spks):
    cfm = mb.CFM(160, 80, CFM_PARAMS, DEC, n_spks=n_spks, spk_emb_dim=64)
    sd = weights.estimator_state_dict(c_in, seed=3)
    assert set(sd) == set(cfm.state_dict()) and len(sd) == 180
    cfm.load_state_dict(sd, strict=True)
    n_params = sum(p.numel() for p in cfm.parameters())
    assert n_params == (11_008_848 if c_in == 160 else 11_139_920)


def test_generator_state_dict_layouts():
    gen = mb.Generator(mb.AttrDict(mb.HIFIGAN_V1))
    assert sum(p.numel() for p in gen.parameters()) == 13_936_130
    sd = weights.generator_state_dict(5, weight_norm=True)
    assert set(sd) == set(gen.state_dict()) and len(sd) == 234
    gen.load_state_dict(sd, strict=True)
    gen.remove_weight_norm()
    assert len(gen.state_dict()) == 156
    gen.load_state_dict(weights.generator_state_dict(5, weight_norm=False), strict=True)


def test_unsupported_configurations_raise_at_construction():
    with pytest.raises(NotImplementedError):
        mb.CFM(160, 80, CFM_PARAMS, dict(DEC, channels=[128, 128]))
    with pytest.raises(NotImplementedError):
        mb.CFM(160, 80, CFM_PARAMS, dict(DEC, act_fn="snake"))
    with pytest.raises(NotImplementedError):
        mb.CFM(160, 80, CFM_PARAMS, dict(DEC, down_block_type="conformer"))
    with pytest.raises(ValueError):  # decoder.py:341
        mb.CFM(160, 80, CFM_PARAMS, dict(DEC, down_block_type="nope"))
    with pytest.raises(NotImplementedError):
        mb.Generator(mb.AttrDict(dict(mb.HIFIGAN_V1, resblock="2")))


@pytest.mark.reference
def test_key_order_and_shapes_match_live_reference():
    from oracle import ref_shims
    CFM, Generator, AttrDict, v1, _ = ref_shims.load_reference()
    ref = CFM(160, 80, CFM_PARAMS, DEC, n_spks=109, spk_emb_dim=64)
    ours = mb.CFM(160, 80, CFM_PARAMS, DEC, n_spks=109, spk_emb_dim=64)
    assert list(ref.state_dict()) == list(ours.state_dict())
    assert all(ref.state_dict()[k].shape == ours.state_dict()[k].shape for k in ref.state_dict())
    ours.load_state_dict(ref.state_dict(), strict=True)
    rg, og = Generator(AttrDict(v1)), mb.Generator(AttrDict(v1))
    assert list(rg.state_dict()) == list(og.state_dict())
    og.load_state_dict(rg.state_dict(), strict=True)
    rg.remove_weight_norm()
    og.remove_weight_norm()
    assert list(rg.state_dict()) == list(og.state_dict())
