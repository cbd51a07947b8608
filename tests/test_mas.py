"""Monotonic alignment search: oracle/mas.py against the reference's own Cython code (compiled into
oracle/_ref by oracle/build_ref.py, and the fixtures it generated), and the CUDA kernel against both."""
import os

import numpy as np
import pytest
import torch

from oracle import mas as omas
from oracle.build_ref import load_ref_mas

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _fx(name):
    fx = np.load(os.path.join(GOLDEN, name + ".npz"))
    value, xl, yl = fx["value"], fx["x_lengths"], fx["y_lengths"]
    mask = np.zeros_like(value)
    for i, (nx, ny) in enumerate(zip(xl, yl)):
        mask[i, :nx, :ny] = 1.0
    return value, mask, fx["path"].astype(np.int32)


def _case(b, tx, ty, seed, square=False):
    rng = np.random.default_rng(seed)
    xl = rng.integers(1, tx + 1, size=b)
    xl[0] = tx
    yl = np.array([ty if i == 0 else int(rng.integers(xl[i], ty + 1)) for i in range(b)])
    if square:
        yl[-1] = xl[-1]                      # t_x == t_y: the path is forced onto the diagonal
    value = (rng.standard_normal((b, tx, ty)) * 40.0 - 120.0).astype(np.float32)
    mask = np.zeros((b, tx, ty), dtype=np.float32)
    for i in range(b):
        mask[i, :xl[i], :yl[i]] = 1.0
    return value, mask, xl, yl


def _ref_path(core, value, mask):
    v = (value * mask).astype(np.float32)
    path = np.zeros_like(v).astype(np.int32)
    core.maximum_path_c(path, v, mask.sum(1)[:, 0].astype(np.int32), mask.sum(2)[:, 0].astype(np.int32))
    return path


@pytest.mark.parametrize("name", ["mas_ragged", "mas_long"])
def test_oracle_matches_reference_golden(name):
    value, mask, path = _fx(name)
    assert np.array_equal(omas.maximum_path(value, mask), path)


def test_oracle_matches_compiled_reference():
    core = load_ref_mas()
    if core is None:
        pytest.skip("oracle/_ref has not been built (python -m oracle.build_ref; needs /root/reference)")
    for seed, (b, tx, ty) in enumerate([(3, 30, 90), (2, 64, 64), (4, 7, 200), (1, 1, 1)]):
        value, mask, _, _ = _case(b, tx, ty, seed, square=True)
        assert np.array_equal(omas.maximum_path(value, mask), _ref_path(core, value, mask))


def _check_properties(path, xl, yl):
    """Size-independent properties of a monotonic alignment: one token per frame, every token used when t_x <= t_y,
    token index non-decreasing and advancing by at most one per frame, from token 0 to token t_x - 1."""
    for i in range(path.shape[0]):
        p = path[i, :xl[i], :yl[i]]
        assert path[i].sum() == yl[i] and p.sum() == yl[i]
        assert (p.sum(0) == 1).all()
        tok = p.argmax(0)
        d = np.diff(tok)
        assert tok[0] == 0 and tok[-1] == xl[i] - 1 and ((d == 0) | (d == 1)).all()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["mas_ragged", "mas_long"])
def test_cuda_matches_reference_golden(name):
    from matcha_b200 import monotonic_align
    value, mask, path = _fx(name)
    out = monotonic_align.maximum_path(torch.from_numpy(value).cuda(), torch.from_numpy(mask).cuda())
    assert out.is_cuda and out.dtype == torch.float32
    assert np.array_equal(out.cpu().numpy().astype(np.int32), path)


@pytest.mark.gpu
@pytest.mark.parametrize("b,tx,ty", [(5, 40, 130), (3, 64, 64), (2, 1, 9), (4, 33, 257), (2, 300, 1000)])
def test_cuda_matches_oracle_and_compiled_reference(b, tx, ty):
    from matcha_b200 import monotonic_align
    value, mask, xl, yl = _case(b, tx, ty, seed=tx + ty, square=True)
    out = monotonic_align.maximum_path(torch.from_numpy(value).cuda(), torch.from_numpy(mask).cuda()).cpu().numpy().astype(np.int32)
    assert np.array_equal(out, omas.maximum_path(value, mask))
    core = load_ref_mas()
    if core is not None:
        assert np.array_equal(out, _ref_path(core, value, mask))
    _check_properties(out, xl, yl)


@pytest.mark.gpu
def test_cuda_full_size_properties_and_dtype():
    """A training batch at BASELINE cfg5's scale (B = 32, up to 2001 tokens x 3004 frames is cfg4's synthesis length):
    properties only, plus the half-precision dtype round trip of the wrapper."""
    from matcha_b200 import monotonic_align
    value, mask, xl, yl = _case(6, 2001, 3004, seed=5)
    out = monotonic_align.maximum_path(torch.from_numpy(value).cuda().half(), torch.from_numpy(mask).cuda().half())
    assert out.dtype == torch.float16
    _check_properties(out.float().cpu().numpy().astype(np.int32), xl, yl)
