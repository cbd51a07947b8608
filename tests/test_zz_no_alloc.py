"""`mb200_maximum_path_static`: the allocation-free variant of the monotonic alignment search (scratch in a static device
array inside the library; no counterpart in the reference).

Called through the C ABI (ctypes binding in matcha_b200/_lib.py) and compared with the existing entry point
`mb200_maximum_path` on the same inputs: the paths are 0/1 matrices from identical fp32 arithmetic, so the bar is equality
(the bar tests/test_mas.py holds the kernel to).  Two calls in a row, the second on a smaller, differently shaped batch: the
static array still holds the first call's decision bits, which must not leak into the second result.  This file is named to be
collected last (pytest -x stops at the first failure); it passes alone in a fresh process and after the
other tests."""
import numpy as np
import pytest
import torch

from matcha_b200 import _lib

pytestmark = pytest.mark.gpu


def _case(b, tx, ty, seed):
    rng = np.random.default_rng(seed)
    xl = rng.integers(1, tx + 1, size=b)
    xl[0] = tx
    yl = np.array([ty if i == 0 else int(rng.integers(xl[i], ty + 1)) for i in range(b)])
    value = (rng.standard_normal((b, tx, ty)) * 40.0 - 120.0).astype(np.float32)
    mask = np.zeros((b, tx, ty), dtype=np.float32)
    for i in range(b):
        mask[i, :xl[i], :yl[i]] = 1.0
    return torch.from_numpy(value).cuda(), torch.from_numpy(mask).cuda()


def test_static_scratch_entry_point_matches_the_allocating_one_twice_in_a_row():
    cases = [_case(8, 170, 512, 1), _case(3, 45, 97, 2)]    # a training-sized batch, then a smaller ragged one
    want = [_lib.maximum_path(v, m) for v, m in cases]      # existing entry point (torch-allocated workspace)
    outs = [torch.full_like(v, 7.0) for v, _ in cases]      # caller-owned outputs, poisoned
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    alloc_before = torch.cuda.memory_allocated()
    for (v, m), out in zip(cases, outs):                    # two calls in a row, nothing in between
        assert _lib.maximum_path_static(v, m, out) is out
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == alloc_before    # torch allocated nothing ...
    assert torch.cuda.mem_get_info()[0] == free_before      # ... and neither did the library (no cudaMalloc below the ABI)
    for out, ref in zip(outs, want):
        assert torch.equal(out, ref)
    # once more on the first batch: the array now holds the smaller call's bits
    again = torch.empty_like(cases[0][0])
    _lib.maximum_path_static(*cases[0], again)
    assert torch.equal(again, want[0])
    assert float(want[0].sum()) == float(sum(int(m[i, 0].sum()) for _, m in cases[:1] for i in range(m.shape[0])))  # one cell per frame


def test_static_scratch_entry_point_rejects_shapes_beyond_its_limit():
    b, tx, ty = 16, 667, 3200                               # 16 * 3200 * 21 * 4 = 4 300 800 bytes > 4 MiB = 4 194 304
    assert b * ty * ((tx + 31) // 32) * 4 > _lib.MAS_STATIC_SCRATCH_BYTES
    v = torch.zeros(b, tx, ty, device="cuda")
    with pytest.raises(RuntimeError, match="exceed the static limit"):
        _lib.maximum_path_static(v, torch.ones_like(v), torch.empty_like(v))
    with pytest.raises(ValueError):
        _lib.maximum_path_static(v.half(), torch.ones_like(v), torch.empty_like(v))
