"""TEST INFRASTRUCTURE ONLY - CPU restatement of the reference's monotonic alignment search.

Follows matcha/utils/monotonic_align/core.pyx:11-47 (`maximum_path_each`, `maximum_path_c`) and the wrapper
matcha/utils/monotonic_align/__init__.py:7-22.  The column recurrence is vectorised over tokens with numpy fp32
arithmetic (column y reads column y-1 only, so this is the same fp32 operation per cell); back-tracking is the
reference's loop.  Pinned against the reference's own Cython code compiled into oracle/_ref (oracle/build_ref.py) and
against the fixtures that build generated (tests/golden/mas_*.npz): tests/test_mas.py.  Only tests may import this.
"""
import numpy as np

MAX_NEG = np.float32(-1e9)


def maximum_path_each(value, t_x, t_y):
    """value (Tx, Ty) fp32, modified in place like the reference; returns the int32 0/1 path."""
    path = np.zeros(value.shape, dtype=np.int32)
    for y in range(t_y):
        lo, hi = max(0, t_x + y - t_y), min(t_x, y + 1)
        if hi <= lo:
            continue
        x = np.arange(lo, hi)
        if y == 0:
            v_cur = np.full(hi - lo, MAX_NEG, dtype=np.float32)          # only x == 0 == y is in range
            v_prev = np.zeros(hi - lo, dtype=np.float32)
        else:
            v_cur = np.where(x == y, MAX_NEG, value[x, y - 1]).astype(np.float32)
            v_prev = np.where(x == 0, MAX_NEG, value[np.maximum(x - 1, 0), y - 1]).astype(np.float32)
        value[lo:hi, y] = np.maximum(v_cur, v_prev) + value[lo:hi, y]
    index = t_x - 1
    for y in range(t_y - 1, -1, -1):
        path[index, y] = 1
        if index != 0 and (index == y or value[index, y - 1] < value[index - 1, y - 1]):
            index -= 1
    return path


def maximum_path(value, mask):
    """numpy (B, Tx, Ty) value and 0/1 mask -> int32 path, as monotonic_align/__init__.py:7-22."""
    value = (np.asarray(value, dtype=np.float32) * np.asarray(mask, dtype=np.float32)).astype(np.float32)
    mask = np.asarray(mask)
    t_xs = mask.sum(1)[:, 0].astype(np.int32)
    t_ys = mask.sum(2)[:, 0].astype(np.int32)
    return np.stack([maximum_path_each(value[b], int(t_xs[b]), int(t_ys[b])) for b in range(value.shape[0])])
