"""Host logic of the image mosaic: packing of narrow maps for the 64-pixel tiles.
Runs without a GPU: dau_plan_mosaic / dau_mosaic_to_real are pure host functions of the C ABI."""
import ctypes as C

import numpy as np
import pytest

from dau_conv import _lib


def plan(N, H, W, gap):
    out = (C.c_int * 10)()
    _lib.check(_lib.lib.dau_plan_mosaic(N, H, W, gap, out))
    return out, dict(zip(("kx", "ky", "px", "py", "W", "H", "Wv", "Hv", "Nv", "gap"), list(out)))


def to_real(q, N, nv, vy, vx):
    n, y, x = C.c_int(-1), C.c_int(-1), C.c_int(-1)
    ok = _lib.lib.dau_mosaic_to_real(q, N, nv, vy, vx, C.byref(n), C.byref(y), C.byref(x))
    return (n.value, y.value, x.value) if ok else None


@pytest.mark.parametrize("N,H,W,gap", [(256, 13, 13, 4), (32, 13, 13, 4), (256, 27, 27, 4), (256, 28, 28, 8),
                                       (9, 13, 11, 6), (33, 6, 5, 4), (1, 8, 8, 4), (128, 64, 64, 4), (3, 65, 8, 4)])
def test_every_real_pixel_appears_exactly_once_and_gaps_separate_images(N, H, W, gap):
    q, m = plan(N, H, W, gap)
    assert m["px"] == W + gap and m["py"] == H + gap
    assert m["Wv"] == m["kx"] * m["px"] - gap and m["Hv"] == m["ky"] * m["py"] - gap
    cells = m["kx"] * m["ky"]
    assert m["Nv"] == -(-N // cells)
    assert cells == 1 or 2 * cells <= N                     # never pads the batch by more than a pair
    seen = np.zeros((N, H, W), np.int32)
    owner = np.full((m["Nv"], m["Hv"], m["Wv"]), -1, np.int64)
    for nv in range(m["Nv"]):
        for vy in range(m["Hv"]):
            for vx in range(m["Wv"]):
                r = to_real(q, N, nv, vy, vx)
                if r is not None:
                    n, y, x = r
                    assert 0 <= n < N and 0 <= y < H and 0 <= x < W
                    seen[n, y, x] += 1
                    owner[nv, vy, vx] = n
    assert (seen == 1).all()
    # two different images are never closer than `gap` pixels in either direction
    for nv in range(m["Nv"]):
        o = owner[nv]
        for d in range(1, gap + 1):
            a, b = o[:, :-d], o[:, d:]
            assert not ((a >= 0) & (b >= 0) & (a != b)).any()
            a, b = o[:-d, :], o[d:, :]
            assert not ((a >= 0) & (b >= 0) & (a != b)).any()


def test_wide_images_keep_the_identity_and_narrow_ones_are_packed():
    _, m = plan(128, 64, 64, 4)
    assert (m["kx"], m["ky"], m["Wv"], m["Hv"], m["Nv"]) == (1, 1, 64, 64, 128)
    _, m = plan(256, 13, 13, 4)                              # 4 x 2 images fill a 64 x 32 tile to 66 %
    assert (m["kx"], m["ky"], m["Wv"], m["Hv"]) == (4, 2, 64, 30)
    _, m = plan(256, 28, 28, 8)
    assert (m["kx"], m["Wv"]) == (2, 64)
    _, m = plan(2, 13, 13, 4)                                # too few images to pack
    assert (m["kx"], m["ky"]) == (1, 1)


def test_out_of_range_queries_and_bad_sizes():
    q, m = plan(16, 13, 13, 4)
    assert to_real(q, 16, 0, m["Hv"], 0) is None and to_real(q, 16, m["Nv"], 0, 0) is None
    assert to_real(q, 16, 0, 13, 0) is None                  # first gap row
    out = (C.c_int * 10)()
    assert _lib.lib.dau_plan_mosaic(0, 13, 13, 4, out) != 0
    assert b"positive" in _lib.lib.dau_last_error()


@pytest.mark.parametrize("N,H,W,gap", [(256, 13, 13, 4), (32, 13, 13, 4), (256, 27, 27, 4), (256, 28, 28, 8), (7, 27, 27, 4),
                                       (9, 13, 11, 6), (33, 6, 5, 4), (1, 8, 8, 4), (128, 64, 64, 4), (3, 65, 8, 4),
                                       (5, 3, 70, 4), (4, 1, 1, 4), (16, 16, 16, 4), (2, 130, 129, 16)])
def test_gradient_row_strips_cover_every_real_row_once_and_no_gap_row(N, H, W, gap):
    """dau_plan_strips (the gradient kernel's own item -> origin map, StripGeom in dau_kernels.cuh): H // 4 full strips
    of 4 rows + one tail strip of H % 4 rows per cell row; gap rows belong to no strip; the 64-column tiles cover the
    mosaic width; items are numbered the way the strip-major Xk layout is written (full items first)."""
    q, m = plan(N, H, W, gap)
    counts = (C.c_longlong * 2)()
    _lib.check(_lib.lib.dau_plan_strips(N, H, W, gap, counts, None, 0))
    n_full, n_tail = counts[0], counts[1]
    pairs, tiles_x = -(-m["Nv"] // 2), -(-m["Wv"] // 64)
    assert n_full == pairs * m["ky"] * (H // 4) * tiles_x
    assert n_tail == (pairs * m["ky"] * tiles_x if H % 4 else 0)
    items = (C.c_int * (4 * (n_full + n_tail)))()
    _lib.check(_lib.lib.dau_plan_strips(N, H, W, gap, counts, items, n_full + n_tail))
    it = np.array(list(items), np.int64).reshape(-1, 4)
    cover = np.zeros((pairs, m["Hv"], tiles_x), np.int32)
    for k, (pair, y0, rows, tx) in enumerate(it):
        assert 0 <= pair < pairs and 0 <= tx < tiles_x and y0 >= 0 and y0 + rows <= m["Hv"]
        assert rows == (4 if k < n_full else H % 4) and rows > 0
        cover[pair, y0:y0 + rows, tx] += 1
    real_row = np.array([(vy % m["py"]) < H for vy in range(m["Hv"])])
    assert (cover[:, real_row, :] == 1).all()               # every real row of every cell row, once per column tile
    assert (cover[:, ~real_row, :] == 0).all()               # gap rows are nobody's work
    # a strip never crosses a cell row
    for pair, y0, rows, tx in it:
        assert y0 // m["py"] == (y0 + rows - 1) // m["py"]
    with pytest.raises(_lib.DAUError):
        _lib.check(_lib.lib.dau_plan_strips(N, H, W, gap, counts, items, 0 if n_full + n_tail else -1))
