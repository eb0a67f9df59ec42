"""Integer pieces of the path held bit-exact against an integer-only host model (SURVEY.md section 8(d)
parity gate): floor(mu) -> integer displacement (reference dau_conv_forward_core.hpp:2025-2028), the
fraction -> four bilinear factors (:2155-2213), their flipped use on the error in the gradient kernel
(dau_conv_backward_core.hpp:846-894, :2078-2081), the negated offsets of the back-propagation
(base_dau_conv_layer.cu:299-315), the clamp of displacements into the kernel's bucket, and the
ignore_edge_gradients rule (dau_conv_test.py:110-136, base_dau_conv_layer.cpp:763-771).

The model below is written from those definitions in exact arithmetic: an fp32 number is a dyadic
rational, so floor and the fraction are computed with Python integers / fractions.Fraction and only the
final factors are rounded to fp32 the way a chain of fp32 multiplications rounds them.
"""
import ctypes as C
from fractions import Fraction

import numpy as np
import pytest


# ---- host model ------------------------------------------------------------------------------------
def model_edge_rule(extent):
    for patch in (64, 32, 16, 8):
        if extent >= patch:
            return extent % patch == 0
    return False


def _f32(x):
    return float(np.float32(x))


def split_offset(mu, max_off, interp):
    """(integer displacement, fraction as exact Fraction) of one fp32 offset, clamped into the
    displacement bucket [-max_off, max_off - 1] (+ fraction 1 at the upper end = the pixel max_off)."""
    q = Fraction(float(np.float32(mu)))                # exact value of the fp32 number
    fl = q.numerator // q.denominator                  # floor, integer arithmetic
    fr = q - fl                                        # exact; rounded to fp32 once, as the fp32 subtraction rounds it
    if fl < -max_off:
        fl, fr = -max_off, Fraction(0)
    if fl > max_off - 1:
        fl, fr = max_off - 1, Fraction(1 if interp else 0)
    if not interp:
        fr = Fraction(0)
    return int(fl), fr


def bilinear(fx, fy):
    """fp32 factors b00, b01, b10, b11 as fp32 products of fp32 (1 - f) and f."""
    fx32, fy32 = np.float32(float(fx)), np.float32(float(fy))
    ox, oy = np.float32(1) - fx32, np.float32(1) - fy32
    return ox * oy, fx32 * oy, ox * fy32, fx32 * fy32


def model_tables(w, mu1, mu2, max_off, interp, g_active, center=0.0):
    """dict name -> (offx, offy, weights[4]) arrays indexed [s, g, f] (before the halo is added)."""
    S, G, F = w.shape[1:]
    out = {k: (np.zeros((S, G, F), np.int64), np.zeros((S, G, F), np.int64), np.zeros((S, G, F, 4), np.float32))
           for k in ("fwd", "bwd", "grad")}
    for s in range(S):
        for g in range(G):
            for f in range(F):
                wv = np.float32(w[0, s, g, f])
                m1 = np.float32(mu1[0, s, g, f]) - np.float32(center)
                m2 = np.float32(mu2[0, s, g, f]) - np.float32(center)
                ox, fx = split_offset(m1, max_off, interp)
                oy, fy = split_offset(m2, max_off, interp)
                b = bilinear(fx, fy)
                wa = wv if g < g_active else np.float32(0)
                out["fwd"][0][s, g, f], out["fwd"][1][s, g, f] = ox, oy
                out["fwd"][2][s, g, f] = [wa * v for v in b]
                out["grad"][0][s, g, f], out["grad"][1][s, g, f] = -ox - 1, -oy - 1
                out["grad"][2][s, g, f] = [b[3], b[2], b[1], b[0]]
                bx, gx = split_offset(-m1, max_off, interp)
                by, gy = split_offset(-m2, max_off, interp)
                bb = bilinear(gx, gy)
                out["bwd"][0][s, g, f], out["bwd"][1][s, g, f] = bx, by
                # the kernel multiplies w * (1 - gx) first, then * (1 - gy) (left to right)
                gx32, gy32 = np.float32(float(gx)), np.float32(float(gy))
                out["bwd"][2][s, g, f] = [wv * (np.float32(1) - gx32) * (np.float32(1) - gy32),
                                          wv * gx32 * (np.float32(1) - gy32),
                                          wv * (np.float32(1) - gx32) * gy32, wv * gx32 * gy32]
                del bb
    return out


# ---- CPU: the edge-mask rule for every extent --------------------------------------------------------
def test_edge_mask_rule_for_every_extent():
    import dau_conv
    lib = dau_conv._lib.lib
    for n in range(1, 300):
        assert bool(lib.dau_edge_mask_rule(n)) == model_edge_rule(n), n
    # spot values of the reference's rule: multiples of the LARGEST patch that fits, not of any patch
    assert model_edge_rule(64) and model_edge_rule(128) and not model_edge_rule(96) and not model_edge_rule(48)
    assert model_edge_rule(8) and not model_edge_rule(7) and not model_edge_rule(72) and model_edge_rule(16) and model_edge_rule(32)


def test_split_offset_model_itself():
    assert split_offset(np.float32(-0.0), 4, True) == (0, Fraction(0))
    assert split_offset(np.float32(-1e-30), 4, True)[0] == -1
    assert split_offset(np.float32(3.99), 4, True)[0] == 3
    assert split_offset(np.float32(4.0), 4, True) == (3, Fraction(1))         # the pixel at +4 through tap 1
    assert split_offset(np.float32(4.0), 4, False) == (3, Fraction(0))
    assert split_offset(np.float32(-3.5), 4, True) == (-4, Fraction(1, 2))
    assert split_offset(np.float32(-4.5), 4, True) == (-4, Fraction(0))
    assert split_offset(np.float32(-9.0), 4, True) == (-4, Fraction(0))
    assert split_offset(np.float32(2) - np.float32(2.0 ** -20), 4, True)[0] == 1
    assert split_offset(np.float32(2) + np.float32(2.0 ** -20), 4, True)[0] == 2


# ---- GPU: the three unit tables, bit for bit ---------------------------------------------------------
REC = np.dtype([("w", np.float32, 4), ("ox", np.int32), ("oy", np.int32), ("pad", np.int32, 2)])


def _special_offsets(ks, rng, count):
    """Offsets that sit on every integer boundary the index math has."""
    m = ks // 2
    vals = [-0.0, 0.0, 0.5, -0.5, m - 0.01, -(m - 0.01), float(m), float(-m), m - 1.0, m + 3.5, -(m + 3.5),
            np.nextafter(np.float32(m), np.float32(0)), np.nextafter(np.float32(-m), np.float32(0)),
            1e-30, -1e-30, np.float32(1e-45), -np.float32(1e-45)]
    for k in range(-m, m + 1):
        for d in (2.0 ** -20, 2.0 ** -10):
            vals += [k + d, k - d, float(k)]
    vals = [np.float32(v) for v in vals]
    extra = rng.uniform(-(m + 1), m + 1, max(0, count - len(vals))).astype(np.float32)
    return np.concatenate([np.asarray(vals, np.float32), extra])


@pytest.mark.gpu
@pytest.mark.parametrize("ks", [3, 9, 17, 33, 65])
@pytest.mark.parametrize("interp", [True, False], ids=["interp", "floor_only"])
@pytest.mark.parametrize("centered", [True, False], ids=["centred", "corner"])
def test_unit_tables_bit_exact(ks, interp, centered):
    import torch
    import dau_conv
    from dau_conv import _lib
    dev = torch.device("cuda:0")
    rng = np.random.default_rng(ks * 7 + interp)
    S, G, F = 5, 3, 16
    m = ks // 2
    n = S * G * F
    mu1 = _special_offsets(ks, rng, n)
    mu2 = _special_offsets(ks, rng, n)
    reps = -(-n // len(mu1))
    mu1 = np.tile(mu1, reps)[:n]
    mu2 = np.roll(np.tile(mu2, reps), 7)[:n]
    rng.shuffle(mu2)
    center = 0.0 if centered else float(m)
    mu1 = (mu1 + np.float32(center)).astype(np.float32).reshape(1, S, G, F)
    mu2 = (mu2 + np.float32(center)).astype(np.float32).reshape(1, S, G, F)
    w = rng.standard_normal((1, S, G, F)).astype(np.float32)
    st = _lib.default_settings(num_output=F, number_units_x=1, number_units_y=G, kernel_size=ks, pad=m,
                               use_interpolation=int(interp), number_units_ignore=1,
                               offsets_already_centered=int(centered))
    h = dau_conv.DAULayerHandle(st, dev)
    h.set_sigma_hint(0.5)
    h.reshape(2, S, 24, 40)
    t = [torch.from_numpy(a).to(dev) for a in (w, mu1, mu2)]
    bufs = [torch.zeros(n * 32, dtype=torch.uint8, device=dev) for _ in range(3)]
    halos = (C.c_int * 4)()
    _lib.check(_lib.lib.dau_debug_unit_tables(h._h, t[0].data_ptr(), t[1].data_ptr(), t[2].data_ptr(),
                                              bufs[0].data_ptr(), bufs[1].data_ptr(), bufs[2].data_ptr(), halos,
                                              torch.cuda.current_stream(dev).cuda_stream))
    torch.cuda.synchronize()
    fwd, bwd, grad = [b.cpu().numpy().view(REC) for b in bufs]
    ref = model_tables(w, mu1, mu2, m, interp, G - 1, center)
    hx, hy, ghx, ghy = list(halos)
    assert hx >= m and hy >= m and ghx >= m and ghy >= m
    got = {"fwd": (fwd.reshape(S, F, G).transpose(0, 2, 1), hx, hy),           # [s][f][g] -> [s,g,f]
           "grad": (grad.reshape(S, F, G).transpose(0, 2, 1), ghx, ghy),
           "bwd": (bwd.reshape(F, S, G).transpose(1, 2, 0), hx, hy)}           # [f][s][g] -> [s,g,f]
    for name, (tab, ax, ay) in got.items():
        ox, oy, wt = ref[name]
        assert np.array_equal(tab["ox"].astype(np.int64) - ax, ox), name + " x displacement"
        assert np.array_equal(tab["oy"].astype(np.int64) - ay, oy), name + " y displacement"
        # every haloed offset addresses inside the tile: 0 <= off and off + 1 (second tap) <= 2 * halo
        assert tab["ox"].min() >= 0 and tab["oy"].min() >= 0
        assert tab["ox"].max() + 1 <= 2 * ax and tab["oy"].max() + 1 <= 2 * ay
        assert np.array_equal(tab["w"].view(np.uint32), wt.view(np.uint32)), name + " tap weights (bit pattern)"
        assert not tab["pad"].any()


@pytest.mark.gpu
def test_offset_hint_bucket_clamps_displacements():
    """dau_set_offset_hint(3.9) on a kernel-33 layer: displacements are clamped into the '< 4' bucket."""
    import torch
    import dau_conv
    from dau_conv import _lib
    dev = torch.device("cuda:0")
    S, G, F = 2, 2, 8
    rng = np.random.default_rng(1)
    mu1 = rng.uniform(-15, 15, (1, S, G, F)).astype(np.float32)
    mu2 = rng.uniform(-15, 15, (1, S, G, F)).astype(np.float32)
    w = rng.standard_normal((1, S, G, F)).astype(np.float32)
    st = _lib.default_settings(num_output=F, number_units_x=1, number_units_y=G, kernel_size=33, pad=16)
    h = dau_conv.DAULayerHandle(st, dev)
    h.set_sigma_hint(0.5)
    h.set_offset_hint(3.9)
    h.reshape(2, S, 20, 20)
    t = [torch.from_numpy(a).to(dev) for a in (w, mu1, mu2)]
    buf = torch.zeros(S * G * F * 32, dtype=torch.uint8, device=dev)
    halos = (C.c_int * 4)()
    _lib.check(_lib.lib.dau_debug_unit_tables(h._h, t[0].data_ptr(), t[1].data_ptr(), t[2].data_ptr(),
                                              buf.data_ptr(), None, None, halos, None))
    torch.cuda.synchronize()
    fwd = buf.cpu().numpy().view(REC).reshape(S, F, G).transpose(0, 2, 1)
    ref = model_tables(w, mu1, mu2, 4, True, G)["fwd"]
    assert np.array_equal(fwd["ox"].astype(np.int64) - halos[0], ref[0])
    assert np.array_equal(fwd["oy"].astype(np.int64) - halos[1], ref[1])
    assert ref[0].min() == -4 and ref[0].max() == 3
    assert np.array_equal(fwd["w"].view(np.uint32), ref[2].view(np.uint32))
