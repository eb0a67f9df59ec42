"""The header-only C++ face (include/dau_conv_b200_layer.hpp) of the C ABI: a host subclass written
against the reference's BaseDAUConvLayer<float> contract (base_dau_conv_layer.hpp:177-264) must
compile with a plain host compiler, and LayerSetUp -> Reshape -> Forward_gpu -> Backward_gpu must
give the oracle's numbers.  Tolerance as in test_parity_gpu.py (1e-4 rel / 1e-5 abs)."""
import os
import struct
import subprocess

import numpy as np
import pytest

from oracle import dau_oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
CPP = os.path.join(HERE, "cpp")
EXE = os.path.join(CPP, "layer_adapter_main")


def _build():
    subprocess.check_call(["make", "-C", CPP], stdout=subprocess.DEVNULL)
    return EXE


def test_adapter_header_compiles_without_cuda_headers(tmp_path):
    """The header needs only dau_conv_b200.h: compile a subclass with g++ and no CUDA include path."""
    src = tmp_path / "t.cpp"
    src.write_text(
        '#include "dau_conv_b200_layer.hpp"\n'
        "struct L : dau_b200::BaseDAUConvLayer {\n"
        "  float* param_w() override { return nullptr; } float* param_mu1() override { return nullptr; }\n"
        "  float* param_mu2() override { return nullptr; } float* param_sigma() override { return nullptr; }\n"
        "  float* param_bias() override { return nullptr; } float* param_w_grad() override { return nullptr; }\n"
        "  float* param_mu1_grad() override { return nullptr; } float* param_mu2_grad() override { return nullptr; }\n"
        "  float* param_sigma_grad() override { return nullptr; } float* param_bias_grad() override { return nullptr; }\n"
        "  void* allocate_workspace_mem(size_t) override { return nullptr; } void deallocate_workspace_mem() override {}\n"
        "};\n"
        "int main() { L l; dau_b200::DAUConvSettings s; s.num_output = 4; s.number_units = {2, 2};\n"
        "  try { l.LayerSetUp(s, {1, 2, 3}); } catch (const dau_b200::DAUException& e) { return e.status() == DAU_ERR_INVALID_ARGUMENT ? 0 : 1; }\n"
        "  return 1; }\n")
    lib = os.path.join(os.path.dirname(HERE), "dau-convnet_b200", "dau_conv")
    exe = tmp_path / "t"
    subprocess.check_call(["g++", "-std=c++14", "-Wall", "-Werror", "-I", os.path.join(os.path.dirname(HERE), "include"),
                           str(src), "-o", str(exe), "-L", lib, "-ldau_conv_b200", "-Wl,-rpath," + lib])
    # a malformed bottom shape is refused on the host, before any CUDA call (base_dau_conv_layer.cpp:33-35)
    assert subprocess.call([str(exe)]) == 0


def test_adapter_driver_builds():
    assert os.path.exists(_build())


CASES = [
    # N, S, F, Gx, Gy, H, W, ks, pad, edge, bias, learn_sigma, mu_lr
    (2, 3, 4, 2, 1, 12, 20, 9, 4, 0, 0, 1, 1.0),
    (3, 5, 6, 2, 2, 16, 16, 9, 4, 1, 1, 1, 500.0),
    (2, 4, 3, 1, 2, 32, 40, 17, 8, 0, 0, 0, 10.0),
    (2, 2, 3, 2, 1, 24, 24, 9, 2, 0, 1, 1, 1.0),      # pad < (ks-1)/2: output 20x20
]


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=lambda c: "N%dS%dF%dG%dx%d_%dx%d_k%dp%d" % c[:9])
def test_adapter_matches_oracle(case, tmp_path):
    from test_parity_gpu import close
    N, S, F, Gx, Gy, H, W, ks, pad, edge, bias_term, learn_sigma, mu_lr = case
    G = Gx * Gy
    crop = (ks - 1) // 2 - pad
    Ho, Wo = H - 2 * crop, W - 2 * crop
    rng = np.random.default_rng(4242 + N * 7 + S)
    x = rng.random((N, S, H, W)).astype(np.float32)
    err = rng.standard_normal((N, F, Ho, Wo)).astype(np.float32)
    w = (0.1 * rng.standard_normal((1, S, G, F))).astype(np.float32)
    m = ks // 2 - 0.01
    mu1 = rng.uniform(-m, m, (1, S, G, F)).astype(np.float32)
    mu2 = rng.uniform(-m, m, (1, S, G, F)).astype(np.float32)
    sigma = np.full((1, S, G, F), 0.5, np.float32)
    bias = rng.standard_normal(F).astype(np.float32)
    fin, fout = tmp_path / "in.bin", tmp_path / "out.bin"
    with open(fin, "wb") as f:
        f.write(struct.pack("12i", N, S, F, Gx, Gy, H, W, ks, pad, edge, bias_term, learn_sigma))
        f.write(struct.pack("f", mu_lr))
        for a in (x, err, w, mu1, mu2, sigma, bias):
            f.write(np.ascontiguousarray(a, np.float32).tobytes())
    r = subprocess.run([_build(), str(fin), str(fout)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    raw = open(fout, "rb").read()
    ho, wo = struct.unpack("2i", raw[:8])
    assert (ho, wo) == (Ho, Wo)
    flat = np.frombuffer(raw[8:], np.float32)
    sizes = [N * F * Ho * Wo, N * S * H * W] + [S * G * F] * 4 + [F]
    assert flat.size == sum(sizes)
    parts, o = [], 0
    for s in sizes:
        parts.append(flat[o:o + s]); o += s
    y, dx, dw, dmu1, dmu2, dsigma, dbias = parts

    y_ref = O.forward_cropped(x, w, mu1, mu2, sigma, ks, pad)
    if bias_term:
        y_ref = y_ref + bias[None, :, None, None]
    g_ref = O.backward_cropped(x, err, w, mu1, mu2, sigma, ks, pad, ignore_edge_gradients=bool(edge))
    close(y.reshape(y_ref.shape), y_ref, "top")
    close(dx.reshape(x.shape), g_ref[0], "bottom_error", scale_atol=True)
    close(dw.reshape(w.shape), g_ref[1], "dw", scale_atol=True)
    close(dmu1.reshape(w.shape), g_ref[2] * mu_lr, "dmu1", scale_atol=True)
    close(dmu2.reshape(w.shape), g_ref[3] * mu_lr, "dmu2", scale_atol=True)
    if learn_sigma:
        close(dsigma.reshape(w.shape), g_ref[4], "dsigma", scale_atol=True)
    else:   # enable_sigma_learning(false): the sigma gradient buffer is left untouched (last_k_optional)
        assert not dsigma.any()
    if bias_term:
        close(dbias, err.astype(np.float64).sum(axis=(0, 2, 3)), "dbias", scale_atol=True)
