"""ctypes binding of oracle/_ref/libdau_ref_cuda.so: the REFERENCE's own CUDA path
(BaseDAUConvLayer<float>::Forward_gpu / Backward_gpu) compiled for sm_100a from the sources under
/root/reference by oracle/Makefile.refcuda, behind the plain-cudaMalloc host of oracle/ref_driver.cpp.
Test infrastructure: used by tests/test_reference_cuda.py (parity of this repo's kernels against the
reference's) and tests/ref_cuda_time.py (the reference's kernels timed on the same B200)."""
import ctypes
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "oracle", "_ref", "libdau_ref_cuda.so")


def available():
    return os.path.exists(SO)


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(SO)
        fp, i = ctypes.c_void_p, ctypes.c_int
        _lib.dau_ref_cuda_run.argtypes = [i] * 10 + [ctypes.c_float] + [fp] * 12 + [
            i, ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_size_t),
            ctypes.c_char_p, i, i]
        _lib.dau_ref_cuda_run.restype = i
    return _lib


def run(x, err, w, mu1, mu2, sigma, units, kernel_size=9, ignore_edge_gradients=True, units_ignore=0,
        mu_learning_rate_factor=1.0, iters=1, opt_flags=0):
    """x, err, w, mu1, mu2, sigma: contiguous fp32 CUDA tensors (NCHW, [1,S,G,F]).  Returns
    (y, dx, dw, dmu1, dmu2, dsigma, info) as CUDA tensors; info = fwd/bwd ms per pass and workspace bytes."""
    import torch
    N, S, H, W = x.shape
    F = w.shape[3]
    gx, gy = units
    assert gx * gy == w.shape[2]
    y = torch.empty((N, F, H, W), device=x.device, dtype=torch.float32)
    dx = torch.empty_like(x)
    grads = [torch.zeros_like(w) for _ in range(4)]
    fwd, bwd, ws = ctypes.c_float(0), ctypes.c_float(0), ctypes.c_size_t(0)
    msg = ctypes.create_string_buffer(1024)
    torch.cuda.synchronize()
    rc = lib().dau_ref_cuda_run(N, S, F, gx, gy, H, W, kernel_size, int(ignore_edge_gradients), units_ignore,
                                mu_learning_rate_factor,
                                *[t.data_ptr() for t in (x, err, w, mu1, mu2, sigma, y, dx, *grads)],
                                iters, ctypes.byref(fwd), ctypes.byref(bwd), ctypes.byref(ws), msg, 1024, opt_flags)
    if rc != 0:
        raise RuntimeError("reference CUDA path failed: " + msg.value.decode(errors="replace"))
    torch.cuda.synchronize()
    return (y, dx, *grads, dict(fwd_ms=fwd.value, bwd_ms=bwd.value, workspace_bytes=ws.value))


class ReferenceFault(RuntimeError):
    """The reference build refused the shape (exit 3) or died inside its own kernels (any other exit)."""


def run_isolated(x, err, w, mu1, mu2, sigma, units, tmp_dir, kernel_size=9, ignore_edge_gradients=True,
                 units_ignore=0, mu_learning_rate_factor=1.0, iters=1, timeout=300, opt_flags=0):
    """Same call on NumPy arrays, in a process of its own (tests/ref_cuda_worker.py): the reference build aborts
    the process on some shapes, which must not take the test session with it.  Returns NumPy arrays
    (y, dx, dw, dmu1, dmu2, dsigma, info).  opt_flags: 1 use_interpolation = false, 2 single_dim_kernel,
    4 forbid_positive_dim1."""
    import numpy as np
    tmp_dir = str(tmp_dir)
    np.savez(os.path.join(tmp_dir, "in.npz"), x=x, err=err, w=w, mu1=mu1, mu2=mu2, sigma=sigma,
             meta=np.array([units[0], units[1], kernel_size, int(ignore_edge_gradients), units_ignore, iters, opt_flags]),
             mu_lr=np.float32(mu_learning_rate_factor))
    out = os.path.join(tmp_dir, "out.npz")
    if os.path.exists(out):
        os.remove(out)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "ref_cuda_worker.py"), tmp_dir],
                       capture_output=True, text=True, timeout=timeout)
    if r.returncode != 0 or not os.path.exists(out):
        raise ReferenceFault("exit %d: %s" % (r.returncode, (r.stderr or "").strip()[-400:]))
    z = np.load(out)
    f, b, ws = z["info"]
    return (z["y"], z["dx"], z["dw"], z["dmu1"], z["dmu2"], z["dsigma"],
            dict(fwd_ms=float(f), bwd_ms=float(b), workspace_bytes=int(ws)))


def kernels_isolated(sigma, flags, tmp_dir, timeout=120):
    """The reference's DAUKernelCompute::get_kernels for one unit (w = 1, mu = 0) at `sigma`, run in a process of
    its own: array [6, Kb, Kb] = blur, d_w, d_mu1, d_mu2, d_sigma, d_error.  flags: 1 single_dimension_kernel,
    2 forbid_positive_dim1, 4 no unit normalisation, 8 square unit normalisation."""
    import numpy as np
    out = os.path.join(str(tmp_dir), "kernels_%g_%d.npy" % (sigma, flags))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "ref_cuda_worker.py"), "--kernels",
                        repr(float(sigma)), str(int(flags)), out], capture_output=True, text=True, timeout=timeout)
    if r.returncode != 0 or not os.path.exists(out):
        raise ReferenceFault("exit %d: %s" % (r.returncode, (r.stderr or "").strip()[-400:]))
    return np.load(out)
