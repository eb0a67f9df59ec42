"""Runs ONE call of the reference's CUDA layer (oracle/_ref/libdau_ref_cuda.so) in a process of its own.
The reference build aborts the process on shapes outside what its authors tested (an exception out of a
destructor after a faulting kernel), so tests never load it in-process: tests/ref_cuda.py run_isolated()
spawns this file.  No torch here: device memory through libcudart via ctypes.
usage: python tests/ref_cuda_worker.py <dir>   (reads <dir>/in.npz, writes <dir>/out.npz)
       python tests/ref_cuda_worker.py --kernels <sigma> <flags> <out.npy>
       python tests/ref_cuda_worker.py --bench <bench.py config> <images> <iters> <out.json>   (timing only)"""
import ctypes
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(os.path.dirname(HERE), "oracle", "_ref", "libdau_ref_cuda.so")


def kernels_main(sigma, flags, out_path):
    """--kernels <sigma> <flags> <out.npy>: the reference's DAUKernelCompute filters for one unit."""
    lib = ctypes.CDLL(SO)
    lib.dau_ref_cuda_kernels.argtypes = [ctypes.c_float, ctypes.c_int, ctypes.POINTER(ctypes.c_int), ctypes.c_void_p,
                                         ctypes.c_int, ctypes.c_char_p, ctypes.c_int]
    buf = np.zeros(6 * 64 * 64, np.float32)
    kb = ctypes.c_int(0)
    msg = ctypes.create_string_buffer(1024)
    rc = lib.dau_ref_cuda_kernels(sigma, flags, ctypes.byref(kb), buf.ctypes.data_as(ctypes.c_void_p), buf.size, msg, 1024)
    if rc != 0:
        sys.stderr.write("reference DAUKernelCompute refused: " + msg.value.decode(errors="replace") + "\n")
        sys.exit(3)
    k = kb.value
    np.save(out_path, buf[:6 * k * k].reshape(6, k, k))


def bench_inputs(cfg, n_images):
    """bench.py's synthetic tensors for one of its configurations (no file round trip)."""
    sys.path.insert(0, os.path.dirname(HERE))
    import bench
    N, S, F, units, H, W, ks, m = bench.CONFIGS[cfg]
    x, e, w, mu1, mu2 = bench.make_inputs_numpy(cfg, n_images, 1234)
    return dict(x=x, err=e, w=w, mu1=mu1, mu2=mu2, sigma=np.full(w.shape, 0.5, np.float32),
                meta=np.array([units[0], units[1], ks, 1, 0, 1, 0]), mu_lr=np.float32(1.0))


def main():
    if sys.argv[1] == "--kernels":
        return kernels_main(float(sys.argv[2]), int(sys.argv[3]), sys.argv[4])
    bench_out = None
    if sys.argv[1] == "--bench":
        z = bench_inputs(sys.argv[2], int(sys.argv[3]))
        z["meta"][5] = int(sys.argv[4])
        bench_out = sys.argv[5]
        d = None
    else:
        d = sys.argv[1]
        z = np.load(os.path.join(d, "in.npz"))
    rt = ctypes.CDLL("libcudart.so.12")
    rt.cudaMalloc.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_size_t]
    rt.cudaMemcpy.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
    rt.cudaMemset.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t]

    def ck(rc, what):
        if rc != 0:
            raise RuntimeError("%s failed with CUDA error %d" % (what, rc))

    def dev(nbytes):
        p = ctypes.c_void_p()
        ck(rt.cudaMalloc(ctypes.byref(p), max(nbytes, 4)), "cudaMalloc")
        ck(rt.cudaMemset(p, 0, max(nbytes, 4)), "cudaMemset")
        return p

    def up(a):
        a = np.ascontiguousarray(a, np.float32)
        p = dev(a.nbytes)
        ck(rt.cudaMemcpy(p, a.ctypes.data_as(ctypes.c_void_p), a.nbytes, 1), "cudaMemcpy H2D")
        return p

    def down(p, shape):
        a = np.empty(shape, np.float32)
        ck(rt.cudaMemcpy(a.ctypes.data_as(ctypes.c_void_p), p, a.nbytes, 2), "cudaMemcpy D2H")
        return a

    x, err, w = z["x"], z["err"], z["w"]
    N, S, H, W = x.shape
    F = w.shape[3]
    gx, gy, ks, edge, ignore, iters, opt_flags = [int(v) for v in z["meta"]]
    px, pe, pw, pm1, pm2, ps = [up(z[k]) for k in ("x", "err", "w", "mu1", "mu2", "sigma")]
    py, pdx = dev(4 * N * F * H * W), dev(x.nbytes)
    pg = [dev(w.nbytes) for _ in range(4)]
    lib = ctypes.CDLL(SO)
    fp, i = ctypes.c_void_p, ctypes.c_int
    lib.dau_ref_cuda_run.argtypes = [i] * 10 + [ctypes.c_float] + [fp] * 12 + [
        i, ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_float), ctypes.POINTER(ctypes.c_size_t),
        ctypes.c_char_p, i, i]
    fwd, bwd, ws = ctypes.c_float(0), ctypes.c_float(0), ctypes.c_size_t(0)
    msg = ctypes.create_string_buffer(1024)
    rc = lib.dau_ref_cuda_run(N, S, F, gx, gy, H, W, ks, edge, ignore, float(z["mu_lr"]),
                              px, pe, pw, pm1, pm2, ps, py, pdx, *pg, iters,
                              ctypes.byref(fwd), ctypes.byref(bwd), ctypes.byref(ws), msg, 1024, opt_flags)
    if rc != 0:
        sys.stderr.write("reference CUDA path refused: " + msg.value.decode(errors="replace") + "\n")
        sys.exit(3)
    if bench_out is not None:
        import json
        with open(bench_out, "w") as f:
            json.dump({"fwd_ms": fwd.value, "bwd_ms": bwd.value, "workspace_bytes": int(ws.value), "iters": iters}, f)
        return
    np.savez(os.path.join(d, "out.npz"), y=down(py, (N, F, H, W)), dx=down(pdx, x.shape),
             dw=down(pg[0], w.shape), dmu1=down(pg[1], w.shape), dmu2=down(pg[2], w.shape),
             dsigma=down(pg[3], w.shape), info=np.array([fwd.value, bwd.value, float(ws.value)]))


if __name__ == "__main__":
    main()
