"""The DAUConv / DAUConvGrad operators on top of the C ABI, with the TF op signatures.

Mirrors (reference, plugins/tensorflow):
  DAUConv(input, weights, mu1, mu2, sigma) -> output      src/dau_conv_op.cpp:22-84, 150-324
  DAUConvGrad(grad, input, weights, mu1, mu2, sigma)
      -> (grad_input, grad_weights, grad_mu1, grad_mu2, grad_sigma)   src/dau_conv_grad_op.cpp:18-49, 115-318
  gradient registration                                   dau_conv/_dau_conv_grad_op.py:14-60

torch is used for device memory and streams only; all arithmetic happens in
libdau_conv_b200.so.  Tensors are NCHW fp32 on a CUDA device, parameters [1,S,G,F].
"""
import ctypes as C

import torch

from . import _lib

# TF op attributes and their defaults (dau_conv_op.cpp:29-48)
OP_ATTRS = dict(
    number_units_x=2, number_units_y=2, number_units_ignore=0, num_output=64, kernel_size=9, pad=4,
    stride=1, unit_normalization=True, square_unit_normalization=False, mean_iteration_step=1,
    sigma_iteration_step=1, component_border_bound=1.0, sigma_lower_bound=0.3, merge_iteration_step=0,
    merge_threshold=1, unit_testing=False, mu_learning_rate_factor=1.0, single_dim_kernel=False,
    forbid_positive_dim1=False, use_interpolation=True,
)
# beyond the TF op: settings of BaseDAUConvLayer the fused layer path uses (DAUConvSettings.bias_term,
# base_dau_conv_layer.hpp:112)
EXT_ATTRS = dict(bias_term=False)


def _settings_from_attrs(attrs):
    a = dict(OP_ATTRS)
    a.update(EXT_ATTRS)
    for k, v in attrs.items():
        if k not in a:
            raise TypeError("DAUConv: unknown attribute %r" % k)
        a[k] = v
    s = _lib.default_settings(
        num_output=int(a["num_output"]), number_units_x=int(a["number_units_x"]),
        number_units_y=int(a["number_units_y"]), number_units_ignore=int(a["number_units_ignore"]),
        kernel_size=int(a["kernel_size"]), pad=int(a["pad"]), stride=int(a["stride"]),
        unit_normalization=int(bool(a["unit_normalization"])),
        square_unit_normalization=int(bool(a["square_unit_normalization"])),
        mean_iteration_step=int(a["mean_iteration_step"]), sigma_iteration_step=int(a["sigma_iteration_step"]),
        component_border_bound=float(a["component_border_bound"]), sigma_lower_bound=float(a["sigma_lower_bound"]),
        merge_iteration_step=int(a["merge_iteration_step"]), merge_threshold=float(a["merge_threshold"]),
        ignore_edge_gradients=int(bool(a["unit_testing"])),
        mu_learning_rate_factor=float(a["mu_learning_rate_factor"]),
        single_dimension_kernel=int(bool(a["single_dim_kernel"])),
        forbid_positive_dim1=int(bool(a["forbid_positive_dim1"])),
        use_interpolation=int(bool(a["use_interpolation"])),
        bias_term=int(bool(a["bias_term"])),
        unit_bounds_guard=0,           # the TF op disables the guard (dau_conv_op.cpp:283)
        learn_sigma=1,                 # enable_sigma_learning(true) (dau_conv_grad_op.cpp:272)
    )
    return s, tuple(sorted(a.items()))


class DAULayerHandle(object):
    """One dau_handle_t (= one BaseDAUConvLayer object) with its caller-owned workspace."""

    def __init__(self, settings, device):
        self.settings = settings
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("the DAU layer runs on CUDA devices only (no CPU fallback)")
        self._h = C.c_void_p()
        idx = self.device.index if self.device.index is not None else torch.cuda.current_device()
        _lib.check(_lib.lib.dau_create(C.byref(settings), idx, C.byref(self._h)))
        self._shape = None
        self._hint = None
        self._off_hint = -1.0
        self._ws = None
        self.fwd_bytes = self.bwd_bytes = 0

    def __del__(self):
        try:
            if self._h:
                _lib.lib.dau_destroy(self._h)
                self._h = C.c_void_p()
        except Exception:
            pass

    def reshape(self, N, S, H, W):
        # the layer plans its layouts (mosaic gap) from the shape AND the blur width, i.e. sigma
        if self._shape != (N, S, H, W, self._hint, self._off_hint):
            f, b = C.c_size_t(0), C.c_size_t(0)
            _lib.check(_lib.lib.dau_reshape(self._h, N, S, H, W, C.byref(f), C.byref(b)))
            self.fwd_bytes, self.bwd_bytes = f.value, b.value
            self._shape = (N, S, H, W, self._hint, self._off_hint)
            ho, wo = C.c_int(0), C.c_int(0)
            _lib.check(_lib.lib.dau_get_output_size(self._h, C.byref(ho), C.byref(wo)))
            self.out_hw = (ho.value, wo.value)      # == (H, W) unless pad < (kernel_size-1)/2

    def workspace(self, nbytes):
        if self._ws is None or self._ws.numel() < nbytes:
            self._ws = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        return self._ws

    def set_sigma_hint(self, value):
        """Tell the layer sigma's value (it is one number per layer, base_dau_conv_layer.hpp:150):
        without it every call reads sigma[0] back from the device with a blocking copy."""
        value = float(value)
        if value != self._hint:
            _lib.check(_lib.lib.dau_set_sigma_hint(self._h, value))
            self._hint = value

    def set_offset_hint(self, max_abs_offset):
        """Upper bound of |mu| (None / negative: unknown): the layer then plans for the displacement
        bucket < 4 / 8 / 16 / 32 that holds it instead of kernel_size/2 (dau_conv_op.cpp:222-256)."""
        value = -1.0 if max_abs_offset is None else float(max_abs_offset)
        if value != self._off_hint:
            _lib.check(_lib.lib.dau_set_offset_hint(self._h, value))
            self._off_hint = value

    def enable_timing(self, enable=True):
        _lib.check(_lib.lib.dau_enable_timing(self._h, int(enable)))

    def timing(self, which):
        buf = (C.c_float * 16)()
        n = C.c_int(0)
        _lib.check(_lib.lib.dau_get_timing(self._h, which, buf, 16, C.byref(n)))
        return {_lib.lib.dau_timing_name(which, i).decode(): buf[i] for i in range(n.value)}

    def launch_count(self):
        return int(_lib.lib.dau_launch_count(self._h))

    # -- the two calls of the contract ---------------------------------------------------------
    def take_nan_flag(self):
        """True if a mu was NaN in any call since the last query (the TF op raises instead,
        dau_conv_op.cpp:258-262); blocks on the current stream."""
        found = C.c_int(0)
        _lib.check(_lib.lib.dau_take_nan_flag(self._h, C.byref(found),
                                              torch.cuda.current_stream(self.device).cuda_stream))
        return bool(found.value)

    def set_activation(self, activation):
        """dau_set_activation: DAU_ACT_NONE (0) or DAU_ACT_RELU (1), fused into forward / backward."""
        _lib.check(_lib.lib.dau_set_activation(self._h, int(activation)))

    def forward(self, x, w, mu1, mu2, sigma, bias=None, out=None, activation=_lib.ACT_NONE):
        N, S, H, W = x.shape
        self.reshape(N, S, H, W)
        self.set_activation(activation)
        F = self.settings.num_output
        if out is None:
            out = torch.empty((N, F) + self.out_hw, dtype=torch.float32, device=x.device)
        ws = self.workspace(self.fwd_bytes)
        stream = torch.cuda.current_stream(x.device).cuda_stream
        _lib.check(_lib.lib.dau_forward(
            self._h, x.data_ptr(), w.data_ptr(), mu1.data_ptr(), mu2.data_ptr(), sigma.data_ptr(),
            bias.data_ptr() if bias is not None else None, out.data_ptr(),
            ws.data_ptr(), ws.numel(), stream))
        return out

    def backward(self, x, top_error, w, mu1, mu2, sigma, flags=_lib.GRAD_ALL, grads=None, dbias=None,
                 top_data=None, activation=_lib.ACT_NONE):
        """Returns (dx, dw, dmu1, dmu2, dsigma); param grads are accumulated into `grads` if given.
        With a fused activation, `top_data` is the forward output and `top_error` the gradient
        w.r.t. that (activated) output: dau_backward_act."""
        N, S, H, W = x.shape
        self.reshape(N, S, H, W)
        self.set_activation(activation)
        if tuple(top_error.shape) != (N, self.settings.num_output) + self.out_hw:
            raise RuntimeError("DAUConvGrad: grad has shape %s, expected %s" % (
                tuple(top_error.shape), (N, self.settings.num_output) + self.out_hw))
        dx = torch.empty_like(x) if flags & _lib.GRAD_INPUT else None
        if grads is None:
            grads = [torch.zeros_like(w) for _ in range(4)]     # the TF op pre-zeroes (grad_op.cpp:202-205)
        ws = self.workspace(self.bwd_bytes)
        stream = torch.cuda.current_stream(x.device).cuda_stream
        tail = (w.data_ptr(), mu1.data_ptr(), mu2.data_ptr(),
                sigma.data_ptr(), dx.data_ptr() if dx is not None else None,
                grads[0].data_ptr(), grads[1].data_ptr(), grads[2].data_ptr(), grads[3].data_ptr(),
                dbias.data_ptr() if dbias is not None else None, flags, ws.data_ptr(), ws.numel(), stream)
        if activation != _lib.ACT_NONE:
            _lib.check(_lib.lib.dau_backward_act(self._h, x.data_ptr(), top_data.data_ptr(), top_error.data_ptr(), *tail))
        else:
            _lib.check(_lib.lib.dau_backward(self._h, x.data_ptr(), top_error.data_ptr(), *tail))
        return (dx,) + tuple(grads)


    def train_step_host(self, x_host, err_host, w, mu1, mu2, sigma, sigma_value, y_host=None, dx_host=None,
                        grads=None, chunk_images=16):
        """dau_train_step_host: forward + backward on HOST tensors (pin them), copies overlapped
        with the kernels chunk by chunk; parameter gradients accumulate into `grads` (device)."""
        N, S, H, W = x_host.shape
        for t in (x_host, err_host, y_host, dx_host):
            if t is not None and (t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous()):
                raise RuntimeError("train_step_host: activations must be contiguous float32 host tensors")
        if grads is None:
            grads = [torch.zeros_like(w) for _ in range(4)]
        stream = torch.cuda.current_stream(self.device).cuda_stream
        _lib.check(_lib.lib.dau_train_step_host(
            self._h, N, S, H, W, int(chunk_images), float(sigma_value), x_host.data_ptr(), err_host.data_ptr(),
            w.data_ptr(), mu1.data_ptr(), mu2.data_ptr(), sigma.data_ptr(),
            y_host.data_ptr() if y_host is not None else None, dx_host.data_ptr() if dx_host is not None else None,
            grads[0].data_ptr(), grads[1].data_ptr(), grads[2].data_ptr(), grads[3].data_ptr(), stream))
        return grads


class GraphedStep(object):
    """forward + backward of one handle captured ONCE in a CUDA graph and replayed: for layers whose whole
    step is shorter than the launch latency of its dozen kernels (the reference's own speed-test shape,
    plugins/tensorflow/tests/dau_conv_test.py:504-628, or BASELINE config C1).  Every call of the C ABI is
    stream-ordered and allocation-free once sigma is known (set_sigma_hint), so the capture is plain stream
    capture; the graph is keyed on the tensors it was captured with -- write new data INTO them.
    Parameter gradients are zeroed inside the graph (the op's contract: it accumulates into pre-zeroed buffers)."""

    def __init__(self, handle, x, top_error, w, mu1, mu2, sigma, flags=_lib.GRAD_ALL):
        self.h = handle
        self.x, self.e, self.w, self.mu1, self.mu2, self.sigma = x, top_error, w, mu1, mu2, sigma
        N, S, H, W = x.shape
        handle.reshape(N, S, H, W)
        self.y = torch.empty((N, handle.settings.num_output) + handle.out_hw, dtype=torch.float32, device=x.device)
        self.flat = torch.zeros((4,) + tuple(w.shape), dtype=torch.float32, device=x.device)
        self.grads = [self.flat[k] for k in range(4)]
        self.flags = flags
        handle.workspace(max(handle.fwd_bytes, handle.bwd_bytes))
        side = torch.cuda.Stream(x.device)
        side.wait_stream(torch.cuda.current_stream(x.device))
        with torch.cuda.stream(side):                    # warm-up outside the capture (lazy attribute / table setup)
            for _ in range(2):
                self._eager()
        torch.cuda.current_stream(x.device).wait_stream(side)
        torch.cuda.synchronize(x.device)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._eager()

    def _eager(self):
        self.h.forward(self.x, self.w, self.mu1, self.mu2, self.sigma, out=self.y)
        self.flat.zero_()
        out = self.h.backward(self.x, self.e, self.w, self.mu1, self.mu2, self.sigma, flags=self.flags, grads=self.grads)
        self.dx = out[0]
        return out

    def replay(self):
        """Runs the captured step; returns (y, dx, dw, dmu1, dmu2, dsigma) -- the SAME tensors every time."""
        self.graph.replay()
        return (self.y, self.dx) + tuple(self.grads)


_HANDLES = {}


def get_handle(device, **attrs):
    settings, key = _settings_from_attrs(attrs)
    dev = torch.device(device)
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    k = (key, idx)
    h = _HANDLES.get(k)
    if h is None:
        h = DAULayerHandle(settings, torch.device("cuda", idx))
        _HANDLES[k] = h
    return h


def _check_inputs(x, w, mu1, mu2, sigma, attrs):
    for name, t in (("input", x), ("weights", w), ("mu1", mu1), ("mu2", mu2), ("sigma", sigma)):
        if not t.is_cuda or t.dtype != torch.float32:
            raise RuntimeError("DAUConv: %s must be a float32 CUDA tensor" % name)
        if t.dim() != 4:
            raise RuntimeError("DAUConv: %s must have rank 4" % name)      # dau_conv_op.cpp:49-84
    num_output = attrs.get("num_output", OP_ATTRS["num_output"])
    G = attrs.get("number_units_x", 2) * attrs.get("number_units_y", 2)
    want = (1, x.shape[1], G, num_output)
    for name, t in (("weights", w), ("mu1", mu1), ("mu2", mu2), ("sigma", sigma)):
        if tuple(t.shape) != want:
            raise RuntimeError("DAUConv: %s has shape %s, expected %s" % (name, tuple(t.shape), want))


def _param_flags(attrs):
    """params_propagate_down of the reference's grad op: with single_dim_kernel it passes false for mu2
    and returns zeros for it (dau_conv_grad_op.cpp:291-292)."""
    f = _lib.GRAD_W | _lib.GRAD_MU1 | _lib.GRAD_SIGMA
    if not attrs.get("single_dim_kernel", False):
        f |= _lib.GRAD_MU2
    return f


def _sigma_of(sigma, sigma_value):
    # the TF op builds its layer from the live sigma (one number per layer); a caller that knows
    # the value passes it and spares the blocking device read
    return float(sigma.reshape(-1)[0]) if sigma_value is None else float(sigma_value)


def dau_conv_op(input, weights, mu1, mu2, sigma, sigma_value=None, check_nan=False, max_offset=None, **attrs):
    """TF op DAUConv: output[N,F,H,W] (forward only, no autograd).  check_nan=True reproduces the op's
    FAILED_PRECONDITION on NaN offsets (dau_conv_op.cpp:258-262) at the price of one host sync."""
    _check_inputs(input, weights, mu1, mu2, sigma, attrs)
    h = get_handle(input.device, **attrs)
    h.set_sigma_hint(_sigma_of(sigma, sigma_value))
    h.set_offset_hint(max_offset)
    out = h.forward(input.contiguous(), weights.contiguous(), mu1.contiguous(), mu2.contiguous(),
                    sigma.contiguous())
    if check_nan and h.take_nan_flag():
        raise RuntimeError("DAUConv: mu1 or mu2 contains NaN values")
    return out


def dau_conv_grad_op(grad, input, weights, mu1, mu2, sigma, sigma_value=None, max_offset=None, **attrs):
    """TF op DAUConvGrad: (grad_input, grad_weights, grad_mu1, grad_mu2, grad_sigma)."""
    _check_inputs(input, weights, mu1, mu2, sigma, attrs)
    h = get_handle(input.device, **attrs)
    h.set_sigma_hint(_sigma_of(sigma, sigma_value))
    h.set_offset_hint(max_offset)
    return h.backward(input.contiguous(), grad.contiguous(), weights.contiguous(), mu1.contiguous(),
                      mu2.contiguous(), sigma.contiguous(), flags=_lib.GRAD_INPUT | _param_flags(attrs))


class _DAUConvFunction(torch.autograd.Function):
    """@ops.RegisterGradient("DAUConv") equivalent (_dau_conv_grad_op.py:14-60)."""

    @staticmethod
    def forward(ctx, input, weights, mu1, mu2, sigma, attrs, sigma_value, max_offset=None):
        ctx.attrs, ctx.max_offset = attrs, max_offset
        ctx.sigma_value = _sigma_of(sigma, sigma_value)
        ctx.save_for_backward(input, weights, mu1, mu2, sigma)
        return dau_conv_op(input, weights, mu1, mu2, sigma, sigma_value=ctx.sigma_value, max_offset=max_offset,
                           **attrs)

    @staticmethod
    def backward(ctx, grad):
        input, weights, mu1, mu2, sigma = ctx.saved_tensors
        flags = 0
        need = ctx.needs_input_grad
        if need[0]:
            flags |= _lib.GRAD_INPUT
        if any(need[1:5]):
            flags |= _param_flags(ctx.attrs)
        _check_inputs(input, weights, mu1, mu2, sigma, ctx.attrs)
        h = get_handle(input.device, **ctx.attrs)
        h.set_sigma_hint(ctx.sigma_value)
        h.set_offset_hint(ctx.max_offset)
        dx, dw, dm1, dm2, dsg = h.backward(input.contiguous(), grad.contiguous(), weights.contiguous(),
                                           mu1.contiguous(), mu2.contiguous(), sigma.contiguous(), flags=flags)
        return (dx if need[0] else None, dw if need[1] else None, dm1 if need[2] else None,
                dm2 if need[3] else None, dsg if need[4] else None, None, None, None)


def dau_conv(input, weights, mu1, mu2, sigma, sigma_value=None, max_offset=None, **attrs):
    """Differentiable DAUConv (forward + registered gradient).  max_offset: an upper bound of |mu| the
    caller knows (plans the displacement bucket from it, see DAULayerHandle.set_offset_hint)."""
    return _DAUConvFunction.apply(input, weights, mu1, mu2, sigma, attrs, sigma_value, max_offset)


class _DAUConvFusedFunction(torch.autograd.Function):
    """DAUConv followed by the bias add and activation of the Python layer (dau_conv.py:500-526),
    both fused into the operator's kernels: the epilogue of the forward gather stores
    act(y + bias), the backward pass masks the incoming gradient as it loads it."""

    @staticmethod
    def forward(ctx, input, weights, mu1, mu2, sigma, bias, attrs, sigma_value, activation, max_offset=None):
        ctx.attrs, ctx.activation, ctx.max_offset = attrs, activation, max_offset
        ctx.sigma_value = _sigma_of(sigma, sigma_value)
        _check_inputs(input, weights, mu1, mu2, sigma, attrs)
        h = get_handle(input.device, **attrs)
        h.set_sigma_hint(ctx.sigma_value)
        h.set_offset_hint(max_offset)
        out = h.forward(input.contiguous(), weights.contiguous(), mu1.contiguous(), mu2.contiguous(),
                        sigma.contiguous(), bias=bias.contiguous() if bias is not None else None,
                        activation=activation)
        ctx.save_for_backward(input, weights, mu1, mu2, sigma, out)
        return out

    @staticmethod
    def backward(ctx, grad):
        input, weights, mu1, mu2, sigma, out = ctx.saved_tensors
        need = ctx.needs_input_grad
        flags = 0
        if need[0]:
            flags |= _lib.GRAD_INPUT
        if any(need[1:5]):
            flags |= _param_flags(ctx.attrs)
        dbias = None
        if need[5]:
            flags |= _lib.GRAD_BIAS
            dbias = torch.zeros(weights.shape[-1], dtype=torch.float32, device=grad.device)
        h = get_handle(input.device, **ctx.attrs)
        h.set_sigma_hint(ctx.sigma_value)
        h.set_offset_hint(ctx.max_offset)
        dx, dw, dm1, dm2, dsg = h.backward(input.contiguous(), grad.contiguous(), weights.contiguous(),
                                           mu1.contiguous(), mu2.contiguous(), sigma.contiguous(), flags=flags,
                                           dbias=dbias, top_data=out, activation=ctx.activation)
        return (dx if need[0] else None, dw if need[1] else None, dm1 if need[2] else None,
                dm2 if need[3] else None, dsg if need[4] else None, dbias, None, None, None, None)


def dau_conv_fused(input, weights, mu1, mu2, sigma, bias=None, activation=None, sigma_value=None, max_offset=None,
                   **attrs):
    """Differentiable act(DAUConv(input) + bias) in ONE pass of the operator (SURVEY section 8
    row f3).  activation: None or "relu"."""
    if activation not in (None, "relu"):
        raise ValueError("dau_conv_fused: activation must be None or 'relu'")
    attrs = dict(attrs, bias_term=bias is not None)
    act = _lib.ACT_RELU if activation == "relu" else _lib.ACT_NONE
    return _DAUConvFusedFunction.apply(input, weights, mu1, mu2, sigma, bias, attrs, sigma_value, act, max_offset)
