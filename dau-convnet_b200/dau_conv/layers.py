"""DAUConv2d / DAUConv1d / dau_conv2d: the layer API of the reference's Python package on top of
the B200 operator.

Mirrors plugins/tensorflow/dau_conv/dau_conv.py of the reference:
  DAUGridMean, ZeroNLast        :24-110   (initialisers)
  _DAUConvolution2d.__call__    :183-218  (mu clipping, attribute set handed to the op)
  DAUConv2d                     :221-558  (variables [1,S,G,F], sigma (1,) tiled, odd-G padding,
                                           stride by slicing, bias, activation)
  DAUConv1d                     :560-571
  dau_conv2d                    :580-795  (functional wrapper: normaliser or bias, activation)
Names, argument order and defaults are the reference's; tensors are torch tensors (NCHW, CUDA).
"""
import math

import numpy as np
import torch
from torch import nn

from . import functional as Fn


def dau_grid_mean(shape, dau_units, max_value, dau_unit_axis=2, exact_grid=False):
    """DAUGridMean (dau_conv.py:24-74): unit offsets on a regular grid in [-max_value, max_value].

    shape is [1, S, G, F]; dau_unit_axis 2 -> mu1 (x), 1 -> mu2 (y).  The reference builds the
    values on a [S, units0, units1, F] view and replicates them with tf.tile, but indexes the
    replication list with the NUMBER of units instead of the axis (:61-62); for mu2 the grid
    therefore comes out interleaved.  The default reproduces that exactly (same initial values
    as the reference); exact_grid=True gives the intended axis-aligned grid."""
    assert len(shape) == 4, "DAUGridMean requires rank 4: [1, input_channels, mu1*mu2, out_filters]"
    separated = shape[2] != dau_units[0] * dau_units[1]
    view = list(shape) if separated else [shape[1], dau_units[0], dau_units[1], shape[-1]]
    n = view[dau_unit_axis]
    span = 2 * max_value + 1
    vals = np.arange(n) * span / float(n) + (-0.5 + span / float(2 * n)) - max_value
    vshape = [1, 1, 1, 1]
    vshape[dau_unit_axis] = n
    vals = vals.reshape(vshape).astype(np.float32)
    def tiled(axis_to_keep):
        reps = list(view)
        reps[axis_to_keep] = 1
        return np.tile(vals, reps)
    out = None
    if not exact_grid and n < 4:                          # reference: tile_rep_shape[num_units] = 1
        out = tiled(n)
        if out.size != int(np.prod(view)):                # the reference fails for this unit layout
            out = None
    if out is None:
        out = tiled(dau_unit_axis)
    if not separated:
        out = out.reshape(1, view[0], view[1] * view[2], view[3])
    return torch.from_numpy(np.ascontiguousarray(out))


def zero_n_last(t, last_num_to_zero, axis):
    """ZeroNLast (dau_conv.py:76-110): zero the last N entries along `axis` (dummy units)."""
    if last_num_to_zero > 0:
        idx = [slice(None)] * t.dim()
        idx[axis] = slice(t.shape[axis] - last_num_to_zero, None)
        t[tuple(idx)] = 0
    return t


class DAUConv2d(nn.Module):
    """Displaced-aggregation-unit convolution layer (dau_conv.py:221-558).

    Parameters `weights`, `mu1`, `mu2` have shape [1, S, G, F]; `sigma` is ONE number (shape (1,),
    trainable only with dau_sigma_trainable) tiled to [1, S, G, F] for the op (:418-430)."""
    DAU_UNITS_GROUP = 2      # the reference's kernels process units in pairs (:224)

    def __init__(self, filters, dau_units, max_kernel_size, strides=1, data_format="channels_first",
                 activation=None, use_bias=True, weight_initializer=None, mu1_initializer=None,
                 mu2_initializer=None, sigma_initializer=None, bias_initializer=None,
                 mu_learning_rate_factor=500, dau_unit_border_bound=0.01, dau_unit_single_dim=False,
                 dau_aggregation_forbid_positive_dim1=False, dau_sigma_trainable=False,
                 dau_mu_interpolation=True, unit_testing=False, in_channels=None, dynamic_kernel_size=False):
        super().__init__()
        if data_format != "channels_first":
            raise ValueError("Only `channels_first` supported, i.e., NCHW format.")      # :374-376
        if isinstance(dau_units, int):
            dau_units = (dau_units, dau_units)
        self.in_channels, self.filters = None, int(filters)
        self.dau_units = (int(dau_units[0]), int(dau_units[1]))
        self.max_kernel_size = int(max_kernel_size)
        self.padding = int(math.floor(self.max_kernel_size / 2.0))
        self.strides = int(strides)
        self.activation, self.use_bias = activation, use_bias
        self.mu_learning_rate_factor = mu_learning_rate_factor
        self.dau_unit_border_bound = dau_unit_border_bound
        self.dau_unit_single_dim = dau_unit_single_dim
        self.dau_aggregation_forbid_positive_dim1 = dau_aggregation_forbid_positive_dim1
        self.dau_sigma_trainable = dau_sigma_trainable
        self.dau_mu_interpolation = dau_mu_interpolation
        self.unit_testing = unit_testing
        # True: read max|mu| back on every call and plan the displacement bucket that holds it, what the
        # reference's op does (dau_conv_op.cpp:222-256); False (default): plan for max_kernel_size, no sync
        self.dynamic_kernel_size = bool(dynamic_kernel_size)
        # odd unit counts get dummy units with zero weight (:311-327)
        self.num_dau_units_all = self.dau_units[0] * self.dau_units[1]
        self.num_dau_units_ignore = 0
        if self.num_dau_units_all % self.DAU_UNITS_GROUP != 0:
            new_n = int(math.ceil(self.num_dau_units_all / float(self.DAU_UNITS_GROUP)) * self.DAU_UNITS_GROUP)
            self.num_dau_units_ignore = new_n - self.num_dau_units_all
            if self.dau_units[0] < self.dau_units[1]:
                self.dau_units = (self.dau_units[0] + self.num_dau_units_ignore, self.dau_units[1])
            else:
                self.dau_units = (self.dau_units[0], self.dau_units[1] + self.num_dau_units_ignore)
            if self.dau_units[0] * self.dau_units[1] != new_n:
                # the reference's rule only balances for 1 x 1 units; keep G = new_n for the others
                self.dau_units = (1, new_n) if dau_units[0] < dau_units[1] else (new_n, 1)
            self.num_dau_units_all = new_n
        self._init = dict(weight=weight_initializer, mu1=mu1_initializer, mu2=mu2_initializer,
                          sigma=sigma_initializer, bias=bias_initializer)
        self.mean_max_allowed_offset = math.floor(self.max_kernel_size / 2.0) - self.dau_unit_border_bound
        self.weights = self.mu1 = self.mu2 = self.sigma = self.bias = None
        self._sigma_value = None             # host copy of sigma: spares the op a device read per call
        self._sigma_key = None               # (id, version) of the sigma parameter the copy was read from
        if in_channels is not None:
            self.build(in_channels)

    def get_dau_variable_shape(self, in_channels):
        return (1, int(in_channels), self.num_dau_units_all, self.filters)       # :392-399

    def build(self, in_channels, device=None):
        """Creates the variables (the reference does this in build(), :441-491, once the number of
        input channels is known); called by the first forward when in_channels was not given."""
        self.in_channels = int(in_channels)
        shape = self.get_dau_variable_shape(in_channels)
        max_mu = math.floor(self.max_kernel_size / 2.0) - 1
        ini = self._init
        w = ini["weight"](shape) if ini["weight"] else 0.1 * torch.randn(shape)             # stddev 0.1 (:232)
        w = zero_n_last(w.clone(), self.num_dau_units_ignore, axis=2)
        mu1 = ini["mu1"](shape) if ini["mu1"] else dau_grid_mean(shape, self.dau_units, max_mu, 2)
        mu2 = ini["mu2"](shape) if ini["mu2"] else dau_grid_mean(shape, self.dau_units, max_mu, 1)
        sigma = ini["sigma"]((1,)) if ini["sigma"] else torch.full((1,), 0.5)               # :296-297
        self.weights = nn.Parameter(w.float().to(device))
        self.mu1 = nn.Parameter(mu1.float().reshape(shape).to(device))
        self.mu2 = nn.Parameter(mu2.float().reshape(shape).to(device))
        self.sigma = nn.Parameter(sigma.float().to(device), requires_grad=bool(self.dau_sigma_trainable))
        if self.use_bias:
            b = ini["bias"]((self.filters,)) if ini["bias"] else torch.zeros(self.filters)
            self.bias = nn.Parameter(b.float().to(device))
        return self

    def set_dau_variables_manually(self, w=None, mu1=None, mu2=None, sigma=None):
        """dau_conv.py:357-372: install caller-made tensors of shape get_dau_variable_shape()."""
        for name, t in (("weights", w), ("mu1", mu1), ("mu2", mu2)):
            if t is not None:
                if self.in_channels is None:
                    self.build(t.shape[1], device=t.device)
                if tuple(t.shape) != self.get_dau_variable_shape(self.in_channels):
                    raise ValueError("Shape mismatch for variable `dau_%s`" % name)         # :446-458
                setattr(self, name, nn.Parameter(t.float()))
        if sigma is not None:
            self.sigma = nn.Parameter(sigma.float().reshape(-1)[:1].clone(), requires_grad=bool(self.dau_sigma_trainable))
            self._sigma_value = None

    def _op_attrs(self):
        # the attribute set _DAUConvolution2d hands to the op (:197-210)
        return dict(num_output=self.filters, number_units_x=self.dau_units[0], number_units_y=self.dau_units[1],
                    number_units_ignore=self.num_dau_units_ignore, kernel_size=self.max_kernel_size,
                    pad=self.padding, component_border_bound=self.dau_unit_border_bound,
                    sigma_lower_bound=0.01, mu_learning_rate_factor=self.mu_learning_rate_factor,
                    single_dim_kernel=self.dau_unit_single_dim,
                    forbid_positive_dim1=self.dau_aggregation_forbid_positive_dim1,
                    use_interpolation=self.dau_mu_interpolation, unit_testing=self.unit_testing)

    def forward(self, inputs):
        if inputs.dim() != 4:
            raise ValueError("DAU convolution not supported for input with rank %d" % inputs.dim())
        if self.weights is None:
            self.build(inputs.shape[1], device=inputs.device)
        if inputs.shape[1] != self.in_channels:
            raise ValueError("expected %d input channels, got %d" % (self.in_channels, inputs.shape[1]))
        lim = self.mean_max_allowed_offset
        mu1 = self.mu1.clamp(-lim, lim)                       # :185-186 (clipping is part of the graph)
        mu2 = self.mu2.clamp(-lim, lim)
        sigma = self.sigma.reshape(1, 1, 1, 1).expand(self.weights.shape)      # tiled view (:418-430)
        # host copy of sigma, refreshed whenever the parameter object or its version counter changed
        # (optimizer step, load_state_dict, in-place edit): one blocking read per change, none otherwise
        key = (id(self.sigma), self.sigma._version)
        if self._sigma_value is None or self._sigma_key != key:
            self._sigma_value = float(self.sigma.detach().reshape(-1)[0])
            self._sigma_key = key
        # bias add and ReLU (:500-526) are element-wise, so they commute with the stride slicing and
        # run inside the operator's own kernels; any other activation is applied afterwards
        relu = self.activation in (torch.relu, nn.functional.relu) or isinstance(self.activation, nn.ReLU)
        out = Fn.dau_conv_fused(inputs, self.weights, mu1, mu2, sigma, bias=self.bias if self.use_bias else None,
                                activation="relu" if relu else None, sigma_value=self._sigma_value,
                                max_offset=(float(torch.maximum(mu1.detach().abs().max(), mu2.detach().abs().max()))
                                            if self.dynamic_kernel_size else None),
                                **self._op_attrs())
        if self.strides > 1:                                  # emulated by slicing (:496-498)
            out = out[:, :, ::self.strides, ::self.strides]
        return self.activation(out) if (self.activation is not None and not relu) else out


class DAUConv1d(DAUConv2d):
    """dau_conv.py:560-571: single-dimension kernel, mu2 pinned to zero."""

    def __init__(self, filters, dau_units, max_kernel_size, **kwargs):
        kwargs.pop("mu2_initializer", None)
        super().__init__(filters, dau_units, max_kernel_size, mu2_initializer=torch.zeros,
                         dau_unit_single_dim=True, **kwargs)

    def forward(self, inputs):
        if self.mu2 is not None:
            with torch.no_grad():
                self.mu2.zero_()                              # mu2_constraint = zeros_like (:565-566)
        return super().forward(inputs)


def dau_conv2d(inputs, filters, dau_units, max_kernel_size, stride=1, mu_learning_rate_factor=500,
               data_format=None, activation_fn=torch.relu, normalizer_fn=None, normalizer_params=None,
               weights_initializer=None, mu1_initializer=None, mu2_initializer=None, sigma_initializer=None,
               biases_initializer=None, dau_unit_border_bound=0.01, dau_sigma_trainable=False,
               dau_mu_interpolation=True, layer=None):
    """Functional form (dau_conv.py:580-795): builds (or re-uses, via `layer`) a DAUConv2d without
    activation, then applies the normaliser INSTEAD of the bias when one is given, then the
    activation.  Returns (outputs, layer) so the caller can keep the variables."""
    if data_format not in (None, "NCHW"):
        raise ValueError("Invalid data_format: %r" % (data_format,))
    if inputs.dim() != 4:
        raise ValueError("DAU convolution not supported for input with rank %d" % inputs.dim())
    if layer is None:
        layer = DAUConv2d(filters, dau_units, max_kernel_size, strides=stride, activation=None,
                          use_bias=(normalizer_fn is None) and biases_initializer is not False,
                          weight_initializer=weights_initializer, mu1_initializer=mu1_initializer,
                          mu2_initializer=mu2_initializer, sigma_initializer=sigma_initializer,
                          bias_initializer=biases_initializer or None,
                          mu_learning_rate_factor=mu_learning_rate_factor,
                          dau_unit_border_bound=dau_unit_border_bound, dau_sigma_trainable=dau_sigma_trainable,
                          dau_mu_interpolation=dau_mu_interpolation)
        layer.build(inputs.shape[1], device=inputs.device)
    out = layer(inputs)
    if normalizer_fn is not None:
        out = normalizer_fn(out, **(normalizer_params or {}))
    if activation_fn is not None:
        out = activation_fn(out)
    return out, layer
