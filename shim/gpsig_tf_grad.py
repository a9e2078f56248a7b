"""Gradient registration of the TensorFlow shim ops (shim/gpsig_tf_ops.cc), the counterpart of
gpsig/covariance_op/_untrunc_cov_grad.py:25-77 (``@ops.RegisterGradient("UntruncCov")``).

OPTIONAL, needs TensorFlow (absent from this image; see shim/build_shim.py).  The C ABI's adjoints take the
coordinates already divided by the lengthscales and return gradients with respect to those, so the division
(gpsig/kernels.py:358-359) is chained here.
"""
import os

import tensorflow as tf
from tensorflow.python.framework import ops

_here = os.path.dirname(os.path.abspath(__file__))
gpsig_ops = tf.load_op_library(os.path.join(_here, 'gpsig_tf_ops.so'))     # like kernels_pde.py:17-22


def _trunc_attrs(op):
    return {k: op.get_attr(k) for k in ('base', 'num_levels', 'order', 'difference', 'base_p0', 'base_p1')}


@ops.RegisterGradient("GpsigTruncGram")
def _gpsig_trunc_gram_grad(op, grad_levels):
    x, x2, ls = op.inputs
    xs, x2s = x / ls, x2 / ls
    gx, gx2 = gpsig_ops.gpsig_trunc_gram_grad(xs, x2s, grad_levels, **_trunc_attrs(op))
    # d/dx = g / ls;  d/dls = -sum(g * x) / ls^2 over both inputs
    gls = -(tf.reduce_sum(gx * xs, axis=[0, 1]) + tf.reduce_sum(gx2 * x2s, axis=[0, 1])) / ls
    return gx / ls, gx2 / ls, gls


@ops.RegisterGradient("GpsigPdeDiag")
def _gpsig_pde_diag_grad(op, grad_kdiag):
    x, ls = op.inputs
    xs = x / ls
    gx = gpsig_ops.gpsig_pde_diag_grad(xs, grad_kdiag, order=op.get_attr('order'), rev_solver=1)
    return gx / ls, -tf.reduce_sum(gx * xs, axis=[0, 1]) / ls
