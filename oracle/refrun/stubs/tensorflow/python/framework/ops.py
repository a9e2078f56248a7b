"""``tensorflow.python.framework.ops`` as far as gpsig/kernels_pde.py and covariance_op/_untrunc_cov_grad.py touch it."""
import contextlib


def RegisterGradient(name):
    def deco(fn):
        return fn
    return deco


@contextlib.contextmanager
def name_scope(name, default_name=None, values=None):
    yield name or default_name
