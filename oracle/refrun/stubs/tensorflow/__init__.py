"""Eager numpy stand-in for the TensorFlow-1.15 ops the reference's truncated signature-kernel path calls
(test infrastructure, see oracle/refrun/__init__.py).  Every op evaluates at once in float64 numpy with TF's semantics."""
import numpy as np

float64, float32, int32, int64 = np.float64, np.float32, np.int32, np.int64
newaxis = None


class TensorShape(tuple):
    @property
    def ndims(self):
        return len(self)

    def as_list(self):
        return list(self)


class Tensor:
    """Opaque to numpy like a graph tensor (``np.asarray([[M]])`` must give an OBJECT array holding M,
    signature_algs.py:58; numpy scalars must defer to its reflected operators), array arithmetic inside."""
    __array_ufunc__ = None
    __array_priority__ = 1000

    def __init__(self, a):
        self.a = a.a if isinstance(a, Tensor) else np.asarray(a)

    @property
    def shape(self):
        return TensorShape(self.a.shape)

    @property
    def dtype(self):
        return self.a.dtype

    def get_shape(self):
        return self.shape

    def __len__(self):
        raise TypeError('len is not well defined for symbolic Tensors')

    def __iter__(self):
        raise TypeError('Tensor objects are only iterable when eager execution is enabled')

    def __getitem__(self, idx):
        if not isinstance(idx, tuple):
            idx = (idx,)
        return Tensor(self.a[tuple(_a(i) if isinstance(i, Tensor) else i for i in idx)])

    def __neg__(self):
        return Tensor(-self.a)

    def __repr__(self):
        return 'Tensor(%r)' % (self.a,)


def _binop(name, f, reflected=False):
    def op(self, other):
        x, y = self.a, _a(other)
        return Tensor(f(y, x) if reflected else f(x, y))
    op.__name__ = name
    setattr(Tensor, name, op)


for _n, _f in (('add', np.add), ('sub', np.subtract), ('mul', np.multiply), ('truediv', np.true_divide), ('pow', np.power),
               ('floordiv', np.floor_divide), ('mod', np.mod)):
    _binop('__%s__' % _n, _f)
    _binop('__r%s__' % _n, _f, reflected=True)
for _n, _f in (('lt', np.less), ('le', np.less_equal), ('gt', np.greater), ('ge', np.greater_equal)):
    _binop('__%s__' % _n, _f)


def _a(x):
    """numpy value of a tensor / array / scalar / (nested) list of them"""
    if isinstance(x, Tensor):
        return x.a
    if isinstance(x, (list, tuple)):
        return np.asarray([_a(v) for v in x])
    return np.asarray(x)


def _t(x, dtype=None):
    return Tensor(np.asarray(_a(x), dtype=dtype))


def constant(x, dtype=None):
    return _t(x, dtype)


convert_to_tensor = constant


def shape(x):
    return np.asarray(_a(x).shape, dtype=np.int64)      # indexable, unstackable integers


def _dims(s):
    return tuple(int(v) for v in np.atleast_1d(_a(s)))


def reshape(x, s):
    return _t(np.reshape(_a(x), _dims(s)))


def cast(x, dtype):
    return _t(_a(x).astype(dtype))


def ones(s, dtype=np.float64):
    return _t(np.ones(_dims(s), dtype=dtype))


def zeros(s, dtype=np.float64):
    return _t(np.zeros(_dims(s), dtype=dtype))


def fill(dims, value):
    return _t(np.full(_dims(dims), _a(value)))


def ones_like(x):
    return _t(np.ones_like(_a(x)))


def eye(n, dtype=np.float64):
    return _t(np.eye(int(n), dtype=dtype))


def range(limit, dtype=np.float64):          # only the one-argument form is used (lags.py:56)
    return _t(np.arange(int(_a(limit)), dtype=dtype))


def reduce_sum(x, axis=None, keepdims=False, keep_dims=None):
    if keep_dims is not None:
        keepdims = keep_dims
    if isinstance(axis, (list, np.ndarray)):
        axis = tuple(int(a) for a in axis)
    return _t(np.sum(_a(x), axis=axis, keepdims=keepdims))


def add_n(values):
    out = _a(values[0]).copy()
    for v in values[1:]:
        out = out + _a(v)
    return _t(out)


def cumsum(x, axis=0, exclusive=False, reverse=False):
    a = _a(x)
    if reverse:
        a = np.flip(a, axis=axis)
    c = np.cumsum(a, axis=axis)
    if exclusive:                              # shifted by one along the axis, zero first
        c = c - a
    if reverse:
        c = np.flip(c, axis=axis)
    return _t(c)


def matmul(a, b, transpose_a=False, transpose_b=False):
    a, b = _a(a), _a(b)
    if transpose_a:
        a = np.swapaxes(a, -1, -2)
    if transpose_b:
        b = np.swapaxes(b, -1, -2)
    return _t(np.matmul(a, b))


def transpose(x, perm=None):
    return _t(np.transpose(_a(x), perm))


def stack(values, axis=0):
    return _t(np.stack([_a(v) for v in values], axis=axis))


def unstack(x, num=None, axis=0):
    a = _a(x)
    parts = [np.take(a, i, axis=axis) for i in np.arange(a.shape[axis])]
    return [p.item() if (p.ndim == 0 and np.issubdtype(p.dtype, np.integer)) else _t(p) for p in parts]


def concat(values, axis):
    return _t(np.concatenate([_a(v) for v in values], axis=axis))


def tile(x, multiples):
    return _t(np.tile(_a(x), _dims(multiples)))


def gather(params, indices, axis=0):
    return _t(np.take(_a(params), _a(indices), axis=axis))


def matrix_diag_part(x):
    return _t(np.diagonal(_a(x), axis1=-2, axis2=-1).copy())


def where(cond, x, y):
    return _t(np.where(_a(cond), _a(x), _a(y)))


def argmax(x, axis=None):
    return np.argmax(_a(x), axis=axis)


def argmin(x, axis=None):
    return np.argmin(_a(x), axis=axis)


def maximum(x, y):
    return _t(np.maximum(_a(x), _a(y)))


def minimum(x, y):
    return _t(np.minimum(_a(x), _a(y)))


def _unary(f):
    return lambda x: _t(f(_a(x)))


square, sqrt, exp, log, cos, sin, floor, ceil, abs = (_unary(f) for f in (np.square, np.sqrt, np.exp, np.log, np.cos, np.sin, np.floor, np.ceil, np.abs))


def repeat(x, repeats, axis=None):
    return _t(np.repeat(_a(x), int(_a(repeats)), axis=axis))


def reverse(x, axis, name=None):
    return _t(np.flip(_a(x), axis=tuple(int(v) for v in np.atleast_1d(_a(axis)))))


class math:
    @staticmethod
    def pow(x, y):
        return _t(np.power(_a(x), _a(y)))


class dtypes:
    cast = staticmethod(cast)


# graph plumbing of the Cython path (kernels_pde.py:444-463): py_func evaluates at once, gradients are not registered
def py_func(func, inp, Tout, stateful=True, name=None):
    outs = func(*[_a(v) if isinstance(v, Tensor) else v for v in inp])
    return [_t(np.asarray(o)) for o in outs]


def RegisterGradient(name):
    def deco(fn):
        return fn
    return deco


class _Graph:
    import contextlib as _cl

    @_cl.contextmanager
    def gradient_override_map(self, mapping):
        yield


def get_default_graph():
    return _Graph()


def load_op_library(path):
    """The compiled TF custom op (covariance_op/untrunc_cov_op_gpu.so) does not exist here; nothing on the Cython path calls it."""
    return None


def random_uniform(*args, **kwargs):
    raise NotImplementedError('low-rank (randomised) paths are out of scope of this stand-in')


class _Contrib:
    stateless = None


contrib = _Contrib()
