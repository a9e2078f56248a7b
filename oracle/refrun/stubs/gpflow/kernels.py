import numpy as np
import tensorflow as tf


class Kernel:
    """gpflow/kernels.py (1.5.1): ``input_dim``, ``active_dims`` and the column slicing of ``_slice``."""

    def __init__(self, input_dim, active_dims=None, name=None):
        self.input_dim = int(input_dim)
        self.name = name
        if active_dims is None:
            self.active_dims = slice(input_dim)
        elif isinstance(active_dims, slice):
            self.active_dims = active_dims
        else:
            self.active_dims = np.array(active_dims, dtype=np.int32)

    def _slice(self, X, X2):
        if isinstance(self.active_dims, slice):
            X = X[..., self.active_dims]
            if X2 is not None:
                X2 = X2[..., self.active_dims]
        else:
            X = tf.gather(X, self.active_dims, axis=-1)
            if X2 is not None:
                X2 = tf.gather(X2, self.active_dims, axis=-1)
        return X, X2


class Combination(Kernel):
    pass


class Sum(Combination):
    pass


class Product(Combination):
    pass
