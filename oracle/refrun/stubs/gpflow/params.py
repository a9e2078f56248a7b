import numpy as np
import tensorflow as tf


def Parameter(value, transform=None, dtype=None, trainable=True, name=None):
    """The constrained value as a tensor (what ``params_as_tensors`` hands to the kernel's methods)."""
    return tf.Tensor(np.array(value, dtype=dtype if dtype is not None else np.float64))


class ParamList(list):
    pass
