import numpy as np

float_type = np.float64
int_type = np.int32
jitter = 1e-6           # gpflow 1.5.1 default numerics.jitter_level


class dtypes:            # ``settings.dtypes.float_type`` (inducing_variables.py:61)
    float_type = float_type
    int_type = int_type
