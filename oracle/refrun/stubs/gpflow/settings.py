import numpy as np

float_type = np.float64
int_type = np.int32
jitter = 1e-6           # gpflow 1.5.1 default numerics.jitter_level
