"""Global numerics settings, mirroring the two GPflow 1.5 settings the reference
reads on the hot path (``settings.float_type`` kernels.py:66,130;
``settings.jitter`` kernels.py:432,464,610,689).  GPflow's defaults are float64
and 1e-6."""
import torch

float_type = torch.float64
jitter = 1e-6
_device = None


def device():
    """Device used by the numpy ``compute_*`` wrappers (current CUDA device)."""
    if _device is not None:
        return torch.device(_device)
    if not torch.cuda.is_available():
        raise RuntimeError("gpsig_b200: no CUDA device is available and there is no CPU path")
    return torch.device('cuda', torch.cuda.current_device())


def set_device(dev):
    global _device
    _device = dev
