"""ctypes binding of ``libgpsig_b200.so`` (C ABI in ``include/gpsig_b200.h``).

There is NO fallback: if the CUDA library has not been built, importing this
module raises.  torch is used only for device memory and streams; the raw
device pointers and the current stream handle are what crosses the boundary.
"""
import ctypes
import os

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, os.environ.get('GPSIG_LIB_NAME', 'libgpsig_b200.so'))

BASE_IDS = {'linear': 0, 'rbf': 1, 'cosine': 2, 'poly': 3, 'mix': 4,
            'matern12': 5, 'matern32': 6, 'matern52': 7, 'spectral': 8}
SPECTRAL_FAMILIES = {'rbf': 0, 'exp': 1}
PDE_DIAG, PDE_GRAM = 0, 1


class TruncDesc(ctypes.Structure):
    _fields_ = [('base', ctypes.c_int), ('num_levels', ctypes.c_int), ('order', ctypes.c_int),
                ('difference', ctypes.c_int), ('base_params', ctypes.c_double * 4),
                ('spectral', ctypes.c_void_p), ('spectral_q', ctypes.c_int), ('spectral_family', ctypes.c_int)]


class Epilogue(ctypes.Structure):
    _fields_ = [('norm_x', ctypes.c_void_p), ('norm_x2', ctypes.c_void_p), ('level_scale', ctypes.c_void_p),
                ('sym_jitter', ctypes.c_double), ('return_levels', ctypes.c_int), ('upper_only', ctypes.c_int)]


class GpsigError(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "gpsig_b200: %s is missing. Build it with `python gpsig_b200/build.py` "
            "(there is no CPU or PyTorch fallback for the signature-kernel ops)." % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    vp, i, sz, dbl = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t, ctypes.c_double
    lib.gpsig_last_error_string.restype = ctypes.c_char_p
    lib.gpsig_version.restype = i
    lib.gpsig_trunc_gram_workspace_bytes.restype = sz
    lib.gpsig_trunc_gram_workspace_bytes.argtypes = [i] * 6
    lib.gpsig_trunc_diag_workspace_bytes.restype = sz
    lib.gpsig_trunc_diag_workspace_bytes.argtypes = [i] * 4
    lib.gpsig_tens_vs_seq_workspace_bytes.restype = sz
    lib.gpsig_tens_vs_seq_workspace_bytes.argtypes = [i] * 7
    lib.gpsig_pde_workspace_bytes.restype = sz
    lib.gpsig_pde_workspace_bytes.argtypes = [i] * 7
    for sfx in ('f32', 'f64'):
        f = getattr(lib, 'gpsig_trunc_gram_' + sfx)
        f.restype = i
        f.argtypes = [ctypes.POINTER(TruncDesc), vp, i, i, vp, i, i, i, vp, i, i, ctypes.POINTER(Epilogue), vp, vp, sz, vp]
        f = getattr(lib, 'gpsig_trunc_diag_' + sfx)
        f.restype = i
        f.argtypes = [ctypes.POINTER(TruncDesc), vp, i, i, i, vp, vp, vp, sz, vp]
        f = getattr(lib, 'gpsig_tens_vs_seq_' + sfx)
        f.restype = i
        f.argtypes = [ctypes.POINTER(TruncDesc), vp, i, i, vp, i, i, i, vp, i, i, ctypes.POINTER(Epilogue), vp, vp, sz, vp]
        f = getattr(lib, 'gpsig_tens_tens_' + sfx)
        f.restype = i
        f.argtypes = [ctypes.POINTER(TruncDesc), vp, i, i, i, vp, ctypes.POINTER(Epilogue), vp, vp]
    lib.gpsig_trunc_diag_static_f64.restype = i          # no-allocation variant (csrc/static_scratch.cu)
    lib.gpsig_trunc_diag_static_f64.argtypes = [ctypes.POINTER(TruncDesc), vp, i, i, i, vp, vp, vp]
    lib.gpsig_trunc_vjp_workspace_bytes.restype = sz
    lib.gpsig_trunc_vjp_workspace_bytes.argtypes = [i] * 6
    lib.gpsig_trunc_vjp_workspace_bytes_order.restype = sz
    lib.gpsig_trunc_vjp_workspace_bytes_order.argtypes = [i] * 7
    lib.gpsig_tens_vs_seq_vjp_workspace_bytes.restype = sz
    lib.gpsig_tens_vs_seq_vjp_workspace_bytes.argtypes = [i] * 5
    lib.gpsig_tens_vs_seq_lifted_gd_workspace_bytes.restype = sz
    lib.gpsig_tens_vs_seq_lifted_gd_workspace_bytes.argtypes = [i] * 7
    for sfx in ('f32', 'f64'):
        f = getattr(lib, 'gpsig_trunc_vjp_' + sfx)
        f.restype = i
        f.argtypes = [ctypes.POINTER(TruncDesc), vp, i, i, vp, i, i, i, i, i, i, vp, vp, vp, vp, sz, vp]
        f = getattr(lib, 'gpsig_tens_vs_seq_vjp_' + sfx)
        f.restype = i
        f.argtypes = [ctypes.POINTER(TruncDesc), vp, i, i, vp, i, i, i, i, i, vp, vp, vp, vp, sz, vp]
        f = getattr(lib, 'gpsig_tens_vs_seq_lifted_gd_' + sfx)
        f.restype = i
        f.argtypes = [ctypes.POINTER(TruncDesc), vp, i, i, vp, i, i, i, i, i, vp, vp, vp, sz, vp]
        f = getattr(lib, 'gpsig_tens_tens_vjp_' + sfx)
        f.restype = i
        f.argtypes = [ctypes.POINTER(TruncDesc), vp, i, i, i, vp, vp, vp]
    lib.gpsig_launch_count.restype = ctypes.c_ulonglong
    lib.gpsig_profile_enable.argtypes = [i]
    lib.gpsig_profile_enable.restype = None
    lib.gpsig_profile_last_kernel_ms.restype = dbl
    lib.gpsig_fma_peak.restype = i
    lib.gpsig_fma_peak.argtypes = [i, i, i, vp, vp]
    lib.gpsig_pde_f64.restype = i
    lib.gpsig_pde_f64.argtypes = [vp, i, i, vp, i, i, i, vp, i, i, i, i, dbl, vp, vp, sz, vp]
    for sfx_ in ('f64', 'f32'):
        fn = getattr(lib, 'gpsig_mirror_lower_' + sfx_)
        fn.restype = i
        fn.argtypes = [vp, i, i, vp]
    lib.gpsig_pde_kdiag_vjp_workspace_bytes.restype = sz
    lib.gpsig_pde_kdiag_vjp_workspace_bytes.argtypes = [i, i, i, i]
    lib.gpsig_pde_kdiag_vjp_f64.restype = i
    lib.gpsig_pde_kdiag_vjp_f64.argtypes = [vp, i, i, i, i, i, i, i, vp, vp, vp, sz, vp]
    return lib


lib = _load()

_ERR = {-1: ValueError, -2: NotImplementedError, -3: GpsigError, -4: GpsigError}


def check(rc):
    if rc != 0:
        msg = lib.gpsig_last_error_string().decode()
        raise _ERR.get(rc, GpsigError)("gpsig_b200: " + msg)


def sfx(dtype):
    if dtype == torch.float64:
        return 'f64'
    if dtype == torch.float32:
        return 'f32'
    raise TypeError("gpsig_b200 kernels take float32 or float64 tensors, got %s" % dtype)


def ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def stream_handle(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def require_cuda(t, name):
    if not t.is_cuda:
        raise RuntimeError("gpsig_b200: %s must be a CUDA tensor (no CPU path exists)" % name)
    if not t.is_contiguous():
        raise RuntimeError("gpsig_b200: %s must be contiguous" % name)


def workspace(nbytes, device):
    return torch.empty(max(int(nbytes), 16), dtype=torch.uint8, device=device)


def make_desc(base, num_levels, order, difference, base_params=()):
    d = TruncDesc()
    d.base = BASE_IDS[base]
    d.num_levels = int(num_levels)
    d.order = int(order)
    d.difference = int(bool(difference))
    if base == 'spectral':
        # base_params = (family, block): block = float64 CUDA tensor alpha[Q] ++ omega[Q, d] ++ gamma[Q, d], Q = block[0] size
        family, q, block = base_params
        if not (torch.is_tensor(block) and block.is_cuda and block.dtype == torch.float64 and block.is_contiguous()):
            raise TypeError("gpsig_b200: the spectral parameter block must be a contiguous float64 CUDA tensor")
        d.spectral = block.data_ptr()
        d.spectral_q = int(q)
        d.spectral_family = SPECTRAL_FAMILIES[family]
        d._keep = block            # the descriptor only holds a raw pointer
        base_params = ()
    bp = list(base_params) + [0.0] * 4
    for k in range(4):
        d.base_params[k] = float(bp[k])
    return d


def make_epilogue(norm_x=None, norm_x2=None, level_scale=None, sym_jitter=0.0, return_levels=True, upper_only=False):
    e = Epilogue()
    e.norm_x = None if norm_x is None else norm_x.data_ptr()
    e.norm_x2 = None if norm_x2 is None else norm_x2.data_ptr()
    e.level_scale = None if level_scale is None else level_scale.data_ptr()
    e.sym_jitter = float(sym_jitter)
    e.return_levels = int(bool(return_levels))
    e.upper_only = int(bool(upper_only))
    return e
