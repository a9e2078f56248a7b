"""Build the reference's own Cython PDE solver, and its own CUDA PDE kernel, into ``oracle/_ref/``.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Compiles ``/root/reference/gpsig/sigKer_fast.pyx`` *where it lies* (no source
is copied into this repository; the generated C file and the extension module
land in the git-ignored ``oracle/_ref/``).  Recipe, recorded in SURVEY.md 8c:
Cython 3 needs the ``cpow=True`` directive (otherwise ``2**n`` is typed double
and ``pyx:23-24`` is rejected), ``language_level=2``, and the system gcc for
OpenMP (the default gcc wrapper cannot find ``libgomp.spec``).

    python oracle/build_ref.py          # no-op when the reference is absent

``build_cuda()`` does the same for the reference's CUDA op kernel
(``gpsig/covariance_op/untrunc_cov_op_gpu.cu``: plain CUDA, no TensorFlow headers):
``nvcc -arch=sm_100a`` on the file where it lies plus the C-ABI doorway
``oracle/csrc/untrunc_cov_ref_shim.cu`` -> ``oracle/_ref/libuntrunc_cov_ref.so``.  It is the bar
SURVEY.md sets for the PDE kernel ("this kernel re-compiled on the same box"); ``bench.py``
times it when the file is there (``other_configs.c3_vs_reference_cuda``).

The GPU box has no ``/root/reference``; it uses the prebuilt module that
travels with the repository snapshot.
"""
import glob
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, '_ref')
PYX = '/root/reference/gpsig/sigKer_fast.pyx'


def built_module():
    hits = glob.glob(os.path.join(OUT, 'sigKer_fast*.so'))
    return hits[0] if hits else None


def build(force=False):
    if built_module() and not force:
        return built_module()
    if not os.path.exists(PYX):
        return None
    import numpy as np
    os.makedirs(OUT, exist_ok=True)
    c_file = os.path.join(OUT, 'sigKer_fast.c')
    subprocess.check_call([sys.executable, '-m', 'cython', '-2', '-X', 'cpow=True',
                           '-o', c_file, PYX])
    ext = sysconfig.get_config_var('EXT_SUFFIX')
    so = os.path.join(OUT, 'sigKer_fast' + ext)
    gcc = '/usr/bin/gcc' if os.path.exists('/usr/bin/gcc') else 'gcc'
    subprocess.check_call([gcc, '-shared', '-fPIC', '-O2', '-fopenmp', '-fwrapv',
                           '-I', sysconfig.get_paths()['include'], '-I', np.get_include(),
                           c_file, '-o', so])
    os.remove(c_file)   # generated from the reference source: keep only the binary
    return so


CU = '/root/reference/gpsig/covariance_op/untrunc_cov_op_gpu.cu'
CU_SO = os.path.join(OUT, 'libuntrunc_cov_ref.so')


def build_cuda(force=False):
    """The reference's UntruncCovKernel compiled for sm_100a (None when the reference or nvcc is absent)."""
    if os.path.exists(CU_SO) and not force:
        return CU_SO
    nvcc = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')
    if not os.path.exists(CU) or not os.path.exists(nvcc):
        return None
    os.makedirs(OUT, exist_ok=True)
    subprocess.check_call([nvcc, '-gencode', 'arch=compute_100a,code=sm_100a', '-O3', '-lineinfo', '-shared',
                           '-Xcompiler', '-fPIC', CU, os.path.join(HERE, 'csrc', 'untrunc_cov_ref_shim.cu'),
                           '-o', CU_SO, '-lcudart_static', '-lpthread', '-ldl', '-lrt'])
    return CU_SO


def load():
    """Import the compiled reference module, or return None if it was never built."""
    so = built_module()
    if so is None:
        return None
    import importlib.util
    spec = importlib.util.spec_from_file_location('sigKer_fast', so)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == '__main__':
    print(build(force='--force' in sys.argv))
    print(build_cuda(force='--force' in sys.argv))
