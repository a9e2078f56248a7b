"""Builds ``gpsig_b200/libgpsig_b200.so`` (sm_100a only) with nvcc, in-tree.

    python gpsig_b200/build.py [--force] [--fast]     (run as a script: importing the package needs the built library)

The library has no PyTorch dependency: it is a plain C-ABI shared object
(``include/gpsig_b200.h``).  One object file per (dtype, base-kernel
specialisation, padded state dimension) so the template instantiations
compile in parallel.  ``--fast`` / ``GPSIG_FAST_BUILD=1`` builds only the
shapes of the BASELINE configs (developer iteration).
"""
import hashlib
import os
import re
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, 'csrc')
LIB = os.path.join(HERE, os.environ.get('GPSIG_LIB_NAME', 'libgpsig_b200.so'))   # variants: developer experiments only
EXTRA = os.environ.get('GPSIG_EXTRA_DEFS', '').split()
OBJ = os.path.join(HERE, 'build', os.path.basename(LIB))
NVCC = os.environ.get('NVCC', '/usr/local/cuda/bin/nvcc')
ARCH = ['-gencode', 'arch=compute_100a,code=sm_100a']
FLAGS = ARCH + ['-O3', '-std=c++17', '-lineinfo', '-Xcompiler', '-fPIC', 
                '--expt-relaxed-constexpr']


def _deps():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(('.cuh', '.cu'))) + \
        [os.path.join(HERE, '..', 'include', 'gpsig_b200.h')]


def _stamp(extra):
    h = hashlib.sha1()
    for f in _deps():
        h.update(open(f, 'rb').read())
    h.update(' '.join(FLAGS + extra + EXTRA + [LIB]).encode())
    return h.hexdigest()


_INC = re.compile(r'^\s*#\s*include\s+"([^"]+)"', re.M)


def _closure(src, seen=None):
    """Transitive closure of the quoted includes of one translation unit."""
    seen = set() if seen is None else seen
    src = os.path.normpath(src)
    if src in seen or not os.path.exists(src):
        return seen
    seen.add(src)
    for inc in _INC.findall(open(src).read()):
        _closure(os.path.join(os.path.dirname(src), inc), seen)
    return seen


def _unit_stamp(u):
    name, src, extra = u
    h = hashlib.sha1()
    for f in sorted(_closure(src)):
        h.update(open(f, 'rb').read())
    h.update(' '.join(FLAGS + extra + [name]).encode())
    return h.hexdigest()


def _units(fast):
    extra = (['-DGPSIG_FAST_BUILD'] if fast else []) + EXTRA
    units = [('abi', os.path.join(CSRC, 'abi.cu'), extra),
             ('pde_vjp', os.path.join(CSRC, 'pde_vjp.cu'), extra),
             ('static_scratch', os.path.join(CSRC, 'static_scratch.cu'), extra),
             ('lifted', os.path.join(CSRC, 'lifted.cu'), extra),
             ('tens_tens', os.path.join(CSRC, 'tens_tens.cu'), extra)]
    for t in ('float', 'double'):
        for op in (0, 1, 2):
            for dp in (4, 8, 16):
                units.append(('trunc_%s_%d_%d' % (t, op, dp), os.path.join(CSRC, 'trunc_inst.cu'),
                              extra + ['-DGPSIG_T=%s' % t, '-DGPSIG_OP=%d' % op, '-DGPSIG_DP=%d' % dp]))
    return units


def build(force=False, fast=None, verbose=False):
    if fast is None:
        fast = os.environ.get('GPSIG_FAST_BUILD', '0') == '1'
    os.makedirs(OBJ, exist_ok=True)
    stamp_file = os.path.join(OBJ, 'stamp')
    stamp = _stamp(['fast' if fast else 'full'])
    if not force and os.path.exists(LIB) and os.path.exists(stamp_file) and open(stamp_file).read() == stamp:
        return LIB
    units = _units(fast)

    def compile_one(u):
        name, src, extra = u
        obj = os.path.join(OBJ, name + '.o')
        ustamp = _unit_stamp(u)
        if not force and os.path.exists(obj) and os.path.exists(obj + '.stamp') and open(obj + '.stamp').read() == ustamp:
            return obj
        cmd = [NVCC] + FLAGS + extra + ['-I', CSRC, '-c', src, '-o', obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError('nvcc failed for %s:\n%s\n%s' % (name, ' '.join(cmd), r.stderr))
        if verbose and r.stderr.strip():
            print(r.stderr)
        open(obj + '.stamp', 'w').write(ustamp)
        return obj

    with ThreadPoolExecutor(max_workers=os.cpu_count() or 4) as ex:
        objs = list(ex.map(compile_one, units))
    cmd = [NVCC] + ARCH + ['-shared', '-o', LIB] + objs + ['-lcudart_static', '-lpthread', '-ldl', '-lrt']
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError('link failed:\n%s' % r.stderr)
    open(stamp_file, 'w').write(stamp)
    return LIB


if __name__ == '__main__':
    print(build(force='--force' in sys.argv, fast=True if '--fast' in sys.argv else None, verbose='-v' in sys.argv))
