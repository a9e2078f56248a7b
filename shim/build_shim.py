"""Builds shim/gpsig_tf_ops.so when TensorFlow is installed; otherwise reports that it is absent (exit 0).

    python shim/build_shim.py            # g++ -shared with tf.sysconfig compile/link flags, linked to libgpsig_b200.so
    python shim/build_shim.py --check    # syntax check only, against the header stubs under tests/tf_stub (what the
                                         # CPU test-suite runs: this image has no TensorFlow)
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SRC = os.path.join(HERE, 'gpsig_tf_ops.cc')


def check():
    cmd = ['g++', '-std=c++17', '-fsyntax-only', '-Wall', '-I', os.path.join(ROOT, 'tests', 'tf_stub'), '-I', os.path.join(ROOT, 'include'), '-I', '/usr/local/cuda/include', SRC]
    return subprocess.run(cmd, capture_output=True, text=True)


def build():
    try:
        import tensorflow as tf
    except ImportError:
        print("TensorFlow is not installed: the shim is not built (nothing else depends on it)")
        return None
    out = os.path.join(HERE, 'gpsig_tf_ops.so')
    cmd = (['g++', '-std=c++17', '-shared', '-fPIC', '-O2', SRC, '-o', out, '-I', os.path.join(ROOT, 'include'), '-I', '/usr/local/cuda/include', '-DGOOGLE_CUDA=1']
           + tf.sysconfig.get_compile_flags() + tf.sysconfig.get_link_flags()
           + ['-L', os.path.join(ROOT, 'gpsig_b200'), '-lgpsig_b200', '-Wl,-rpath,' + os.path.join(ROOT, 'gpsig_b200')])
    subprocess.check_call(cmd)
    return out


if __name__ == '__main__':
    if '--check' in sys.argv:
        r = check()
        print(r.stderr or 'syntax ok')
        sys.exit(r.returncode)
    print(build())
