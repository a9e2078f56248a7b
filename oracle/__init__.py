"""CPU oracle for the signature-kernel Gram hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline
legs may import it.  The product (``gpsig_b200``) never does, and fails
loudly when its CUDA library is missing.

Contents
--------
``trunc_np``    numpy restatement of the reference's TensorFlow graphs
                (``gpsig/signature_algs.py`` + the ``gpsig/kernels.py`` glue).
``pde_np``      numpy restatement of ``gpsig/sigKer_fast.pyx`` generalised to
                rectangular (x_i, y_j) pairs.
``sigs_np``     brute-force truncated signatures (Chen's identity): the
                independent known-answer used by the reference's notebooks
                (``notebooks/signature_kernel.ipynb`` cells 11-29, via esig).
``csrc/``       plain-C restatement of the same algorithms (OpenMP over
                pairs) used as the timed CPU baseline ("port").
``_ref/``       build outputs of the reference's own ``sigKer_fast.pyx``
                compiled from ``/root/reference`` (git-ignored).

Parity pinning status
---------------------
* PDE path: PINNED against the reference's own Cython solver run in the build
  container (``tests/golden/pde_ref_*.npz`` were produced by
  ``oracle/_ref/sigKer_fast`` -- see ``tests/golden/make_golden.py``).
* Truncated path: PINNED against the reference's own Python code.  TensorFlow 1.15 /
  GPflow 1.5.1 cannot be installed here (Python 3.12, no network), so
  ``oracle/refrun`` executes ``gpsig/kernels.py`` + ``signature_algs.py`` +
  ``lags.py`` where they lie under ``/root/reference`` on minimal EAGER numpy
  stand-ins for the two dependencies (the ~30 TF ops the path calls, with TF's
  semantics; ``Parameter`` / decorators / ``Kernel`` of GPflow); the outputs are the
  committed fixtures ``tests/golden/trunc_ref.npz`` (``make_golden.py --trunc-ref``).
  The numpy restatement and the C port reproduce them to 1e-13, the CUDA path to
  1e-10 / 1e-5.  Independent of that: the brute-force definition of the first-order
  kernel and the signature known answers of the reference notebooks.
"""
