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
* Truncated path: the reference executes inside TensorFlow 1.15 / GPflow 1.5
  which cannot be installed here (Python 3.12, no network) and the reference
  holds no golden vectors or tests.  It is pinned by the reference notebooks'
  known-answer properties (signature inner products) instead; in the
  golden-vector sense it is "parity unpinned".
"""
