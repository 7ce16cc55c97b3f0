"""CPU oracle for the CGNN hot path.  TEST INFRASTRUCTURE ONLY.

This package restates, on the CPU (numpy / torch-CPU, fp64 by default), the
arithmetic of the reference's generator + MMD + Adam path:

* ``oracle.mmd``      <- Code/cgnn/utils/Loss.py:10-32
* ``oracle.adam``     <- tf.train.AdamOptimizer (TF 1.x, third-party, un-vendored;
                         call sites Code/cgnn/GNN.py:74-75, CGNN.py:97-99)
* ``oracle.gnn``      <- Code/cgnn/GNN.py:21-166 (pairwise generator, loops)
* ``oracle.cgnn``     <- Code/cgnn/CGNN.py:24-195 and CGNN_confounders.py:26-216
* ``oracle.tf1_shim`` <- a tiny graph-mode TensorFlow-1 API emulator (torch-backed)
                         that lets the reference's OWN, unmodified Python files be
                         executed in this container to pin the restatement.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` legs may import it.  The product (``cgnn_b200``) never does:
it fails loudly when the CUDA library is missing.

Parity pinning status: the reference ships no golden vectors, no tests and no
seeds, and its arithmetic lives in TensorFlow 1.x which is absent here.  The
restatement is pinned as far as this container allows: against the reference's
own Python source executed over ``oracle.tf1_shim`` (fixtures in
``tests/golden/``, generator ``oracle/make_golden.py``).  TensorFlow's own
kernels were never run => "parity unpinned" at the TF-kernel level.
"""
GAMMAS = (0.005, 0.05, 0.25, 0.5, 1.0, 5.0, 50.0)   # Loss.py:10
