"""CPU oracle for the QP-assembly hot path -- TEST INFRASTRUCTURE ONLY.

Nothing in the product package (``mpc_interface_b200``) may import this.
Only ``tests/``, ``__graft_entry__.smoke()`` and the CPU-baseline legs of
``bench.py`` use it, and only as the checker / the timed CPU baseline.

Parity status: PINNED.  ``oracle.qp_oracle`` is checked (tests/test_oracle.py)
against (a) the reference's own golden pickle ``python/tests/LIP_matrices``
(a copy of its arrays is committed as ``tests/golden/lip_matrices.npz``),
(b) the identities / shapes the reference's unit tests assert, and (c) golden
vectors produced by importing the unmodified reference in the build container
(``tests/golden/make_goldens.py`` -> ``tests/golden/*.npz``).
"""
