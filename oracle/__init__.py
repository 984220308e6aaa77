"""CPU oracle for the planar-flow free-energy hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it, and there only as the checker or as
the timed CPU baseline -- never as the thing shipped.  The product path
(``vi-normflows_b200/``) raises when ``libnfcore.so`` or a CUDA device is
missing; it never falls back to this code.

Contents
--------
``planar_np``     NumPy (float64) restatement of the reference *forward* math,
                  one function per reference function, each citing the
                  reference file:line it follows.
``planar_torch``  the same op sequence in torch float64 so that
                  ``torch.autograd`` can stand in for HIPS ``autograd`` 1.3
                  (environment.yml:9), which is absent from /root/reference
                  and cannot be installed here.
``refshim``       import shim that lets the *unmodified* reference modules be
                  imported in the build container (``/root/reference`` does
                  not exist on the GPU box).  Used only by ``make_golden.py``
                  and by CPU tests that skip when the reference is absent.
``make_golden``   generates ``tests/golden/*.npz`` by running the reference
                  itself; the script and its outputs are committed.

Parity pinning status
---------------------
* forward values: PINNED -- every restated function is checked against the
  reference's own code run in this container (tests/test_oracle_vs_reference.py)
  and against committed golden vectors generated from it (tests/golden/).
  The only fixture the reference itself ships for this path
  (normflows/normflows/test_gmm.py:6-11 inputs) is golden file ``gmm_kat.npz``.
* gradients: the reference obtains them from HIPS autograd 1.3, which is not
  present; no reference test pins any gradient.  The gradient oracle is torch
  float64 autograd over the op-for-op restatement, cross-checked against
  central differences of the *reference's own forward* (tests/golden/ holds
  those finite-difference vectors too).  Read: "gradient parity pinned to the
  reference forward by finite differences, not to autograd 1.3 itself".
"""
