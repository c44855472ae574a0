"""
oracle/ -- CPU restatement of graphistician's Gibbs hot path.  TEST INFRASTRUCTURE ONLY.

This package is the checker, never the product: only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import it.  Nothing under
``graphistician_b200/`` imports or links it; the product path fails loudly when the CUDA
library is missing.

Every function cites the reference file:line it restates (paths relative to the upstream
repository root, i.e. ``/root/reference`` in the build container).

Pinning status
--------------
The reference is Python 2 and depends on un-vendored packages (pybasicbayes, pypolyagamma,
hips, autograd), so it cannot be imported as a whole.  What *is* pinned:

* ``internals/utils.py`` and ``internals/distributions.py`` parse under Python 3 and are
  imported unmodified (3-alias NumPy shim) by ``tests/golden/make_golden.py``; the truncated
  normal / Beta / Bernoulli / Gaussian-negentropy values they produce are committed in
  ``tests/golden/ref_internals.npz`` and the oracle is checked against them.
* The known-answer values of ``test/test_distributions.py:14,28,32,36`` (the reference's only
  asserts) are reproduced in ``tests/test_oracle_golden.py``.
* For the SBM / eigenmodel / latent-distance sweeps the reference has no golden vector and cannot
  run: **parity unpinned by the reference**.  There the oracle is held to (i) literal loop
  transcriptions (T0, ``*_loops`` functions) cross-checked against the vectorised forms (T1), and
  (ii) golden vectors generated from T0 and committed under ``tests/golden/``.
* Third-party semantics (pybasicbayes NIW draw, ``sample_discrete_from_log``, hips HMC,
  pypolyagamma's Devroye sampler) are restated from their published algorithms; flagged
  "recollection" in each docstring.
"""
