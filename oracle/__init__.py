"""CPU oracle for the NeRF hot path -- TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it, and there only as the checker or the
timed CPU baseline.  The product path (``nerf_b200``) never imports this
package and fails loudly when its CUDA library is missing.

Parity status
-------------
* ``oracle.head`` (HEAD ``nerf/model.py``, the only generation whose source is
  in the reference tree) is pinned: ``tests/golden/make_golden.py`` imports the
  real reference from ``/root/reference`` and stores its outputs; the
  restatement is checked against those vectors.
* ``oracle.g1`` (the "v1.2" generation that ``examples/nerf.pth`` and the
  example notebook belong to) has NO source in the reference tree.  It is a
  restatement from evidence pinned by (a) the
  checkpoint's key set / shapes (strict ``load_state_dict``), (b) the notebook's
  embedded "Iteration: 40000" render (>= 32 dB) and (c) the shared HEAD
  functions (ray generation, world transform, compositing, chunking) which ARE
  pinned by the real reference.  Bit-level details of G1 beyond that are
  "parity unpinned" (SURVEY.md A.4).
"""
