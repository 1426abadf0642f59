"""CPU oracle for the JointEncoding hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import this package, and only as the checker or the timed CPU baseline.  Nothing in
``mneslam_b200/`` imports it; the product path fails loudly when its CUDA library is missing.

Parity status (see DESIGN.md "Oracle"):
* sampling, tri-plane bilinear encoding, decoder MLPs, sdf2weights compositing and all losses are
  PINNED: ``oracle/make_golden.py`` imports the reference's own ``model/`` package (in the build
  container, from ``/root/reference``) and the fixtures it wrote live in ``tests/golden/``;
  ``tests/test_oracle_golden.py`` checks this restatement against them.
* OneBlob (``oracle/oneblob.py``) and the multi-resolution hash grid (``oracle/hashgrid.py``) restate the
  published algorithm of the un-vendored third-party dependency tiny-cuda-nn; the reference holds no
  test or golden vector for them and tiny-cuda-nn cannot be installed here: PARITY UNPINNED for
  these two functions.
* The on-device batch draw (``oracle/draw.py``) restates this repository's own keyed permutation: the reference draws with
  Python's ``random.sample``, whose stream cannot be reproduced on the device; the property (distinct rows) is what is tested.
"""
