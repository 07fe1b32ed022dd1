"""CPU oracle for the invert hot path -- TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs
of ``bench.py`` may import it, and only as the checker or the timed CPU baseline.
The product (``ska-sdp-continuum-imaging-pipeline_b200/``) never imports it and
fails loudly when its CUDA library is missing.

Pinning status
--------------
* ``oracle.tiling``  (T1-T5, reference ``src/ska_sdp_cip/uvw_tiling/tiling_plan.py:29-181``,
  ``tile.py:83-115``): PINNED.  It is checked against fixtures under
  ``tests/golden/`` that were produced by importing the reference's own
  ``tiling_plan.py`` / ``tile.py`` in the build container
  (``tools/make_golden.py``), bit for bit.
* ``oracle.stokes`` (P1/P2/P4/P5, reference ``src/ska_sdp_cip/invert.py:72-116,149,200-209``)
  and ``oracle.partition`` (P7, ``measurement_set.py:361-391``): PINNED the same
  way (reference functions imported with stubbed third-party modules) and by the
  reference's golden partition bounds
  (``tests/test_measurement_set_partition_indices.py:33-62``).
* ``oracle.dft`` / ``oracle.wstack`` (G0-G4): **PARITY UNPINNED** against the
  reference's numerics.  The arithmetic of the reference lives in the third
  party wheel ``ducc0 == 0.34.0`` (``poetry.lock:421-422``), reached through one
  call site (``src/ska_sdp_cip/invert.py:170-183``); its source is not under
  ``/root/reference``, it is not installed and cannot be fetched.  The reference
  tests pin no dirty-image values (``tests/test_invert_measurement_set.py:16-17``
  checks type and shape only).  The oracle therefore restates the measurement
  equation ms2dirty is documented to approximate (direct DFT, float64) and the
  published improved-w-stacking algorithm (Arras et al. 2021, Ye et al. 2022) and
  anchors parity on (i) the DFT to the requested epsilon and (ii) the one
  property the reference does test, chunk invariance to 1e-5
  (``tests/test_dask_invert_measurement_set.py:31-34``).
"""
