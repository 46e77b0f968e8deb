"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the AutoBub3hs per-event image hot path.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs may import it, and only as the checker / the timed CPU arm.

Two restatements live here:

* ``oracle.cv2_oracle``  -- function-by-function restatement of the reference's
  hot path that calls the *same* OpenCV entry points through ``cv2`` (the
  reference delegates almost all arithmetic to OpenCV, which is not vendored in
  /root/reference and is unpinned there; the pin chosen here is cv2 4.13.0).
* ``oracle/abub_ref.c``   -- a cv2-free plain-C restatement of the integer math
  (ProcessFrame, histograms, significance trigger, Welford, box blur, Otsu,
  external-contour tracing + TC89_L1, polygon moments), itself checked against
  the cv2 oracle in ``tests/``; loaded through ``oracle.cref`` (ctypes).

PARITY PIN: the reference has no tests, golden vectors or fixtures for this path
 and cannot be compiled here (needs OpenCV C++ + Boost
headers).  The pin is therefore "cv2 4.13.0 driven exactly as the reference
drives OpenCV"; golden vectors under tests/golden/ are generated from
``oracle.cv2_oracle`` by ``tests/golden/make_golden.py``.  "parity unpinned"
with respect to reference-owned vectors.
"""
