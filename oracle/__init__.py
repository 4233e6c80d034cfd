"""TEST INFRASTRUCTURE ONLY — CPU oracle for the VINS-Mono visual front-end.

Nothing under ``oracle/`` is part of the product path.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import it, and only as the checker / timed CPU baseline.

Modules
  spec.py          numpy restatement of the OpenCV arithmetic the reference calls
                   (SURVEY.md §8 a-spec S1..S6), pinned against cv2 4.13.0 in tests.
  cv_reference.py  line-by-line restatement of /root/reference/src/frontend/feature_tracker.cpp
                   driving the same five OpenCV calls through cv2 — "the reference's own
                   OpenCV CPU tracker" named by BASELINE.json north_star.
  camera.py        liftProjective for PINHOLE / KANNALA_BRANDT / MEI / SCARAMUZZA (float64).

Parity status: the reference's own tests hold NO golden vectors for this path
 => "parity unpinned by reference tests"; the oracle is pinned
instead against cv2 4.13.0 (the OpenCV build in this image) on seeded inputs, and the
vectors it generated are committed under tests/golden/ with their generating script.
"""
