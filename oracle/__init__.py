"""CPU oracle for the RGP hot path — TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is product code.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it, and there only as the checker or the
timed CPU baseline.  The product path (``recurrent_gaze_prediction_b200``)
never imports this package and fails loudly when its CUDA library is missing.

Parity status: the reference ships no tests, golden vectors or fixtures and
TensorFlow 1.x cannot be installed here, so the oracle cannot be checked
against a *running* TF graph.  It is pinned two ways instead:

* ``oracle/metrics_oracle.py`` is checked against the reference's own
  ``evaluation_metrics.py`` imported unmodified in the build container
  (``oracle/gen_golden_metrics.py`` -> ``tests/golden/metrics_*.npz``).
* ``oracle/model_oracle.py`` is checked against the reference's own graph
  builders (``models/gaze_grcn.py`` etc., imported unmodified) executed over an
  eager NumPy stand-in for the TF1 ops (``oracle/tf1_eager_shim.py``;
  ``oracle/gen_golden_model.py`` -> ``tests/golden/model_*.npz``).  The wiring
  is therefore the reference's; the semantics of each TF op are restated from
  the TF1 documentation -> "parity unpinned at the TF-library boundary".
"""
