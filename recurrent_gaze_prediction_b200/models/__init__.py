"""Host-side mirror of the reference's ``models/`` package for the recurrent gaze-prediction path: the same class /
method / flag names (``models/base.py``, ``gaze_rnn.py``, ``gaze_grcn.py``, ``gaze_grcn77.py``, ``gaze_lstm.py``,
``train_gaze.py``, ``evaluate_gaze.py``), with every numeric site routed to ``librgp_sm100.so``."""
