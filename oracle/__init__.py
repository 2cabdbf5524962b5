"""CPU oracle for the Silent-Speech clip-classification hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package
(``silent_speech_b200``) imports, calls or links anything from here.  The only
legitimate importers are ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` (as the checker or
as the CPU arm being timed -- never as the thing that is shipped).

Parity status: **pinned by execution, not by reference tests.**  The reference
(davdwan21/Silent-Speech) holds no tests, golden vectors or fixtures for this
path, so the restatement here is pinned against outputs
of the reference's own code, imported from ``/root/reference`` in the build
container by ``oracle/make_golden.py`` and committed under ``tests/golden/``
(the reference cannot travel to the GPU box).

Modules
-------
ref_features   numpy restatement of ``extract_feature`` / ``mouth_width_px``
               (live_infer_official.py:141-169, record_landmarks_official.py:52-100)
ref_model      torch-CPU restatement of ``BiGRUClassifier.forward`` and the
               train step using the same ATen call sites as the reference
               (train_model_official.py:209-310, 402-405, 433-439)
ref_explicit   float64 explicit-math restatement (no nn.GRU / conv modules):
               the "truth" used to bound fp32 noise
ref_roi        numpy/cv2-free restatement of ``crop_roi_gray`` / ``crop_roi``
               (live_infer_official.py:171-187, record_landmarks_official.py:102-118)
"""
