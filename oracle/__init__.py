"""CPU oracle for the EgoScaler trajectory-generation hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is product code: only
``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl
reference`` legs of ``bench.py`` may import it, and there only as the checker
or the timed CPU baseline -- never as the thing shipped.  The product package
(``egoscaler_b200``) must not import from here.

Contents
--------
``traj_vlm.py``        plain-PyTorch fp32 restatement of the reference model
                       (PointBERT encoder -> projector -> splice -> LLaMA ->
                       lm_head, loss slice, manual KV-cache greedy decode).
``traj_tokenizer.py``  NumPy restatement of the trajectory (de)quantiser and
                       the de-normalising scalers.
``ref_shim.py``        imports the *real* reference from ``/root/reference``
                       (dev container only) -- used by ``make_golden.py``.
``make_golden.py``     generates ``tests/golden/*.npz`` from the real reference.

Parity pinning: the reference ships no tests/golden vectors for this path
(SURVEY.md section 8c), so the restatement is pinned against outputs of the
imported reference itself, generated here by ``make_golden.py`` and committed
under ``tests/golden/`` (torch/transformers versions recorded in the files).
"""
