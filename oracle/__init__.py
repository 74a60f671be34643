"""CPU oracle for the AWD-LSTM hot path.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import anything from this
package.  The product path (``pytorch-zappa-serverless_b200/``) never does: it
fails loudly when the CUDA library is missing.

Parity pin: the reference ships no tests or golden vectors, so
the oracle is pinned against outputs of the reference module itself, run in
the build container by ``oracle/make_golden.py`` (which imports
``/root/reference/pytorch_models/awd_lstm.py`` and the ``sample_model`` loop
of ``/root/reference/main.py``) and committed under ``tests/golden/``.
``tests/test_oracle_golden.py`` checks the restatement against those files.

Third-party arithmetic: every number on this path is computed by PyTorch
(pinned ``torch-nightly==1.0.0.dev20181105``, /root/reference/requirements.txt:38-39),
which is not vendored in the reference.  ``awd_numpy`` restates the published
definitions of ``F.embedding``, ``nn.LSTM`` (gate order i,f,g,o), ``nn.Linear``
at the reference's call sites; ``awd_torch_cpu`` restates the same call
sequence with torch's CPU operators and is what the CPU baseline times.
"""
