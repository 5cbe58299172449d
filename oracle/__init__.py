"""oracle/ -- TEST INFRASTRUCTURE, NOT PRODUCT.

A CPU (torch-CPU / numpy) restatement of the reference's batched speech-generation hot path
(coezbek/Zonos: zonos/model.py generate, zonos/backbone/_torch.py, zonos/sampling.py,
zonos/codebook_pattern.py, zonos/autoencoder.py decode -> transformers DacModel.decode).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package, and only as the checker or the timed CPU baseline.  zonos_b200/ never imports it; the
product path fails loudly if the CUDA library is missing.

Parity pin: every function here is checked against golden vectors produced by running the *unmodified*
reference on CPU in the build container (oracle/gen_golden.py -> tests/golden/*.pt, checked by
tests/test_oracle_golden.py), and -- when /root/reference is present -- against the live reference
(tests/test_oracle_vs_reference.py).  The reference itself ships no tests/golden vectors.
The hybrid (mamba_ssm) backbone is NOT covered: parity unpinned for that row.
"""
