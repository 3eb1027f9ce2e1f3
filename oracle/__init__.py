"""TEST INFRASTRUCTURE ONLY.

CPU restatement of mjrl's policy-update hot path, used as the parity checker.
Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline /
`--impl reference` legs may import this package.  Nothing under `mjrl_b200/`
imports it; the product path fails loudly when the CUDA library is missing.
"""
