"""Bench / test kit: the fixtures (tests/golden), the synthetic simulator state and the Workload object that bench.py, the
tests, __graft_entry__.smoke() and the tools share.  Not part of the product package (pbhc_b200/): nothing in there imports it,
except the optional build-time warm-up of the specialised kernels for the configs the repo ships fixtures for."""
