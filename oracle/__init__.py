"""TEST INFRASTRUCTURE ONLY.

CPU restatements of the reference's algorithms for the POD + Galerkin-projection hot path.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package, and only as the checker or as the timed CPU baseline. The product path
(ithaca-fv_b200/) never imports it and fails loudly when its CUDA library is missing.
"""
