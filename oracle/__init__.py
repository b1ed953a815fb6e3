"""oracle/ -- TEST INFRASTRUCTURE ONLY.

CPU restatements of the reference's hot path (log-mel extraction, pathway packing, SlowFast
ResNet-50 forward/backward) used as the parity checker.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may
import anything from here; the product package (``asf_b200``) never does and fails loudly when
its CUDA extension is missing.
"""
