"""CPU oracle for the MAT-AS hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``mat_dcml_b200/`` may import this
package; only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` do, and there only as the checker or
as the timed CPU restatement, never as the product.

The reference (brithbutter/mat-dcml) is pure Python, so the oracle is Python
too (torch-CPU for the fp32 model arithmetic, numpy/float64 loops for the DCML
simulator).  Parity status: PINNED -- every module here is checked in
``tests/test_oracle_golden.py`` against fixtures under ``tests/golden/`` that
``tests/golden/make_golden.py`` produced by running the unmodified reference
from ``/root/reference`` in the build container (the reference ships no golden
vectors or tests of its own).
"""
