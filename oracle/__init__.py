"""CPU oracle for the quantized-linear hot path.  TEST INFRASTRUCTURE ONLY.

This package restates, on the CPU with plain torch ops, the *torch reference
implementations* that the reference repository's own test-suite uses to judge
its SYCL kernels (``/root/reference/tests/*.py``).  Each function cites the
reference file:line it follows.

Rules (enforced by tests/test_no_oracle_in_product.py):
  * only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
    ``--impl reference`` legs of ``bench.py`` may import this package;
  * nothing under ``vllm_xpu_kernels_b200/`` may import it -- the product path
    is the CUDA extension and fails loudly when that is missing.

Parity status: PINNED.  ``tests/golden/make_golden.py`` imports the
reference's own Python reference functions from ``/root/reference/tests`` (with
the XPU ``.so`` imports stubbed out and ``device="xpu"`` redirected to the
CPU), runs them on seeded inputs and commits input/output pairs under
``tests/golden/``.  ``tests/test_oracle_golden.py`` checks every oracle
function bit-for-bit against those fixtures.

The reference has no CPU *kernel* implementation of these ops (its device code
is SYCL + oneDNN, neither buildable here), so there is no ``oracle/_ref``.
"""

from . import formats, quant, gemm, cache  # noqa: F401
