"""CPU oracle for the quantization hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is part of the product: only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl
reference`` legs may import it.  The product path (``quantized.pytorch_b200``)
never does and raises if the CUDA library is missing.

Parity pin: the reference ships no tests or golden vectors of its own
(SURVEY.md section 4), so the oracle is pinned against outputs of the unmodified
reference executed in the build container (``oracle/gen_golden.py`` ->
``tests/golden/*.npz``).
"""
