"""CPU oracle for the VQ-VAE hot path of dendisuhubdy/neural_sound_generation.

TEST INFRASTRUCTURE ONLY.  Nothing under ``neural_sound_generation_b200/`` may
import this package: it is the checker, never the thing measured or shipped.
Allowed importers: ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs.

Parity status: the reference ships NO tests, golden vectors or known-answer
fixtures for this path (``/root/reference/test/test_{models,train,...}.py`` are
zero-byte files).  The oracle is therefore pinned the other
way the task allows: against outputs of the reference itself, generated in the
authoring container by importing ``/root/reference/src/{vector_quantization,
models,train}.py`` unmodified (``oracle/gen_golden.py``) and committed as
fixtures under ``tests/golden/``.  ``tests/test_oracle_golden.py`` checks every
function here against those fixtures.  The EMA codebook update is an extension
that does not exist in the reference: *parity unpinned* (see ``ema_update``).
"""
