"""CPU oracle for the fused rollout path -- TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is product code.  Only ``tests/``, ``__graft_entry__.smoke()``
and the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it, and there
only as the checker (or as the timed CPU baseline), never as the thing shipped.

Parity status: PINNED against the reference itself.  ``oracle/pin_reference.py`` imports the
unmodified reference modules from ``/root/reference`` (scratch-layout recipe),
checks that every function restated in ``rollout_ref.py`` is bit-identical to the reference
module on the shipped checkpoints and validation pickles, and writes the golden vectors under
``tests/golden/``.  The reference holds no tests or golden vectors of its own, and
its multi-step loop exists only for the missing CasADi model, so the closed-loop semantics are
the ones defined in DESIGN.md ("Rollout semantics"), driven through the reference's own modules.
"""
