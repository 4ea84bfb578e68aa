"""CPU oracle for the PointDSC spatial-consistency hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``pointdsc_b200/`` may import this
package.  Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline``
/ ``--impl reference`` legs of ``bench.py`` use it, and there only as the
checker or the timed CPU baseline, never as the thing shipped.

Parity pin: the restatement in ``oracle/restate.py`` is validated against the
reference's own code (``/root/reference/models/PointDSC.py`` imported read-only
by ``oracle/ref_harness.py``) in the build container, and the outputs of that
true reference are committed as golden vectors under ``tests/golden/`` by
``oracle/gen_golden.py``.  The reference has no tests or golden vectors of its
own, so executing it is the pin.
"""
