"""TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference's inference hot path.

Nothing under ``oracle/`` is product code: only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it.
"""
