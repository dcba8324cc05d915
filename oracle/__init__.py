"""CPU oracle for the ACT DETRVAE hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs may import it, and there only as
the checker / the timed CPU baseline, never as the path that is shipped.

Parity status: the reference (jie0530/act-plus-plus) ships no tests, golden
vectors or fixtures for this path, so the oracle is
pinned against *outputs of the reference itself run in the build container*:
``oracle/make_golden.py`` imports the unmodified reference modules from
``/root/reference`` on CPU, loads the oracle's seeded state_dict into them,
asserts that this restatement reproduces the reference's outputs, and writes
the small golden vectors under ``tests/golden/``.
"""
