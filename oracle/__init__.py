"""Parity checkers for flashgmm_b200 -- TEST INFRASTRUCTURE ONLY.

Nothing under ``flashgmm_b200/`` may import this package. Allowed importers:
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs.

* :mod:`oracle.port` -- ctypes binding of ``oracle/gmm_oracle.c`` (the CPU
  restatement, "port").
* :mod:`oracle.ref`  -- loader / worker-process client for the reference's own
  compiled extension in ``oracle/_ref`` ("reference").
"""
