"""CPU oracle for the hot path. TEST INFRASTRUCTURE ONLY: imported by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, never by the
product package. Parity status and pinning: see oracle/ref_numpy.py."""
