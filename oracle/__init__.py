"""CPU oracle for the mantis hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package.  See oracle/mantis_oracle.h for the parity pin.
"""
