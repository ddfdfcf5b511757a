"""CPU oracle for the sparse-linear hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this package.  See oracle/oracle.h for the parity status.
"""
from .oracle import *  # noqa: F401,F403
