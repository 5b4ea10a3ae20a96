"""CPU oracle of the ihm2 NMPC hot path -- TEST INFRASTRUCTURE ONLY (parity unpinned).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  See oracle/ihm2_oracle.cpp for what it restates.
"""
from .lib import Oracle, build_oracle  # noqa: F401
