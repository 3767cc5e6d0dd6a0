"""CPU oracle for the CoLoC front-end hot path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  Nothing under coloc_b200/ does (tests/test_boundary.py greps for it).
"""
from .oracle import *  # noqa: F401,F403
