"""CPU oracle for the RVQ path -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package.  The product (audiocraft_b200/) must never do so.
"""
from .rvq_oracle import (  # noqa: F401
    build, cnorm, encode, decode, forward, dot, set_threads, get_threads, LIB_PATH,
)
