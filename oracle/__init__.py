"""TEST INFRASTRUCTURE ONLY -- CPU oracle of the Filter scoring pass.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this package.  PARITY UNPINNED (see oracle/score_ref.h).
score_ref.c holds R1 / R2 (C); score_py.py holds R0, a third restatement in plain Python for small cases.
"""
from .ref import build, score_faithful, score_packed, to_upper, typemask  # noqa: F401
