"""Test oracle for the DIS hot path.  TEST INFRASTRUCTURE ONLY — see oracle/dis_oracle.c.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package.  The product package never does.
"""
from .binding import Oracle, RefViz, build, load_port, load_ref, load_refviz  # noqa: F401
