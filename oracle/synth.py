"""Kept for the tests' imports: the synthetic-sequence generator lives in /synthdata.py (it produces INPUT data for
bench.py, the tools and the tests; it is not part of the oracle)."""
from synthdata import *  # noqa: F401,F403
from synthdata import CALIB, Sequence, make_sequence, projection_matrices  # noqa: F401
