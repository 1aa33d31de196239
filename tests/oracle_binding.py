"""Kept for the tests' imports: the binding lives with the oracle (oracle/binding.py)."""
from oracle.binding import *  # noqa: F401,F403
from oracle.binding import load, load_ref, build_oracle, build_ref, MATH_CANONICAL, MATH_LIBM  # noqa: F401
