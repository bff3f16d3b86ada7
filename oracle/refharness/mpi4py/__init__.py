"""Thread-backed stand-in for mpi4py (test infrastructure; see MPI.py)."""
from . import MPI  # noqa: F401
