from .mpi_ops import *  # noqa: F401,F403
