"""Entry module of a rank process (kept separate from ``mpi_worker`` so ``python -m`` does not
re-execute a module the package already imported)."""
import sys

from raydp_b200.mpi.mpi_worker import main

if __name__ == "__main__":
    sys.exit(main())
