"""Shim for the reference's ``src/distributed_evaluator.py`` command line."""
from ewdml.cli.distributed_evaluator import main

if __name__ == "__main__":
    main()
