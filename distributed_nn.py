"""Thin shim so ``python -m torch.distributed.run ... distributed_nn.py <flags>`` works from the
repo root exactly like the reference's ``src/distributed_nn.py``."""
from ewdml.cli.distributed_nn import main

if __name__ == "__main__":
    main()
