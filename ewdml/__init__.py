"""ewdml -- Blackwell-native compressed-gradient data-parallel training.

Capability map to the reference (AnirudhKaushik10/Efficient-Workers-in-Distributed-Machine-Learning):

* ``ewdml.cli.distributed_nn``      <- ``PyTorch-parameter-server/src/distributed_nn.py``
* ``ewdml.cli.distributed_evaluator`` <- ``src/distributed_evaluator.py``
* ``ewdml.cli.allreduce_mnist``     <- ``horvod_pytorch.py`` (Horovod allreduce trainer)
* ``ewdml.compress``                <- ``src/Compresssor/{qsgd,TopK}.py``, ``horovod_compression.py``
* ``ewdml.optim``                   <- ``src/optim/{sgd,adam}.py`` (``step(grads=...)``)
* ``ewdml.models``                  <- ``src/model_ops/{lenet,vgg,resnet}.py``
* ``ewdml.parallel``                <- master/worker roles, collectives (gloo plumbing + native NVLink kernels)
* ``ewdml.runtime``                 <- what the reference delegated to gloo / Horovod / Open MPI
"""
__version__ = "0.1.0"

from .config import TrainConfig, registry  # noqa: F401
