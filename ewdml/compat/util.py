"""``from util import build_model, prepare_data`` (src/util.py)."""
from ewdml.config import TrainConfig
from ewdml.data import prepare_data as _prepare
from ewdml.models import build_model  # noqa: F401


def prepare_data(args, shard_rank: int = 0, num_shards: int = 1):
    """``args`` is the argparse namespace of distributed_nn.py (needs .dataset, .batch_size, .test_batch_size)."""
    cfg = TrainConfig(dataset=args.dataset, batch_size=args.batch_size, test_batch_size=getattr(args, "test_batch_size", 500),
                      data=getattr(args, "data", "auto"), data_root=getattr(args, "data_root", "."),
                      synthetic_size=getattr(args, "synthetic_size", 4096))
    return _prepare(cfg, shard_rank=shard_rank, num_shards=num_shards)
