"""kwargs dictionaries of the reference's entry point (distributed_nn.py:89-121) -> TrainConfig."""
from ewdml.config import TrainConfig

_MAP = {"batch_size": "batch_size", "learning_rate": "lr", "max_epochs": "epochs", "momentum": "momentum", "network": "network",
        "comm_method": "comm_type", "eval_freq": "eval_freq", "train_dir": "train_dir", "max_steps": "max_steps",
        "compress_grad": "compress_grad", "gather_type": "gather_type"}


def config_from_kwargs(kw: dict, **extra) -> TrainConfig:
    vals = {dst: kw[src] for src, dst in _MAP.items() if src in kw}
    net = vals.get("network", "LeNet")
    vals.setdefault("dataset", "MNIST" if net in ("LeNet", "MnistNet") else "Cifar10")
    # the reference wires kill_threshold <- num_aggregate on the server (distributed_nn.py:97): not replicated
    if "timeout_threshold" in kw:
        vals["kill_threshold"] = float(kw["timeout_threshold"])
    elif "kill_threshold" in kw and "rank" in kw:
        vals["kill_threshold"] = float(kw["kill_threshold"])
    vals.update(extra)
    return TrainConfig(**vals).validate()
