"""Typed configuration + environment-override registry.

Replaces the reference's flag plumbing (argparse flags repacked into two kwargs dicts,
``PyTorch-parameter-server/src/distributed_nn.py:24-121``) and the role the MCA variable system
plays in the vendored Open MPI (``openmpi-4.1.0/opal/mca/base/mca_base_var.c``: default < file <
env < command line).  Here: dataclass default < ``EWDML_<NAME>`` environment variable < CLI flag.

Quirks of the reference that are deliberately NOT reproduced:
``kill_threshold`` is not wired to ``--num-aggregate``; ``--enable-gpu`` is a real boolean flag
that also accepts the reference's ``--enable-gpu=<anything>`` spelling.
"""
from __future__ import annotations

import argparse
import dataclasses
import os
from dataclasses import dataclass, field
from typing import Any, Dict, Optional

ENV_PREFIX = "EWDML_"

NETWORKS = ("LeNet", "ResNet18", "ResNet34", "ResNet50", "ResNet101", "ResNet152", "VGG11",
            "VGG13", "VGG16", "VGG19", "MnistNet")
DATASETS = ("MNIST", "Cifar10", "Cifar100", "SVHN")
COMPRESSORS = ("none", "qsgd", "topk", "topk-qsgd")
EXCHANGES = ("ps", "allgather", "allreduce")
# The six "methods" of the report (Report.zip!main.tex:80-119); resolved by TrainConfig.resolved_*() below.
METHODS = (0, 1, 2, 3, 4, 5, 6)


def _str2bool(v: Any) -> bool:
    if isinstance(v, bool):
        return v
    if v is None:
        return True
    s = str(v).strip().lower()
    # The reference declares ``--enable-gpu`` with type=bool, so "" is False and any other
    # string is True (distributed_nn.py:67).  Keep "" -> False, but parse words sanely.
    if s in ("", "0", "false", "no", "off", "none"):
        return False
    return True


@dataclass
class TrainConfig:
    # ---- reference flag surface (distributed_nn.py:30-70), same names and defaults ----
    batch_size: int = 128
    test_batch_size: int = 500
    epochs: int = 100
    max_steps: int = 10000
    lr: float = 0.01
    momentum: float = 0.5
    no_cuda: bool = False
    seed: int = 1
    log_interval: int = 10
    network: str = "LeNet"
    mode: str = "normal"            # normal | kill (straggler cut-off) | async (asynchronous parameter server)
    kill_threshold: float = 7.0
    dataset: str = "MNIST"
    comm_type: str = "Bcast"
    num_aggregate: int = 5
    eval_freq: int = 50
    train_dir: str = "output/models/"
    compress_grad: str = "compress"
    gather_type: str = "gather"
    enable_gpu: bool = False
    local_rank: Optional[int] = None
    # ---- new flags ----
    method: int = 0                 # 0 = derive from the flags below; 1..6 = report's methods
    compressor: str = ""            # "" = derive from --compress-grad / --method
    qsgd_levels: int = 128          # qsgd.py:9
    topk_ratio: float = 0.5         # qsgd.py:10 (slides use 0.4)
    bucket_size: int = 0            # bytes per exchange bucket; 0 = the whole model in one launch (per-tensor norms / top-k either way)
    sync_period: int = 1            # 20 in the report's method 6
    exchange: str = "ps"            # ps | allgather | allreduce
    pull: str = ""                  # "" = derive; weights | grads
    compress_pull: bool = False     # method 4/5: compress server->worker direction too
    dtype: str = "fp32"             # fp32 | bf16 (compute dtype on GPU)
    resume: str = ""                # path to a checkpoint (state_dict or sidecar) to resume from
    data: str = "auto"              # auto | synthetic | real
    data_root: str = "."            # where ./mnist_data, ./cifar10_data live for --data real
    synthetic_size: int = 4096      # samples in the synthetic training set
    backend: str = "auto"           # auto | gloo | nccl | native
    compute: str = "torch"          # torch | native (hand-written sm_100a kernels)
    weight_decay: float = 0.0
    nesterov: bool = False
    optimizer: str = "sgd"          # sgd | adam
    error_feedback: bool = False    # residual accumulation for top-k (off = reference semantics)
    metrics_file: str = ""          # JSONL per-step records
    legacy_checkpoint_name: bool = True  # keep "model_step_" (same file every time)
    fault_inject_step: int = -1     # test hook: this rank exits at step k
    fault_inject_rank: int = -1
    straggler_rank: int = -1        # test hook: this rank sleeps straggler_delay_s before every step
    straggler_delay_s: float = 0.0
    timeout_s: float = 300.0        # collective timeout for the watchdog
    shard_data: bool = True         # per-rank sharding (reference has none; fix per Appendix A)
    overlap: bool = False           # launch each bucket's exchange from backward (autograd hooks) on a side stream
    graphs: bool = False            # --backend native: GpuTrainer fast path, whole step replayed from one CUDA graph

    # ------------------------------------------------------------------
    def resolved_compressor(self) -> str:
        if self.compressor:
            return self.compressor
        if self.method in (1, 3):
            return "none"
        if self.method in (2, 4):
            return "qsgd"
        if self.method in (5, 6):
            return "topk-qsgd"
        return "none" if self.compress_grad == "none" else "qsgd"

    def resolved_pull(self) -> str:
        if self.pull:
            return self.pull
        return "weights" if self.method in (1, 2) else "grads"

    def resolved_compress_pull(self) -> bool:
        return self.compress_pull or self.method in (4, 5, 6)

    def resolved_sync_period(self) -> int:
        if self.method == 6 and self.sync_period == 1:
            return 20
        return max(1, self.sync_period)

    def num_classes(self) -> int:
        return 100 if self.dataset == "Cifar100" else 10

    def validate(self) -> "TrainConfig":
        if self.network not in NETWORKS:
            raise ValueError(f"--network must be one of {NETWORKS}, got {self.network!r}")
        if self.dataset not in DATASETS:
            raise ValueError(f"--dataset must be one of {DATASETS}, got {self.dataset!r}")
        if self.resolved_compressor() not in COMPRESSORS:
            raise ValueError(f"--compressor must be one of {COMPRESSORS}")
        if self.exchange not in EXCHANGES:
            raise ValueError(f"--exchange must be one of {EXCHANGES}")
        if self.mode not in ("normal", "kill", "async"):
            raise ValueError("--mode must be normal, kill or async")
        if self.method not in METHODS:
            raise ValueError(f"--method must be in {METHODS}")
        if not (1 <= self.qsgd_levels <= 32768):
            raise ValueError("--qsgd-levels must be in [1, 32768]")
        if not (0.0 < self.topk_ratio <= 1.0):
            raise ValueError("--topk-ratio must be in (0, 1]")
        if self.overlap and self.error_feedback:
            # the residual is folded into the gradient after backward, the overlapped exchange starts during it
            raise ValueError("--overlap and --error-feedback cannot be combined")
        if self.batch_size < 1 or self.max_steps < 0 or self.sync_period < 0:
            raise ValueError("--batch-size must be >= 1, --max-steps and --sync-period >= 0")
        if self.lr < 0 or not (0.0 <= self.momentum < 1.0):
            raise ValueError("--lr must be >= 0 and --momentum in [0, 1)")
        return self

    def as_dict(self) -> Dict[str, Any]:
        return dataclasses.asdict(self)


# ----------------------------------------------------------------------------------------
# Registry of tunables with env override (the MCA-var analogue).  Kernel tunables that are
# not CLI flags register themselves here so `python -m ewdml.cli.info` can list them.
# ----------------------------------------------------------------------------------------
@dataclass
class _Var:
    name: str
    default: Any
    typ: type
    help: str
    source: str = "default"
    value: Any = None


class Registry:
    """Small typed variable registry: default < env (EWDML_*) < explicit set()."""

    def __init__(self) -> None:
        self._vars: Dict[str, _Var] = {}

    def register(self, name: str, default: Any, help: str = "", typ: Optional[type] = None) -> Any:
        typ = typ or type(default)
        var = _Var(name=name, default=default, typ=typ, help=help, value=default)
        env = os.environ.get(ENV_PREFIX + name.upper())
        if env is not None:
            var.value = _str2bool(env) if typ is bool else typ(env)
            var.source = "env"
        self._vars[name] = var
        return var.value

    def get(self, name: str) -> Any:
        return self._vars[name].value

    def set(self, name: str, value: Any, source: str = "api") -> None:
        var = self._vars[name]
        var.value = _str2bool(value) if var.typ is bool else var.typ(value)
        var.source = source

    def dump(self) -> Dict[str, Dict[str, Any]]:
        return {k: {"value": v.value, "default": v.default, "source": v.source, "help": v.help}
                for k, v in sorted(self._vars.items())}


registry = Registry()
registry.register("spin_limit", 1 << 26, "device-side flag poll cap before the error word is set: 4096 back-to-back polls, then ~1 us per poll (2^26 = about a minute)")
registry.register("allreduce_algo", "auto", "fused dense allreduce (K4): auto | oneshot | twoshot | nvls (auto = nvls when the buffer has a multicast mapping)")
registry.register("heap", "auto", "symmetric heap transport: auto (VMM + NVLS multicast when every rank supports it) | vmm | ipc (cudaMalloc + CUDA IPC)")
registry.register("multicast", True, "use NVLS multicast stores / in-switch reduction when the heap provides multicast mappings")
registry.register("coll_allreduce_algo", "auto", "collective engine allreduce: auto | oneshot | twoshot (read by the library)")
registry.register("coll_oneshot_max_bytes", 1 << 18, "collective engine: one-shot allreduce at or below this size (read by the library)")
registry.register("gemm_bn", 0, "experimental GEMM tile width: 256 = 128x256 tiles where N % 256 == 0 (read by the library)")
registry.register("gemm_persist", 0, "experimental persistent GEMM kernel with two TMEM accumulators: 1 = on (read by the library)")
registry.register("trace", False, "print one line per native collective call (libompitrace analogue)")
registry.register("nvtx", False, "emit NVTX ranges around exchange phases")
registry.register("debug_poison", False, "poison receive slots after consumption (debug build aid)")


def _field_flag(name: str) -> str:
    # reference spells --local_rank with an underscore (distributed_nn.py:69)
    return "--local_rank" if name == "local_rank" else "--" + name.replace("_", "-")


def build_parser(description: str = "ewdml distributed trainer") -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description=description)
    for f in dataclasses.fields(TrainConfig):
        flag = _field_flag(f.name)
        default = f.default
        env = os.environ.get(ENV_PREFIX + f.name.upper())
        if f.type in ("bool", bool):
            if env is not None:
                default = _str2bool(env)
            p.add_argument(flag, nargs="?", const=True, default=default, type=_str2bool,
                           help=f"(bool) default {default}")
        elif f.type in ("Optional[int]",):
            p.add_argument(flag, type=int, default=int(env) if env is not None else default)
        else:
            typ = {"int": int, "float": float, "str": str}[f.type] if isinstance(f.type, str) else f.type
            if env is not None:
                default = typ(env)
            p.add_argument(flag, type=typ, default=default, help=f"default {default}")
    return p


def parse_args(argv=None, description: str = "ewdml distributed trainer") -> TrainConfig:
    ns = build_parser(description).parse_args(argv)
    cfg = TrainConfig(**{f.name: getattr(ns, f.name) for f in dataclasses.fields(TrainConfig)})
    return cfg.validate()
