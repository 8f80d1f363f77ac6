"""``python -m ewdml.cli.info``: what ``ompi_info`` is to Open MPI -- build, tunables and their sources."""
import json
import os

from ..config import ENV_PREFIX, TrainConfig, registry
from ..runtime import lib


def main():
    info = {"library": lib.lib_path(), "library_present": lib.available(), "env_prefix": ENV_PREFIX,
            "tunables": registry.dump(), "flags": TrainConfig().as_dict(),
            "env_overrides": {k: v for k, v in os.environ.items() if k.startswith(ENV_PREFIX)}}
    if lib.available():
        l = lib.load()
        info["library_version"] = l.ewdml_version()
        info["counters"] = {n: int(l.ewdml_counter_get(i)) for i, n in enumerate(
            ["launches", "send_calls", "recv_calls", "allreduce_calls", "bytes_sent", "bytes_recv", "compute_launches"])}
    try:
        import torch
        info["torch"] = torch.__version__
        info["cuda_devices"] = torch.cuda.device_count() if torch.cuda.is_available() else 0
        if lib.available() and torch.cuda.is_available():
            # what the symmetric heap can use on this device: VMM allocations, POSIX-fd handles, NVLS multicast objects
            import ctypes as C
            q = (C.c_int * 3)()
            lib.load().ewdml_vmm_query(q)
            info["transport"] = {"vmm": bool(q[0]), "multicast": bool(q[1]), "posix_fd_handles": bool(q[2]),
                                 "sm_count": int(lib.load().ewdml_sm_count())}
    except Exception:
        pass
    print(json.dumps(info, indent=1, default=str))


if __name__ == "__main__":
    main()
