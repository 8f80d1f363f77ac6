"""``python -m ewdml.cli.launch --nproc N -- <command ...>``: start N local ranks under the native
launcher + watchdog (``ewdml_launch``), the single-node stand-in for ``mpirun`` / ``horovodrun`` /
``torch.distributed.launch`` (``src/run_pytorch_single.sh:1-3``, ``Horovod all reduce.ipynb`` cells 9-10).
Sets RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT for every rank and tears the whole job
down when one rank exits non-zero instead of leaving the others blocked in a collective.  With ``--nnodes N
--node-rank i`` (one launcher per node, same ``--master-addr``) ranks are numbered ``i * nproc + local_rank`` in a
world of ``N * nproc``."""
import argparse
import os
import sys

from ..runtime import lib


def main(argv=None):
    p = argparse.ArgumentParser(description=__doc__)
    p.add_argument("--nproc", "-np", type=int, required=True)
    p.add_argument("--master-addr", default="127.0.0.1")
    p.add_argument("--master-port", type=int, default=29500)
    p.add_argument("--grace-seconds", type=float, default=5.0)
    p.add_argument("--nnodes", type=int, default=int(os.environ.get("NNODES", "1")), help="number of nodes running this launcher")
    p.add_argument("--node-rank", type=int, default=int(os.environ.get("NODE_RANK", "0")), help="index of this node")
    p.add_argument("cmd", nargs=argparse.REMAINDER)
    a = p.parse_args(argv)
    cmd = a.cmd[1:] if a.cmd and a.cmd[0] == "--" else a.cmd
    if not cmd:
        p.error("no command given")
    os.environ["NNODES"], os.environ["NODE_RANK"] = str(a.nnodes), str(a.node_rank)
    status, failed = lib.launch(a.nproc, cmd, a.master_addr, a.master_port, a.grace_seconds)
    if status != 0:
        sys.stderr.write(f"[ewdml.launch] job failed: rank {failed} exited with status {status}\n")
    sys.exit(status)


if __name__ == "__main__":
    main()
