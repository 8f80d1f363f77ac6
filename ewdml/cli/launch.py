"""``python -m ewdml.cli.launch --nproc N -- <command ...>``: start N local ranks under the native
launcher + watchdog (``ewdml_launch``), the single-node stand-in for ``mpirun`` / ``horovodrun`` /
``torch.distributed.launch`` (``src/run_pytorch_single.sh:1-3``, ``Horovod all reduce.ipynb`` cells 9-10).
Sets RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT for every rank and tears the whole job
down when one rank exits non-zero instead of leaving the others blocked in a collective.  With ``--nnodes N
--node-rank i`` (one launcher per node, same ``--master-addr``) ranks are numbered ``i * nproc + local_rank`` in a
world of ``N * nproc``.

``--max-restarts R --checkpoint-dir DIR`` is the reference's recovery story (``Report.zip!main.tex:90``: "restart from the
periodically saved model") done by the launcher instead of by hand: when the job fails it is started again, up to R times,
with ``--resume <newest checkpoint in DIR that has its .state sidecar>`` added to the command (``EWDML_RESTART_COUNT`` tells the
ranks which attempt they are).  A job that was interrupted by a signal to the launcher is not restarted."""
import argparse
import glob
import os
import signal
import sys

from ..runtime import lib


def latest_checkpoint(directory: str) -> str:
    """Newest ``model_step_*`` file in ``directory`` that has a ``.state`` sidecar ('' if there is none)."""
    best, best_t = "", -1.0
    for side in glob.glob(os.path.join(directory, "model_step_*.state")):
        path = side[:-len(".state")]
        if os.path.isfile(path) and os.path.getmtime(side) > best_t:
            best, best_t = path, os.path.getmtime(side)
    return best


def with_resume(cmd, checkpoint: str):
    """``cmd`` with ``--resume <checkpoint>`` (replacing a --resume that is already there, in either spelling)."""
    out, skip = [], False
    for arg in cmd:
        if skip:
            skip = False
            continue
        if arg == "--resume":
            skip = True
            continue
        if arg.startswith("--resume="):
            continue
        out.append(arg)
    return out + ["--resume", checkpoint]


def main(argv=None):
    p = argparse.ArgumentParser(description=__doc__)
    p.add_argument("--nproc", "-np", type=int, required=True)
    p.add_argument("--master-addr", default="127.0.0.1")
    p.add_argument("--master-port", type=int, default=29500)
    p.add_argument("--grace-seconds", type=float, default=5.0)
    p.add_argument("--nnodes", type=int, default=int(os.environ.get("NNODES", "1")), help="number of nodes running this launcher")
    p.add_argument("--node-rank", type=int, default=int(os.environ.get("NODE_RANK", "0")), help="index of this node")
    p.add_argument("--max-restarts", type=int, default=0, help="start a failed job again this many times")
    p.add_argument("--checkpoint-dir", default="", help="the job's --train-dir: a restart resumes from the newest checkpoint in it")
    p.add_argument("cmd", nargs=argparse.REMAINDER)
    a = p.parse_args(argv)
    cmd = a.cmd[1:] if a.cmd and a.cmd[0] == "--" else a.cmd
    if not cmd:
        p.error("no command given")
    os.environ["NNODES"], os.environ["NODE_RANK"] = str(a.nnodes), str(a.node_rank)
    interrupted = {128 + signal.SIGINT, 128 + signal.SIGTERM, 128 + signal.SIGHUP}
    attempt = 0
    while True:
        os.environ["EWDML_RESTART_COUNT"] = str(attempt)
        run_cmd = cmd
        if attempt > 0 and a.checkpoint_dir:
            ck = latest_checkpoint(a.checkpoint_dir)
            if ck:
                run_cmd = with_resume(cmd, ck)
            sys.stderr.write(f"[ewdml.launch] restart {attempt}/{a.max_restarts}: "
                             + (f"resuming from {ck}\n" if ck else f"no checkpoint in {a.checkpoint_dir}, starting over\n"))
        # a fresh port per attempt: the previous rendezvous socket may still be in TIME_WAIT
        status, failed = lib.launch(a.nproc, run_cmd, a.master_addr, a.master_port + attempt, a.grace_seconds)
        if status == 0:
            break
        sys.stderr.write(f"[ewdml.launch] job failed: rank {failed} exited with status {status}\n")
        if attempt >= a.max_restarts or status in interrupted:
            break
        attempt += 1
    sys.exit(status)


if __name__ == "__main__":
    main()
