#!/usr/bin/env python3
"""Cluster helper: the host-side capability of the reference's ``tools/pytorch_ec2.py`` without the cloud API.

``pytorch_ec2.py`` (975 lines, Python 2, boto3 + paramiko) mixes two jobs: renting EC2 spot instances, and then
driving a set of hosts (``get_hosts``, ``run_command``, ``kill_all_python``, ``setup_nfs``, ... ``:938-951``).  The
first has no equivalent on a machine without network or cloud credentials; the second is what this tool does, over
plain ``ssh``/``rsync`` and a hosts file (``tools/hosts``: one ``host [slots]`` per line; the pseudo host ``local``
runs commands on this machine without ssh, which is also how the tests exercise the tool).

    python tools/cluster.py get_hosts
    python tools/cluster.py check                      # reachability + GPUs per host
    python tools/cluster.py sync --dest /opt/ewdml     # copy this repo to every host (update_git_dir.sh / setup_nfs)
    python tools/cluster.py run_command -- nvidia-smi -L
    python tools/cluster.py launch --nproc-per-node 8 -- python distributed_nn.py --network VGG11 ...
    python tools/cluster.py status | kill_all

``launch`` starts one ``python -m ewdml.cli.launch`` (native launcher + watchdog) per host with the
``NODE_RANK / NNODES / MASTER_ADDR / MASTER_PORT`` contract and records ``host pid`` pairs in ``tools/pids``;
``kill_all`` signals exactly those process groups (the reference's ``killall.sh:7-10`` runs ``killall python`` on
every host, which also kills everything else the user is running).  Note that the peer-memory exchange of this
framework is single-node (one NVSwitch domain); multi-host launches are for the gloo/NCCL paths.
"""
from __future__ import annotations

import argparse
import os
import shlex
import subprocess
import sys
from typing import List, Tuple

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)


def read_hosts(path: str) -> List[Tuple[str, int]]:
    hosts = []
    if os.path.isfile(path):
        for line in open(path):
            line = line.split("#")[0].strip()
            if line:
                parts = line.split()
                hosts.append((parts[0], int(parts[1]) if len(parts) > 1 else 0))
    return hosts


def remote(host: str, command: str, dry: bool, background: bool = False) -> subprocess.CompletedProcess:
    """Run ``command`` through bash on ``host`` (``local`` = this machine)."""
    if host == "local":
        argv = ["bash", "-c", command]
    else:
        argv = ["ssh", "-o", "BatchMode=yes", "-o", "StrictHostKeyChecking=accept-new", host, command]
    if dry:
        print("+ " + " ".join(shlex.quote(a) for a in argv))
        return subprocess.CompletedProcess(argv, 0, "", "")
    return subprocess.run(argv, capture_output=True, text=True)


def cmd_get_hosts(a, hosts):
    for h, s in hosts:
        print(h if not s else f"{h} slots={s}")
    return 0


def cmd_check(a, hosts):
    bad = 0
    for h, _ in hosts:
        r = remote(h, "hostname && (nvidia-smi -L 2>/dev/null | wc -l)", a.dry_run)
        ok = r.returncode == 0
        bad += not ok
        print(f"{h}: {'ok' if ok else 'UNREACHABLE'} {' '.join(r.stdout.split())}")
    return 1 if bad else 0


def cmd_run(a, hosts):
    command = " ".join(shlex.quote(c) for c in a.cmd)
    rc = 0
    for h, _ in hosts:
        r = remote(h, command, a.dry_run)
        for line in (r.stdout + r.stderr).splitlines():
            print(f"[{h}] {line}")
        rc = rc or r.returncode
    return rc


def cmd_sync(a, hosts):
    rc = 0
    for h, _ in hosts:
        if h == "local":
            argv = ["rsync", "-a", "--delete", "--exclude", ".git", "--exclude", "gpurun_out", REPO + "/", a.dest + "/"]
        else:
            argv = ["rsync", "-az", "--delete", "--exclude", ".git", "--exclude", "gpurun_out", REPO + "/", f"{h}:{a.dest}/"]
        if a.dry_run:
            print("+ " + " ".join(shlex.quote(x) for x in argv))
            continue
        rc = rc or subprocess.run(argv).returncode
    return rc


def cmd_launch(a, hosts):
    if not hosts:
        print("no hosts", file=sys.stderr)
        return 2
    master = a.master_addr or ("127.0.0.1" if hosts[0][0] == "local" else hosts[0][0])
    inner = " ".join(shlex.quote(c) for c in a.cmd)
    lines = []
    for node, (h, slots) in enumerate(hosts):
        nproc = a.nproc_per_node or slots or 1
        env = f"NODE_RANK={node} NNODES={len(hosts)} EWDML_NODE_RANK={node}"
        # `cd X || exit; cmd &`: only the launcher itself is backgrounded (an `&&` list would keep a subshell -- and the
        # ssh channel's pipes -- alive for the whole job)
        launch = (f"cd {shlex.quote(a.workdir)} || exit 1; {env} setsid nohup {shlex.quote(a.python)} -m ewdml.cli.launch "
                  f"-np {nproc} --master-addr {master} --master-port {a.master_port} -- {inner} "
                  f"> {shlex.quote(a.log)}.{node} 2>&1 < /dev/null & echo $!")
        r = remote(h, launch, a.dry_run)
        if r.returncode != 0:
            print(f"[{h}] launch failed: {r.stderr.strip()}", file=sys.stderr)
            return r.returncode
        pid = r.stdout.strip().splitlines()[-1] if r.stdout.strip() else "0"
        lines.append(f"{h} {pid}")
        print(f"[{h}] node {node}: pid {pid}, log {a.log}.{node}")
    if not a.dry_run:
        with open(a.pids, "w") as f:
            f.write("\n".join(lines) + "\n")
    return 0


def _pids(a):
    return [tuple(l.split()) for l in open(a.pids).read().splitlines() if l.strip()] if os.path.isfile(a.pids) else []


def cmd_status(a, hosts):
    alive = 0
    for h, pid in _pids(a):
        r = remote(h, f"kill -0 {int(pid)} 2>/dev/null && echo running || echo exited", a.dry_run)
        state = r.stdout.strip() or "?"
        alive += state == "running"
        print(f"{h} {pid} {state}")
    return 0 if alive == 0 else 3


def cmd_kill_all(a, hosts):
    for h, pid in _pids(a):
        # the launcher was started with setsid: its pid is the process-group id of the whole job on that host
        remote(h, f"kill -TERM -- -{int(pid)} 2>/dev/null || kill -TERM {int(pid)} 2>/dev/null; true", a.dry_run)
        print(f"{h}: signalled process group {pid}")
    if not a.dry_run and os.path.isfile(a.pids):
        os.remove(a.pids)
    return 0


def main(argv=None):
    p = argparse.ArgumentParser(description="ewdml cluster helper (hosts file + ssh)")
    p.add_argument("--hosts", default=os.path.join(HERE, "hosts"))
    p.add_argument("--pids", default=os.path.join(HERE, "pids"))
    p.add_argument("--dry-run", action="store_true", help="print the ssh/rsync commands instead of running them")
    sub = p.add_subparsers(dest="command", required=True)
    sub.add_parser("get_hosts")
    sub.add_parser("check")
    sub.add_parser("status")
    sub.add_parser("kill_all")
    s = sub.add_parser("sync"); s.add_argument("--dest", required=True)
    r = sub.add_parser("run_command"); r.add_argument("cmd", nargs=argparse.REMAINDER)
    l = sub.add_parser("launch")
    l.add_argument("--nproc-per-node", type=int, default=0, help="default: the host's slots=, else 1")
    l.add_argument("--master-addr", default="")
    l.add_argument("--master-port", type=int, default=29500)
    l.add_argument("--workdir", default=REPO)
    l.add_argument("--python", default=sys.executable)
    l.add_argument("--log", default="/tmp/ewdml_job.log")
    l.add_argument("cmd", nargs=argparse.REMAINDER)
    a = p.parse_args(argv)
    if getattr(a, "cmd", None) and a.cmd[0] == "--":
        a.cmd = a.cmd[1:]
    hosts = read_hosts(a.hosts)
    fn = {"get_hosts": cmd_get_hosts, "check": cmd_check, "run_command": cmd_run, "sync": cmd_sync, "launch": cmd_launch,
          "status": cmd_status, "kill_all": cmd_kill_all}[a.command]
    return fn(a, hosts)


if __name__ == "__main__":
    sys.exit(main())
