#!/usr/bin/env python3
"""Cloud half of the cluster tooling: rent / list / release EC2 instances and turn them into a hosts file.

The reference's ``tools/pytorch_ec2.py`` (Python 2, boto3 + paramiko) launches 1 master + N worker spot instances from a
self-interpolating configuration dictionary (``:12-91``), waits until they run, writes their addresses to ``hosts`` /
``hosts_address`` (``get_hosts``, ``:938-951``) and terminates them again.  This file is that capability for
``tools/cluster.py``: it only talks to the cloud API; everything that happens ON the hosts afterwards (sync, run_command,
launch, kill_all) is ``cluster.py`` over ssh.

    python tools/cloud.py --config tools/cloud.example.json plan          # resolved configuration + the API requests, offline
    python tools/cloud.py --config my.json launch                         # request the instances, wait, write tools/hosts
    python tools/cloud.py --config my.json hosts                          # re-read the running instances into tools/hosts
    python tools/cloud.py --config my.json terminate

Configuration values may refer to each other with ``%(key)s`` (resolved repeatedly until nothing changes), like the
reference's ``Cfg`` class.  ``boto3`` is imported only by the commands that need the network; on a machine without it
(or without credentials) ``plan`` and ``--dry-run`` still show exactly what would be requested.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time
from typing import Any, Dict, List

HERE = os.path.dirname(os.path.abspath(__file__))

DEFAULTS: Dict[str, Any] = {
    "name": "ewdml",                       # tag that marks this job's instances (used by hosts / terminate)
    "region": "us-west-2",
    "availability_zone": "%(region)sb",
    "image_id": "",                        # an AMI with CUDA 12.9 + PyTorch; required for launch
    "key_name": "",                        # EC2 key pair for ssh
    "ssh_username": "ubuntu",
    "n_masters": 1,
    "n_workers": 1,
    "master_type": "p6-b200.48xlarge",     # 8 x B200; the framework's peer-memory exchange is single-node
    "worker_type": "%(master_type)s",
    "method": "spot",                      # spot | on-demand
    "spot_price": "",                      # string, as the API wants it; empty = market price cap
    "security_group_ids": [],
    "subnet_id": "",
    "slots_per_host": 8,                   # GPUs per host, written into the hosts file
    "hosts_file": os.path.join(HERE, "hosts"),
    "wait_timeout_s": 900,
}


def resolve(cfg: Dict[str, Any]) -> Dict[str, Any]:
    """``%(key)s`` interpolation between values, iterated to a fixed point (cycles are reported)."""
    out = dict(cfg)
    for _ in range(32):
        changed = False
        flat = {k: v for k, v in out.items() if isinstance(v, (str, int, float))}
        for k, v in out.items():
            if isinstance(v, str) and "%(" in v:
                new = v % flat
                changed |= new != v
                out[k] = new
        if not changed:
            return out
    raise ValueError("configuration values refer to each other in a cycle")


def load_config(path: str) -> Dict[str, Any]:
    cfg = dict(DEFAULTS)
    if path:
        with open(path) as f:
            user = json.load(f)
        unknown = sorted(set(user) - set(DEFAULTS))
        if unknown:
            raise SystemExit(f"unknown configuration keys: {unknown}")
        cfg.update(user)
    return resolve(cfg)


def requests_for(cfg: Dict[str, Any]) -> List[Dict[str, Any]]:
    """The ``run_instances`` requests: one for the master(s), one for the workers (same shape the API takes)."""
    reqs = []
    for role, count, itype in (("master", cfg["n_masters"], cfg["master_type"]), ("worker", cfg["n_workers"], cfg["worker_type"])):
        if int(count) <= 0:
            continue
        r: Dict[str, Any] = {
            "ImageId": cfg["image_id"], "InstanceType": itype, "MinCount": int(count), "MaxCount": int(count),
            "KeyName": cfg["key_name"], "Placement": {"AvailabilityZone": cfg["availability_zone"]},
            "TagSpecifications": [{"ResourceType": "instance", "Tags": [{"Key": "ewdml-job", "Value": cfg["name"]},
                                                                          {"Key": "ewdml-role", "Value": role}]}],
        }
        if cfg["security_group_ids"]:
            r["SecurityGroupIds"] = list(cfg["security_group_ids"])
        if cfg["subnet_id"]:
            r["SubnetId"] = cfg["subnet_id"]
        if cfg["method"] == "spot":
            opts: Dict[str, Any] = {"SpotInstanceType": "one-time"}
            if cfg["spot_price"]:
                opts["MaxPrice"] = str(cfg["spot_price"])
            r["InstanceMarketOptions"] = {"MarketType": "spot", "SpotOptions": opts}
        reqs.append(r)
    return reqs


def _ec2(cfg):
    try:
        import boto3
    except ImportError:
        raise SystemExit("boto3 is not installed on this machine: `plan` and --dry-run work offline; launch / hosts / terminate "
                         "need boto3 and AWS credentials")
    return boto3.client("ec2", region_name=cfg["region"])


def _job_instances(client, cfg, states=("pending", "running")) -> List[Dict[str, Any]]:
    res = client.describe_instances(Filters=[{"Name": "tag:ewdml-job", "Values": [cfg["name"]]},
                                             {"Name": "instance-state-name", "Values": list(states)}])
    out = []
    for r in res.get("Reservations", []):
        for i in r.get("Instances", []):
            tags = {t["Key"]: t["Value"] for t in i.get("Tags", [])}
            out.append({"id": i["InstanceId"], "state": i["State"]["Name"], "role": tags.get("ewdml-role", "worker"),
                        "private_ip": i.get("PrivateIpAddress", ""), "public_ip": i.get("PublicIpAddress", "")})
    return sorted(out, key=lambda x: (x["role"] != "master", x["private_ip"]))


def write_hosts(instances: List[Dict[str, Any]], cfg: Dict[str, Any], use_public: bool = False) -> str:
    """``tools/hosts`` in the format ``cluster.py`` reads: master first (it becomes MASTER_ADDR / node rank 0)."""
    lines = [f"# written by tools/cloud.py for job {cfg['name']}"]
    for i in instances:
        ip = i["public_ip"] if use_public and i["public_ip"] else i["private_ip"]
        lines.append(f"{cfg['ssh_username']}@{ip} {int(cfg['slots_per_host'])}   # {i['role']} {i['id']}")
    with open(cfg["hosts_file"], "w") as f:
        f.write("\n".join(lines) + "\n")
    return cfg["hosts_file"]


def cmd_plan(a, cfg):
    print(json.dumps({"config": cfg, "run_instances": requests_for(cfg)}, indent=1))
    return 0


def cmd_launch(a, cfg):
    if not cfg["image_id"] or not cfg["key_name"]:
        raise SystemExit("launch needs image_id and key_name in the configuration")
    reqs = requests_for(cfg)
    if a.dry_run:
        for r in reqs:
            print("+ ec2.run_instances(" + json.dumps(r) + ")")
        return 0
    client = _ec2(cfg)
    for r in reqs:
        client.run_instances(**r)
    want = int(cfg["n_masters"]) + int(cfg["n_workers"])
    deadline = time.time() + float(cfg["wait_timeout_s"])
    while time.time() < deadline:
        inst = _job_instances(client, cfg)
        running = [i for i in inst if i["state"] == "running" and i["private_ip"]]
        print(f"{len(running)} / {want} instances running", flush=True)
        if len(running) >= want:
            print("hosts file:", write_hosts(running, cfg, a.public))
            return 0
        time.sleep(10)
    raise SystemExit("timed out waiting for the instances; `terminate` releases what was started")


def cmd_hosts(a, cfg):
    if a.dry_run:
        print("+ ec2.describe_instances(tag:ewdml-job = %s)" % cfg["name"])
        return 0
    inst = [i for i in _job_instances(_ec2(cfg), cfg) if i["state"] == "running"]
    print("hosts file:", write_hosts(inst, cfg, a.public), f"({len(inst)} hosts)")
    return 0


def cmd_terminate(a, cfg):
    if a.dry_run:
        print("+ ec2.terminate_instances(every instance tagged ewdml-job = %s)" % cfg["name"])
        return 0
    client = _ec2(cfg)
    ids = [i["id"] for i in _job_instances(client, cfg, states=("pending", "running", "stopping", "stopped"))]
    if ids:
        client.terminate_instances(InstanceIds=ids)
    print(f"terminated {len(ids)} instances")
    return 0


def main(argv=None):
    p = argparse.ArgumentParser(description="ewdml cloud helper (EC2): launch / hosts / terminate; what runs on the hosts is tools/cluster.py")
    p.add_argument("--config", default="", help="JSON file; keys as in DEFAULTS, values may use %%(key)s")
    p.add_argument("--dry-run", action="store_true", help="print the API requests instead of sending them")
    p.add_argument("--public", action="store_true", help="write public instead of private addresses into the hosts file")
    p.add_argument("command", choices=["plan", "launch", "hosts", "terminate"])
    a = p.parse_args(argv)
    cfg = load_config(a.config)
    return {"plan": cmd_plan, "launch": cmd_launch, "hosts": cmd_hosts, "terminate": cmd_terminate}[a.command](a, cfg)


if __name__ == "__main__":
    sys.exit(main())
