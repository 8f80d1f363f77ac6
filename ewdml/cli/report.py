"""``python -m ewdml.cli.report run.jsonl [other.jsonl ...]``: the table of the reference's report from ``--metrics-file`` output.

The report compares its six methods by top-1 accuracy, minutes spent communicating vs computing, average MB per iteration
and end-to-end minutes (``Report.zip!main.tex:143-158``); its authors collected those by hand from the per-step log lines
(``distributed_worker.py:230-231``).  The trainers here write one JSON object per rank and step (``--metrics-file``); this
tool reduces one or more of those files to the same columns, one row per file, as a text table or ``--json``.

Per file: the slowest worker defines the wall time (the job ends when it does); communication / compute time are that
worker's summed per-step phases; MB per iteration is the mean of (sent + received) bytes per worker and step; accuracy is
the mean top-1 over the last ``--tail`` steps of all workers; "steps to A %" is the first step at which the mean top-1 of
a sliding window of ``--tail`` steps reaches ``--target`` (the report's "epochs to converge", in steps).
"""
from __future__ import annotations

import argparse
import json
import sys
from collections import defaultdict
from typing import Dict, List


def load(path: str) -> List[dict]:
    rows = []
    with open(path) as f:
        for line in f:
            line = line.strip()
            if line.startswith("{"):
                try:
                    rows.append(json.loads(line))
                except json.JSONDecodeError:
                    pass          # a line cut short by a killed job
    return rows


def summarize(rows: List[dict], tail: int = 10, target: float = 0.0) -> Dict[str, float]:
    if not rows:
        return {"steps": 0, "workers": 0}
    by_rank: Dict[int, List[dict]] = defaultdict(list)
    for r in rows:
        by_rank[int(r.get("rank", 0))].append(r)
    for v in by_rank.values():
        v.sort(key=lambda r: r["step"])
    totals = {rank: sum(r.get("ms_compute", 0.0) + r.get("ms_send", 0.0) + r.get("ms_recv", 0.0) for r in v)
              for rank, v in by_rank.items()}
    slow = max(totals, key=totals.get)
    sv = by_rank[slow]
    comm_ms = sum(r.get("ms_send", 0.0) + r.get("ms_recv", 0.0) for r in sv)
    comp_ms = sum(r.get("ms_compute", 0.0) for r in sv)
    nsteps = max(len(v) for v in by_rank.values())
    mb = [(r.get("bytes_sent", 0) + r.get("bytes_recv", 0)) / 1e6 for v in by_rank.values() for r in v]
    # mean top-1 per step over the workers, then windows of `tail` steps
    per_step: Dict[int, List[float]] = defaultdict(list)
    for v in by_rank.values():
        for r in v:
            if "prec1" in r:
                per_step[int(r["step"])].append(float(r["prec1"]))
    steps = sorted(per_step)
    acc = [sum(per_step[s]) / len(per_step[s]) for s in steps]
    tail = max(1, min(tail, len(acc))) if acc else 1
    final_acc = sum(acc[-tail:]) / tail if acc else float("nan")
    reached = None
    if target > 0 and acc:
        run = 0.0
        for i, a in enumerate(acc):
            run += a
            if i >= tail:
                run -= acc[i - tail]
            if i >= tail - 1 and run / tail >= target:
                reached = steps[i]
                break
    wall_s = None
    if all("time" in r for r in sv) and len(sv) > 1:
        wall_s = sv[-1]["time"] - sv[0]["time"]
    ratios = [r["compression_ratio"] for v in by_rank.values() for r in v if r.get("compression_ratio")]
    return {
        "workers": len(by_rank), "steps": nsteps,
        "top1_final": final_acc, "loss_final": sum(r.get("loss", 0.0) for r in sv[-tail:]) / min(tail, len(sv)),
        "comm_minutes": comm_ms / 60e3, "compute_minutes": comp_ms / 60e3, "total_minutes": (comm_ms + comp_ms) / 60e3,
        "wall_minutes": None if wall_s is None else wall_s / 60.0,
        "mb_per_iteration": sum(mb) / max(1, len(mb)),
        "compression_ratio": sum(ratios) / len(ratios) if ratios else 1.0,
        "comm_fraction": comm_ms / max(1e-9, comm_ms + comp_ms),
        "steps_to_target": reached,
    }


def main(argv=None):
    p = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    p.add_argument("files", nargs="+", help="--metrics-file outputs, one per run (a label can be given as label=path)")
    p.add_argument("--tail", type=int, default=10, help="steps averaged for the final accuracy / the convergence window")
    p.add_argument("--target", type=float, default=0.0, help="top-1 %% for the 'steps to target' column (0 = off)")
    p.add_argument("--json", action="store_true")
    a = p.parse_args(argv)
    out = []
    for spec in a.files:
        label, _, path = spec.rpartition("=")
        label = label or path
        s = summarize(load(path), a.tail, a.target)
        s["run"] = label
        out.append(s)
    if a.json:
        print(json.dumps(out, indent=1))
        return 0
    base = out[0].get("mb_per_iteration") or 0.0
    hdr = f"{'run':<28}{'workers':>8}{'steps':>7}{'top-1 %':>9}{'comm min':>10}{'comp min':>10}{'total min':>10}{'MB/iter':>10}{'vs first':>9}"
    if a.target > 0:
        hdr += f"{'steps to ' + format(a.target, 'g') + ' %':>16}"
    print(hdr)
    for s in out:
        if not s.get("steps"):
            print(f"{s['run']:<28}  (no records)")
            continue
        rel = base / s["mb_per_iteration"] if s["mb_per_iteration"] else float("inf")
        line = (f"{s['run'][-28:]:<28}{s['workers']:>8}{s['steps']:>7}{s['top1_final']:>9.2f}{s['comm_minutes']:>10.3f}"
                f"{s['compute_minutes']:>10.3f}{s['total_minutes']:>10.3f}{s['mb_per_iteration']:>10.3f}{rel:>8.1f}x")
        if a.target > 0:
            line += f"{(s['steps_to_target'] if s['steps_to_target'] is not None else '-'):>16}"
        print(line)
    return 0


if __name__ == "__main__":
    sys.exit(main())
