#!/usr/bin/env python
"""Print a one-line summary of the last JSON line of each bench output file given on the command line."""
import json, sys
for f in sys.argv[1:]:
    try:
        d = [json.loads(l) for l in open(f) if l.startswith("{")][-1]
    except Exception as e:  # noqa: BLE001
        print(f, "no json:", e)
        continue
    if "unavailable" in d:
        print(f, "unavailable:", d["unavailable"])
        continue
    ex = {k: (round(v, 3) if isinstance(v, float) else v) for k, v in (d.get("exchange") or {}).items()} if isinstance(d.get("exchange"), dict) else None
    print(f.split("/")[-1], d.get("impl"), round(d["value"]), "img/s", round(d["ms_per_step"], 3), "ms  e2e",
          round(d["e2e"]["value"]) if d.get("e2e") else None, ex, "err", d.get("device_error"))
