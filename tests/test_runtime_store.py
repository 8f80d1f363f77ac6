"""Native TCP rendezvous store (csrc/runtime/store.cpp, the PMIx-modex analogue): set / blocking get / barrier / all_gather
across processes, on CPU."""
import json
import os
import socket
import time

import torch.multiprocessing as mp


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _proc(rank, world, port, out_dir):
    from ewdml.runtime import lib as L
    if rank == 2:
        time.sleep(0.3)                           # a late joiner: the others block in get / barrier until it arrives
    st = L.Store("127.0.0.1", port, rank, world, timeout_s=30.0)
    st.set(f"key/{rank}", bytes([rank]) * (100 + rank))
    got = {r: st.get(f"key/{r}") for r in range(world)}          # blocks until every key exists
    st.barrier("phase1")
    handles = st.all_gather("handles", f"ipc-handle-of-{rank}".encode().ljust(64, b"\0"))
    big = os.urandom(40000) if rank == 0 else b""
    if rank == 0:
        st.set("big", big)
    blob = st.get("big", cap=1 << 17)
    st.barrier("phase2")
    json.dump({"lens": {str(r): len(v) for r, v in got.items()}, "first": {str(r): v[0] for r, v in got.items()},
               "handles": [h.rstrip(b"\0").decode() for h in handles], "big": len(blob)},
              open(os.path.join(out_dir, f"r{rank}.json"), "w"))
    st.close()


def test_store_set_get_barrier_all_gather(tmp_path):
    world = 3
    mp.spawn(_proc, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        res = json.load(open(tmp_path / f"r{r}.json"))
        assert res["lens"] == {"0": 100, "1": 101, "2": 102} and res["first"] == {"0": 0, "1": 1, "2": 2}
        assert res["handles"] == [f"ipc-handle-of-{i}" for i in range(world)]
        assert res["big"] == 40000
