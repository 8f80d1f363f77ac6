#!/usr/bin/env python
"""How does K2 (wait + decode + average + SGD) scale with the number of senders?  One GPU, emulated ranks (LocalPeerGroup: every
"peer" mailbox is local memory), ResNet-50's real segment table, top-k 1 % + QSGD.  W senders push into rank 0's mailbox one after
another, then ONE k_recv_apply over W sources is timed with CUDA events (L2 flushed before it).  Prints one JSON line per W."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ewdml.models import build_model
from ewdml.parallel.flat import FlatModel
from ewdml.parallel.native_exchange import ExchangePlan, Mailbox, make_codec
from ewdml.parallel.peer import LocalPeerGroup
from ewdml.runtime import lib as L

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
model = build_model(sys.argv[1] if len(sys.argv) > 1 else "ResNet50", 10)
flat = FlatModel(model, dev, channels_last=True, with_bf16=True)
codec = make_codec("topk-qsgd", 128, 0.01)
plan = ExchangePlan.for_flat(flat, codec)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for W in (1, 2, 4, 8):
    lpg = LocalPeerGroup(W, dev)
    mbs = [Mailbox(b, W, plan.slot_bytes) for b in lpg.alloc(f"mb{W}", Mailbox.nbytes(W, plan.slot_bytes))]
    grads = [torch.randn(flat.total, device=dev) * 0.01 for _ in range(W)]
    times, sends = [], []
    for q in range(1, 8):
        par = q & 1
        e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        e[0].record()
        for r in range(W):
            plan.send(grads[r], [(mbs[0].slot(0, par, r), mbs[0].flag(0, par, r))], q, 7, q, r)
        e[1].record()
        flush.zero_()
        e[2].record()
        plan.recv_apply([(mbs[0].slot(0, par, r), mbs[0].flag(0, par, r)) for r in range(W)], q, float(W), L.OUT_SGD,
                        sgd=__import__("ewdml.parallel.native_exchange", fromlist=["make_sgd"]).make_sgd(type("C", (), dict(lr=0.01, momentum=0.9, weight_decay=0.0, nesterov=False))(), False),
                        param=flat.param, momentum=flat.momentum, param_bf16=flat.param_bf16)
        e[3].record()
        torch.cuda.synchronize()
        if q > 3:
            times.append(e[2].elapsed_time(e[3])); sends.append(e[0].elapsed_time(e[1]) / W)
    print(json.dumps({"model": sys.argv[1] if len(sys.argv) > 1 else "ResNet50", "senders": W, "ntiles": plan.ntiles, "nseg": plan.nseg,
                      "k2_ms": round(sorted(times)[len(times) // 2], 4), "k1_ms_per_sender": round(sorted(sends)[len(sends) // 2], 4),
                      "payload_bytes": plan.stats()["payload_bytes"]}), flush=True)
