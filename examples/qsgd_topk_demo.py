"""The unit demos of the reference's ``QSGD and topk Sparsification.ipynb`` (cells 0, 2, 3), as a script.

* QSGD round trip of a 3-element tensor with s = 64 (cell 0): the reference prints a level of 128 > s because a huge
  norm rounds in fp32 and nothing clamps; here levels are clamped to s and a zero-norm tensor encodes to zeros.
* TopK round trip of a 4x3 tensor at ratio 0.4 (cell 2).
* Payload accounting (cell 3): the reference measures ``sys.getsizeof(tensor.storage())`` = payload + 72 bytes of
  Python object overhead; here the byte counts are those of the packed wire format.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch

from ewdml.compress import Compression, QSGDCompressor, TopKCompressor
from ewdml.parallel.flat import Segment
from ewdml.parallel.host_codec import HostCodec


def main():
    torch.manual_seed(0)
    t = torch.tensor([1.0e9, 2.0e9, 6.5573e10])
    q = QSGDCompressor(quantum_num=64)
    levels, norm = q.compress(t)
    print("QSGD  compress  :", levels.tolist(), "norm %.4e" % float(norm))
    print("QSGD  decompress:", q.decompress((levels, norm)).tolist())
    print("QSGD  zero norm :", q.compress(torch.zeros(4))[0].tolist())

    m = torch.arange(12, dtype=torch.float32).reshape(4, 3) - 5
    k = TopKCompressor(0.4)
    (vals, idx), ctx = k.compress(m)
    print("TopK  values/idx:", vals.tolist(), idx.tolist())
    print("TopK  decompress:\n", k.decompress((vals, idx), ctx))

    hq, hctx = Compression.qsgd.compress(t)            # Horovod convention: compress(t) -> (q, ctx)
    print("hvd   round trip:", Compression.qsgd.decompress(hq, hctx).tolist())

    n = 100000
    g = torch.randn(n)
    seg = [Segment("w", 0, n, torch.Size([n]), False)]
    for comp, ratio in (("none", 1.0), ("qsgd", 1.0), ("topk", 0.4), ("topk-qsgd", 0.4), ("topk-qsgd", 0.01)):
        codec = HostCodec(seg, comp, 128, ratio)
        msg = codec.encode(g, step=0)
        err = (codec.decode(msg) - g).norm() / g.norm()
        print(f"wire  {comp:9s} ratio {ratio:<4}: {msg.numel():7d} bytes ({4 * n / msg.numel():5.1f}x), relative error {float(err):.3f}")


if __name__ == "__main__":
    main()
