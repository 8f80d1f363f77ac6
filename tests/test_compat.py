"""Reference-style imports and keyword dictionaries (ewdml/compat): a driver written the way the reference's
distributed_nn.py is (kwargs_master / kwargs_worker, start_updated / train_updated) runs unchanged."""
import argparse
import json
import os
import socket
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

COMPAT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "ewdml", "compat")


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _main(rank, world_size, port, out_dir):
    sys.path.insert(0, COMPAT)
    from distributed_worker import DistributedWorker
    from sync_replicas_master_nn import SyncReplicasMaster_NN
    from util import prepare_data
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.set_num_threads(1)
    dist.init_process_group(backend="gloo", world_size=world_size, rank=rank)
    args = argparse.Namespace(batch_size=16, lr=0.02, epochs=1, momentum=0.9, network="LeNet", comm_type="Bcast", num_aggregate=5,
                              kill_threshold=7.0, eval_freq=3, train_dir=out_dir + "/", max_steps=5, compress_grad="compress",
                              gather_type="gather", dataset="MNIST", test_batch_size=64, data="synthetic", synthetic_size=256)
    device = torch.device("cpu")
    common = {'world_size': world_size, 'batch_size': args.batch_size, 'learning_rate': args.lr, 'max_epochs': args.epochs,
              'momentum': args.momentum, 'network': args.network, 'comm_method': args.comm_type, 'eval_freq': args.eval_freq,
              'train_dir': args.train_dir, 'max_steps': args.max_steps, 'compress_grad': args.compress_grad,
              'gather_type': args.gather_type, 'device': device}
    if rank == 0:
        master = SyncReplicasMaster_NN(kill_threshold=args.num_aggregate, timeout_threshold=args.kill_threshold, **common)
        master.build_model(num_classes=10)
        master.start_updated()
        res = {"cur_step": master.cur_step}
    else:
        train_loader, test_loader = prepare_data(args, shard_rank=rank - 1, num_shards=world_size - 1)
        worker = DistributedWorker(rank=rank, kill_threshold=args.kill_threshold, **common)
        worker.build_model(num_classes=10)
        hist = worker.train_updated(train_loader=train_loader, test_loader=test_loader)
        res = {"next_step": worker.next_step, "losses": [h["loss"] for h in hist]}
    json.dump(res, open(os.path.join(out_dir, f"r{rank}.json"), "w"))
    dist.barrier()
    dist.destroy_process_group()


def test_reference_style_driver_runs(tmp_path):
    mp.spawn(_main, args=(3, _free_port(), str(tmp_path)), nprocs=3, join=True)
    res = [json.load(open(tmp_path / f"r{r}.json")) for r in range(3)]
    assert res[0]["cur_step"] == 5 and res[1]["next_step"] == 5 and len(res[2]["losses"]) == 5
    assert os.path.isfile(tmp_path / "model_step_")


def test_reference_style_helper_classes():
    sys.path.insert(0, COMPAT)
    from Compresssor.qsgd import QSGDCompressor
    from Compresssor.TopK import TopKCompressor, desparsify, sparsify
    from distributed_worker import ModelBuffer, accuracy
    from model_ops.lenet import LeNet
    from optim.sgd import SGD
    from sync_replicas_master_nn import GradientAccumulator, update_params_dist_version
    net = LeNet()
    mb, ga = ModelBuffer(net), GradientAccumulator(net, world_size=3)
    assert [tuple(b.shape) for b in mb.recv_buf] == [tuple(p.shape) for p in net.parameters()]
    ga.gradient_aggregator[0][1].fill_(2.0)
    ga.meset_everything()
    assert float(ga.gradient_aggregator[0][1].abs().sum()) == 0.0 and len(ga.gradient_aggregator[4]) == 3
    p = torch.ones(3)
    assert torch.allclose(update_params_dist_version(p, torch.ones(3), 0.5), torch.full((3,), 0.5))
    q, norm = QSGDCompressor().compress(torch.tensor([3.0, -4.0]))
    assert float(norm) == 5.0 and q.abs().max() <= 128
    vals, idx = sparsify(torch.tensor([1.0, -5.0, 2.0, 0.5]), 0.5)
    assert sorted(idx.tolist()) == [1, 2] and torch.equal(desparsify((vals, idx), 4).abs(), torch.tensor([0.0, 5.0, 2.0, 0.0]))
    opt = SGD(net.parameters(), lr=0.1, momentum=0.9)
    opt.step(grads=[torch.zeros_like(p) for p in net.parameters()])
    assert accuracy(torch.eye(4), torch.arange(4), topk=(1,))[0].item() == 100.0
    assert TopKCompressor(0.5).compress(torch.arange(4.0))[1] == (4, torch.Size([4]))
