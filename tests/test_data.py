"""Data helpers: device feeder (CPU mode), whole-dataset tensors for real datasets, batched CIFAR augmentation."""
import numpy as np
import torch

from ewdml.data import DeviceFeeder, SyntheticImageDataset, dataset_tensors, gpu_augment


def test_feeder_shards_are_disjoint_and_epochs_cover_the_shard():
    ds = SyntheticImageDataset("Cifar10", 96)
    ds.data[:, 0, 0, 0] = torch.arange(96, dtype=torch.float32)              # tag every sample
    seen = []
    for rank in range(3):
        f = DeviceFeeder(ds.data, ds.targets, 8, torch.device("cpu"), shard_rank=rank, num_shards=3, seed=5)
        ids = []
        for _ in range(f.steps_per_epoch):
            x, y = f.next(); f.mark_consumed()
            assert x.shape == (8, 3, 32, 32) and y.shape == (8,)
            ids += x[:, 0, 0, 0].long().tolist()
        assert sorted(ids) == list(range(rank, 96, 3))                       # one epoch = the whole shard, once
        seen += ids
        x2, _ = f.next()                                                     # next epoch reshuffles
        assert f.h2d_bytes_per_step == 8 * (3 * 32 * 32 * 4 + 8)
    assert sorted(seen) == list(range(96))


def test_dataset_tensors_for_torchvision_style_datasets():
    class FakeCifar:
        data = np.full((5, 32, 32, 3), 255, dtype=np.uint8)
        targets = [0, 1, 2, 3, 4]

    class FakeMnist:
        data = torch.zeros(4, 28, 28, dtype=torch.uint8)
        targets = torch.tensor([9, 8, 7, 6])

    class FakeSvhn:
        data = np.zeros((3, 3, 32, 32), dtype=np.uint8)
        labels = np.array([1, 2, 3])

    x, y = dataset_tensors(FakeCifar(), "Cifar10")
    assert x.shape == (5, 3, 32, 32) and y.tolist() == [0, 1, 2, 3, 4]
    mean, std = 125.3 / 255, 63.0 / 255
    assert torch.allclose(x[:, 0], torch.full((5, 32, 32), (1.0 - mean) / std), atol=1e-5)
    x, y = dataset_tensors(FakeMnist(), "MNIST")
    assert x.shape == (4, 1, 28, 28) and torch.allclose(x, torch.full_like(x, -0.1307 / 0.3081), atol=1e-5) and y.dtype == torch.int64
    x, y = dataset_tensors(FakeSvhn(), "SVHN")
    assert x.shape == (3, 3, 32, 32) and y.tolist() == [1, 2, 3]
    syn = SyntheticImageDataset("MNIST", 16)
    assert dataset_tensors(syn, "MNIST")[0] is syn.data


def test_batched_augmentation_is_a_crop_of_the_reflect_padded_image_optionally_flipped():
    torch.manual_seed(0)
    x = torch.arange(4 * 3 * 32 * 32, dtype=torch.float32).view(4, 3, 32, 32)
    y = gpu_augment(x)
    assert y.shape == x.shape and y.is_contiguous(memory_format=torch.channels_last)
    xp = torch.nn.functional.pad(x, (4, 4, 4, 4), mode="reflect")
    kinds = set()
    for n in range(4):
        hits = [(r, c, f) for r in range(9) for c in range(9) for f in (False, True)
                if torch.equal(y[n], xp[n, :, r:r + 32, c:c + 32].flip(2) if f else xp[n, :, r:r + 32, c:c + 32])]
        assert hits, "not a crop of the padded image"
        kinds.add(hits[0][2])
    assert torch.equal(gpu_augment(x[:, :, :, :], pad=0).flip(3).flip(3).abs().sum() > 0, torch.tensor(True))


def test_data_prepare_reports_missing_datasets_without_failing(monkeypatch, capsys):
    """src/data/data_prepare.py downloads the four datasets; offline, the tool names what is missing and exits 0."""
    import pytest
    from torchvision import datasets
    from ewdml.cli import data_prepare

    def offline(*a, **k):
        raise RuntimeError("no network")

    class Ready:
        def __init__(self, *a, **k):
            pass

    monkeypatch.setattr(datasets, "MNIST", Ready)
    for name in ("CIFAR10", "CIFAR100", "SVHN"):
        monkeypatch.setattr(datasets, name, offline)
    with pytest.raises(SystemExit) as e:
        data_prepare.main()
    assert e.value.code == 0
    out = capsys.readouterr().out
    assert "MNIST: ready" in out and "missing: CIFAR10, CIFAR100, SVHN" in out
