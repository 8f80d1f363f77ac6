"""``python -m ewdml.cli.data_prepare``: fetch MNIST / CIFAR-10 / CIFAR-100 / SVHN once before a parallel launch
(``src/data/data_prepare.py``).  Without network access it reports which datasets are missing; training then uses
the synthetic datasets of the same shape (``--data auto``)."""
import sys


def main():
    from torchvision import datasets
    jobs = [("MNIST", lambda: datasets.MNIST("./mnist_data", train=True, download=True)),
            ("CIFAR10", lambda: datasets.CIFAR10("./cifar10_data", train=True, download=True)),
            ("CIFAR100", lambda: datasets.CIFAR100("./cifar100_data", train=True, download=True)),
            ("SVHN", lambda: datasets.SVHN("./svhn_data", split="train", download=True))]
    missing = []
    for name, fn in jobs:
        try:
            fn(); print(f"{name}: ready")
        except Exception as e:  # no network on the build box
            missing.append(name); print(f"{name}: unavailable ({type(e).__name__}); synthetic data will be used")
    if missing:
        print("missing: " + ", ".join(missing))
    sys.exit(0)       # a missing dataset is not an error: --data auto falls back to the synthetic one of the same shape


if __name__ == "__main__":
    main()
