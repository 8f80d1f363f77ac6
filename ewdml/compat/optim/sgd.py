from ewdml.optim.sgd import SGD  # noqa: F401
