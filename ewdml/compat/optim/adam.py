from ewdml.optim.adam import Adam  # noqa: F401
