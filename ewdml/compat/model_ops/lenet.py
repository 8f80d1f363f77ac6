from ewdml.models.lenet import LeNet  # noqa: F401
