from ewdml.models.resnet import *  # noqa: F401,F403
