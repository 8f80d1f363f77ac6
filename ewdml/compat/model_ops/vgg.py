from ewdml.models.vgg import *  # noqa: F401,F403
