from ewdml.compress.qsgd import QSGDCompressor  # noqa: F401
