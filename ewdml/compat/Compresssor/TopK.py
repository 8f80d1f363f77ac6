from ewdml.compress.topk import TopKCompressor, sparsify, desparsify  # noqa: F401
