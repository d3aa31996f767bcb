from . import checkpoint, data, logging, stats  # noqa: F401
