from . import checkpoint, data, logging, stats, tracing  # noqa: F401
