from .sgd import SGD, lookup_optimizer  # noqa: F401
