from . import strategy, comm  # noqa: F401
