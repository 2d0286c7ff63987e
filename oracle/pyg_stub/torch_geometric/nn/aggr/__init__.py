from . import scaler  # noqa: F401
