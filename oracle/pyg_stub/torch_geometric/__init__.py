"""TEST-ONLY stand-in for torch_geometric 2.2.0 (see ../README.md)."""
__version__ = "2.2.0-stub"
from . import typing, utils, data, nn  # noqa: F401,E402
