from . import _abi, scene  # noqa: F401
