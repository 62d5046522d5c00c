"""Operators: native sm_100a kernels (``native``) and their plain-PyTorch statements (``reference``)."""
from . import reference  # noqa: F401
from . import native  # noqa: F401
