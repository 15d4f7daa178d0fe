"""Ops layer: raw kernels (``kernels``), autograd wrappers (``functional``), fp32 oracles
(``reference``) and the ctypes loader for the sm_100a library (``_lib``)."""
from . import _lib, kernels, reference  # noqa: F401
from .reference import ACT_BY_NAME, ACT_GELU_ERF, ACT_GELU_TANH, ACT_NONE, ACT_RELU  # noqa: F401
