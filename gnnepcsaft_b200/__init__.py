"""gnnepcsaft_b200 -- B200 (sm_100a) message-passing hot path of wildsonbbl/gnnepcsaft.

Public surface mirrors the reference (gnnepcsaft/train/models.py): ``GNNePCSAFT``, ``create_model``,
``get_conv``, ``get_global_pool`` and the PyG/ogb-compatible modules in ``gnnepcsaft_b200.nn``.
All compute goes through ``libgnnx_sm100.so`` (C ABI in ``include/gnnx.h``); there is no CPU path.
"""

from .models import GNNePCSAFT, create_model, get_conv, get_global_pool  # noqa: F401
from .ops import set_matmul_precision  # noqa: F401

__all__ = ["GNNePCSAFT", "create_model", "get_conv", "get_global_pool", "set_matmul_precision"]
