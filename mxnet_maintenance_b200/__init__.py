"""mxnet_maintenance_b200 — B200-native gradient-sync + update path for MXNet 1.x.

Importing the package registers the ``b200`` kvstore with ``KVStoreBase`` (so
``mx.kv.create('b200')`` / ``create('b200')`` work) and binds libb200kv.so; it raises if the CUDA
library has not been built — there is no CPU fallback.
"""
from . import _lib
from ._lib import B200KVError
from ._mx_shim import KVStoreBase, create, HAVE_MXNET
from .kvstore import B200, RowSparse
from . import optimizer, ops, shapes

__all__ = ["B200", "RowSparse", "B200KVError", "KVStoreBase", "create", "optimizer", "ops", "shapes"]
