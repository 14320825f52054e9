"""The plugin interface the store registers with (MXNet-absent stand-in: see below).

Inside a real MXNet process this is ``mxnet.kvstore.KVStoreBase`` itself, so that
``mx.kv.create('b200')`` and ``gluon.Trainer(kvstore=...)`` find the class.  MXNet cannot be
imported in this image (no libmxnet.so), so otherwise an API-identical restatement
of python/mxnet/kvstore/base.py:74-243,404-454 is used: same abstract methods, same
``KVStoreBase.register`` (class name lower-cased, warning on override) and same ``create(name)``
lookup (case-insensitive, instantiated with no arguments)."""
import warnings

try:  # pragma: no cover - exercised only inside an MXNet installation
    from mxnet.kvstore.base import KVStoreBase, create  # noqa: F401
    HAVE_MXNET = True
except Exception:
    HAVE_MXNET = False

    class KVStoreBase(object):
        """An abstract key-value store interface for data parallel training (base.py:74-215)."""

        def broadcast(self, key, value, out, priority=0):
            raise NotImplementedError()

        def pushpull(self, key, value, out=None, priority=0):
            raise NotImplementedError()

        def set_optimizer(self, optimizer):
            raise NotImplementedError()

        OPTIMIZER = 'optimizer'

        def is_capable(self, capability):
            raise NotImplementedError()

        def save_optimizer_states(self, fname, dump_optimizer=False):
            raise NotImplementedError()

        def load_optimizer_states(self, fname):
            raise NotImplementedError()

        @property
        def type(self):
            raise NotImplementedError()

        @property
        def rank(self):
            raise NotImplementedError()

        @property
        def num_workers(self):
            raise NotImplementedError()

        kv_registry = {}

        @staticmethod
        def register(klass):
            """base.py:219-243"""
            assert isinstance(klass, type)
            name = klass.__name__.lower()
            if name in KVStoreBase.kv_registry:
                warnings.warn('WARNING: New kvstore %s.%s is overriding existing kvstore %s.%s' % (
                    klass.__module__, klass.__name__, KVStoreBase.kv_registry[name].__module__,
                    KVStoreBase.kv_registry[name].__name__))
            KVStoreBase.kv_registry[name] = klass
            return klass

    def create(name='local'):
        """base.py:404-454 — registry lookup only (there is no native MXKVStoreCreate here)."""
        if not isinstance(name, str):
            raise TypeError('name must be a string')
        name = name.lower()
        if name in KVStoreBase.kv_registry:
            return KVStoreBase.kv_registry[name]()
        raise ValueError("unknown kvstore type %r (registered: %s)" % (
            name, sorted(KVStoreBase.kv_registry)))
