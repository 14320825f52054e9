"""Host-side optimizer plumbing of the store: which scalars reach the fused update each step.

Original code of this package: ``kind_of`` (which fused kernel an optimizer object maps to),
``resolve_step`` (the per-step lr / wd / update-count resolution the kvstore updater performs,
python/mxnet/optimizer/optimizer.py:2103-2124 + :615-617,1331-1338,1609-1619) and ``Updater`` (the
object ``Trainer.save_states/load_states`` expect behind ``kvstore._updater``).  All three call the
USER's optimizer object's own ``_update_count`` / ``_get_lrs`` / ``_get_wds``.

``SGD`` / ``Adam`` / ``LAMB`` / ``AdaGrad`` / ``create`` exported here are MXNet's own classes when
MXNet is importable; otherwise the stand-ins of ``_mx_shim_optimizer`` (see its header).
"""
import math
import pickle

try:  # pragma: no cover - only inside an MXNet installation
    from mxnet.optimizer import Optimizer, SGD, Adam, LAMB, AdaGrad, create  # noqa: F401
except Exception:
    from ._mx_shim_optimizer import Optimizer, SGD, Adam, LAMB, AdaGrad, create  # noqa: F401

__all__ = ["Optimizer", "SGD", "Adam", "LAMB", "AdaGrad", "create", "Updater", "resolve_step", "kind_of"]


def kind_of(optimizer):
    """'sgd' | 'adam' | 'lamb' for any object that looks like the reference classes."""
    for klass in type(optimizer).__mro__:
        n = klass.__name__.lower()
        if n in ("sgd", "adam", "lamb", "adagrad"):
            return n
    raise ValueError("the b200 kvstore runs SGD, Adam, LAMB and AdaGrad on the store; got %s (use "
                     "update_on_kvstore=False for other optimizers)" % type(optimizer).__name__)


def resolve_step(optimizer, indices):
    """What reaches the fused op for one update of `indices`: (lrs, wds, steps) as Python floats.

    Follows the aggregated branch of Updater.__call__ -> update_multi_precision
    (optimizer.py:2103-2124, :615-617 for SGD, :1331-1338 for LAMB, :1609-1619 for Adam):
    counters first, then the scheduler at the new num_update, then multipliers.  Adam folds the
    bias correction into lr in double precision."""
    indices = list(indices)
    if hasattr(optimizer, "_set_current_context"):
        optimizer._set_current_context(0)
    optimizer._update_count(indices)
    lrs = optimizer._get_lrs(indices)
    wds = optimizer._get_wds(indices)
    steps = [optimizer._index_update_count[i] for i in indices]
    if kind_of(optimizer) == "adam":
        b1, b2 = optimizer.beta1, optimizer.beta2
        for i, t in enumerate(steps):
            coef1 = 1. - b1 ** t
            coef2 = 1. - b2 ** t
            lrs[i] *= math.sqrt(coef2) / coef1
    return lrs, wds, steps


class Updater(object):
    """What Trainer.save_states/load_states expect behind ``kvstore._updater``
    (python/mxnet/gluon/trainer.py:468-527; optimizer.py:2069-2160): ``.optimizer``, ``.states``,
    ``get_states`` / ``set_states`` with the reference's pickle layout.  The tensors themselves live
    in the store; ``states`` is filled on demand by the kvstore (``_collect_states``)."""

    def __init__(self, optimizer):
        self.optimizer = optimizer
        self.states = {}
        self.states_synced = {}

    def get_states(self, dump_optimizer=False):
        return pickle.dumps((self.states, self.optimizer) if dump_optimizer else self.states)

    def set_states(self, states):
        states = pickle.loads(states)
        if isinstance(states, tuple) and len(states) == 2:
            self.states, self.optimizer = states
        else:
            self.states = states
        self.states_synced = dict.fromkeys(self.states.keys(), False)
