"""Host-side optimizer front-end: the scalar plumbing of ``mxnet.optimizer`` for SGD / Adam / LAMB.

In a real MXNet process the user's own ``mx.optimizer.SGD/Adam/LAMB`` instance is handed to
``kvstore.set_optimizer`` and the store calls ITS ``_update_count`` / ``_get_lrs`` / ``_get_wds``
(pure-Python methods), so schedules and multipliers behave exactly as in the reference.  MXNet
cannot be imported in this image, so the classes below restate that logic with the same
constructor arguments and attribute names (python/mxnet/optimizer/optimizer.py:103-141 ``__init__``,
:411-508 counters / lr / wd, :526-600 SGD, :1250-1268 LAMB, :1590-1607 Adam).  They hold NO tensor
math: every update runs in the CUDA library.
"""
import math
import pickle

__all__ = ["Optimizer", "SGD", "Adam", "LAMB", "AdaGrad", "create", "Updater", "resolve_step"]


class Optimizer(object):
    opt_registry = {}

    def __init__(self, rescale_grad=1., param_idx2name=None, wd=0., clip_gradient=None,
                 learning_rate=None, lr_scheduler=None, sym=None, begin_num_update=0,
                 multi_precision=False, param_dict=None):
        self.rescale_grad = rescale_grad
        self.lr_scheduler = lr_scheduler
        if self.lr_scheduler is None and learning_rate is None:
            learning_rate = 0.01
        self.lr = learning_rate
        if self.lr_scheduler is not None and learning_rate is not None:
            if self.lr_scheduler.base_lr != learning_rate:
                self.lr_scheduler.base_lr = learning_rate
        self.wd = wd
        self.lr_mult = {}
        self.wd_mult = {}
        self.begin_num_update = begin_num_update
        self.num_update = begin_num_update
        self._all_index_update_counts = {0: {}}
        self._index_update_count = self._all_index_update_counts[0]
        self.clip_gradient = clip_gradient
        self.multi_precision = multi_precision
        self.aggregate_num = 0
        self.idx2name = dict(param_idx2name or {})
        self.param_dict = param_dict if param_dict else {}

    @staticmethod
    def register(klass):
        Optimizer.opt_registry[klass.__name__.lower()] = klass
        return klass

    @property
    def learning_rate(self):
        if self.lr_scheduler is not None:
            return self.lr_scheduler(self.num_update)
        return self.lr

    def set_learning_rate(self, lr):
        if self.lr_scheduler is not None:
            raise UserWarning("LRScheduler of the optimizer has already been defined.")
        self.lr = lr

    def set_lr_mult(self, args_lr_mult):
        self.lr_mult = dict(args_lr_mult)

    def set_wd_mult(self, args_wd_mult):
        # optimizer.py:375-397: names not ending in _weight/_gamma get wd_mult 0 when names are known
        self.wd_mult = {}
        for n in self.idx2name.values():
            if not (n.endswith('_weight') or n.endswith('_gamma')):
                self.wd_mult[n] = 0.0
        self.wd_mult.update(args_wd_mult)

    def _set_current_context(self, device_id):
        if device_id not in self._all_index_update_counts:
            self._all_index_update_counts[device_id] = {}
        self._index_update_count = self._all_index_update_counts[device_id]

    def _update_count(self, index):
        if not isinstance(index, (list, tuple)):
            index = [index]
        for idx in index:
            if idx not in self._index_update_count:
                self._index_update_count[idx] = self.begin_num_update
            self._index_update_count[idx] += 1
            self.num_update = max(self._index_update_count[idx], self.num_update)

    def _get_lrs(self, indices):
        lr = self.lr_scheduler(self.num_update) if self.lr_scheduler is not None else self.lr
        lrs = [lr for _ in indices]
        for i, index in enumerate(indices):
            if index in self.param_dict:
                lrs[i] *= self.param_dict[index].lr_mult
            elif index in self.lr_mult:
                lrs[i] *= self.lr_mult[index]
            elif index in self.idx2name:
                lrs[i] *= self.lr_mult.get(self.idx2name[index], 1.0)
        return lrs

    def _get_lr(self, index):
        return self._get_lrs([index])[0]

    def _get_wds(self, indices):
        wds = [self.wd for _ in indices]
        for i, index in enumerate(indices):
            if index in self.param_dict:
                wds[i] *= self.param_dict[index].wd_mult
            elif index in self.wd_mult:
                wds[i] *= self.wd_mult[index]
            elif index in self.idx2name:
                wds[i] *= self.wd_mult.get(self.idx2name[index], 1.0)
        return wds

    def _get_wd(self, index):
        return self._get_wds([index])[0]


@Optimizer.register
class SGD(Optimizer):
    def __init__(self, momentum=0.0, lazy_update=True, **kwargs):
        super(SGD, self).__init__(**kwargs)
        self.momentum = momentum
        self.lazy_update = lazy_update
        self.aggregate_num = 4


@Optimizer.register
class Adam(Optimizer):
    def __init__(self, learning_rate=0.001, beta1=0.9, beta2=0.999, epsilon=1e-8, lazy_update=True,
                 **kwargs):
        super(Adam, self).__init__(learning_rate=learning_rate, **kwargs)
        self.beta1, self.beta2, self.epsilon, self.lazy_update = beta1, beta2, epsilon, lazy_update


@Optimizer.register
class LAMB(Optimizer):
    def __init__(self, learning_rate=0.001, beta1=0.9, beta2=0.999, epsilon=1e-6, lower_bound=None,
                 upper_bound=None, bias_correction=True, **kwargs):
        super(LAMB, self).__init__(learning_rate=learning_rate, **kwargs)
        self.beta1, self.beta2, self.epsilon = beta1, beta2, epsilon
        self.lower_bound, self.upper_bound, self.bias_correction = lower_bound, upper_bound, bias_correction
        self.aggregate_num = 45


@Optimizer.register
class AdaGrad(Optimizer):
    """optimizer.py:1633-1687 (``eps`` is the history stabiliser ``float_stable_eps``)."""

    def __init__(self, eps=1e-7, **kwargs):
        super(AdaGrad, self).__init__(**kwargs)
        self.float_stable_eps = eps


def create(name, **kwargs):
    """mx.optimizer.create (optimizer.py:169-199)."""
    if isinstance(name, Optimizer):
        return name
    return Optimizer.opt_registry[name.lower()](**kwargs)


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
