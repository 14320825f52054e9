"""MXNet-absent stand-in for the USER's optimizer object — NOT part of the product path.

In a real MXNet process the user hands an ``mx.optimizer.SGD/Adam/LAMB/AdaGrad`` instance to
``kvstore.set_optimizer`` and the store calls that object's own pure-Python ``_update_count`` /
``_get_lrs`` / ``_get_wds``.  MXNet cannot be imported in this image (no libmxnet.so), so the tests
and bench.py need an object with the same constructor arguments, attribute names and scalar
bookkeeping.  The method bodies below therefore restate — deliberately line for line, they ARE the
contract — python/mxnet/optimizer/optimizer.py:
  ``Optimizer.__init__`` :103-141, ``register`` :143-167, ``learning_rate`` / ``set_learning_rate``
  :335-356, ``set_lr_mult`` / ``set_wd_mult`` :358-397, ``_set_current_context`` :399-409,
  ``_update_count`` :411-427, ``_get_lrs`` / ``_get_lr`` :429-470, ``_get_wds`` / ``_get_wd`` :472-508,
  ``SGD.__init__`` :563-567, ``LAMB.__init__`` :1253-1262, ``Adam.__init__`` :1590-1597,
  ``AdaGrad.__init__`` :1653-1655, ``create`` :169-199.
They hold no tensor math (every update runs in libb200kv.so).  mxnet_maintenance_b200.optimizer
re-exports them only when ``import mxnet`` fails.
"""

__all__ = ["Optimizer", "SGD", "Adam", "LAMB", "AdaGrad", "create"]


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
