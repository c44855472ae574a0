"""Old-API network objects and base classes.

The reference imports these from ``graphistician.old_abstractions`` (stochastic_block_model.py:12-13,
eigenmodel.py:1319-1320), a module that is missing from its tree; their contract is reconstructed from the use
sites: a network exposes ``.A`` (N x N), ``.W`` (N x N x B) and the mean-field expectations
``.E_A .E_W .E_WWT``.  Here a network additionally owns its device-resident copy (``device()``), uploaded once
and reused by every ``resample(network)`` call until ``invalidate()``.
"""
import numpy as np

from .device import DeviceNetwork


class WeightedDirectedNetwork(object):
    """Abstract network: subclasses provide A, W and their expectations."""
    A = W = None

    @property
    def E_A(self):
        return self.A

    @property
    def E_W(self):
        return self.W

    @property
    def E_WWT(self):
        return np.einsum("ijb,ijc->ijbc", self.W, self.W)


class FixedGaussianNetwork(WeightedDirectedNetwork):
    """Deterministic data (stochastic_block_model.py:163,224): E_A = A, E_W = W, E_WWT = W W^T per edge."""

    def __init__(self, A, W):
        A = np.asarray(A)
        assert A.ndim == 2 and A.shape[0] == A.shape[1], "A must be N x N"
        self.A = A
        self.N = A.shape[0]
        if W is not None:
            W = np.asarray(W)
            if W.ndim == 2:
                W = W[:, :, None]
            assert W.shape[:2] == A.shape, "W must be N x N x B"
        self.W = W
        self.B = 0 if W is None else W.shape[2]
        self._dev = {}
        self._stale = set()

    def device(self, dtype="f32", ctx=None):
        """Device-resident copy (uploaded on first use; see SURVEY.md 8b 'Ownership': data is never mutated)."""
        key = (dtype, id(ctx))
        if key in self._stale:
            self._stale.discard(key)
            self._dev[key].upload(self.A, self.W)
        if key not in self._dev:
            comm = getattr(self, "_comm", None)
            if comm is not None and comm.world > 1:
                from .sharding import sharded_device_network
                self._dev[key] = sharded_device_network(self.A, self.W, comm, dtype=dtype, ctx=ctx)
            else:
                self._dev[key] = DeviceNetwork(self.A, self.W, dtype=dtype, ctx=ctx)
        return self._dev[key]

    def shard(self, comm):
        """Row-shard the device copy over the ranks of ``comm`` (graphistician_b200.sharding.Comm): each rank then
        uploads only its rows, and every device pass of ``model.resample(network)`` becomes collective."""
        self._drop()
        self._comm = comm
        return self

    def _drop(self):
        """Free every device copy."""
        for d in self._dev.values():
            d.close()
        self._dev = {}
        self._stale = set()
        self.__dict__.pop("_sum_A", None)

    def invalidate(self):
        """Call after mutating A or W in place: the next ``device()`` copies the host arrays to the device again.
        Unsharded copies keep their device allocations (and the model-side state bound to them) and are refilled by
        ``DeviceNetwork.upload``; sharded copies are dropped and rebuilt."""
        from .device import DeviceNetwork
        self.__dict__.pop("_sum_A", None)          # host-side caches of A go stale with it
        for key, d in list(self._dev.items()):
            if type(d) is DeviceNetwork and d.nrows == d.N:
                self._stale.add(key)
            else:
                d.close()
                del self._dev[key]

    def __deepcopy__(self, memo):
        return FixedGaussianNetwork(self.A.copy(), None if self.W is None else self.W.copy())


def as_network(obj):
    """Accept what the reference's callers pass: a network object, an (A, W) tuple, or a bare ndarray
    (test/test_logistic_eigenmodel_geweke.py:41)."""
    if isinstance(obj, WeightedDirectedNetwork):
        return obj
    if isinstance(obj, (tuple, list)) and len(obj) == 2:
        return FixedGaussianNetwork(obj[0], obj[1])
    return FixedGaussianNetwork(np.asarray(obj), None)


class NetworkDistribution(object):
    """Old-API base (stochastic_block_model.py:18): log_probability = log_likelihood + log_prior."""

    def log_prior(self):
        raise NotImplementedError

    def log_likelihood(self, network):
        raise NotImplementedError

    def log_probability(self, network):
        return self.log_likelihood(network) + self.log_prior()

    def copy_sample(self):
        """pybasicbayes ``GibbsSampling.copy_sample`` = deep copy (test/test_logistic_eigenmodel_geweke.py:47)."""
        import copy
        return copy.deepcopy(self)


class GaussianWeightedNetworkDistribution(object):
    """Mixin: models exposing P (N x N), Mu (N x N x B), Sigma (N x N x B x B)."""
    pass


# ------------------------------------------------------------------------------------------------------
# new-API callers pass bare arrays -- ``adjacency.resample(A)``, ``weights.resample((A, W))``
# (networks.py:69-71).  The device copy is found again through a small, BOUNDED identity cache: one network
# object per host array A (upgraded in place when a W shows up for the same A).  The cache owns its entries
# (a network holds A, so the id cannot be reused while the entry lives) and evicts the least recently used one,
# freeing its device copies, once more than ``MAX_CACHED_NETWORKS`` distinct arrays have been seen -- a Geweke or
# AIS loop that draws a fresh A every iteration therefore keeps a flat device footprint.
# ------------------------------------------------------------------------------------------------------
import collections

MAX_CACHED_NETWORKS = 4
_NETWORKS = collections.OrderedDict()


def network_for(A, W=None):
    """The (cached) network object that owns the device copy of the host arrays A (N x N) and W (N x N x B).
    Data is treated as immutable, as the reference treats it (SURVEY.md 8b 'Ownership'); after an in-place edit
    of A or W call ``network_for(A, W).invalidate()``."""
    if isinstance(A, WeightedDirectedNetwork):
        return A
    A = np.asarray(A)
    key = id(A)
    net = _NETWORKS.get(key)
    if net is not None and net.A is A:
        _NETWORKS.move_to_end(key)
        if W is not None and (net.W is None or net._W_id != id(W)):
            net._drop()
            net.__init__(A, W)
            net._W_id = id(W)
        return net
    net = FixedGaussianNetwork(A, W)
    net._W_id = id(W) if W is not None else None
    _NETWORKS[key] = net
    while len(_NETWORKS) > MAX_CACHED_NETWORKS:
        _, old = _NETWORKS.popitem(last=False)
        old._drop()
    return net


def drop_cached_networks():
    """Free every cached network's device copies (the cache is rebuilt on demand)."""
    while _NETWORKS:
        _, old = _NETWORKS.popitem(last=False)
        old._drop()
