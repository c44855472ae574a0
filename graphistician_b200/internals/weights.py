"""Old-API NIW weight model (internals/weights.py:62-101 in the reference).  The ``W[A==1]`` gather and its
reduction run on the device; this object holds the NIW state and draws (mu, Sigma) from statistics."""
import numpy as np

from .gaussian import Gaussian


class GaussianWeights(Gaussian):
    def __init__(self, B, mu_0=None, Sigma_0=None, nu_0=None, kappa_0=None):
        # defaults: internals/weights.py:69-79
        self.B = B
        super(GaussianWeights, self).__init__(
            mu_0=np.zeros(B) if mu_0 is None else mu_0,
            sigma_0=np.eye(B) if Sigma_0 is None else Sigma_0,
            nu_0=B + 2 if nu_0 is None else nu_0,
            kappa_0=1.0 if kappa_0 is None else kappa_0)

    def log_prior(self):
        return 0                                  # internals/weights.py:84-86 (TODO in the reference)

    def resample(self, networks=[], rng=np.random):
        """internals/weights.py:88-101: NIW draw given W where A == 1, over one network or a list."""
        from ..old_abstractions import as_network
        if not isinstance(networks, list):
            networks = [networks]
        n, sw, sww = 0, np.zeros(self.B), np.zeros((self.B, self.B))
        for net in networks:
            net = as_network(net)
            dev = net.device()
            ne, s1, s2 = dev.block_stats(np.zeros(net.N, np.int32), 1)
            d = dev.diag_stats(np.zeros(net.N, np.int64), 1)
            n += int(ne[0, 0] + d["d_edge"][0])
            sw += s1[0, 0] + d["d_sum_w"][0]
            sww += s2[0, 0] + d["d_sum_ww"][0]
        return self.resample_from_stats(n, sw, sww, rng)
