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

    # ---- mean field (internals/weights.py:105-132, 200-221) -------------------------------------------------
    def meanfieldupdate_from_stats(self, n, sum_w, sum_ww, stepsize=1.0, minibatchfrac=1.0):
        """`meanfieldupdate` (:105-132; stepsize == 1) / `svi_step` (:134-156) given the weighted statistics
        sum_ij weights[i,j] * (1, E_W, E_WWT) -- the N^2 reduction itself runs on the device (gk_mf_sbm_soft_stats)."""
        stats = self._statistics_matrix(n, sum_w, sum_ww)
        target = self.natural_hypparam + stats / minibatchfrac
        self.mf_natural_hypparam = (1 - stepsize) * self.mf_natural_hypparam + stepsize * target

    def mf_expected_mu(self):
        return self.mu_mf

    def mf_expected_Sigma(self):
        return self.sigma_mf / (self.nu_mf - self.D - 1)

    def mf_expected_mumuT(self):
        E_mu = self.mu_mf
        return self.mf_expected_Sigma() + np.outer(E_mu, E_mu)

    def mf_expected_Sigma_inv(self):
        return self.nu_mf * np.linalg.inv(self.sigma_mf)

    def mf_expected_logdet_Sigma(self):
        return -self._loglmbdatilde()

    def resample_from_mf(self):
        self._resample_from_mf()

    def expected_log_likelihood(self, network):
        """internals/weights.py:174-198: sum over ALL N^2 ordered pairs (no A mask, as the reference writes it) of
        Gaussian().negentropy(E_x = E_W[n1, n2], E_xxT = E_WWT[n1, n2], ...).sum().  The per-pair term is linear in
        (1, E_W, E_WWT), so the N^2 Python loop becomes three sums over the network's arrays (chunked NumPy: a host
        diagnostic, not a sweep)."""
        E_W = getattr(network, "E_W", None)
        E_W = np.asarray(network.W if E_W is None else E_W)
        N = E_W.shape[0]
        assert E_W.shape == (N, N, self.B)
        E_WWT = getattr(network, "E_WWT", None)
        c0, c1, C2 = self.expected_log_likelihood_coefficients()
        s1 = np.zeros(self.B); s2 = np.zeros((self.B, self.B))
        for r0 in range(0, N, 256):
            w = E_W[r0:r0 + 256].astype(np.float64)
            s1 += w.sum(axis=(0, 1))
            s2 += (np.einsum("ijb,ijc->bc", w, w) if E_WWT is None
                   else np.asarray(E_WWT[r0:r0 + 256], dtype=np.float64).sum(axis=(0, 1)))
        return float(N * N * c0 + c1.dot(s1) + np.sum(C2 * s2))

    def expected_log_likelihood_coefficients(self):
        """Coefficients (c0, c1[B], C2[B,B]) of E_q[ln N(w | mu, Sigma)] = c0 + c1.w + w^T C2 w as the reference's
        Gaussian.negentropy writes it (internals/distributions.py:278-308; its -0.5*log(2*pi) constant kept)."""
        E_Sinv = self.mf_expected_Sigma_inv()
        c0 = -0.5 * np.log(2 * np.pi) - 0.5 * self.mf_expected_logdet_Sigma() \
            - 0.5 * np.trace(E_Sinv.dot(self.mf_expected_mumuT()))
        return c0, E_Sinv.dot(self.mf_expected_mu()), -0.5 * E_Sinv
