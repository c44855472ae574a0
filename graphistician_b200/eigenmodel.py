"""Eigenmodel (Hoff 2008) with a logistic link and Polya-gamma augmentation -- old-API class shell.

Mirrors ``graphistician.eigenmodel`` (``_EigenmodelBase`` :14-135, ``_LogisticEigenmodelBase`` :776-820,
``_GibbsLogisticEigenmodel`` :822-965, ``LogisticEigenmodel`` :1313).  Every O(N^2) step of
``resample(network)`` runs in libgk:

  Omega  <- gk_eig_pg_draw      K3   (replaces ppg.pgdrawvpar over Mu.ravel(), :859-874)
  F      <- gk_eig_resample_F   K4+K5 (replaces the N^2 Python loop of :876-908, same serial semantics)
  mu_0, lmbda: O(1)/O(D^2) host draws (np.random.normal / np.random.multivariate_normal, as the reference calls
               them) from pair statistics the feature sweep returns (K6), :910-965.

RNG: the reference seeds pypolyagamma's private generators from ``np.random`` (:825-827); here one 31-bit seed
is drawn from ``np.random`` at construction and keys the device's counter-based generator, so
``np.random.seed(s)`` makes a run reproducible.  The per-node Gaussian draws use N x D standard normals taken
from ``np.random`` once per sweep with the fixed map f = L^-T (L^-1 h + z) instead of N calls to
``np.random.multivariate_normal`` (:908) -- same conditional, different sample path (DESIGN.md "Parity").
"""
import copy

import numpy as np
from scipy.stats import norm

from .old_abstractions import as_network, NetworkDistribution
from .device import DeviceEigenmodel


def logistic(x):
    """internals/utils.py:5-6."""
    return 1.0 / (1 + np.exp(-x))


class _EigenmodelBase(NetworkDistribution):
    """eigenmodel.py:14-135."""

    def __init__(self, N, D, p=0.5, sigma_mu0=1.0, sigma_F=1.0, lmbda=None, mu_lmbda=0, sigma_lmbda=1.0,
                 dtype="f32"):
        self.N, self.D = N, D
        assert p >= 0 and p <= 1.0, "p must be a baseline probability of connection"
        self.p = p
        self.mu_mu_0 = norm(0, 1).ppf(p)
        self.mu_0 = 0
        assert sigma_mu0 >= 0 and sigma_F >= 0 and sigma_lmbda >= 0, "q,r,s must be positive variances"
        self.sigma_mu0, self.sigma_F = sigma_mu0, sigma_F
        self.F = np.zeros((N, D))
        self.dtype = dtype                      # device storage of Omega: "f32" (default) or "f64" (parity mode)
        if lmbda is not None:
            assert lmbda.shape == (self.D,)
            self.lmbda_given = True
            self.lmbda = lmbda
        else:
            assert np.isscalar(mu_lmbda) or mu_lmbda.shape == (self.D,)
            self.lmbda_given = False
            self.mu_lmbda = mu_lmbda * np.ones(self.D)
            self.sigma_lmbda = sigma_lmbda
            self.lmbda = self.mu_lmbda.copy()
        self._dev = None

    @property
    def Mu(self):
        """eigenmodel.py:75-83 (N x N on the host: reporting only, never on the sweep path)."""
        Mu = self.mu_0 + (self.F * self.lmbda[None, :]).dot(self.F.T)
        np.fill_diagonal(Mu, self.mu_0)
        return Mu

    def log_prior(self):
        """eigenmodel.py:92-108."""
        lp = -0.5 * (self.F * self.F / self.sigma_F).sum()
        lp += -0.5 * (self.mu_0 - self.mu_mu_0) ** 2 / self.sigma_mu0
        if not self.lmbda_given:
            lp += -0.5 * (self.lmbda * self.lmbda / self.sigma_lmbda).sum()
        return lp

    _link = 0

    def log_likelihood(self, x):
        """eigenmodel.py:110-129, fused on the device (K9): P is never materialised."""
        network = as_network(x)
        assert network.A.shape == (self.N, self.N)
        dev = self._device_state(network)
        dev.set_params(self.F, self.mu_0, self.lmbda)
        return dev.log_likelihood(self._link)

    def log_probability(self, A):
        return self.log_likelihood(A) + self.log_prior()

    def rvs(self, size=[]):
        """eigenmodel.py:134-148."""
        return np.random.rand(self.N, self.N) < self.P

    def compute_optimal_rotation(self, F_true, lmbda_true):
        """eigenmodel.py:151-174."""
        from scipy.linalg import orthogonal_procrustes
        return orthogonal_procrustes(self.F / np.sqrt(abs(self.lmbda[None, :])),
                                     F_true / np.sqrt(abs(lmbda_true[None, :])))[0]

    # ---- device state ----------------------------------------------------------------------------------
    def _device_state(self, network):
        net = network.device(self.dtype)
        if self._dev is None or self._dev[0] is not net:
            if self._dev is not None:
                self._dev[1].close()
            self._dev = (net, DeviceEigenmodel(net, self.D))
        return self._dev[1]

    def __deepcopy__(self, memo):
        cls = self.__class__
        new = cls.__new__(cls)
        memo[id(self)] = new
        for k, v in self.__dict__.items():
            setattr(new, k, None if k == "_dev" else copy.deepcopy(v, memo))
        return new


class _LogisticEigenmodelBase(_EigenmodelBase):
    """eigenmodel.py:776-820."""
    _link = 0

    @property
    def P(self):
        return np.clip(logistic(self.Mu), 1e-16, 1 - 1e-16)

    def kappa(self, A):
        return A - 0.5


class _GibbsLogisticEigenmodel(_LogisticEigenmodelBase):
    def __init__(self, N, D, **kwargs):
        """eigenmodel.py:823-834."""
        super(_GibbsLogisticEigenmodel, self).__init__(N, D, **kwargs)
        self._pg_seed = int(np.random.randint(0, 2 ** 31 - 1))
        self._pg_calls = 0
        self.F = np.sqrt(self.sigma_F) * np.random.randn(self.N, self.D)
        self.mu_0 = self.mu_mu_0 + np.sqrt(self.sigma_mu0) * np.random.randn()
        self.resample()

    def resample(self, networks=[], omega=None, normals=None, want_conditionals=False):
        """eigenmodel.py:836-857.  ``omega`` (N x N) / ``normals`` (N x D) inject the random draws (parity tests);
        ``want_conditionals`` keeps prec_n / h_n of every node in ``self._last_conditionals``."""
        if not isinstance(networks, list):
            networks = [networks]
        if len(networks) == 0:
            return self._resample_prior()
        if len(networks) > 1:
            raise NotImplementedError("this build resamples given one network (the reference's demos and tests pass one)")
        network = as_network(networks[0])
        dev = self._device_state(network)
        dev.set_params(self.F, self.mu_0, self.lmbda)
        # Omega (K3)
        if omega is None:
            sum_omega = dev.pg_draw((self._pg_seed << 20) + self._pg_calls)
            self._pg_calls += 1
        else:
            omega = np.asarray(omega, dtype=np.float64)
            S = omega + omega.T
            if dev.comm is not None:                     # row-sharded: this rank's rows only
                S = S[dev.local.row0:dev.local.row0 + dev.local.nrows]
            dev.set_S(S)
            sum_omega = omega.sum()
        # F (K4 + K5); the pair statistics for mu_0 and lmbda come back with it (K6)
        z = np.random.randn(self.N, self.D) if normals is None else normals
        self.F, st, prec, h = dev.resample_F(self.sigma_F, z, want_conditionals)
        self._last_conditionals = dict(prec=prec, h=h, sum_omega=sum_omega, stats=st)
        self._resample_mu_0(network, sum_omega, st)
        if not self.lmbda_given:
            self._resample_lmbda(st)

    def _resample_prior(self):
        """``resample()`` with no data (the constructor's call, :834): every sum is empty."""
        cov = self.sigma_F * np.eye(self.D)
        for n in range(self.N):
            self.F[n, :] = np.random.multivariate_normal(np.zeros(self.D), cov)
        self.mu_0 = np.random.normal(self.mu_mu_0, np.sqrt(self.sigma_mu0))
        if not self.lmbda_given:
            self.lmbda = np.random.multivariate_normal(self.mu_lmbda, self.sigma_lmbda * np.eye(self.D))

    def _mu0_conditional(self, network, sum_omega, st):
        """eigenmodel.py:914-927: (post_prec, post_mean_dot_prec)."""
        if not hasattr(network, "_sum_A"):
            network._sum_A = float(np.count_nonzero(network.A))
        post_prec = 1.0 / self.sigma_mu0 + sum_omega
        sum_kappa = network._sum_A - 0.5 * self.N * self.N
        h = self.mu_mu_0 / self.sigma_mu0 + sum_kappa - float(np.dot(self.lmbda, st["hS"]))
        return post_prec, h

    def _resample_mu_0(self, network, sum_omega, st):
        """eigenmodel.py:910-932."""
        post_prec, h = self._mu0_conditional(network, sum_omega, st)
        post_var = 1.0 / post_prec
        self._last_conditionals["mu0"] = (post_prec, h)
        self.mu_0 = np.random.normal(post_var * h, np.sqrt(post_var))

    def _resample_lmbda(self, st):
        """eigenmodel.py:934-965 from the pair statistics (uses the NEW mu_0, as :945 does)."""
        post_prec = 1.0 / self.sigma_lmbda * np.eye(self.D) + st["Pq"]
        h = self.mu_lmbda / self.sigma_lmbda + st["hK"] - self.mu_0 * st["hS"]
        self._last_conditionals["lmbda"] = (post_prec, h)
        post_cov = np.linalg.inv(post_prec)
        self.lmbda = np.random.multivariate_normal(post_cov.dot(h), post_cov)


class LogisticEigenmodel(_GibbsLogisticEigenmodel):
    """eigenmodel.py:1313 (Gibbs part; the mean-field mixin is a later row of SURVEY.md 8a)."""
    pass


# ------------------------------------------------------------------------------------------------------
# probit link (eigenmodel.py:222-379, 772)
# ------------------------------------------------------------------------------------------------------
class _ProbitEigenmodelBase(_EigenmodelBase):
    """eigenmodel.py:222-236."""
    _link = 1

    @property
    def P(self):
        from scipy.special import erf
        return np.clip(0.5 * (1 + erf(self.Mu / np.sqrt(2.0))), 1e-16, 1 - 1e-16)


class _GibbsProbitEigenmodel(_ProbitEigenmodelBase):
    """eigenmodel.py:238-379.  Every O(N^2) step of ``resample(A)`` runs in libgk:

      Z      <- gk_eig_tn_draw      K7    (replaces sample_truncnorm over N x N, :268-287)
      F      <- gk_eig_resample_F   K4+K5 in probit mode (S == 2, h from Z - mu_0; replaces the N^2 loop of :289-321)
      mu_0, lmbda: host draws from the pair statistics of the feature sweep (:323-379)."""

    def __init__(self, N, D, **kwargs):
        super(_GibbsProbitEigenmodel, self).__init__(N, D, **kwargs)
        self._tn_seed = int(np.random.randint(0, 2 ** 31 - 1))
        self._tn_calls = 0
        self.resample()

    def resample(self, data=[], uniforms=None, normals=None, want_conditionals=False):
        """eigenmodel.py:244-266.  ``uniforms`` (N x N) / ``normals`` (N x D) inject the random draws (parity tests)."""
        if not isinstance(data, list):
            data = [data]
        if len(data) == 0:
            return self._resample_prior()
        if len(data) > 1:
            raise NotImplementedError("this build resamples given one network (the reference's demos and tests pass one)")
        network = as_network(data[0])
        dev = self._device_state(network)
        dev.set_params(self.F, self.mu_0, self.lmbda)
        if uniforms is None:
            sum_z = dev.tn_draw((self._tn_seed << 20) + self._tn_calls)
            self._tn_calls += 1
        else:
            from .device import truncnorm_draw
            Z = truncnorm_draw(self.Mu, network.A, uniforms, dev.ctx)
            np.fill_diagonal(Z, 0.0)
            sum_z = Z.sum()
            Zs = Z + Z.T
            if dev.comm is not None:
                Zs = Zs[dev.local.row0:dev.local.row0 + dev.local.nrows]
            dev.set_Zsym(Zs)
        z = np.random.randn(self.N, self.D) if normals is None else normals
        self.F, st, prec, h = dev.resample_F(self.sigma_F, z, want_conditionals)
        self._last_conditionals = dict(prec=prec, h=h, sum_z=sum_z, stats=st)
        # mu_0 (:323-347): Zcent = Z - F diag(lmbda) F^T off the diagonal; sum_{n1 != n2} f1 Lambda f2 = lmbda . hS
        post_prec = 1.0 / self.sigma_mu0 + self.N ** 2 - self.N
        hm = self.mu_mu_0 / self.sigma_mu0 + sum_z - float(np.dot(self.lmbda, st["hS"]))
        self._last_conditionals["mu0"] = (post_prec, hm)
        self.mu_0 = np.random.normal(hm / post_prec, np.sqrt(1.0 / post_prec))
        if not self.lmbda_given:
            # lmbda (:349-379) with the NEW mu_0: zcent pair sums = hK - mu_0 hS
            pp = 1.0 / self.sigma_lmbda * np.eye(self.D) + st["Pq"]
            hl = self.mu_lmbda / self.sigma_lmbda + st["hK"] - self.mu_0 * st["hS"]
            self._last_conditionals["lmbda"] = (pp, hl)
            cov = np.linalg.inv(pp)
            self.lmbda = np.random.multivariate_normal(cov.dot(hl), cov)

    def _resample_prior(self):
        """``resample()`` with no data (the constructor's call, :242): every data sum is empty."""
        cov = self.sigma_F * np.eye(self.D)
        for n in range(self.N):
            self.F[n, :] = np.random.multivariate_normal(np.zeros(self.D), cov)
        self.mu_0 = np.random.normal(self.mu_mu_0, np.sqrt(self.sigma_mu0))
        if not self.lmbda_given:
            self.lmbda = np.random.multivariate_normal(self.mu_lmbda, self.sigma_lmbda * np.eye(self.D))


class ProbitEigenmodel(_GibbsProbitEigenmodel):
    """eigenmodel.py:772 (Gibbs part)."""
    pass


Eigenmodel = ProbitEigenmodel        # the name test/test_probit_eigenmodel_geweke.py:8 imports
