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

    def _edge_total(self, network):
        """sum of A over all N^2 entries, counted on the device with the upload it belongs to (summed over row shards)."""
        net = network.device(self.dtype)
        if getattr(net, "comm", None) is not None:
            return float(net.comm.allreduce_np(np.array([float(net.local.edge_count())]))[0])
        return float(net.edge_count())

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
        post_prec = 1.0 / self.sigma_mu0 + sum_omega
        sum_kappa = self._edge_total(network) - 0.5 * self.N * self.N
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


class _MeanFieldLogisticEigenModel(_LogisticEigenmodelBase):
    """eigenmodel.py:968-1311.  Fixed networks (E_A = A).  The N^2 work runs in libgk:
      _meanfieldupdate_Omega  <- gk_eig_mf_omega      (:1055-1064; N_samples activations per element, on the device)
      _meanfieldupdate_F      <- gk_eig_mf_update_F   (:1090-1124; serial in n)
      mu0 / lmbda sums        <- the omega kernel's tile sums, gk_eig_mf_pair_stats (:1066-1088, 1126-1157)
    `literal_h_features` (default True) keeps a reference quirk: _meanfieldupdate_F builds h_n from `self.F` (the Gibbs
    sample, :1113-1114) rather than from mf_mu_F; False uses mf_mu_F, following the update."""
    literal_h_features = True
    mf_n_samples = 1000                      # :1061

    def _mf_init(self):
        """:969-985 (called by the concrete class after the Gibbs constructor)."""
        N, D = self.N, self.D
        self.mf_mu_Z = np.zeros((N, N)) if N <= 4096 else None     # never read by the reference's updates
        self.mf_mu_mu0 = self.mu_mu_0
        self.mf_sigma_mu0 = self.sigma_mu0
        self.mf_mu_F = np.sqrt(self.sigma_F) * np.random.randn(N, D)
        self.mf_Sigma_F = np.tile((self.sigma_F * np.eye(D))[None, :, :], [N, 1, 1])
        if self.lmbda_given:
            self.mf_mu_lmbda = self.lmbda.copy()
            self.mf_Sigma_lmbda = 1e-16 * np.eye(D)
        else:
            self.mf_mu_lmbda = self.mu_lmbda.copy()
            self.mf_Sigma_lmbda = self.sigma_lmbda * np.eye(D)
        self._mf_seed = int(np.random.randint(0, 2 ** 31 - 1))
        self._mf_calls = 0

    def init_with_gibbs(self, gibbs_model):
        """:988-996"""
        self.mf_mu_lmbda = gibbs_model.lmbda.copy()
        self.mf_Sigma_lmbda = 1e-4 * np.eye(self.D)
        self.mf_mu_mu0 = gibbs_model.mu_0
        self.mf_sigma_mu0 = 1e-4
        self.mf_mu_F = gibbs_model.F.copy()
        self.mf_Sigma_F = np.tile((1e-4 * np.eye(self.D))[None, :, :], [self.N, 1, 1])

    # ---- the updates ----------------------------------------------------------------------------------------
    def meanfieldupdate(self, network, stepsize_mu0=1.0):
        """:998-1022 (order: Omega, mu0, F, lmbda)."""
        network = as_network(network)
        sums = self._meanfieldupdate_Omega(network)
        self._meanfieldupdate_mu0(network, sums, stepsize=stepsize_mu0)
        self._meanfieldupdate_F(network)
        if not self.lmbda_given:
            self._meanfieldupdate_lmbda(network)

    def svi_step(self, network, minibatchfrac, stepsize):
        """:1024-1053: the stepsize reaches the mu0 update only."""
        self.meanfieldupdate(network, stepsize_mu0=stepsize)

    def _meanfieldupdate_Omega(self, network):
        """:1055-1064.  Leaves S = E_O + E_O^T on the device; returns (sum E_O, sum_{n != m} E_O * E_FdotF)."""
        dev = self._device_state(network)
        self._mf_calls += 1
        return dev.mf_omega(self.mf_mu_F, self.mf_Sigma_F, self.mf_expected_llT(), self.mf_mu_lmbda, self.mf_mu_mu0,
                            self.mf_sigma_mu0, self.mf_n_samples, (self._mf_seed + 7919 * self._mf_calls) % (2 ** 63))

    def _sum_kappa(self, network):
        """sum of kappa(E_A) = A - 1/2 over all N^2 entries (:1083)."""
        return self._edge_total(network) - 0.5 * self.N * self.N

    def _meanfieldupdate_mu0(self, network, sums, stepsize=1.0):
        """:1066-1088"""
        sum_EO, sum_EOF = sums
        post_prec = 1.0 / self.sigma_mu0 + sum_EO
        post_mean_dot_prec = self.mu_mu_0 / self.sigma_mu0 + self._sum_kappa(network) - sum_EOF
        self.mf_sigma_mu0 = (1 - stepsize) * self.mf_sigma_mu0 + stepsize * 1.0 / post_prec
        self.mf_mu_mu0 = (1 - stepsize) * self.mf_mu_mu0 + stepsize * self.mf_sigma_mu0 * post_mean_dot_prec

    def _meanfieldupdate_F(self, network, stepsize=1.0, want_conditionals=False):
        """:1090-1124"""
        dev = self._device_state(network)
        F_lin = self.F if self.literal_h_features else None
        self.mf_mu_F, self.mf_Sigma_F, prec, h = dev.mf_update_F(
            self.mf_mu_F, self.mf_Sigma_F, F_lin, self.mf_expected_llT(), self.mf_mu_lmbda, self.mf_mu_mu0, self.sigma_F,
            stepsize, want_conditionals)
        assert np.all(np.isfinite(self.mf_mu_F)) and np.all(np.isfinite(self.mf_Sigma_F))
        return prec, h

    def _meanfieldupdate_lmbda(self, network, stepsize=1.0):
        """:1126-1157"""
        dev = self._device_state(network)
        st = dev.mf_pair_stats(self.mf_mu_F, self.mf_Sigma_F)
        post_prec = 1.0 / self.sigma_lmbda * np.eye(self.D) + st["Pq"]
        post_mean_dot_prec = self.mu_lmbda / self.sigma_lmbda + st["hK"] - self.mf_mu_mu0 * st["hS"]
        self.mf_Sigma_lmbda = (1.0 - stepsize) * self.mf_Sigma_lmbda + stepsize * np.linalg.inv(post_prec)
        self.mf_mu_lmbda = (1.0 - stepsize) * self.mf_mu_lmbda + stepsize * self.mf_Sigma_lmbda.dot(post_mean_dot_prec)

    def get_vlb(self):
        """:1262-1298.  Prior and entropy terms of q(mu0), q(F), q(lmbda) -- N * D^2 host arithmetic, no N^2 work; the
        reference leaves the likelihood term out (the entropy of q(omega) is unknown).  Gaussian.negentropy's constant is
        -0.5 log 2 pi whatever D (internals/distributions.py:302), kept."""
        D = self.D
        c = -0.5 * np.log(2 * np.pi)

        def cross(mu, sig_iso, E_x, E_xxT):           # Gaussian(mu, sig_iso I).negentropy(E_x, E_xxT)
            return c - 0.5 * D * np.log(sig_iso) - 0.5 * np.trace(E_xxT) / sig_iso + E_x.dot(mu) / sig_iso - 0.5 * mu.dot(mu) / sig_iso

        def own(mu, Sigma):                           # Gaussian(mu, Sigma).negentropy()
            Si = np.linalg.inv(Sigma)
            mm = np.outer(mu, mu)
            return c - 0.5 * np.linalg.slogdet(Sigma)[1] - 0.5 * np.trace(Si.dot(Sigma + mm)) + mu.dot(Si).dot(mu) - 0.5 * np.trace(Si.dot(mm))

        s0, m0 = self.sigma_mu0, self.mu_mu_0
        vlb = c - 0.5 * np.log(s0) - 0.5 / s0 * self.mf_expected_mu0sq() + self.mf_mu_mu0 * m0 / s0 - 0.5 * m0 ** 2 / s0
        qs, qm = self.mf_sigma_mu0, self.mf_mu_mu0
        vlb -= c - 0.5 * np.log(qs) - 0.5 / qs * (qs + qm ** 2) + qm * qm / qs - 0.5 * qm ** 2 / qs
        zero = np.zeros(D)
        for n in range(self.N):
            vlb += cross(zero, self.sigma_F, self.mf_mu_F[n], self.mf_expected_ffT(n))
            vlb -= own(self.mf_mu_F[n], self.mf_Sigma_F[n])
        if not self.lmbda_given:
            vlb += cross(np.asarray(self.mu_lmbda, float), self.sigma_lmbda, self.mf_mu_lmbda, self.mf_expected_llT())
            vlb -= own(self.mf_mu_lmbda, self.mf_Sigma_lmbda)
        return float(vlb)

    # ---- expectations (:1159-1248; N x N on the host: reporting only) --------------------------------------------
    def mf_expected_mu(self):
        E_mu = self.mf_mu_mu0 + (self.mf_mu_F * self.mf_mu_lmbda[None, :]).dot(self.mf_mu_F.T)
        np.fill_diagonal(E_mu, self.mf_mu_mu0)
        return E_mu

    def mf_expected_mu0sq(self):
        return self.mf_sigma_mu0 + self.mf_mu_mu0 ** 2

    def mf_expected_ffT(self, n):
        return self.mf_Sigma_F[n, :, :] + np.outer(self.mf_mu_F[n, :], self.mf_mu_F[n, :])

    def mf_expected_llT(self):
        return self.mf_Sigma_lmbda + np.outer(self.mf_mu_lmbda, self.mf_mu_lmbda)

    def mf_expected_musq(self):
        N, D = self.N, self.D
        E = np.zeros((N, N)) + self.mf_expected_mu0sq()
        E += 2 * self.mf_mu_mu0 * (self.mf_mu_F * self.mf_mu_lmbda[None, :]).dot(self.mf_mu_F.T)
        ffT = (self.mf_Sigma_F + self.mf_mu_F[:, :, None] * self.mf_mu_F[:, None, :]).reshape(N, D * D)
        E += (ffT * self.mf_expected_llT().reshape(D * D)[None, :]).dot(ffT.T)
        np.fill_diagonal(E, self.mf_expected_mu0sq())
        return E

    def mf_variance_mu(self):
        return self.mf_expected_musq() - self.mf_expected_mu() ** 2

    def mf_sample_mus(self, N_samples=10):
        E_mu = self.mf_expected_mu()[None, :, :]
        std_mu = np.sqrt(self.mf_variance_mu())[None, :, :]
        return E_mu + std_mu * np.random.randn(N_samples, self.N, self.N)

    def mf_expected_log_p(self):
        Ps = np.clip(logistic(self.mf_sample_mus()), 1e-16, 1 - 1e-16)
        return np.log(Ps).mean(0)

    def mf_expected_log_notp(self):
        Ps = np.clip(logistic(self.mf_sample_mus()), 1e-16, 1 - 1e-16)
        return np.log(1 - Ps).mean(0)

    def expected_log_likelihood(self, network):
        A = as_network(network).A
        return (A * self.mf_expected_log_p() + (1 - A) * self.mf_expected_log_notp()).sum()

    def resample_from_mf(self):
        """:1300-1311"""
        for n in range(self.N):
            self.F[n, :] = np.random.multivariate_normal(self.mf_mu_F[n, :], self.mf_Sigma_F[n, :, :])
        self.mu_0 = np.random.normal(self.mf_mu_mu0, np.sqrt(self.mf_sigma_mu0))
        if not self.lmbda_given:
            self.lmbda = np.random.multivariate_normal(self.mf_mu_lmbda, self.mf_Sigma_lmbda)


class LogisticEigenmodel(_GibbsLogisticEigenmodel, _MeanFieldLogisticEigenModel):
    """eigenmodel.py:1313."""

    def __init__(self, N, D, **kwargs):
        super(LogisticEigenmodel, self).__init__(N, D, **kwargs)
        self._mf_init()


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


class _MeanFieldProbitEigenModel(_ProbitEigenmodelBase):
    """eigenmodel.py:382-770.  Fixed networks (mf_A = A).  The N^2 work runs in libgk:
      _meanfieldupdate_Z + mf_expected_Z  <- gk_eig_mf_probit_Z   (:570-585, 519-534; E[Z] per tile, never an N x N host array)
      _meanfieldupdate_F                  <- gk_eig_mf_update_F   (:610-641, serial in n; probit link: S == 2, h from E[Z] - E[mu_0])
      mu0 / lmbda sums                    <- the E[Z] kernel's tile sums, gk_eig_mf_pair_stats (:587-608, 643-674)
    `literal_truncnorm_normalizer` (default True) keeps a reference quirk: expected_truncnorm normalises by
    normal_cdf(ub) - normal_cdf(lb) of the RAW bounds (internals/utils.py:93), i.e. by 1/2 whatever mu; False uses the
    truncated-normal mean."""
    literal_truncnorm_normalizer = True

    def _mf_init(self):
        """:383-399 (called by the concrete class after the Gibbs constructor)."""
        N, D = self.N, self.D
        self.mf_mu_mu0 = self.mu_mu_0
        self.mf_sigma_mu0 = self.sigma_mu0
        self.mf_mu_F = np.sqrt(self.sigma_F) * np.random.randn(N, D)
        self.mf_Sigma_F = np.tile((self.sigma_F * np.eye(D))[None, :, :], [N, 1, 1])
        if self.lmbda_given:
            self.mf_mu_lmbda = self.lmbda.copy()
            self.mf_Sigma_lmbda = 1e-16 * np.eye(D)
        else:
            self.mf_mu_lmbda = self.mu_lmbda.copy()
            self.mf_Sigma_lmbda = self.sigma_lmbda * np.eye(D)

    init_with_gibbs = _MeanFieldLogisticEigenModel.init_with_gibbs            # :402-410, same body as :988-996

    # ---- expectations (:472-568; N x N on the host: reporting only) ------------------------------------------------
    mf_expected_mu = _MeanFieldLogisticEigenModel.mf_expected_mu
    mf_expected_musq = _MeanFieldLogisticEigenModel.mf_expected_musq
    mf_variance_mu = _MeanFieldLogisticEigenModel.mf_variance_mu
    mf_expected_mu0sq = _MeanFieldLogisticEigenModel.mf_expected_mu0sq
    mf_expected_ffT = _MeanFieldLogisticEigenModel.mf_expected_ffT
    mf_expected_llT = _MeanFieldLogisticEigenModel.mf_expected_llT
    resample_from_mf = _MeanFieldLogisticEigenModel.resample_from_mf          # :758-769

    _mf_Z_snapshot = None

    @property
    def mf_mu_Z(self):
        """The N x N array _meanfieldupdate_Z sets (:570-585), rebuilt on demand from the (E[mu_0], E[F], E[lmbda]) it
        was computed from -- the device recomputes it per tile and never stores it.  Zeros before the first update (:388)."""
        if self._mf_Z_snapshot is None:
            return np.zeros((self.N, self.N))
        mu0, F, lam = self._mf_Z_snapshot
        Mu = mu0 + (F * lam[None, :]).dot(F.T)
        np.fill_diagonal(Mu, mu0)
        return Mu

    def _truncnorm_mean(self, mu, edge):
        phi = np.exp(-0.5 * mu * mu) / np.sqrt(2 * np.pi)
        if self.literal_truncnorm_normalizer:
            return np.where(edge, mu + 2.0 * phi, mu - 2.0 * phi)
        from scipy.special import erfc
        cdf = 0.5 * erfc(np.where(edge, -mu, mu) / np.sqrt(2.0))
        r = phi / np.maximum(cdf, 1e-300)
        return np.where(edge, mu + r, mu - r)

    def mf_expected_Z(self, A):
        """:519-534 for a fixed network (host, N x N: reporting / tests only)."""
        return self._truncnorm_mean(self.mf_mu_Z, np.asarray(A) != 0)

    def mf_expected_log_p(self):
        from scipy.special import erfc
        return np.log(1.0 - 0.5 * erfc(self.mf_mu_Z / np.sqrt(2.0)))          # :440-447

    def mf_expected_log_notp(self):
        from scipy.special import erfc
        return np.log(0.5 * erfc(self.mf_mu_Z / np.sqrt(2.0)))                # :449-455

    def expected_log_likelihood(self, network):
        A = as_network(network).A
        return (A * self.mf_expected_log_p() + (1 - A) * self.mf_expected_log_notp()).sum()

    # ---- the updates --------------------------------------------------------------------------------------------
    def meanfieldupdate(self, network):
        """:412-438 (order: Z, mu0, F, lmbda).  E[Z] is computed once, from the q(mu_0), q(F), q(lmbda) at entry, and is
        held fixed through the three updates, as the reference does (mf_mu_Z is set at :570-585 and not touched again)."""
        network = as_network(network)
        sum_EZ = self._meanfieldupdate_Z(network)
        self._meanfieldupdate_mu0(sum_EZ)
        self._meanfieldupdate_F(network)
        if not self.lmbda_given:
            self._meanfieldupdate_lmbda(network)

    def _meanfieldupdate_Z(self, network):
        """:570-585 + :519-534.  Leaves E[Z] + E[Z]^T on the device; returns the sum of E[Z] over n != m."""
        dev = self._device_state(as_network(network))
        self._mf_Z_snapshot = (float(self.mf_mu_mu0), self.mf_mu_F.copy(), self.mf_mu_lmbda.copy())
        return dev.mf_probit_Z(self.mf_mu_F, self.mf_mu_lmbda, self.mf_mu_mu0, not self.literal_truncnorm_normalizer)

    def _meanfieldupdate_mu0(self, sum_EZ):
        """:587-608: Zcent = E[Z] - E[F] diag(E[lmbda]) E[F]^T off the diagonal."""
        F, lam = self.mf_mu_F, self.mf_mu_lmbda
        cs = F.sum(0)
        sum_FLF = float(((cs * lam) * cs).sum() - ((F * lam) * F).sum())      # sum_{n != m} f_n^T Lambda f_m
        post_prec = 1.0 / self.sigma_mu0 + self.N ** 2 - self.N
        post_mean_dot_prec = self.mu_mu_0 / self.sigma_mu0 + sum_EZ - sum_FLF
        self.mf_sigma_mu0 = 1.0 / post_prec
        self.mf_mu_mu0 = self.mf_sigma_mu0 * post_mean_dot_prec

    def _meanfieldupdate_F(self, network, want_conditionals=False):
        """:610-641 (h built from mf_mu_F as it is updated, :632-633)."""
        dev = self._device_state(network)
        self.mf_mu_F, self.mf_Sigma_F, prec, h = dev.mf_update_F(
            self.mf_mu_F, self.mf_Sigma_F, None, self.mf_expected_llT(), self.mf_mu_lmbda, self.mf_mu_mu0, self.sigma_F,
            1.0, want_conditionals)
        assert np.all(np.isfinite(self.mf_mu_F)) and np.all(np.isfinite(self.mf_Sigma_F))
        return prec, h

    def _meanfieldupdate_lmbda(self, network):
        """:643-674"""
        dev = self._device_state(network)
        st = dev.mf_pair_stats(self.mf_mu_F, self.mf_Sigma_F)
        post_prec = 1.0 / self.sigma_lmbda * np.eye(self.D) + st["Pq"]
        post_mean_dot_prec = self.mu_lmbda / self.sigma_lmbda + st["hK"] - self.mf_mu_mu0 * st["hS"]
        self.mf_Sigma_lmbda = np.linalg.inv(post_prec)
        self.mf_mu_lmbda = self.mf_Sigma_lmbda.dot(post_mean_dot_prec)

    def get_vlb(self, network=None):
        """:688-756.  The prior / entropy terms of q(mu0), q(F), q(lmbda) are the logistic class's (:1262-1298, same code);
        with a network, the E[ln p(z | mu)] - E[ln q(z)] terms of :694-727 are added (N x N host arithmetic, as in the
        reference: get_vlb is a diagnostic, not on the sweep path)."""
        vlb = _MeanFieldLogisticEigenModel.get_vlb(self)
        if network is None:
            return vlb
        from scipy.special import erfc
        A = np.asarray(as_network(network).A) != 0
        mu = self.mf_mu_Z
        phi = np.exp(-0.5 * mu * mu) / np.sqrt(2 * np.pi)
        Zn = np.where(A, 1.0 - 0.5 * erfc(mu / np.sqrt(2.0)), 0.5 * erfc(mu / np.sqrt(2.0)))     # P(z in its half line)
        Zn = np.maximum(Zn, 1e-300)
        E_z = np.where(A, mu + phi / Zn, mu - phi / Zn)
        E_zsq = 1.0 + mu * E_z                                # truncated at 0: E[z^2] = 1 + mu E[z]
        E_mu, E_musq = self.mf_expected_mu(), self.mf_expected_musq()
        c = -0.5 * np.log(2 * np.pi)
        cross = c - 0.5 * E_zsq + E_z * E_mu - 0.5 * E_musq   # ScalarGaussian().negentropy(E_x, E_xsq, E_mu, E_musq), sigmasq = 1
        own = c - 0.5 * (E_zsq - 2 * mu * E_z + mu * mu) - np.log(Zn)          # TruncatedScalarGaussian(mu, 1).negentropy()
        return float(vlb + (cross - own).sum())


class ProbitEigenmodel(_GibbsProbitEigenmodel, _MeanFieldProbitEigenModel):
    """eigenmodel.py:772."""

    def __init__(self, N, D, **kwargs):
        super(ProbitEigenmodel, self).__init__(N, D, **kwargs)
        self._mf_init()


Eigenmodel = ProbitEigenmodel        # the name test/test_probit_eigenmodel_geweke.py:8 imports
