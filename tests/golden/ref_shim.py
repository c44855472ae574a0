"""
ref_shim.py -- load the REFERENCE'S OWN source files from /root/reference under Python 3 and execute
their method bodies.  Build-container only (the GPU box has no /root/reference); used by
tests/golden/make_golden.py to write the ``ref_*.npz`` fixtures that pin oracle/ and the CUDA path.

Nothing is copied into the repo: every file is read where it lies, rewritten IN MEMORY by three mechanical,
value-preserving Python-2 -> Python-3 rewrites, compiled and executed:

  1. tuple parameters      ``def f(self, (A,W)):``          -> ``def f(self, _t0): (A,W) = _t0``
                           ``(n,A,f)=(None,None,None)``      -> ``_t0=(None,None,None)`` + unpack
  2. print statements      ``print "x", y``                  -> ``print("x", y)``
  3. boolean ``True - eye``  (adjacency.py:628; NumPy >= 1.13 refuses bool subtraction; the old result is
                           the logical NOT)                  -> ``~eye``

plus ``builtins.xrange = range`` and the NumPy aliases removed in NumPy 2 (np.Inf, np.int, np.float, np.bool).

Third-party packages that are absent here are replaced by stubs restating their published
behaviour -- each stub says so:

  pybasicbayes.abstractions     empty base classes (Distribution, GibbsSampling, MeanField, MeanFieldSVI, ...)
  pybasicbayes.distributions    Gaussian (NIW), GaussianFixedMean, GaussianFixedCov      (oracle.niw recollection)
  pybasicbayes.util.stats       sample_discrete_from_log  (records the lp it is handed; uniform injectable)
  pypolyagamma                  PyPolyaGamma / pgdrawvpar  (omegas injectable; else oracle/pg.c)
  hips.inference.hmc            hmc  (oracle.latent.hmc recollection; momentum / log-uniform from np.random or injected)
  hips.inference.*sample        discrete_sample, log_sum_exp_sample
  autograd                      autograd.numpy = numpy; grad = complex-step derivative (exact to round-off for
                                the analytic functions the reference differentiates)
  matplotlib, scipy.misc.logsumexp, numpy.core.umath_tests.inner1d   trivial aliases
  graphistician.old_abstractions   the module missing from the reference tree: minimal
                                network containers and empty base classes

The hooks in ``REC`` let a driver read the CONDITIONALS the reference's code computed (the lp vector given to
sample_discrete_from_log, the matrix given to np.linalg.inv, the mean / covariance given to
np.random.multivariate_normal) and inject the random draws.
"""
import builtins
import importlib.util
import os
import re
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF_ROOT = "/root/reference"
REF = os.path.join(REF_ROOT, "graphistician")
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


class _Rec(object):
    """Recording / injection hooks shared by the stubs."""

    def __init__(self):
        self.reset()

    def reset(self):
        self.lp = []              # lp vectors handed to sample_discrete_from_log / log_sum_exp_sample
        self.uniforms = None      # list of uniforms to consume (None -> np.random)
        self.inv_in = []          # matrices handed to np.linalg.inv (while `patched_numpy` is active)
        self.mvn = []             # (mean, cov) handed to np.random.multivariate_normal
        self.mvn_z = None         # list of standard-normal vectors to consume (None -> real draw)
        self.omegas = None        # array returned by pgdrawvpar (None -> C sampler)
        self.hmc_draws = None     # list of (p0, log_u) to consume (None -> np.random)
        self.hmc_log = []         # per hmc call: dict(step, accepted, log_ratio)
        self.gauss_data = []      # data handed to Gaussian.resample
        self.normal = []          # (mean, std) handed to np.random.normal
        self.normal_z = None


REC = _Rec()


# --------------------------------------------------------------------------------------------------
# source rewrites
# --------------------------------------------------------------------------------------------------
_DEF = re.compile(r"^(\s*)def\s+(\w+)\((.*)\):\s*$")


def _split_args(s):
    out, depth, cur = [], 0, ""
    for ch in s:
        if ch in "([{":
            depth += 1
        elif ch in ")]}":
            depth -= 1
        if ch == "," and depth == 0:
            out.append(cur); cur = ""
        else:
            cur += ch
    if cur.strip():
        out.append(cur)
    return out


def py2to3(src):
    lines = src.split("\n")
    out = []
    for line in lines:
        m = _DEF.match(line)
        if m and re.search(r"[(,]\s*\(", m.group(3) and "(" + m.group(3)):
            indent, name, args = m.groups()
            new_args, unpack = [], []
            for i, a in enumerate(_split_args(args)):
                a = a.strip()
                if a.startswith("("):
                    close = a.index(")")
                    tup, rest = a[:close + 1], a[close + 1:].strip()
                    nm = "_t%d" % i
                    new_args.append(nm + rest)
                    unpack.append("%s    %s = %s" % (indent, tup, nm))
                else:
                    new_args.append(a)
            out.append("%sdef %s(%s):" % (indent, name, ", ".join(new_args)))
            out.extend(unpack)
            continue
        m = re.match(r"^(\s*)print\s+(.*)$", line)
        if m and not m.group(2).startswith("("):
            out.append("%sprint(%s)" % (m.group(1), m.group(2).rstrip()))
            continue
        line = line.replace("True - np.eye(", "~np.eye(")
        out.append(line)
    return "\n".join(out)


# --------------------------------------------------------------------------------------------------
# stubs
# --------------------------------------------------------------------------------------------------
def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    parent, _, child = name.rpartition(".")
    if parent:
        if parent not in sys.modules:
            _mod(parent)
        setattr(sys.modules[parent], child, m)
    if not hasattr(m, "__path__"):
        m.__path__ = []
    return m


def _sample_discrete_from_log(p_log, axis=0, dtype=np.int32):
    """pybasicbayes.util.stats.sample_discrete_from_log (recollection): cumsum of exp(lp - max), one uniform
    scaled by the total, count of strictly smaller cumulative values."""
    p_log = np.asarray(p_log, dtype=np.float64)
    REC.lp.append(p_log.copy())
    cum = np.exp(p_log - p_log.max()).cumsum()
    u = REC.uniforms.pop(0) if REC.uniforms is not None else np.random.random()
    return dtype(np.sum(u * cum[-1] > cum))


def _log_sum_exp_sample(lnp):
    """hips.inference.log_sum_exp_sample (recollection): normalise with log-sum-exp, cdf = cumsum(p),
    first index with cdf >= u  (one np.random.rand())."""
    lnp = np.asarray(lnp, dtype=np.float64).ravel()
    REC.lp.append(lnp.copy())
    mx = lnp.max()
    p = np.exp(lnp - (mx + np.log(np.sum(np.exp(lnp - mx)))))
    u = REC.uniforms.pop(0) if REC.uniforms is not None else np.random.rand()
    cdf = np.cumsum(p)
    idx = int(np.searchsorted(cdf, u, side="left"))
    return min(idx, len(p) - 1)


def _discrete_sample(p, N=1):
    """hips.inference.discrete_sample (recollection): inverse cdf of a normalised vector."""
    p = np.asarray(p, dtype=np.float64).ravel()
    cdf = np.cumsum(p / p.sum())
    out = [min(int(np.searchsorted(cdf, REC.uniforms.pop(0) if REC.uniforms is not None else np.random.rand(),
                                   side="left")), len(p) - 1) for _ in range(N)]
    return out[0] if N == 1 else np.array(out)


def _install_pybasicbayes():
    from oracle import niw

    class Distribution(object):
        pass

    names = ["Distribution", "BayesianDistribution", "GibbsSampling", "MeanField", "MeanFieldSVI", "Collapsed",
             "MaxLikelihood", "MAP", "Model", "ModelGibbsSampling", "ModelMeanField"]
    classes = {n: type(n, (object,), {}) for n in names}
    _mod("pybasicbayes")
    _mod("pybasicbayes.abstractions", **classes)
    _mod("pybasicbayes.util")
    _mod("pybasicbayes.util.stats", sample_discrete_from_log=_sample_discrete_from_log,
         sample_niw=niw.sample_niw, sample_invwishart=niw.sample_invwishart)

    class Gaussian(object):
        """pybasicbayes.distributions.Gaussian (recollection, ~2015 natural-parameter version): NIW prior
        (mu_0, sigma_0, kappa_0, nu_0); mean-field factor (mu_mf, sigma_mf, kappa_mf, nu_mf)."""

        def __init__(self, mu=None, sigma=None, mu_0=None, sigma_0=None, kappa_0=None, nu_0=None):
            self.mu, self.sigma = mu, sigma
            self.mu_0 = self.mu_mf = mu_0
            self.sigma_0 = self.sigma_mf = sigma_0
            self.kappa_0 = self.kappa_mf = kappa_0
            self.nu_0 = self.nu_mf = nu_0
            if mu is None and sigma is None and not any(v is None for v in (mu_0, sigma_0, kappa_0, nu_0)):
                self.resample()                                   # initialise from the prior

        @property
        def D(self):
            return (self.mu if self.mu is not None else self.mu_0).shape[0]

        @property
        def sigma_chol(self):
            return np.linalg.cholesky(self.sigma)

        @staticmethod
        def _standard_to_natural(mu_mf, sigma_mf, kappa_mf, nu_mf):
            D = sigma_mf.shape[0]
            out = np.zeros((D + 2, D + 2))
            out[:D, :D] = sigma_mf + kappa_mf * np.outer(mu_mf, mu_mf)
            out[:D, -2] = out[-2, :D] = kappa_mf * mu_mf
            out[-2, -2] = kappa_mf
            out[-1, -1] = nu_mf + 2 + D
            return out

        @staticmethod
        def _natural_to_standard(nat):
            D = nat.shape[0] - 2
            A, b, kappa, nu = nat[:D, :D], nat[:D, -2], nat[-2, -2], nat[-1, -1]
            mu = b / kappa
            return mu, A - np.outer(b, mu), kappa, nu - 2 - D

        @property
        def natural_hypparam(self):
            return self._standard_to_natural(self.mu_0, self.sigma_0, self.kappa_0, self.nu_0)

        @property
        def mf_natural_hypparam(self):
            return self._standard_to_natural(self.mu_mf, self.sigma_mf, self.kappa_mf, self.nu_mf)

        @mf_natural_hypparam.setter
        def mf_natural_hypparam(self, nat):
            self.mu_mf, self.sigma_mf, self.kappa_mf, self.nu_mf = self._natural_to_standard(nat)
            self.mu, self.sigma = self.mu_mf.copy(), self.sigma_mf / (self.nu_mf - self.D - 1)

        def resample(self, data=[]):
            data = np.asarray(data, dtype=np.float64).reshape(-1, self.mu_0.shape[0]) if len(data) else []
            REC.gauss_data.append(np.array(data, copy=True))
            self.mu, self.sigma = niw.niw_resample(self.mu_0, self.sigma_0, self.kappa_0, self.nu_0,
                                                   data=data if len(data) else None)
            self.mu_mf, self.sigma_mf = self.mu, self.sigma
            return self

        def log_likelihood(self, x):
            x = np.asarray(x, dtype=np.float64)
            r = niw.gaussian_log_likelihood(np.reshape(x, (-1, self.D)), self.mu, self.sigma)
            return r

        def rvs(self, size=None):
            size = 1 if size is None else size
            size = size + (self.D,) if isinstance(size, tuple) else (int(size), self.D)
            return self.mu + np.random.normal(size=size).dot(self.sigma_chol.T)

        def _loglmbdatilde(self):
            from oracle import mf_sbm
            return mf_sbm.loglmbdatilde(self.sigma_mf, self.nu_mf)

        def _resample_from_mf(self):
            self.mu, self.sigma = niw.sample_niw(self.mu_mf, self.sigma_mf, self.kappa_mf, self.nu_mf)
            return self

        def get_vlb(self):
            from oracle import mf_sbm
            return mf_sbm.niw_vlb(self.mu_0, self.sigma_0, self.kappa_0, self.nu_0,
                                  self.mu_mf, self.sigma_mf, self.kappa_mf, self.nu_mf)

    class GaussianFixedMean(object):
        """pybasicbayes.distributions.GaussianFixedMean (recollection): Sigma ~ IW(lmbda_0, nu_0), mean fixed."""

        def __init__(self, mu=None, sigma=None, nu_0=None, lmbda_0=None):
            self.mu, self.sigma, self.nu_0, self.lmbda_0 = mu, sigma, nu_0, lmbda_0
            if sigma is None and nu_0 is not None and lmbda_0 is not None:
                self.resample()

        @property
        def D(self):
            return self.mu.shape[0]

        def resample(self, data=[]):
            n = len(data)
            if n:
                d = np.asarray(data, dtype=np.float64).reshape(-1, self.D) - self.mu
                S = d.T.dot(d)
            else:
                S = 0.0
            self.sigma = niw.sample_invwishart(self.lmbda_0 + S, self.nu_0 + n)
            return self

        def log_likelihood(self, x):
            return niw.gaussian_log_likelihood(np.reshape(np.asarray(x, dtype=np.float64), (-1, self.D)), self.mu, self.sigma)

    class GaussianFixedCov(object):
        """pybasicbayes.distributions.GaussianFixedCov (recollection): mu ~ N(mu_0, sigma_0), covariance fixed."""

        def __init__(self, mu=None, sigma=None, mu_0=None, sigma_0=None):
            self.mu, self.sigma, self.mu_0, self.sigma_0 = mu, sigma, mu_0, sigma_0
            if mu is None and mu_0 is not None and sigma_0 is not None:
                self.resample()

        @property
        def D(self):
            return self.sigma.shape[0]

        def resample(self, data=[]):
            J0 = np.linalg.inv(self.sigma_0)
            h = J0.dot(self.mu_0)
            J = J0.copy()
            if len(data):
                x = np.asarray(data, dtype=np.float64).reshape(-1, self.D)
                Ji = np.linalg.inv(self.sigma)
                J = J + x.shape[0] * Ji
                h = h + Ji.dot(x.sum(0))
            cov = np.linalg.inv(J)
            self.mu = np.random.multivariate_normal(cov.dot(h), cov)
            return self

        def log_likelihood(self, x):
            return niw.gaussian_log_likelihood(np.reshape(np.asarray(x, dtype=np.float64), (-1, self.D)), self.mu, self.sigma)

    _mod("pybasicbayes.distributions", Gaussian=Gaussian, GaussianFixedMean=GaussianFixedMean,
         GaussianFixedCov=GaussianFixedCov)


def _install_pypolyagamma():
    class PyPolyaGamma(object):
        def __init__(self, seed=0):
            self.seed = int(seed)

    def pgdrawvpar(ppgs, b, z, out):
        """pypolyagamma.pgdrawvpar: out[i] ~ PG(b[i], z[i]).  Injected omegas when REC.omegas is set, else the
        C restatement of the Devroye sampler (oracle/pg.c)."""
        if REC.omegas is not None:
            out[:] = np.asarray(REC.omegas, dtype=np.float64).ravel()
            return
        from oracle import pg
        out[:] = pg.pgdraw(np.asarray(z, dtype=np.float64), seed=ppgs[0].seed)

    _mod("pypolyagamma", PyPolyaGamma=PyPolyaGamma, pgdrawvpar=pgdrawvpar, get_omp_num_threads=lambda: 1)


def _grad(f):
    """autograd.grad replaced by the complex-step derivative d f / d x_i = Im f(x + i h e_i) / h, h = 1e-30:
    exact to round-off for the analytic expressions the reference differentiates (exp, log, sums, products)."""
    def g(x):
        x = np.asarray(x, dtype=np.float64)
        out = np.zeros(x.shape)
        h = 1e-30
        xc = x.astype(np.complex128)
        it = np.nditer(x, flags=["multi_index"]) if x.ndim else None
        if x.ndim == 0:
            return np.array(np.imag(f(xc + 1j * h)) / h)
        for _ in it:
            i = it.multi_index
            xp = xc.copy(); xp[i] += 1j * h
            out[i] = np.imag(f(xp)) / h
        return out
    return g


def _hmc(U, grad_U, step_sz, n_steps, q_curr, adaptive_step_sz=False, tgt_accept_rate=0.9,
         avg_accept_time_const=0.95, avg_accept_rate=0.9, min_step_sz=0.00005, max_step_sz=1.0,
         negative_log_prob=True):
    """hips.inference.hmc.hmc (recollection) through oracle.latent.hmc; momentum then log-uniform from
    np.random (hips' order: randn at entry, rand at the accept test) unless injected."""
    from oracle import latent
    q_curr = np.asarray(q_curr, dtype=np.float64)
    if REC.hmc_draws is not None:
        p0, log_u = REC.hmc_draws.pop(0)
    else:
        p0 = np.random.randn(*np.shape(q_curr))
        log_u = np.log(np.random.rand())
    Uf = lambda q: float(np.real(U(q)))
    r = latent.hmc(Uf, grad_U, step_sz, n_steps, q_curr, (np.asarray(p0, dtype=np.float64).reshape(q_curr.shape), log_u),
                   negative_log_prob=negative_log_prob, avg_accept_rate=avg_accept_rate,
                   adaptive_step_sz=adaptive_step_sz, tgt_accept_rate=tgt_accept_rate,
                   avg_accept_time_const=avg_accept_time_const, min_step_sz=min_step_sz, max_step_sz=max_step_sz)
    REC.hmc_log.append(dict(step=step_sz, out=r))
    return r


def install_stubs():
    for name, val in (("Inf", np.inf), ("int", int), ("float", float), ("bool", bool)):
        if not hasattr(np, name):
            setattr(np, name, val)
    builtins.xrange = range
    _install_pybasicbayes()
    _install_pypolyagamma()
    import scipy.special
    m = _mod("scipy.misc", logsumexp=scipy.special.logsumexp)
    import scipy
    scipy.misc = m
    _mod("numpy.core.umath_tests", inner1d=lambda a, b: np.einsum("...i,...i->...", a, b))
    _mod("autograd", grad=_grad)
    sys.modules["autograd.numpy"] = np
    sys.modules["autograd"].numpy = np
    _mod("hips")
    _mod("hips.inference")
    _mod("hips.inference.hmc", hmc=_hmc)
    _mod("hips.inference.discrete_sample", discrete_sample=_discrete_sample)
    _mod("hips.inference.log_sum_exp_sample", log_sum_exp_sample=_log_sum_exp_sample)
    _mod("hips.plotting")
    _mod("hips.plotting.colormaps", white_to_color_cmap=lambda *a, **k: None)
    plt = _mod("matplotlib.pyplot")
    sys.modules["matplotlib"].pyplot = plt
    _mod("matplotlib.cm")
    _mod("matplotlib.patches")


class _Network(object):
    """graphistician.old_abstractions.FixedGaussianNetwork -- MISSING from the reference tree; reconstructed
    from its use sites (stochastic_block_model.py:163,224,331; internals/weights.py:94-96,114):
    .A, .W and the expectations of a fixed network E_A = A, E_W = W, E_WWT = W W^T."""

    def __init__(self, A, W):
        self.A, self.W = A, W

    @property
    def E_A(self):
        return self.A

    @property
    def E_W(self):
        return self.W

    @property
    def E_WWT(self):
        return self.W[:, :, :, None] * self.W[:, :, None, :]


def load_source(modname, relpath, extra_globals=None):
    """Read a reference file, apply the rewrites, execute it as module `modname`."""
    path = os.path.join(REF_ROOT, relpath)
    src = py2to3(open(path).read())
    m = types.ModuleType(modname)
    m.__file__ = path
    if extra_globals:
        m.__dict__.update(extra_globals)
    sys.modules[modname] = m
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", SyntaxWarning)
        code = compile(src, path, "exec")
    exec(code, m.__dict__)
    return m


_LOADED = {}


def load_reference():
    """Returns a namespace of executed reference modules: utils, distributions, iweights (internals/weights.py),
    abstractions, adjacency, weights, networks, sbm (stochastic_block_model.py), eigenmodel, graph_models."""
    if _LOADED:
        return types.SimpleNamespace(**_LOADED)
    assert os.path.isdir(REF), "the reference tree is only present in the build container"
    install_stubs()
    pkg = _mod("graphistician"); pkg.__path__ = [REF]
    sub = _mod("graphistician.internals"); sub.__path__ = [os.path.join(REF, "internals")]
    base = type("Base", (object,), {})
    _mod("graphistician.old_abstractions", NetworkDistribution=type("NetworkDistribution", (object,), {}),
         GaussianWeightedNetworkDistribution=type("GaussianWeightedNetworkDistribution", (object,), {}),
         WeightedDirectedNetwork=_Network, FixedGaussianNetwork=_Network)
    L = _LOADED
    L["utils"] = load_source("graphistician.internals.utils", "graphistician/internals/utils.py")
    L["distributions"] = load_source("graphistician.internals.distributions", "graphistician/internals/distributions.py")
    L["iweights"] = load_source("graphistician.internals.weights", "graphistician/internals/weights.py")
    sub.utils, sub.distributions, sub.weights = L["utils"], L["distributions"], L["iweights"]
    # Python-2 implicit relative imports inside the package: `from abstractions import ...`, `import adjacency`
    sys.modules["internals"] = sub
    sys.modules["internals.utils"] = L["utils"]
    L["abstractions"] = load_source("graphistician.abstractions", "graphistician/abstractions.py")
    sys.modules["abstractions"] = L["abstractions"]
    L["adjacency"] = load_source("graphistician.adjacency", "graphistician/adjacency.py")
    sys.modules["adjacency"] = L["adjacency"]
    L["weights"] = load_source("graphistician.weights", "graphistician/weights.py")
    sys.modules["weights"] = L["weights"]
    L["networks"] = load_source("graphistician.networks", "graphistician/networks.py")
    L["sbm"] = load_source("graphistician.stochastic_block_model", "graphistician/stochastic_block_model.py")
    # eigenmodel.py:1319-1320 names two base classes that are never imported (NameError at import in the reference
    # itself): pre-seed them so the module body executes; GaussianWeightedEigenmodel is not exercised here.
    L["eigenmodel"] = load_source("graphistician.eigenmodel", "graphistician/eigenmodel.py",
                                  dict(FactorizedWeightedNetworkDistribution=base,
                                       GaussianWeightedNetworkDistribution=type("Base2", (object,), {})))
    L["graph_models"] = load_source("graphistician.internals.graph_models", "graphistician/internals/graph_models.py")
    L["Network"] = _Network
    return types.SimpleNamespace(**L)


class patched_numpy(object):
    """Context manager: record what the reference's code hands to np.linalg.inv / np.random.multivariate_normal /
    np.random.normal, and (optionally) replace the draws by the fixed maps of the parity protocol:
    x = mean + L^-T z with inv(cov) = L L^T  (the map oracle.eigenmodel.gaussian_from_prec uses); normal: mean + std z."""

    def __enter__(self):
        self._inv, self._mvn, self._nrm = np.linalg.inv, np.random.multivariate_normal, np.random.normal

        def inv(a):
            REC.inv_in.append(np.array(a, copy=True))
            return self._inv(a)

        def mvn(mean, cov, *a, **k):
            REC.mvn.append((np.array(mean, copy=True), np.array(cov, copy=True)))
            if REC.mvn_z is None:
                return self._mvn(mean, cov, *a, **k)
            z = np.asarray(REC.mvn_z.pop(0), dtype=np.float64)
            prec = REC.inv_in[-1]                       # the reference inverts post_prec right before it draws
            Lc = np.linalg.cholesky(prec)
            import scipy.linalg
            return np.asarray(mean) + scipy.linalg.solve_triangular(Lc.T, z, lower=False)

        def nrm(loc=0.0, scale=1.0, size=None):
            if size is None and np.ndim(loc) == 0:
                REC.normal.append((float(loc), float(scale)))
                if REC.normal_z is not None:
                    return float(loc) + float(scale) * float(REC.normal_z.pop(0))
            return self._nrm(loc, scale, size)

        np.linalg.inv, np.random.multivariate_normal, np.random.normal = inv, mvn, nrm
        return REC

    def __exit__(self, *exc):
        np.linalg.inv, np.random.multivariate_normal, np.random.normal = self._inv, self._mvn, self._nrm
        return False
