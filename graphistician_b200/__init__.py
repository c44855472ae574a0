"""graphistician_b200 -- B200-native Gibbs engine behind graphistician's model-class API.

Host code is Python; every O(N^2) step runs in libgk.so (hand-written sm_100a CUDA behind the C ABI of
include/gk.h).  There is no CPU fallback."""
from . import _lib
from .device import DeviceNetwork


def __getattr__(name):
    """Top-level model classes (graphistician/__init__.py:2 plus the three the reference's demos import from the package
    root: SURVEY.md App. B), resolved lazily so that importing the package does not need the device library."""
    if name in ("GaussianStochasticBlockModel",):
        from . import stochastic_block_model as m
    elif name in ("LogisticEigenmodel", "ProbitEigenmodel", "Eigenmodel"):
        from . import eigenmodel as m
    elif name in ("GaussianWeightedEigenmodel", "GaussianErdosRenyiFixedSparsity", "FactorizedWeightedNetworkDistribution"):
        from . import weighted_models as m
    elif name in ("FactorizedNetworkDistribution", "GaussianBernoulliNetwork"):
        from . import networks as m
    else:
        raise AttributeError(name)
    return getattr(m, name)


__all__ = ["DeviceNetwork", "_lib", "GaussianStochasticBlockModel", "LogisticEigenmodel", "ProbitEigenmodel",
           "GaussianWeightedEigenmodel", "GaussianErdosRenyiFixedSparsity", "FactorizedNetworkDistribution"]
