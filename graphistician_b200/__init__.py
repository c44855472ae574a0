"""graphistician_b200 -- B200-native Gibbs engine behind graphistician's model-class API.

Host code is Python; every O(N^2) step runs in libgk.so (hand-written sm_100a CUDA behind the C ABI of
include/gk.h).  There is no CPU fallback."""
from . import _lib
from .device import DeviceNetwork

__all__ = ["DeviceNetwork", "_lib"]
