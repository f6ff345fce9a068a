"""fluctus_b200 — B200-native (sm_100a) backend for the wavefront path-tracing hot path of
tigrazone/fluctus.  The product is fluctus_b200/libfluctus_cuda.so (hand-written CUDA kernels
behind the C-ABI of include/fluctus_cuda.h); this package is the thin host-side mirror of the
reference's CLContext / Tracer surface used by the tests, bench.py and the headless renderer."""
from ._lib import FluctusError, LIB_PATH, SIGNATURES, load  # noqa: F401
from .abi import *  # noqa: F401,F403
from .context import CudaContext  # noqa: F401
from .tracer import Tracer  # noqa: F401
from . import scene  # noqa: F401
