"""pbdagcon_b200: B200-native DAGCon consensus hot path.

The product is libdagcon_b200.so (CUDA kernels + C-ABI, include/dagcon_b200.h); this package is
the thin Python host-side mirror used by tests and bench.py.  There is no CPU fallback: every
entry point raises if the CUDA library is missing or no device is present."""
from .api import Consensus, DagconError, load_library, version  # noqa: F401
