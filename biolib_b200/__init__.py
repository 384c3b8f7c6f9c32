"""biolib_b200 -- B200 (sm_100a) implementation of yhhshb/biolib's streaming k-mer path.

ASCII DNA -> 2-bit codes -> k-mers (kmer_view) -> m-mer hashes -> window minimizers (minimizer_view) ->
super-k-mers (super_kmer_view), as hand-written CUDA behind the C ABI of include/biolib_gpu.h.
The package holds only that path: csrc/ (kernels + C ABI), the ctypes binding and the sharding plan used
by the multi-GPU driver. There is no CPU implementation here; oracle/ is test infrastructure.
"""
from . import _capi
from .context import BlgError, Context

__all__ = ["Context", "BlgError", "_capi"]
