"""miraculix_b200 -- B200-native (sm_100a) implementation of the compressed-genotype hot path of
alexfreudenberg/miraculix behind the reference's own C ABI.

* ``csrc/``   hand-written CUDA kernels + the C-ABI (``libmiraculix_b200.so``,
              declarations in ``include/miraculix_b200.h``)
* ``api``     ctypes mirror of the reference's Julia bindings (dgemm_compressed.jl, crossproduct.jl)
* ``build``   in-tree nvcc build for sm_100a
"""
from . import api  # noqa: F401
from .api import (MiraculixError, dgemm_compressed, free_compressed, get_compressed_freq, grm,  # noqa: F401
                  init_compressed, ld, load_library, set_options, snp_crossprod)
