"""spmm_b200 -- B200-native (sm_100a) CSR SpGEMM / SpMV engine behind the call shapes of
wjustin784/spMM's hot path.

  spmm_b200.cusparse.spgemm / spmv / spmm   mirror of cupyx.cusparse (operator boundary)
  spmm_b200.sparse.csr_matrix               mirror of cupyx.scipy.sparse.csr_matrix
  spmm_b200.host                            numpy/host-buffer entry points over the C ABI
  spmm_b200/libb200sp.so                    the C ABI (include/b200sp.h), hand-written CUDA
  spmm_b200/bin/spgemm_from_txt_alg{1,2,3}  CLI drop-ins for the reference drivers (CUEXE=...)

Importing the package loads the native library; it fails loudly if it has not been built.
"""
from . import _lib

_lib.load()

from . import cusparse, host, sparse  # noqa: E402
from .sparse import coo_matrix, csc_matrix, csr_matrix, max_abs_diff  # noqa: E402

__all__ = ["cusparse", "sparse", "host", "csr_matrix", "csc_matrix", "coo_matrix", "max_abs_diff"]
