"""ctypes binding of ``libsqb200.so`` (declarations in ``include/sqb200.h``).

This module is the only place the Python side touches the C ABI.  There is no
CPU fallback: if the library is missing the import fails, and every compute
call needs a CUDA device.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_lib", "libsqb200.so")

c_void_p = ctypes.c_void_p
c_int = ctypes.c_int
c_ll = ctypes.c_longlong
c_ull = ctypes.c_ulonglong
c_double = ctypes.c_double
c_size_t = ctypes.c_size_t

# name -> (restype, argtypes); mirrors include/sqb200.h one to one
SIGNATURES = {
    "sq_last_error": (ctypes.c_char_p, []),
    "sq_abi_version": (c_int, []),
    "sq_tomography_max_fused_dim": (c_int, []),
    "sq_tomography_batched": (c_int, [c_void_p, c_ll, c_int, c_double, c_int, c_int, c_void_p, c_int, c_ull, c_ull,
                                      c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "sq_tomography_workspace_bytes": (c_size_t, [c_ll, c_int]),
    "sq_tomography_batched_ws": (c_int, [c_void_p, c_ll, c_int, c_double, c_int, c_int, c_void_p, c_int, c_ull, c_ull,
                                         c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "sq_tomography_workspace_layout": (c_int, [c_ll, c_int, c_void_p]),
    "sq_tomography_sharded_phase": (c_int, [c_void_p, c_ll, c_int, c_double, c_int, c_int, c_double, c_int, c_ull, c_ull,
                                            c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_int, c_int,
                                            c_void_p]),
    "sq_tomography_batched_spec": (c_int, [c_void_p, c_ll, c_int, c_double, c_int, c_void_p, c_int, c_int, c_ull, c_ull,
                                           c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "sq_tomography_spec_state_init": (c_int, [c_void_p, c_ll, c_void_p]),
    "sq_tomography_batched_spec_fused": (c_int, [c_void_p, c_ll, c_int, c_double, c_int, c_void_p, c_int, c_int, c_ull,
                                                 c_ull, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "sq_tomography_tune": (c_int, [c_int, c_double]),
    "sq_ipe_debug_force_replay": (c_int, [c_int]),
    "sq_qmeans_debug_filter": (c_int, [c_int]),
    "sq_ipe_debug_tail_sample": (c_int, [c_double, c_double, c_ll, c_ull, c_void_p, c_void_p]),
    "sq_qmeans_assign_workspace_bytes": (c_size_t, [c_ll, c_int, c_int]),
    "sq_qmeans_assign": (c_int, [c_void_p, c_void_p, c_ll, c_int, c_void_p, c_int, c_int, c_double, c_ull, c_ull,
                                 c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "sq_qmeans_assign_noalloc": (c_int, [c_void_p, c_void_p, c_ll, c_int, c_void_p, c_int, c_int, c_double, c_ull, c_ull,
                                         c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "sq_profile_assign": (c_int, [c_int, c_void_p, c_void_p]),
    "sq_row_sqnorms": (c_int, [c_void_p, c_ll, c_int, c_void_p, c_void_p]),
    "sq_sum_f64": (c_int, [c_void_p, c_ll, c_void_p, c_void_p, c_size_t, c_void_p]),
    "sq_centroid_workspace_bytes": (c_size_t, [c_ll, c_int, c_int]),
    "sq_centroid_accumulate": (c_int, [c_void_p, c_void_p, c_ll, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                       c_void_p, c_size_t, c_void_p]),
    "sq_centroid_finalize": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "sq_kpp_update": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_ll, c_int, c_void_p, c_void_p]),
    "sq_sqdiff_sum": (c_int, [c_void_p, c_void_p, c_ll, c_void_p, c_void_p]),
    "sq_fejer_sample": (c_int, [c_void_p, c_void_p, c_ll, c_ull, c_ull, c_void_p, c_void_p]),
    "sq_fejer_sample_bruteforce": (c_int, [c_void_p, c_void_p, c_ll, c_ull, c_ull, c_void_p, c_void_p]),
    "sq_truncnorm_add_each": (c_int, [c_void_p, c_ll, c_void_p, c_ull, c_ull, c_void_p, c_void_p]),
    "sq_kernel_matrix_workspace_bytes": (c_size_t, [c_ll, c_int]),
    "sq_kernel_matrix": (c_int, [c_void_p, c_ll, c_void_p, c_int, c_int, c_int, c_double, c_double, c_double,
                                 c_void_p, c_void_p, c_size_t, c_void_p]),
    "sq_truncnorm_add": (c_int, [c_void_p, c_ll, c_double, c_ull, c_ull, c_void_p, c_void_p]),
    "sq_binomial_sample": (c_int, [c_void_p, c_void_p, c_void_p, c_ll, c_ull, c_ull, c_void_p, c_void_p]),
    "sq_mu_norms_workspace_bytes": (c_size_t, [c_ll, c_int, c_int]),
    "sq_project_absmax": (c_int, [c_void_p, c_ll, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "sq_project": (c_int, [c_void_p, c_ll, c_int, c_void_p, c_int, c_void_p, c_void_p, c_void_p]),
    "sq_gram_workspace_bytes": (c_size_t, [c_ll, c_int]),
    "sq_gram_centered": (c_int, [c_void_p, c_ll, c_int, c_void_p, c_void_p, c_void_p, c_size_t, c_void_p]),
    "sq_col_sums_workspace_bytes": (c_size_t, [c_ll, c_int]),
    "sq_col_sums": (c_int, [c_void_p, c_ll, c_int, c_void_p, c_void_p, c_size_t, c_void_p]),
    "sq_sub_rowvec": (c_int, [c_void_p, c_void_p, c_ll, c_int, c_void_p, c_void_p]),
    "sq_mu_norms": (c_int, [c_void_p, c_ll, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p,
                            c_void_p, c_size_t, c_void_p]),
}


class SqError(RuntimeError):
    """A C-ABI call returned a non-zero status."""


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). sqlearn has no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    return lib


lib = _load()


def check(status, what):
    if status != 0:
        msg = lib.sq_last_error()
        raise SqError(f"{what} failed with status {status}: {msg.decode() if msg else ''}")
