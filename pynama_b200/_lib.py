"""ctypes binding of libpynama_b200.so (the C ABI in include/pynama_b200.h).

There is no CPU fallback: if the shared library is missing or no CUDA device
is visible, the first compute call raises.  Host-only helpers (1-D rules,
Lagrange tables, local order) work without a GPU.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpynama_b200.so")

c_int, c_i64, c_dbl = ctypes.c_int, ctypes.c_int64, ctypes.c_double
vp = ctypes.c_void_p
P = ctypes.POINTER


class PynamaError(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise PynamaError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C pynama_b200/csrc`). pynama_b200 has no CPU fallback.")
    return ctypes.CDLL(LIB_PATH)


_lib = None

# name -> argtypes (restype is int unless noted)
_PROTOS = {
    "pyn_version": [],
    "pyn_rule_1d": [c_int, c_int, vp, vp],
    "pyn_lagrange_1d": [c_int, vp, c_int, vp, vp, vp],
    "pyn_local_order": [c_int, c_int, vp],
    "pyn_init": [c_int, P(vp)],
    "pyn_finalize": [vp],
    "pyn_sync": [vp],
    "pyn_timer_start": [vp],
    "pyn_timer_stop": [vp, P(c_dbl)],
    "pyn_launch_count": [vp, P(c_i64)],
    "pyn_flush_l2": [vp],
    "pyn_mem_used": [vp, P(c_i64)],
    "pyn_prof_enable": [vp, c_int],
    "pyn_prof_get": [vp, c_int, P(c_i64), P(c_dbl)],
    "pyn_tune": [ctypes.c_char_p, c_int],
    "pyn_host_alloc": [c_i64, P(vp)],
    "pyn_host_free": [vp],
    "pyn_comm_unique_id": [ctypes.c_char_p, vp],
    "pyn_comm_init": [vp, ctypes.c_char_p, c_int, c_int, vp],
    "pyn_comm_info": [vp, P(c_int), P(c_int), P(c_int), P(c_i64)],
    "pyn_comm_allreduce": [vp, vp, c_int, c_int],
    "pyn_comm_finalize": [vp],
    "pyn_comm_window_export": [vp, vp],
    "pyn_comm_window_attach": [vp, vp],
    "pyn_halo_window_export": [vp, c_int, vp],
    "pyn_halo_window_attach": [vp, vp, vp, vp],
    "pyn_comm_p2p_status": [vp, P(c_int), P(c_int)],
    "pyn_halo_create": [vp, c_i64, c_i64, c_int, vp, vp, vp, vp, vp, P(vp)],
    "pyn_halo_destroy": [vp],
    "pyn_halo_exchange": [vp, vp, c_int],
    "pyn_mat_set_halo": [vp, vp],
    "pyn_mat_set_matfree": [vp, c_int],
    "pyn_vec_set_owned": [vp, c_i64],
    "pyn_mesh_create": [vp, c_int, c_int, c_i64, c_i64, vp, vp, P(vp)],
    "pyn_mesh_destroy": [vp],
    "pyn_mesh_graph_info": [vp, P(c_i64), P(c_i64), P(c_int)],
    "pyn_mesh_set_dof_class": [vp, vp],
    "pyn_mesh_set_global_tables": [vp, c_i64, c_i64, vp, vp],
    "pyn_mesh_set_congruence": [vp, c_dbl, P(c_i64)],
    "pyn_mesh_get_congruence": [vp, P(c_dbl), P(c_i64)],
    "pyn_mesh_phase_ms": [vp, c_int, P(c_dbl)],
    "pyn_ibm_build": [vp, c_i64, c_i64, vp, c_i64, vp, vp, c_dbl, c_dbl, c_dbl, c_dbl, c_int, c_int,
                      P(vp), P(vp), P(vp), P(c_i64), vp],
    "pyn_elem_kle": [vp, c_i64, c_i64, vp, vp, vp],
    "pyn_elem_ops": [vp, c_i64, c_i64, vp, vp, vp, vp],
    "pyn_assemble_kle": [vp, c_int, P(vp)],
    "pyn_assemble_ops": [vp, c_int, P(vp), P(vp), P(vp), P(vp)],
    "pyn_mat_shape": [vp, P(c_i64), P(c_i64), P(c_int), P(c_int)],
    "pyn_mat_nnz": [vp, P(c_i64), P(c_i64)],
    "pyn_mat_export_csr": [vp, vp, vp, vp],
    "pyn_mat_mult": [vp, vp, vp],
    "pyn_mat_mult_axpby": [vp, c_dbl, vp, c_dbl, vp, vp],
    "pyn_mat_add_new": [vp, vp, P(vp)],
    "pyn_mat_get_diagonal": [vp, vp],
    "pyn_mat_scale_rows": [vp, vp],
    "pyn_mat_bytes": [vp, P(c_i64), P(c_i64)],
    "pyn_mat_destroy": [vp],
    "pyn_vec_create": [vp, c_i64, P(vp)],
    "pyn_vec_destroy": [vp],
    "pyn_vec_size": [vp, P(c_i64)],
    "pyn_vec_set": [vp, c_dbl],
    "pyn_vec_copy": [vp, vp],
    "pyn_vec_scale": [vp, c_dbl],
    "pyn_vec_axpy": [vp, c_dbl, vp],
    "pyn_vec_aypx": [vp, c_dbl, vp],
    "pyn_vec_waxpy": [vp, c_dbl, vp, vp],
    "pyn_vec_axpby": [vp, c_dbl, vp, c_dbl],
    "pyn_vec_pointwise_mult": [vp, vp, vp],
    "pyn_vec_reciprocal": [vp],
    "pyn_vec_dot": [vp, vp, P(c_dbl)],
    "pyn_vec_norm2": [vp, P(c_dbl)],
    "pyn_vec_sum": [vp, P(c_dbl)],
    "pyn_vec_maxpy": [vp, c_int, vp, vp],
    "pyn_vec_set_values": [vp, c_i64, vp, vp, c_int],
    "pyn_vec_get_values": [vp, c_i64, vp, vp],
    "pyn_scatter_create": [vp, c_i64, vp, P(vp)],
    "pyn_scatter_apply": [vp, vp, vp, c_int],
    "pyn_scatter_destroy": [vp],
    "pyn_vec_to_host": [vp, vp],
    "pyn_vec_from_host": [vp, vp],
    "pyn_vec_to_host_n": [vp, vp, c_i64],
    "pyn_vec_from_host_n": [vp, vp, c_i64],
    "pyn_vec_device_ptr": [vp, P(vp)],
    "pyn_fp64_peak": [vp, P(c_dbl)],
    "pyn_vec_dot_noalloc": [vp, vp, vp, c_i64, vp, vp],
    "pyn_ksp_create": [vp, vp, c_int, c_int, P(vp)],
    "pyn_ksp_set_tolerances": [vp, c_dbl, c_dbl, c_int, c_int],
    "pyn_ksp_set_fischer_guess": [vp, c_int, c_int],
    "pyn_ksp_set_cg_single_reduction": [vp, c_int],
    "pyn_ksp_hint_iterations": [vp, c_int],
    "pyn_ksp_set_asm_coarse": [vp, c_int],
    "pyn_ksp_solve": [vp, vp, vp, P(c_int), P(c_int), P(c_dbl)],
    "pyn_ksp_destroy": [vp],
    "pyn_vtensv": [vp, vp, c_int],
    "pyn_stress_combine": [vp, vp, c_dbl, c_dbl, c_int],
}

EXPORTS = sorted(list(_PROTOS) + ["pyn_last_error"])


def lib():
    global _lib
    if _lib is None:
        l = _load()
        for name, args in _PROTOS.items():
            fn = getattr(l, name)
            fn.argtypes = args
            fn.restype = c_int
        l.pyn_last_error.argtypes = []
        l.pyn_last_error.restype = ctypes.c_char_p
        _lib = l
    return _lib


def check(rc):
    if rc != 0:
        msg = lib().pyn_last_error()
        raise PynamaError(f"libpynama_b200 error {rc}: {msg.decode() if msg else '?'}")


def call(name, *args):
    check(getattr(lib(), name)(*args))


def ptr(a):
    """void* of a C-contiguous numpy array (None -> NULL)."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(vp)


def as_f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


# ---- host-only helpers --------------------------------------------------------------
def rule_1d(kind, n):
    x, w = np.zeros(n), np.zeros(n)
    call("pyn_rule_1d", {"gauss": 0, "lobatto": 1}[kind], n, ptr(x), ptr(w))
    return x, w


def lagrange_1d(nodes, pts):
    nodes, pts = as_f64(nodes), as_f64(pts)
    h = np.zeros((len(pts), len(nodes)))
    dh = np.zeros_like(h)
    call("pyn_lagrange_1d", len(nodes), ptr(nodes), len(pts), ptr(pts), ptr(h), ptr(dh))
    return h, dh


def local_order(dim, n):
    out = np.zeros((n ** dim, dim), dtype=np.int32)
    call("pyn_local_order", dim, n, ptr(out))
    return out
