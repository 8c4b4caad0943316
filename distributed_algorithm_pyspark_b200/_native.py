"""ctypes binding of ``_lib/libdapb200.so`` (C ABI in ``include/dapb200.h``).

The library is loaded lazily.  On a machine with a GPU a missing or unloadable library is an error
(the CUDA path must be the one that runs -- no silent eager fallback); on a CPU-only machine
``available()`` is False and the torch/NumPy backends are used instead.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

from . import build as _build

MAX_RANKS = 8
FLAG_BYTES = 65536
TC_STATE_WORDS, TC_RUNNING, TC_DONE, TC_GROW = 32, 0, 1, 2

_lib: Optional[C.CDLL] = None
_load_error: Optional[str] = None


class DapbComm(C.Structure):
    _fields_ = [("rank", C.c_int32), ("world", C.c_int32),
                ("heap", C.c_void_p * MAX_RANKS), ("heap_bytes", C.c_uint64), ("mc_heap", C.c_void_p)]


class DapbPrHub(C.Structure):
    _fields_ = [("hub_rows", C.c_void_p), ("hub_col", C.c_void_p), ("chunk_start", C.c_void_p),
                ("chunk_end", C.c_void_p), ("hub_chunk_ptr", C.c_void_p), ("chunk_partial", C.c_void_p),
                ("n_hub", C.c_int32), ("n_chunks", C.c_int32), ("chunk_row", C.c_void_p), ("row_done", C.c_void_p)]


class DapbPrRunArgs(C.Structure):
    _fields_ = [("row_ptr", C.c_void_p), ("col_idx", C.c_void_p), ("inv_outdeg", C.c_void_p), ("blocks", C.c_void_p),
                ("n_blocks", C.c_int32), ("f64", C.c_int32),
                ("nv_local", C.c_int64), ("v0", C.c_int64), ("V", C.c_int64),
                ("base", C.c_double), ("damp", C.c_double),
                ("ranks_local", C.c_void_p), ("c_off", C.c_uint64), ("flag_base", C.c_uint32), ("epoch0", C.c_uint64),
                ("n_iters", C.c_int32), ("it0", C.c_int32), ("wait_first", C.c_int32), ("skip_empty_ok", C.c_int32),
                ("grid", C.c_int32), ("sync", C.c_void_p), ("stats", C.c_void_p), ("hub", DapbPrHub)]


class DapbLrParityArgs(C.Structure):
    _fields_ = [
        ("X", C.c_void_p), ("y", C.c_void_p), ("part_off", C.c_void_p), ("masks", C.c_void_p),
        ("X_test", C.c_void_p), ("y_test", C.c_void_p),
        ("n_local", C.c_int64),
        ("n_test", C.c_int32), ("D", C.c_int32), ("P_local", C.c_int32), ("P_global", C.c_int32),
        ("part_base", C.c_int32), ("T", C.c_int32),
        ("algo", C.c_int32), ("n_local_iters", C.c_int32), ("reg_type", C.c_int32), ("max_part_rows", C.c_int32),
        ("eta", C.c_double), ("lam", C.c_double), ("reg_alpha", C.c_double), ("mu", C.c_double),
        ("zeta", C.c_double), ("alpha", C.c_double), ("beta", C.c_double), ("eps", C.c_double),
        ("w", C.c_void_p), ("ws", C.c_void_p), ("delta_w", C.c_void_p), ("accs", C.c_void_p),
        ("batch_sizes", C.c_void_p), ("w_hist", C.c_void_p),
        ("payload_off", C.c_uint64), ("flag_base", C.c_uint32), ("epoch_base", C.c_uint64),
    ]


class DapbLrStepArgs(C.Structure):
    _fields_ = [
        ("X", C.c_void_p), ("y", C.c_void_p), ("n_rows", C.c_int64), ("row_base", C.c_int64),
        ("D", C.c_int32), ("dtype", C.c_int32),
        ("mask", C.c_void_p), ("seed", C.c_uint64), ("iteration", C.c_uint32), ("threshold24", C.c_uint32),
        ("eps", C.c_float), ("w_fwd", C.c_void_p),
        ("partials", C.c_void_p), ("counts", C.c_void_p), ("ticket", C.c_void_p), ("grid", C.c_int32),
        ("fin", C.c_int32), ("algo", C.c_int32), ("reg_type", C.c_int32), ("use_sum", C.c_int32),
        ("eta", C.c_float), ("lam", C.c_float), ("reg_alpha", C.c_float), ("mu", C.c_float),
        ("zeta", C.c_float), ("alpha", C.c_float), ("beta", C.c_float),
        ("n_replicas_total", C.c_int32),
        ("g_out", C.c_void_p), ("cnt_out", C.c_void_p), ("w", C.c_void_p), ("w_local", C.c_void_p),
        ("delta_w", C.c_void_p), ("reset_replica", C.c_int32),
        ("payload_off", C.c_uint64), ("flag_base", C.c_uint32), ("epoch", C.c_uint64),
        ("cnt_accum", C.c_void_p), ("stats", C.c_void_p),
    ]


LR_STEP_FLAGS = 64
ALGO_ID = {"ssgd": 0, "gd": 1, "ma": 2, "bmuf": 3, "easgd": 4}
REG_ID = {"none": 0, "l2": 1, "l1": 2, "elastic_net": 3}
FIN_GRAD_ONLY, FIN_SSGD, FIN_LOCAL, FIN_LOCAL_SYNC = 0, 1, 2, 3


class NativeError(RuntimeError):
    pass


def _declare(lib: C.CDLL):
    vp, i32, i64, u32, u64 = C.c_void_p, C.c_int32, C.c_int64, C.c_uint32, C.c_uint64
    f32, f64 = C.c_float, C.c_double
    P = C.POINTER
    sig = {
        "dapb_last_error": (C.c_char_p, []),
        "dapb_version": (C.c_int, []),
        "dapb_device_info": (C.c_int, [P(i32), P(i32), P(i32), P(u64)]),
        "dapb_symm_alloc": (C.c_int, [u64, P(vp)]),
        "dapb_symm_free": (C.c_int, [vp]),
        "dapb_ipc_get_handle": (C.c_int, [vp, vp]),
        "dapb_ipc_open_handle": (C.c_int, [vp, P(vp)]),
        "dapb_ipc_close_handle": (C.c_int, [vp]),
        "dapb_comm_selftest": (C.c_int, [vp, u64, i32, i32, i32, vp, vp]),
        "dapb_allreduce_f32": (C.c_int, [P(DapbComm), u64, u32, u64, vp, vp, i32, vp]),
        "dapb_allreduce_f64": (C.c_int, [P(DapbComm), u64, u32, u64, vp, vp, i32, vp]),
        "dapb_barrier": (C.c_int, [P(DapbComm), u32, u64, vp]),
        "dapb_allreduce_stress": (C.c_int, [P(DapbComm), u64, u32, u64, i32, i32, u32, vp, vp]),
        "dapb_allgather": (C.c_int, [P(DapbComm), u64, u32, u64, vp, vp, i64, i32, vp]),
        "dapb_allreduce_twoshot_f32": (C.c_int, [P(DapbComm), u64, u32, u64, vp, vp, i32, vp]),
        "dapb_reduce_scatter_f32": (C.c_int, [P(DapbComm), u64, u32, u64, vp, vp, i32, vp]),
        "dapb_alltoall_u64": (C.c_int, [P(DapbComm), u64, u32, u64, vp, vp, i64, vp]),
        "dapb_lr_parity_run": (C.c_int, [P(DapbLrParityArgs), P(DapbComm), vp]),
        "dapb_lr_parity_payload_bytes": (u64, [i32, i32]),
        "dapb_lr_step_payload_bytes": (u64, [i32, i32]),
        "dapb_lr_step": (C.c_int, [P(DapbLrStepArgs), P(DapbComm), vp]),
        "dapb_lr_step_tc": (C.c_int, [P(DapbLrStepArgs), P(DapbComm), vp]),
        "dapb_philox_mask": (C.c_int, [i64, i64, u64, u32, u32, vp, vp]),
        "dapb_lr_eval": (C.c_int, [vp, vp, i64, i32, i32, vp, f32, vp, vp, vp]),
        "dapb_mc_pi": (C.c_int, [u64, u64, u64, vp, vp]),
        "dapb_mc_pi_allreduce": (C.c_int, [P(DapbComm), u64, u64, u64, vp, vp, vp, u64, u32, u64, vp]),
        "dapb_kmeans_assign": (C.c_int, [vp, i64, i32, i32, vp, i32, vp, vp, vp, vp]),
        "dapb_kmeans_finalize": (C.c_int, [P(DapbComm), vp, vp, vp, i32, i32, i32, u64, u32, u64, vp, i32, vp]),
        "dapb_kmeans_finalize_tc": (C.c_int, [P(DapbComm), vp, vp, vp, i32, i32, u64, u32, u64, vp, i32, vp, vp, C.c_float, vp]),
        "dapb_kmeans_payload_bytes": (u64, [i32, i32, i32, i32]),
        "dapb_kmeans_tc_prepare": (C.c_int, [vp, i32, i32, vp, vp, C.c_float, vp]),
        "dapb_kmeans_assign_tc": (C.c_int, [vp, i64, i32, vp, vp, i32, vp, vp, vp, i32, vp]),
        "dapb_kmeans_assign_tc_f32": (C.c_int, [vp, vp, i64, i32, vp, vp, i32, vp, vp, vp, i32, vp]),
        "dapb_kmeans_assign_tc_dbg": (C.c_int, [vp, i64, i32, vp, vp, i32, vp, vp, vp, i32, vp, vp]),
        "dapb_pagerank_init": (C.c_int, [P(DapbComm), vp, i64, i64, i64, f64, u64, i32, vp]),
        "dapb_pagerank_step": (C.c_int, [P(DapbComm), vp, vp, vp, vp, i32, i64, i64, i64, f32, f32, vp,
                                         u64, u32, u64, vp, i32, i32, i32, vp]),
        "dapb_pagerank_run": (C.c_int, [P(DapbPrRunArgs), P(DapbComm), vp]),
        "dapb_tc_seed": (C.c_int, [vp, i64, vp, u64, vp, vp, i32, i32, vp]),
        "dapb_tc_expand": (C.c_int, [P(DapbComm), vp, vp, vp, i64, vp, u64, vp, vp, u64, u64, vp]),
        "dapb_tc_exchange": (C.c_int, [P(DapbComm), vp, u64, u64, u32, u64, vp]),
        "dapb_tc_drain": (C.c_int, [P(DapbComm), vp, u64, vp, vp, u64, u64, vp]),
        "dapb_tc_closure": (C.c_int, [P(DapbComm), vp, vp, vp, u64, vp, vp, u64, vp, u64, u64, u64, u32, u64, vp]),
        "dapb_gram": (C.c_int, [P(DapbComm), vp, vp, i64, i32, i32, i64, i64, i32, vp, i32, vp, vp, u64, u32, u64, i32, vp]),
        "dapb_gram_tc": (C.c_int, [P(DapbComm), vp, i64, i32, vp, vp, i32, u64, u32, u64, i32, vp]),
        "dapb_gram_tc_payload_bytes": (u64, [i32, i32]),
        "dapb_rows_times_small": (C.c_int, [vp, vp, i64, i32, i32, i64, i32, i32, vp, vp]),
        "dapb_solve_rows": (C.c_int, [vp, f64, vp, i64, i32, i32, vp, vp]),
        "dapb_sq_residual": (C.c_int, [vp, vp, vp, i64, i32, i32, i32, vp, vp]),
        "dapb_cholesky": (C.c_int, [vp, i32, i32, vp, vp]),
        "dapb_cholesky_dbg": (C.c_int, [vp, i32, i32, vp, vp, vp]),
        "dapb_eigh": (C.c_int, [vp, i32, i32, vp, vp, vp, C.c_float, i32, vp, vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name, None)
        if fn is None:
            continue
        fn.restype = res
        fn.argtypes = args


def load(build_if_missing: bool = True) -> C.CDLL:
    global _lib, _load_error
    if _lib is not None:
        return _lib
    path = _build.LIB_PATH
    if not path.exists() and build_if_missing:
        _build.build()
    try:
        _lib = C.CDLL(str(path), mode=C.RTLD_GLOBAL)
    except OSError as e:                                           # pragma: no cover
        _load_error = str(e)
        raise NativeError(f"cannot load {path}: {e}") from e
    _declare(_lib)
    return _lib


def cuda_available() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:                                              # pragma: no cover
        return False


def available() -> bool:
    """True iff the native CUDA path can run here (GPU present and library loadable).
    On a GPU machine a load failure raises instead of returning False."""
    if not cuda_available():
        return False
    load()
    return True


def check(rc: int, what: str = ""):
    if rc != 0:
        msg = load().dapb_last_error().decode("utf-8", "replace")
        raise NativeError(f"{what or 'dapb call'} failed (rc={rc}): {msg}")


def ptr(t) -> int:
    """Device/host pointer of a torch tensor (0 for None)."""
    return 0 if t is None else t.data_ptr()


def stream_ptr(stream=None) -> int:
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


def local_comm(heap_ptr: int, heap_bytes: int) -> DapbComm:
    c = DapbComm()
    c.rank, c.world, c.heap_bytes = 0, 1, heap_bytes
    c.heap[0] = heap_ptr
    return c
