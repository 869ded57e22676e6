"""ctypes view of include/radiomedium_b200.h (librmb200.so).

This is the same C-ABI the Java adapter binds with Panama FFM; Python is used
here only because this image has no JVM.  The library is the product: if it is missing or was
built without CUDA there is nothing to fall back to, and loading fails loudly.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librmb200.so")

OK, ERR_INVALID, ERR_NO_DEVICE, ERR_CUDA, ERR_CAPACITY, ERR_UNIFORMS, ERR_STATE, ERR_UNSUPPORTED = 0, -1, -2, -3, -4, -5, -6, -7
ERR_PEER = -8
PEER_HANDLE_BYTES = 128
MODEL_NULL, MODEL_UDGM, MODEL_UDGM_CONSTANT, MODEL_N2N, MODEL_SINR = 0, 1, 2, 3, 4
NOT_HEARD, INTERFERED, DELIVERED = 0, 1, 2
FLAG_F64_VALUES = 1
FLAG_COMPACT = 2

# every symbol include/radiomedium_b200.h declares
SYMBOLS = ("rmb_version", "rmb_config_default", "rmb_create", "rmb_destroy", "rmb_last_error",
           "rmb_set_nodes", "rmb_set_matrix", "rmb_update_positions", "rmb_update_positions_async", "rmb_update_radio", "rmb_tick",
           "rmb_build_tx_records", "rmb_tick_records", "rmb_tick_resident", "rmb_fetch",
           "rmb_node_info", "rmb_node_info_f64", "rmb_last_timing", "rmb_set_tuning", "rmb_get_stream", "rmb_set_stream",
           "rmb_peer_export", "rmb_peer_open", "rmb_peer_push", "rmb_peer_push_records", "rmb_tick_peers", "rmb_peer_tick",
           "rmb_tick_peers_resident", "rmb_tick_enqueue", "rmb_tick_finish", "rmb_tick_peers_enqueue", "rmb_peer_exchange_tick_enqueue", "rmb_peer_barrier_enqueue", "rmb_tick_draws_begin", "rmb_tick_draws_end",
           "rmb_tick_peers_draws_begin", "rmb_set_shard_base", "rmb_may_draw",
           "rmb_group_create", "rmb_group_destroy", "rmb_group_last_error", "rmb_group_size", "rmb_group_shard",
           "rmb_group_set_nodes", "rmb_group_update_positions", "rmb_group_update_radio", "rmb_group_tick",
           "rmb_group_node_info_f64", "rmb_udgm_tick_static")


class Config(C.Structure):
    _fields_ = [("struct_size", C.c_int32), ("model", C.c_int32), ("device", C.c_int32), ("flags", C.c_int32),
                ("success_ratio_tx", C.c_double), ("success_ratio_rx", C.c_double),
                ("transmission_range", C.c_double), ("constant_range", C.c_double),
                ("pl0_db", C.c_double), ("pl_exponent", C.c_double), ("d0_m", C.c_double),
                ("sensitivity_dbm", C.c_double), ("sinr_threshold_db", C.c_double),
                ("noise_floor_dbm", C.c_double), ("max_pairs", C.c_int64),
                ("rx_base", C.c_int32), ("n_shards", C.c_int32),
                ("half_duplex", C.c_int32), ("reserved", C.c_int32)]


class TxRecord(C.Structure):
    _fields_ = [("x", C.c_double), ("y", C.c_double), ("z", C.c_double), ("power_dbm", C.c_double),
                ("txprob", C.c_double), ("channel", C.c_int32), ("src", C.c_int32)]


class Result(C.Structure):
    _fields_ = [("n_pairs", C.c_int64), ("n_delivered", C.c_int64), ("n_interfered", C.c_int64),
                ("n_decisions", C.c_uint64), ("n_uniforms_used", C.c_int64), ("n_recheck", C.c_int64),
                ("n_tx", C.c_int32), ("near_slots", C.c_int32),
                ("tx", C.POINTER(C.c_int32)), ("rx", C.POINTER(C.c_int32)),
                ("rssi_dbm", C.POINTER(C.c_float)), ("sinr_db", C.POINTER(C.c_float)),
                ("outcome", C.POINTER(C.c_uint8)),
                ("rssi_dbm_f64", C.POINTER(C.c_double)), ("sinr_db_f64", C.POINTER(C.c_double)),
                ("tx_failed", C.POINTER(C.c_uint8)), ("tx_first", C.POINTER(C.c_int32))]


class Timing(C.Structure):
    _fields_ = [("ms_total", C.c_float), ("ms_prepare", C.c_float), ("ms_sweep", C.c_float),
                ("ms_resolve", C.c_float), ("ms_emit", C.c_float),
                ("sweep_launches", C.c_int32), ("kernel_launches", C.c_int32),
                ("recheck_sequential", C.c_int32)]


assert C.sizeof(TxRecord) == 48

_lib = None


def load():
    """dlopen librmb200.so and declare the prototypes.  Raises if the library is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(the CUDA library is the product; there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    H = C.c_void_p
    pd, pi, pu8, pf = C.POINTER(C.c_double), C.POINTER(C.c_int32), C.POINTER(C.c_uint8), C.POINTER(C.c_float)
    L.rmb_version.restype = C.c_int
    L.rmb_config_default.argtypes = [C.POINTER(Config), C.c_int32]
    L.rmb_config_default.restype = None
    L.rmb_create.argtypes = [C.POINTER(Config), C.POINTER(H)]
    L.rmb_destroy.argtypes = [H]
    L.rmb_last_error.argtypes = [H]
    L.rmb_last_error.restype = C.c_char_p
    L.rmb_set_nodes.argtypes = [H, C.c_int32, pd, pd, pd, pd, pi, pu8, pd, pd, pi]
    L.rmb_set_matrix.argtypes = [H, C.c_int32, pd]
    L.rmb_update_positions.argtypes = [H, C.c_int32, pi, pd, pd, pd]
    L.rmb_update_positions_async.argtypes = [H, C.c_int32, pd, pd, pd]
    L.rmb_update_radio.argtypes = [H, C.c_int32, pi, pd, pi, pu8, pd, pd]
    L.rmb_tick.argtypes = [H, C.c_int32, pi, pd, pi, pd, C.c_size_t, C.POINTER(Result)]
    L.rmb_build_tx_records.argtypes = [H, C.c_int32, pi, pd, pi, C.c_void_p]
    L.rmb_tick_records.argtypes = [H, C.c_int32, C.c_void_p, pd, C.c_size_t, C.POINTER(Result)]
    L.rmb_tick_resident.argtypes = [H, C.c_int32, C.c_void_p, C.POINTER(Result)]
    L.rmb_fetch.argtypes = [H, C.POINTER(Result)]
    L.rmb_node_info.argtypes = [H, pf, pu8]
    L.rmb_node_info_f64.argtypes = [H, pd, pu8]
    L.rmb_last_timing.argtypes = [H, C.POINTER(Timing)]
    L.rmb_set_tuning.argtypes = [H, C.c_int32, C.c_int32]
    L.rmb_peer_export.argtypes = [H, C.c_int32, C.c_int32, C.c_int32, C.c_void_p]
    L.rmb_peer_open.argtypes = [H, C.c_void_p]
    L.rmb_peer_push.argtypes = [H, C.c_int32, pi, pd, pi]
    L.rmb_peer_push_records.argtypes = [H, C.c_int32, C.c_void_p]
    L.rmb_tick_peers.argtypes = [H, pd, C.c_size_t, C.POINTER(Result)]
    L.rmb_tick_peers_resident.argtypes = [H, C.POINTER(Result)]
    L.rmb_tick_enqueue.argtypes = [H, C.c_int32, C.c_void_p]
    L.rmb_tick_finish.argtypes = [H, C.POINTER(Result)]
    L.rmb_tick_peers_enqueue.argtypes = [H]
    L.rmb_peer_exchange_tick_enqueue.argtypes = [H, C.c_int32, C.c_void_p]
    L.rmb_peer_barrier_enqueue.argtypes = [H]
    L.rmb_tick_draws_begin.argtypes = [H, C.c_int32, C.c_void_p, C.c_void_p]
    L.rmb_tick_draws_end.argtypes = [H, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, pd, C.c_size_t, C.POINTER(Result)]
    L.rmb_tick_peers_draws_begin.argtypes = [H, C.c_void_p]
    L.rmb_set_shard_base.argtypes = [H, C.c_int32]
    L.rmb_peer_tick.argtypes = [H, C.c_int32, pi, pd, pi, C.POINTER(Result)]
    L.rmb_may_draw.argtypes = [H]
    L.rmb_may_draw.restype = C.c_int32
    L.rmb_udgm_tick_static.argtypes = [C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                       C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_int64, C.c_void_p, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p, C.c_size_t]
    L.rmb_group_create.argtypes = [C.POINTER(Config), C.c_int32, pi, C.c_int32, C.POINTER(H)]
    L.rmb_group_destroy.argtypes = [H]
    L.rmb_group_last_error.argtypes = [H]
    L.rmb_group_last_error.restype = C.c_char_p
    L.rmb_group_size.argtypes = [H]
    L.rmb_group_shard.argtypes = [H, C.c_int32, C.POINTER(H), pi, pi]
    L.rmb_group_set_nodes.argtypes = [H, C.c_int32, pd, pd, pd, pd, pi, pu8, pd, pd, pi]
    L.rmb_group_update_positions.argtypes = [H, C.c_int32, pi, pd, pd, pd]
    L.rmb_group_update_radio.argtypes = [H, C.c_int32, pi, pd, pi, pu8, pd, pd]
    L.rmb_group_tick.argtypes = [H, C.c_int32, pi, pd, pi, pd, C.c_size_t, C.POINTER(Result), pi, pi]
    L.rmb_group_node_info_f64.argtypes = [H, pd, pu8]
    for s in SYMBOLS:
        if s not in ("rmb_config_default", "rmb_last_error", "rmb_group_last_error"):
            getattr(L, s).restype = C.c_int
    _lib = L
    return L
