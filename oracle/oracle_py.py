"""ctypes binding of oracle/liboracle.so -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product (radio-sim_b200/) never does.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liboracle.so")

NOT_HEARD, INTERFERED, DELIVERED = 0, 1, 2
MODEL_NULL, MODEL_UDGM, MODEL_UDGM_CONSTANT, MODEL_N2N, MODEL_SINR = 0, 1, 2, 3, 4


def build(force=False):
    """Compile the oracle with gcc (oracle/Makefile)."""
    srcs = [os.path.join(_HERE, f) for f in ("oracle_math.c", "medium_oracle.c", "oracle_math.h", "medium_oracle.h")]
    if force or not os.path.exists(_LIB) or any(os.path.getmtime(s) > os.path.getmtime(_LIB) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return _LIB


class Params(C.Structure):
    _fields_ = [("model", C.c_int32),
                ("success_ratio_tx", C.c_double), ("success_ratio_rx", C.c_double),
                ("transmission_range", C.c_double), ("constant_range", C.c_double),
                ("matrix", C.POINTER(C.c_double)), ("matrix_n", C.c_int32),
                ("pl0_db", C.c_double), ("pl_exponent", C.c_double), ("d0_m", C.c_double),
                ("sensitivity_dbm", C.c_double), ("sinr_threshold_db", C.c_double),
                ("noise_floor_dbm", C.c_double), ("half_duplex", C.c_int32)]


class Nodes(C.Structure):
    _fields_ = [("n", C.c_int32),
                ("x", C.POINTER(C.c_double)), ("y", C.POINTER(C.c_double)), ("z", C.POINTER(C.c_double)),
                ("channel", C.POINTER(C.c_int32)), ("enabled", C.POINTER(C.c_uint8)),
                ("rxprob", C.POINTER(C.c_double)), ("txprob", C.POINTER(C.c_double)),
                ("id_int", C.POINTER(C.c_int32))]


class Packet(C.Structure):
    _fields_ = [("src", C.c_int32), ("channel", C.c_int32), ("power_dbm", C.c_double)]


class Pair(C.Structure):
    _fields_ = [("tx", C.c_int32), ("rx", C.c_int32), ("rssi_dbm", C.c_double),
                ("sinr_db", C.c_double), ("outcome", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("n_pairs", C.c_size_t), ("n_needed", C.c_size_t), ("n_uniforms_used", C.c_size_t),
                ("n_decisions", C.c_uint64), ("status", C.c_int32)]


class JRandomState(C.Structure):
    _fields_ = [("seed", C.c_uint64)]


PAIR_DTYPE = np.dtype([("tx", "<i4"), ("rx", "<i4"), ("rssi_dbm", "<f8"), ("sinr_db", "<f8"),
                       ("outcome", "<i4"), ("_pad", "<i4")])
assert PAIR_DTYPE.itemsize == C.sizeof(Pair)
PACKET_DTYPE = np.dtype([("src", "<i4"), ("channel", "<i4"), ("power_dbm", "<f8")])
assert PACKET_DTYPE.itemsize == C.sizeof(Packet)

_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB)
        L.orc_log.restype = L.orc_log10.restype = L.orc_exp.restype = C.c_double
        L.orc_log.argtypes = L.orc_log10.argtypes = L.orc_exp.argtypes = [C.c_double]
        L.orc_dbm_to_mw.restype = L.orc_mw_to_dbm.restype = C.c_double
        L.orc_dbm_to_mw.argtypes = L.orc_mw_to_dbm.argtypes = [C.c_double]
        L.orc_distance.restype = C.c_double
        L.orc_distance.argtypes = [C.c_double] * 6
        L.orc_sinr_rx_power_dbm.restype = C.c_double
        L.orc_sinr_rx_power_dbm.argtypes = [C.POINTER(Params), C.c_double, C.c_double]
        L.orc_udgm_rx_probability.restype = C.c_double
        L.orc_udgm_rx_probability.argtypes = [C.POINTER(Params), C.POINTER(Nodes), C.c_int32, C.c_int32]
        L.orc_params_default.argtypes = [C.POINTER(Params), C.c_int32]
        L.orc_jrandom_init.argtypes = [C.POINTER(JRandomState), C.c_int64]
        L.orc_jrandom_next_double.restype = C.c_double
        L.orc_jrandom_next_double.argtypes = [C.POINTER(JRandomState)]
        L.orc_jrandom_next_int_bound.restype = C.c_int32
        L.orc_jrandom_next_int_bound.argtypes = [C.POINTER(JRandomState), C.c_int32]
        L.orc_jrandom_fill_doubles.argtypes = [C.c_int64, C.POINTER(C.c_double), C.c_size_t]
        L.orc_tick.argtypes = [C.POINTER(Params), C.POINTER(Nodes), C.POINTER(Packet), C.c_int32,
                               C.POINTER(C.c_double), C.c_size_t, C.POINTER(Pair), C.c_size_t,
                               C.POINTER(C.c_uint8), C.POINTER(C.c_double), C.c_int, C.POINTER(Stats)]
        _lib = L
    return _lib


def _p(a, ct):
    return a.ctypes.data_as(C.POINTER(ct)) if a is not None else None


def params(model, **kw):
    p = Params()
    lib().orc_params_default(C.byref(p), model)
    keep = None
    for k, v in kw.items():
        if k == "matrix":
            keep = np.ascontiguousarray(v, dtype=np.float64)
            p.matrix = _p(keep, C.c_double)
            p.matrix_n = int(round(keep.size ** 0.5))
        else:
            if not hasattr(p, k):
                raise AttributeError(k)
            setattr(p, k, v)
    p._keep = keep
    return p


class NodeSet:
    """SoA view of Node[] (Simulator.getNodes order)."""

    def __init__(self, x, y, z=None, channel=None, enabled=None, rxprob=None, txprob=None, id_int=None):
        n = len(x)
        f8 = lambda a, d: np.ascontiguousarray(np.full(n, d) if a is None else a, dtype=np.float64)
        self.n = n
        self.x, self.y, self.z = f8(x, 0), f8(y, 0), f8(z, 0.0)
        self.channel = np.ascontiguousarray(np.full(n, 26) if channel is None else channel, dtype=np.int32)
        self.enabled = np.ascontiguousarray(np.ones(n) if enabled is None else enabled, dtype=np.uint8)
        self.rxprob, self.txprob = f8(rxprob, 1.0), f8(txprob, 1.0)
        self.id_int = None if id_int is None else np.ascontiguousarray(id_int, dtype=np.int32)

    def c_struct(self):
        s = Nodes()
        s.n = self.n
        s.x, s.y, s.z = _p(self.x, C.c_double), _p(self.y, C.c_double), _p(self.z, C.c_double)
        s.channel, s.enabled = _p(self.channel, C.c_int32), _p(self.enabled, C.c_uint8)
        s.rxprob, s.txprob = _p(self.rxprob, C.c_double), _p(self.txprob, C.c_double)
        s.id_int = _p(self.id_int, C.c_int32)
        return s


def make_packets(src, power_dbm, channel):
    pk = np.zeros(len(src), dtype=PACKET_DTYPE)
    pk["src"], pk["channel"], pk["power_dbm"] = src, channel, power_dbm
    return pk


def tick(p, nodes, packets, uniforms=None, cap=None, threads=1, want_total=False, want_pairs=True):
    """Run orc_tick.  Returns dict(pairs=structured array in (tx,rx) order, tx_failed, total_dbm, stats)."""
    L = lib()
    T = len(packets)
    packets = np.ascontiguousarray(packets, dtype=PACKET_DTYPE)
    ns = nodes.c_struct()
    u = None if uniforms is None else np.ascontiguousarray(uniforms, dtype=np.float64)
    if cap is None:
        cap = max(1, nodes.n * max(T, 1)) if want_pairs else 0
    out = np.zeros(max(cap, 1), dtype=PAIR_DTYPE) if want_pairs else None
    failed = np.zeros(max(T, 1), dtype=np.uint8)
    total = np.zeros(max(nodes.n, 1), dtype=np.float64) if want_total else None
    st = Stats()
    L.orc_tick(C.byref(p), C.byref(ns), packets.ctypes.data_as(C.POINTER(Packet)), T,
               _p(u, C.c_double), 0 if u is None else len(u),
               None if out is None else out.ctypes.data_as(C.POINTER(Pair)), cap if want_pairs else 0,
               _p(failed, C.c_uint8), _p(total, C.c_double), threads, C.byref(st))
    return dict(pairs=None if out is None else out[:st.n_pairs], tx_failed=failed[:T],
                total_dbm=None if total is None else total[:nodes.n],
                n_pairs=st.n_pairs, n_needed=st.n_needed, n_uniforms_used=st.n_uniforms_used,
                n_decisions=st.n_decisions, status=st.status)


def jrandom_doubles(seed, n):
    out = np.zeros(n, dtype=np.float64)
    lib().orc_jrandom_fill_doubles(seed, _p(out, C.c_double), n)
    return out
