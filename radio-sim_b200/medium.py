"""Host-side handle over the C-ABI: one `Medium` == one `rmb_handle` == one RadioMedium instance.

Mirrors what the reference's media expose (radio-medium/java/se/sics/emul8/radiomedium/
RadioMedium.java:35-45 and the setters of UDGMRadioMedium.java:26-61), batched per tick.
All evaluation happens in librmb200.so on the GPU; this module only marshals numpy arrays.
"""
import ctypes as C
import numpy as np
from . import _abi as A


class MediumError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"rmb error {code}: {msg}")
        self.code = code


def _ptr(a, ct):
    return None if a is None else a.ctypes.data_as(C.POINTER(ct))


def _arr(a, dt, n=None):
    if a is None:
        return None
    a = np.ascontiguousarray(a, dtype=dt)
    if n is not None and a.shape != (n,):
        raise ValueError(f"expected shape ({n},), got {a.shape}")
    return a


_from_memory = C.pythonapi.PyMemoryView_FromMemory
_from_memory.restype, _from_memory.argtypes = C.py_object, [C.c_void_p, C.c_ssize_t, C.c_int]
_PyBUF_WRITE = 0x200


def _column(p, dt, m, copy):
    """numpy view (or copy) of m elements at the ctypes pointer p.  (np.ctypeslib.as_array builds a new ctypes array
    type for every new length -- 45 us a column when the pair count changes from tick to tick.)"""
    if not p or m == 0:
        return np.zeros(0, dtype=dt)
    v = np.frombuffer(_from_memory(C.addressof(p.contents), m * dt.itemsize, _PyBUF_WRITE), dtype=dt)
    return v.copy() if copy else v


class TickResult:
    """Heard (tx, rx) pairs of one tick in generateReceptionEvents call order
    (packets in arrival order, receivers in node-array order; Simulator.java:321-335).
    copy=False: the columns are views of the handle's pinned host memory (valid until the handle's next call), made
    when first touched; copy=True: copies, made here."""

    # column -> (rmb_result field, dtype, length attribute, offset added to the length)
    _COLUMNS = {"tx": ("tx", np.dtype(np.int32), "n_pairs", 0), "rx": ("rx", np.dtype(np.int32), "n_pairs", 0),
                "rssi_dbm": ("rssi_dbm", np.dtype(np.float32), "n_pairs", 0), "sinr_db": ("sinr_db", np.dtype(np.float32), "n_pairs", 0),
                "outcome": ("outcome", np.dtype(np.uint8), "n_pairs", 0),
                "rssi_dbm_f64": ("rssi_dbm_f64", np.dtype(np.float64), "n_pairs", 0),
                "sinr_db_f64": ("sinr_db_f64", np.dtype(np.float64), "n_pairs", 0),
                "tx_failed": ("tx_failed", np.dtype(np.uint8), "n_tx", 0), "tx_first": ("tx_first", np.dtype(np.int32), "n_tx", 1)}

    def __init__(self, r, f64, copy=True):
        self.n_pairs, self.n_delivered, self.n_interfered = int(r.n_pairs), int(r.n_delivered), int(r.n_interfered)
        self.n_decisions, self.n_uniforms_used = int(r.n_decisions), int(r.n_uniforms_used)
        self.n_recheck, self.n_tx, self.near_slots = int(r.n_recheck), int(r.n_tx), int(r.near_slots)
        self._r, self._f64, self._copy = r, f64, copy
        if copy:
            for name in self._COLUMNS:
                getattr(self, name)

    def __getattr__(self, name):                     # reached only for a column not made yet
        spec = TickResult._COLUMNS.get(name)
        if spec is None:
            raise AttributeError(name)
        field, dt, length, extra = spec
        p = getattr(self._r, field)
        if (name.endswith("_f64") and not self._f64) or (name == "tx_first" and not p):
            v = None        # compact form (RMB_FLAG_COMPACT) has tx_first and no tx / sinr columns; f64 columns on request
        else:
            v = _column(p, dt, getattr(self, length) + extra, self._copy)
        setattr(self, name, v)
        return v

    def tx_column(self):
        """The tx column, rebuilt from tx_first when the result is in the compact form."""
        if self.tx_first is None:
            return self.tx
        return np.repeat(np.arange(self.n_tx, dtype=np.int32), np.diff(self.tx_first))


class Medium:
    """A radio medium living on one B200.  model: 'null' | 'udgm' | 'udgm-constant' | 'n2n' | 'sinr'."""

    MODELS = {"null": A.MODEL_NULL, "nullrm": A.MODEL_NULL, "udgm": A.MODEL_UDGM,
              "udgm-constant": A.MODEL_UDGM_CONSTANT, "n2n": A.MODEL_N2N, "n2n-link": A.MODEL_N2N,
              "sinr": A.MODEL_SINR, "sinr-b200": A.MODEL_SINR}

    def __init__(self, model="sinr", device=0, f64_values=False, rx_base=0, n_shards=1, max_pairs=0, compact=False, **params):
        self._lib = A.load()
        cfg = A.Config()
        self._lib.rmb_config_default(C.byref(cfg), self.MODELS[model] if isinstance(model, str) else int(model))
        cfg.device, cfg.rx_base, cfg.n_shards, cfg.max_pairs = device, rx_base, n_shards, max_pairs
        if f64_values:
            cfg.flags |= A.FLAG_F64_VALUES
        if compact:
            cfg.flags |= A.FLAG_COMPACT
        for k, v in params.items():
            if not hasattr(cfg, k):
                raise AttributeError(f"unknown medium parameter {k}")
            setattr(cfg, k, v)
        self.cfg = cfg
        self._f64 = bool(f64_values)
        self._h = C.c_void_p()
        rc = self._lib.rmb_create(C.byref(cfg), C.byref(self._h))
        if rc != A.OK:
            msg = self._lib.rmb_last_error(None).decode()
            self._h = None
            raise MediumError(rc, msg)
        self.n = 0

    # -- lifetime ---------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_h", None):
            self._lib.rmb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _ck(self, rc, allow=()):
        if rc != A.OK and rc not in allow:
            raise MediumError(rc, self._lib.rmb_last_error(self._h).decode())
        return rc

    # -- node state (Simulator.addNode + node-config-set) ----------------------------------------
    def set_nodes(self, x, y, z=None, txpower_dbm=None, channel=None, enabled=None, rxprob=None, txprob=None,
                  id_int=None):
        x = _arr(x, np.float64)
        n = x.shape[0]
        y, z = _arr(y, np.float64, n), _arr(z, np.float64, n)
        txpower_dbm, rxprob, txprob = _arr(txpower_dbm, np.float64, n), _arr(rxprob, np.float64, n), _arr(txprob, np.float64, n)
        channel, id_int, enabled = _arr(channel, np.int32, n), _arr(id_int, np.int32, n), _arr(enabled, np.uint8, n)
        self._ck(self._lib.rmb_set_nodes(self._h, n, _ptr(x, C.c_double), _ptr(y, C.c_double), _ptr(z, C.c_double),
                                         _ptr(txpower_dbm, C.c_double), _ptr(channel, C.c_int32),
                                         _ptr(enabled, C.c_uint8), _ptr(rxprob, C.c_double), _ptr(txprob, C.c_double),
                                         _ptr(id_int, C.c_int32)))
        self.n = n

    def set_matrix(self, matrix):
        m = np.ascontiguousarray(matrix, dtype=np.float64)
        n = int(round(m.size ** 0.5))
        if n * n != m.size:
            raise ValueError("matrix must be square (flat row-major n*n)")
        self._ck(self._lib.rmb_set_matrix(self._h, n, _ptr(m.reshape(-1), C.c_double)))

    def update_positions(self, x, y, z=None, idx=None):
        x = _arr(x, np.float64)
        c = x.shape[0]
        y, z, idx = _arr(y, np.float64, c), _arr(z, np.float64, c), _arr(idx, np.int32, c)
        self._ck(self._lib.rmb_update_positions(self._h, c, _ptr(idx, C.c_int32), _ptr(x, C.c_double),
                                                _ptr(y, C.c_double), _ptr(z, C.c_double)))

    def update_positions_async(self, x, y, z=None):
        """Dense update that does not wait for an enqueued tick: uploaded beside it, swapped in by the next settling call.
        The arrays are used as they are (no copy): keep them untouched until that call returns."""
        if x.dtype != np.float64 or y.dtype != np.float64 or not x.flags.c_contiguous or not y.flags.c_contiguous:
            raise ValueError("update_positions_async takes contiguous float64 arrays (they are read asynchronously)")
        self._ck(self._lib.rmb_update_positions_async(self._h, x.shape[0], _ptr(x, C.c_double), _ptr(y, C.c_double),
                                                      None if z is None else _ptr(z, C.c_double)))

    def update_radio(self, idx=None, txpower_dbm=None, channel=None, enabled=None, rxprob=None, txprob=None):
        fields = [f for f in (txpower_dbm, channel, enabled, rxprob, txprob) if f is not None]
        c = len(idx) if idx is not None else (len(fields[0]) if fields else 0)
        idx = _arr(idx, np.int32, c)
        txpower_dbm, rxprob, txprob = _arr(txpower_dbm, np.float64, c), _arr(rxprob, np.float64, c), _arr(txprob, np.float64, c)
        channel, enabled = _arr(channel, np.int32, c), _arr(enabled, np.uint8, c)
        self._ck(self._lib.rmb_update_radio(self._h, c, _ptr(idx, C.c_int32), _ptr(txpower_dbm, C.c_double),
                                            _ptr(channel, C.c_int32), _ptr(enabled, C.c_uint8),
                                            _ptr(rxprob, C.c_double), _ptr(txprob, C.c_double)))

    # -- the tick (batched RadioMedium.transmit) ----------------------------------------------------
    def tick(self, src, power_dbm=None, channel=None, uniforms=None, copy=True, allow_short_uniforms=False):
        src = _arr(src, np.int32)
        t = src.shape[0]
        power_dbm, channel = _arr(power_dbm, np.float64, t), _arr(channel, np.int32, t)
        uniforms = _arr(uniforms, np.float64)
        r = A.Result()
        rc = self._lib.rmb_tick(self._h, t, _ptr(src, C.c_int32), _ptr(power_dbm, C.c_double), _ptr(channel, C.c_int32),
                                _ptr(uniforms, C.c_double), 0 if uniforms is None else uniforms.shape[0], C.byref(r))
        self._ck(rc, allow=(A.ERR_UNIFORMS,) if allow_short_uniforms else ())
        res = TickResult(r, self._f64, copy)
        res.status = rc
        return res

    def build_tx_records(self, src, d_records, power_dbm=None, channel=None):
        """Write this tick's records for locally owned sources to DEVICE memory (int address)."""
        src = _arr(src, np.int32)
        t = src.shape[0]
        power_dbm, channel = _arr(power_dbm, np.float64, t), _arr(channel, np.int32, t)
        self._ck(self._lib.rmb_build_tx_records(self._h, t, _ptr(src, C.c_int32), _ptr(power_dbm, C.c_double),
                                                _ptr(channel, C.c_int32), C.c_void_p(d_records)))

    def tick_records(self, n_tx, d_records, uniforms=None, copy=True):
        uniforms = _arr(uniforms, np.float64)
        r = A.Result()
        self._ck(self._lib.rmb_tick_records(self._h, n_tx, C.c_void_p(d_records), _ptr(uniforms, C.c_double),
                                            0 if uniforms is None else uniforms.shape[0], C.byref(r)))
        return TickResult(r, self._f64, copy)

    def tick_resident(self, n_tx, d_records):
        r = A.Result()
        self._ck(self._lib.rmb_tick_resident(self._h, n_tx, C.c_void_p(d_records), C.byref(r)))
        return TickResult(r, False, True)

    # -- stochastic reception on sharded receivers (global draw order) ----------------------------------
    def tick_draws_begin(self, n_tx, d_records, d_counts):
        """First half: writes this shard's per-record draw counts (int32[n_tx]) to DEVICE memory d_counts."""
        self._ck(self._lib.rmb_tick_draws_begin(self._h, n_tx, C.c_void_p(d_records), C.c_void_p(d_counts)))

    def tick_peers_draws_begin(self, d_counts):
        self._ck(self._lib.rmb_tick_peers_draws_begin(self._h, C.c_void_p(d_counts)))

    def tick_draws_end(self, d_all_counts, world, rank, uniforms, d_order=0, n_order=0, copy=True, allow_short_uniforms=False):
        """Second half: d_all_counts = int32[world][n_tx] on the device (all shards' counts, rank order); d_order =
        device int32[n_order], the record positions in original packet arrival order (0 = as stored)."""
        uniforms = _arr(uniforms, np.float64)
        r = A.Result()
        rc = self._lib.rmb_tick_draws_end(self._h, C.c_void_p(d_all_counts), world, rank, C.c_void_p(d_order or None), n_order,
                                          _ptr(uniforms, C.c_double),
                                          0 if uniforms is None else uniforms.shape[0], C.byref(r))
        self._ck(rc, allow=(A.ERR_UNIFORMS,) if allow_short_uniforms else ())
        res = TickResult(r, self._f64, copy)
        res.status = rc
        return res

    def tick_enqueue(self, n_tx, d_records):
        """Asynchronous tick on records resident in HBM: returns as soon as the work is on the stream."""
        self._ck(self._lib.rmb_tick_enqueue(self._h, n_tx, C.c_void_p(d_records)))

    def tick_peers_enqueue(self):
        self._ck(self._lib.rmb_tick_peers_enqueue(self._h))

    def peer_exchange_tick_enqueue(self, slots, d_records):
        """Exchange + tick in one step (the preparation kernel pushes the records itself); asynchronous."""
        self._ck(self._lib.rmb_peer_exchange_tick_enqueue(self._h, slots, C.c_void_p(d_records)))

    def may_draw(self):
        """True when a tick on this handle may consume uniforms (some rx / tx probability below 1, or the model draws)."""
        return bool(self._lib.rmb_may_draw(self._h))

    def peer_barrier_enqueue(self):
        """Device-side rendezvous of all ranks' streams (no host synchronisation)."""
        self._ck(self._lib.rmb_peer_barrier_enqueue(self._h))

    def tick_finish(self):
        """Settle an enqueued tick; counts only (fetch() downloads the lists)."""
        r = A.Result()
        self._ck(self._lib.rmb_tick_finish(self._h, C.byref(r)))
        return TickResult(r, False, True)

    def fetch(self, copy=True):
        r = A.Result()
        self._ck(self._lib.rmb_fetch(self._h, C.byref(r)))
        return TickResult(r, self._f64, copy)

    # -- multi-GPU exchange over NVLink peer memory (rmb_peer_*) -------------------------------------
    def peer_export(self, world, rank, max_slots):
        """Allocate this handle's gather buffer; returns the blob (bytes) the peers need to map it."""
        blob = C.create_string_buffer(A.PEER_HANDLE_BYTES)
        self._ck(self._lib.rmb_peer_export(self._h, world, rank, max_slots, blob))
        return blob.raw

    def peer_open(self, blobs):
        """blobs: every rank's peer_export() result, in rank order."""
        raw = b"".join(bytes(b) for b in blobs)
        buf = C.create_string_buffer(raw, len(raw))
        self._ck(self._lib.rmb_peer_open(self._h, buf))

    def peer_push(self, src, power_dbm=None, channel=None):
        """src: this rank's padded local packet list (global node indices, -1 = padding)."""
        src = _arr(src, np.int32)
        t = src.shape[0]
        power_dbm, channel = _arr(power_dbm, np.float64, t), _arr(channel, np.int32, t)
        self._ck(self._lib.rmb_peer_push(self._h, t, _ptr(src, C.c_int32), _ptr(power_dbm, C.c_double), _ptr(channel, C.c_int32)))

    def peer_push_records(self, slots, d_records):
        self._ck(self._lib.rmb_peer_push_records(self._h, slots, C.c_void_p(d_records)))

    def tick_peers(self, uniforms=None, copy=True):
        uniforms = _arr(uniforms, np.float64)
        r = A.Result()
        self._ck(self._lib.rmb_tick_peers(self._h, _ptr(uniforms, C.c_double), 0 if uniforms is None else uniforms.shape[0], C.byref(r)))
        return TickResult(r, self._f64, copy)

    def peer_tick(self, src, power_dbm=None, channel=None, copy=True):
        """peer_push + tick_peers in one call (the exchange inside the tick's first kernel where the medium allows).
        Every rank has to be in this call at the same time: one thread or process per handle."""
        src = _arr(src, np.int32)
        t = src.shape[0]
        power_dbm, channel = _arr(power_dbm, np.float64, t), _arr(channel, np.int32, t)
        r = A.Result()
        self._ck(self._lib.rmb_peer_tick(self._h, t, _ptr(src, C.c_int32), _ptr(power_dbm, C.c_double), _ptr(channel, C.c_int32), C.byref(r)))
        return TickResult(r, self._f64, copy)

    def tick_peers_resident(self):
        r = A.Result()
        self._ck(self._lib.rmb_tick_peers_resident(self._h, C.byref(r)))
        return TickResult(r, False, True)

    # -- node-info (Transciever.getRSSI / getBaseRSSI) ---------------------------------------------
    def node_info(self):
        rssi = np.zeros(max(self.n, 1), dtype=np.float32)
        recv = np.zeros(max(self.n, 1), dtype=np.uint8)
        self._ck(self._lib.rmb_node_info(self._h, _ptr(rssi, C.c_float), _ptr(recv, C.c_uint8)))
        return rssi[:self.n], recv[:self.n]

    def node_info_f64(self):
        """The same with the total in-band power in double (what node-info.rssi carries)."""
        rssi = np.zeros(max(self.n, 1), dtype=np.float64)
        recv = np.zeros(max(self.n, 1), dtype=np.uint8)
        self._ck(self._lib.rmb_node_info_f64(self._h, _ptr(rssi, C.c_double), _ptr(recv, C.c_uint8)))
        return rssi[:self.n], recv[:self.n]

    def timing(self):
        t = A.Timing()
        self._ck(self._lib.rmb_last_timing(self._h, C.byref(t)))
        return {k: getattr(t, k) for k, _ in A.Timing._fields_}

    def stream(self):
        """cudaStream_t (int) the handle's work runs on."""
        p = C.c_void_p()
        self._ck(self._lib.rmb_get_stream(self._h, C.byref(p)))
        return p.value or 0

    def set_stream(self, stream):
        """Run on the caller's CUDA stream (int cudaStream_t; 0/None restores the handle's own)."""
        self._ck(self._lib.rmb_set_stream(self._h, C.c_void_p(stream or None)))

    def set_tuning(self, sweep_variant=0, near_slots=0):
        self._ck(self._lib.rmb_set_tuning(self._h, sweep_variant, near_slots))


class GroupResult:
    """One tick of a `Group`: the shards' lists (`parts`, rank order) and the merged view in the reference's call
    order -- packets in arrival order, receivers in node-array order."""

    def __init__(self, parts, slots, packet_of):
        self.parts, self.slots, self.packet_of = parts, slots, packet_of
        self.n_pairs = sum(p.n_pairs for p in parts)
        self.n_decisions = sum(p.n_decisions for p in parts)
        self.n_delivered = sum(p.n_delivered for p in parts)

    def tx_failed(self):
        """Per packet of the caller's list: the transmission itself failed (tx-loss draw).  Every shard walks the same
        tx draws, so shard 0's column says it for all."""
        p0 = np.asarray(self.parts[0].tx_failed)
        valid = np.flatnonzero(self.packet_of >= 0)
        out = np.zeros(len(valid), dtype=np.uint8)
        out[self.packet_of[valid]] = p0[valid]
        return out

    def merged(self):
        """-> dict(tx, rx, outcome, rssi_dbm): tx = index into the caller's packet list.  (The adapter does not need
        this copy: it walks the shards per packet through tx_first.)"""
        cols = {k: np.concatenate([np.asarray(getattr(p, k)) for p in self.parts]) for k in ("rx", "outcome", "rssi_dbm")}
        pos = np.concatenate([p.tx_column() for p in self.parts])
        tx = self.packet_of[pos] if pos.size else pos.astype(np.int32)
        order = np.lexsort((cols["rx"], tx))
        out = {k: v[order] for k, v in cols.items()}
        out["tx"] = tx[order].astype(np.int32)
        return out


class Group:
    """rmb_group_*: several GPUs behind one object (one process, one worker thread per device inside the library)."""

    def __init__(self, model="sinr", devices=(0,), max_tx_per_tick=16384, f64_values=False, compact=False, **params):
        self._lib = A.load()
        cfg = A.Config()
        self._lib.rmb_config_default(C.byref(cfg), Medium.MODELS[model] if isinstance(model, str) else int(model))
        if f64_values:
            cfg.flags |= A.FLAG_F64_VALUES
        if compact:
            cfg.flags |= A.FLAG_COMPACT
        for k, v in params.items():
            if not hasattr(cfg, k):
                raise AttributeError(f"unknown medium parameter {k}")
            setattr(cfg, k, v)
        self._f64, self.world, self.max_tx = bool(f64_values), len(devices), int(max_tx_per_tick)
        dv = np.asarray(devices, dtype=np.int32)
        self._g = C.c_void_p()
        rc = self._lib.rmb_group_create(C.byref(cfg), self.world, _ptr(dv, C.c_int32), self.max_tx, C.byref(self._g))
        if rc != A.OK:
            self._g = None
            raise MediumError(rc, self._lib.rmb_group_last_error(None).decode())
        self.n = 0

    def close(self):
        if getattr(self, "_g", None):
            self._lib.rmb_group_destroy(self._g)
            self._g = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != A.OK:
            raise MediumError(rc, self._lib.rmb_group_last_error(self._g).decode())

    def set_nodes(self, x, y, z=None, txpower_dbm=None, channel=None, enabled=None, rxprob=None, txprob=None, id_int=None):
        x = _arr(x, np.float64)
        n = x.shape[0]
        y, z = _arr(y, np.float64, n), _arr(z, np.float64, n)
        txpower_dbm, rxprob, txprob = _arr(txpower_dbm, np.float64, n), _arr(rxprob, np.float64, n), _arr(txprob, np.float64, n)
        channel, id_int, enabled = _arr(channel, np.int32, n), _arr(id_int, np.int32, n), _arr(enabled, np.uint8, n)
        self._ck(self._lib.rmb_group_set_nodes(self._g, n, _ptr(x, C.c_double), _ptr(y, C.c_double), _ptr(z, C.c_double),
                                               _ptr(txpower_dbm, C.c_double), _ptr(channel, C.c_int32), _ptr(enabled, C.c_uint8),
                                               _ptr(rxprob, C.c_double), _ptr(txprob, C.c_double), _ptr(id_int, C.c_int32)))
        self.n = n

    def set_matrix(self, matrix):
        """N2N: the matrix medium lives on ONE device (rmb_create refuses RMB_MODEL_N2N with n_shards > 1)."""
        m = np.ascontiguousarray(matrix, dtype=np.float64)
        n = int(round(m.size ** 0.5))
        if n * n != m.size:
            raise ValueError("matrix must be square (flat row-major n*n)")
        h = C.c_void_p()
        self._ck(self._lib.rmb_group_shard(self._g, 0, C.byref(h), None, None))
        rc = self._lib.rmb_set_matrix(h, n, _ptr(m.reshape(-1), C.c_double))
        if rc != A.OK:
            raise MediumError(rc, self._lib.rmb_last_error(h).decode())

    def update_positions(self, x, y, z=None, idx=None):
        x = _arr(x, np.float64)
        c = x.shape[0]
        y, z, idx = _arr(y, np.float64, c), _arr(z, np.float64, c), _arr(idx, np.int32, c)
        self._ck(self._lib.rmb_group_update_positions(self._g, c, _ptr(idx, C.c_int32), _ptr(x, C.c_double), _ptr(y, C.c_double),
                                                      _ptr(z, C.c_double)))

    def update_radio(self, idx, txpower_dbm=None, channel=None, enabled=None, rxprob=None, txprob=None):
        idx = _arr(idx, np.int32)
        c = idx.shape[0]
        txpower_dbm, rxprob, txprob = _arr(txpower_dbm, np.float64, c), _arr(rxprob, np.float64, c), _arr(txprob, np.float64, c)
        channel, enabled = _arr(channel, np.int32, c), _arr(enabled, np.uint8, c)
        self._ck(self._lib.rmb_group_update_radio(self._g, c, _ptr(idx, C.c_int32), _ptr(txpower_dbm, C.c_double), _ptr(channel, C.c_int32),
                                                  _ptr(enabled, C.c_uint8), _ptr(rxprob, C.c_double), _ptr(txprob, C.c_double)))

    def tick(self, src, power_dbm=None, channel=None, uniforms=None, copy=True, allow_short_uniforms=False):
        src = _arr(src, np.int32)
        t = src.shape[0]
        power_dbm, channel = _arr(power_dbm, np.float64, t), _arr(channel, np.int32, t)
        uniforms = _arr(uniforms, np.float64)
        res = (A.Result * self.world)()
        slots = C.c_int32(0)
        packet_of = np.full(self.world * max(self.max_tx, t, 1), -1, dtype=np.int32)
        rc = self._lib.rmb_group_tick(self._g, t, _ptr(src, C.c_int32), _ptr(power_dbm, C.c_double), _ptr(channel, C.c_int32),
                                      _ptr(uniforms, C.c_double), 0 if uniforms is None else uniforms.shape[0],
                                      res, C.byref(slots), _ptr(packet_of, C.c_int32))
        if not (allow_short_uniforms and rc == A.ERR_UNIFORMS):
            self._ck(rc)
        parts = [TickResult(res[r], self._f64, copy) for r in range(self.world)]
        out = GroupResult(parts, slots.value, packet_of[:self.world * slots.value])
        out.status = rc
        out.n_uniforms_used = max(p.n_uniforms_used for p in parts)      # several shards: each carries the global count
        return out

    def node_info_f64(self):
        rssi = np.zeros(max(self.n, 1), dtype=np.float64)
        recv = np.zeros(max(self.n, 1), dtype=np.uint8)
        self._ck(self._lib.rmb_group_node_info_f64(self._g, _ptr(rssi, C.c_double), _ptr(recv, C.c_uint8)))
        return rssi[:self.n], recv[:self.n]
