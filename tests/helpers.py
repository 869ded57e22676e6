"""Shared helpers of the parity tests: deployments and the GPU-vs-oracle comparison."""
import numpy as np

SINR_TOL_DB = 1e-3     # north_star: RSSI and SINR agree within 1e-3 dB; sets bit-exact


def random_deployment(seed, n, side, channels=None, z_spread=0.0):
    rng = np.random.default_rng(seed)
    x, y = rng.uniform(0, side, n), rng.uniform(0, side, n)
    z = rng.uniform(0, z_spread, n) if z_spread else np.zeros(n)
    ch = np.full(n, 26, dtype=np.int32) if channels is None else (11 + (np.arange(n) % channels)).astype(np.int32)
    return x, y, z, ch


def pick_sources(seed, n, t):
    return np.random.default_rng(seed + 1000).permutation(n)[:t].astype(np.int32)


def compare(r, ref, exact_rssi=True, tol=SINR_TOL_DB):
    """GPU TickResult `r` against oracle dict `ref`: index lists and outcomes bit-exact, values in tolerance."""
    p = ref["pairs"]
    assert r.n_decisions == ref["n_decisions"], (r.n_decisions, ref["n_decisions"])
    assert r.n_pairs == ref["n_pairs"], (r.n_pairs, ref["n_pairs"])
    assert np.array_equal(r.tx, p["tx"]), "tx index list differs"
    assert np.array_equal(r.rx, p["rx"]), "rx (neighbour) index list differs"
    assert np.array_equal(r.outcome, p["outcome"]), "delivered/interfered sets differ"
    assert np.array_equal(r.tx_failed, ref["tx_failed"])
    assert r.n_delivered == int((p["outcome"] == 2).sum()) and r.n_interfered == int((p["outcome"] == 1).sum())
    if r.n_pairs:
        assert np.abs(r.rssi_dbm.astype(np.float64) - p["rssi_dbm"]).max() < tol
        fin = np.isfinite(p["sinr_db"])
        if fin.any():
            assert np.abs(r.sinr_db.astype(np.float64)[fin] - p["sinr_db"][fin]).max() < tol
        if r.rssi_dbm_f64 is not None and exact_rssi:
            assert np.array_equal(r.rssi_dbm_f64, p["rssi_dbm"]), "exact-path RSSI not bit-identical"
            if fin.any():
                assert np.abs(r.sinr_db_f64[fin] - p["sinr_db"][fin]).max() < tol
