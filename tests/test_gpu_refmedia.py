"""-m gpu: the reference's own media (Null / UDGM / UDGMConstantLoss / N2N) on the GPU, through the
C-ABI, against the hand-derived known answers and against the oracle on random deployments."""
import json
import os
import numpy as np
import pytest
from helpers import random_deployment, pick_sources, compare

pytestmark = pytest.mark.gpu
KAT = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "kat_reference_media.json")))
OMODEL = {"null": 0, "udgm": 1, "udgm-constant": 2, "n2n": 3}


@pytest.fixture(scope="module")
def rsb():
    import radio_sim_b200
    return radio_sim_b200


@pytest.mark.parametrize("c", KAT["cases"], ids=[c["name"] for c in KAT["cases"]])
def test_known_answers_on_gpu(rsb, c):
    n = len(c["x"])
    with rsb.Medium(c["model"], f64_values=True, **c.get("params", {})) as m:
        m.set_nodes(np.array(c["x"], float), np.array(c["y"], float), channel=c.get("channel"), enabled=c.get("enabled"),
                    rxprob=c.get("rxprob"), txprob=c.get("txprob"), txpower_dbm=c.get("txpower"))
        if "matrix" in c:
            m.set_matrix(np.array(c["matrix"], float))
        u = None
        if "seed" in c:
            from radio_sim_b200.javarandom import JavaRandom
            u = JavaRandom(c["seed"]).next_doubles(64)               # simulator.getRandom() == new Random(seed)
        r = m.tick(np.array(c["src"], dtype=np.int32), channel=c.get("pkt_channel"), uniforms=u)
    got = [[int(t), int(x), int(o)] for t, x, o in zip(r.tx, r.rx, r.outcome)]
    if "draws" in c:
        assert r.n_uniforms_used == c["draws"]
    assert got == c["expect"], c["why"]
    if "tx_failed" in c:
        assert r.tx_failed.tolist() == c["tx_failed"]
    if "rssi" in c:
        assert r.rssi_dbm_f64.tolist() == c["rssi"]
    assert np.isnan(r.sinr_db).all()


def run_ref(rsb, orc, model, x, y, z, ch, src, uniforms=None, rxprob=None, txprob=None, enabled=None, matrix=None, **params):
    with rsb.Medium(model, f64_values=True, **params) as m:
        m.set_nodes(x, y, z, channel=ch, rxprob=rxprob, txprob=txprob, enabled=enabled)
        if matrix is not None:
            m.set_matrix(matrix)
        r = m.tick(src, uniforms=uniforms)
        _, recv = m.node_info()
    kw = dict(params)
    if matrix is not None:
        kw["matrix"] = matrix
    ref = orc.tick(orc.params(OMODEL[model], **kw), orc.NodeSet(x, y, z, channel=ch, rxprob=rxprob, txprob=txprob, enabled=enabled),
                   orc.make_packets(src, 0.0, ch[src]), uniforms=uniforms)
    compare(r, ref)
    if uniforms is not None:
        assert r.n_uniforms_used == ref["n_uniforms_used"]
    heard = np.zeros(len(x), dtype=np.uint8)
    heard[ref["pairs"]["rx"]] = 1
    assert np.array_equal(recv, heard)
    return r, ref


@pytest.mark.parametrize("model", ["null", "udgm", "udgm-constant"])
def test_random_deployment_deterministic(rsb, orc, model):
    x, y, z, ch = random_deployment(21, 3000, 400.0, channels=3, z_spread=5.0)
    en = (np.random.default_rng(22).uniform(size=3000) > 0.05).astype(np.uint8)
    r, _ = run_ref(rsb, orc, model, x, y, z, ch, pick_sources(21, 3000, 60), enabled=en)
    assert r.n_pairs > 0


def test_udgm_points_exactly_on_the_range_boundary(rsb, orc):
    """Pythagorean offsets of length exactly 50 and their one-ulp neighbours around every transmitter."""
    rng = np.random.default_rng(23)
    trip = np.array([[30, 40], [40, 30], [14, 48], [48, 14], [50, 0], [0, 50], [-30, 40], [-14, -48]], float)
    cx, cy = np.round(rng.uniform(100, 900, 40)), np.round(rng.uniform(100, 900, 40))
    xs, ys = [cx], [cy]
    for dx, dy in trip:
        for bump in (0.0, 1.0, -1.0):
            px, py = cx + dx, cy + dy
            if bump:
                py = np.nextafter(py, py + bump * 1e3)
            xs.append(px); ys.append(py)
    x, y = np.concatenate(xs), np.concatenate(ys)
    n = len(x)
    r, ref = run_ref(rsb, orc, "udgm", x, y, np.zeros(n), np.full(n, 26, dtype=np.int32), np.arange(40, dtype=np.int32))
    assert 0 < r.n_pairs < 40 * (n - 1)


def test_udgm_stochastic_draw_order(rsb, orc):
    x, y, z, ch = random_deployment(24, 2000, 300.0)
    rng = np.random.default_rng(24)
    rxprob = np.where(rng.uniform(size=2000) < 0.5, 0.7, 1.0); rxprob[::41] = 0.0
    txprob = np.where(rng.uniform(size=2000) < 0.5, 0.6, 1.0); txprob[::53] = 0.0
    u = orc.jrandom_doubles(7, 200000)
    r, ref = run_ref(rsb, orc, "udgm", x, y, z, ch, pick_sources(24, 2000, 120), uniforms=u, rxprob=rxprob, txprob=txprob,
                     success_ratio_rx=0.8)
    assert ref["n_uniforms_used"] > 1000 and 0 < ref["tx_failed"].sum() < 120


def test_udgm_uniforms_exhausted_reports_error(rsb, orc):
    x, y, z, ch = random_deployment(25, 500, 100.0)
    with rsb.Medium("udgm", success_ratio_rx=0.5) as m:
        m.set_nodes(x, y, z, channel=ch)
        with pytest.raises(rsb.MediumError) as e:
            m.tick(pick_sources(25, 500, 5), uniforms=np.array([0.1, 0.2]))
        assert e.value.code == -5
        u = np.array([0.1, 0.2])
        r = m.tick(pick_sources(25, 500, 5), uniforms=u, allow_short_uniforms=True)
    ref = orc.tick(orc.params(orc.MODEL_UDGM, success_ratio_rx=0.5), orc.NodeSet(x, y, z, channel=ch),
                   orc.make_packets(pick_sources(25, 500, 5), 0.0, 26), uniforms=u)
    assert ref["status"] == -2
    compare(r, ref)          # missing draws read as 0.0 on both sides
    assert r.n_uniforms_used == ref["n_uniforms_used"]


def test_n2n_random_matrix(rsb, orc):
    n = 300
    rng = np.random.default_rng(26)
    mat = rng.choice([0.0, 0.3, 1.0], size=(n, n), p=[0.7, 0.2, 0.1])
    x = np.zeros(n)
    u = orc.jrandom_doubles(9, 100000)
    run_ref(rsb, orc, "n2n", x, x, x, np.full(n, 26, dtype=np.int32), pick_sources(26, n, 50), uniforms=u, matrix=mat.reshape(-1))


def test_errors(rsb):
    with pytest.raises(rsb.MediumError) as e:
        rsb.Medium("sinr", device=99)
    assert e.value.code == -1
    with rsb.Medium("udgm") as m:
        with pytest.raises(rsb.MediumError) as e:
            m.tick(np.array([0], dtype=np.int32))
        assert e.value.code == -6          # tick before set_nodes
        m.set_nodes(np.zeros(4), np.zeros(4))
        with pytest.raises(rsb.MediumError) as e:
            m.tick(np.array([7], dtype=np.int32))
        assert e.value.code == -1          # source index out of range
    with rsb.Medium("null", max_pairs=10) as m:
        m.set_nodes(np.zeros(8), np.zeros(8))
        with pytest.raises(rsb.MediumError) as e:
            m.tick(np.arange(8, dtype=np.int32))
        assert e.value.code == -4          # 56 heard pairs > max_pairs
