"""CPU: a SECOND, independent restatement of the reference's transmit() loops -- pure-Python loops written straight from
the Java (small cases only) -- against the C oracle on random inputs.  It shares no code with oracle/medium_oracle.c:
the draw order, the short-circuits and the filters are spelled out again here, and the generator is the Python port of
java.util.Random.  Two restatements that agree on thousands of random decisions do not pin the oracle to a JVM, but a
transcription slip in either one would show."""
import math
import numpy as np
import pytest

from helpers import random_deployment, pick_sources
from radio_sim_b200.javarandom import JavaRandom

NOT_HEARD, INTERFERED, DELIVERED = 0, 1, 2


def distance(x, y, z, a, b):
    # Position.getDistance (Position.java:56-64)
    dx, dy, dz = x[a] - x[b], y[a] - y[b], z[a] - z[b]
    dx, dy, dz = dx * dx, dy * dy, dz * dz
    return math.sqrt(dx + dy + dz)


def java_transmit(model, x, y, z, channel, enabled, rxprob, txprob, packets, rnd, R=50.0, ratio_rx=1.0, crange=100.0, matrix=None):
    """packets: (src, power, channel) in arrival order -> rows (t, node, outcome, rssi), tx_failed, draws"""
    n = len(x)
    rows, failed, draws = [], [], 0

    def draw():
        nonlocal draws
        draws += 1
        return rnd.next_double()

    for t, (src, power, pch) in enumerate(packets):
        if model == "udgm":                                   # UDGMRadioMedium.java:63-65 (successRatioRx, sic)
            tx_success = ratio_rx * txprob[src]
        elif model == "n2n":                                  # N2NRadioMedium.java:24-26
            tx_success = txprob[src]
        else:
            tx_success = 1.0
        interference = False
        if model in ("udgm", "n2n"):                          # UDGMRadioMedium.java:86-92, N2NRadioMedium.java:44-49
            if tx_success <= 0.0 or (tx_success < 1.0 and draw() > tx_success):
                interference = True
        failed.append(1 if interference else 0)
        for node in range(n):
            if node == src or not enabled[node] or channel[node] != pch:      # UDGMRadioMedium.java:100-102
                continue
            if model == "null":                               # NullRadioMedium.java:63-72
                rows.append((t, node, DELIVERED, power))
                continue
            if model == "udgm-constant":                      # UDGMConstantLossRadioMedium.java:28-31
                if distance(x, y, z, src, node) < crange:
                    rows.append((t, node, DELIVERED, power))
                continue
            if model == "udgm":                               # UDGMRadioMedium.java:67-81
                d = distance(x, y, z, src, node)
                if R == 0.0:
                    rx_success = 0.0
                else:
                    ratio = (d * d) / (R * R)
                    rx_success = 0.0 if ratio > 1.0 else (1.0 - ratio * (1.0 - ratio_rx)) * rxprob[node]
            else:                                             # N2NRadioMedium.java:28-37, ids = index + 1
                sid, did = src + 1, node + 1
                if matrix is None or sid > len(matrix) or did > len(matrix[sid - 1]):
                    rx_success = 0.0
                else:
                    rx_success = matrix[sid - 1][did - 1] * rxprob[node]
            if rx_success <= 0.0:
                continue
            if interference or (rx_success < 1.0 and draw() > rx_success):   # short-circuit = no draw when the tx failed
                rows.append((t, node, INTERFERED, power))
            else:
                rows.append((t, node, DELIVERED, power))
    return rows, failed, draws


def oracle_rows(orc, omodel, x, y, z, ch, en, rxp, txp, packets, seed, **params):
    src = np.array([p[0] for p in packets], dtype=np.int32)
    pw = np.array([p[1] for p in packets], dtype=np.float64)
    pc = np.array([p[2] for p in packets], dtype=np.int32)
    u = orc.jrandom_doubles(seed, 20000)
    r = orc.tick(orc.params(omodel, **params), orc.NodeSet(x, y, z, channel=ch, enabled=en, rxprob=rxp, txprob=txp),
                 orc.make_packets(src, pw, pc), uniforms=u)
    p = r["pairs"]
    rows = [(int(a), int(b), int(c), float(d)) for a, b, c, d in zip(p["tx"], p["rx"], p["outcome"], p["rssi_dbm"])]
    return rows, r["tx_failed"].tolist(), int(r["n_uniforms_used"])


def random_case(seed, n, t, side, lossy):
    rng = np.random.default_rng(seed)
    x, y, z, ch = random_deployment(seed, n, side, channels=2, z_spread=15.0)
    en = (rng.uniform(size=n) > 0.1).astype(np.uint8)
    if lossy:
        rxp = rng.choice([1.0, 1.0, 0.9, 0.5, 0.0], size=n)
        txp = rng.choice([1.0, 1.0, 0.8, 0.3, 0.0], size=n)
    else:
        rxp, txp = np.ones(n), np.ones(n)
    src = pick_sources(seed, n, t)
    src[-1] = src[0]                                           # one node sends twice in the tick
    power = rng.choice([-25.0, -7.0, 0.0], size=t)
    pch = ch[src].copy(); pch[::5] = 11                        # RadioPacket channel override (RadioPacket.java:50,85)
    return x, y, z, ch, en, rxp, txp, [(int(s), float(p), int(c)) for s, p, c in zip(src, power, pch)]


@pytest.mark.parametrize("seed", range(6))
@pytest.mark.parametrize("model", ["null", "udgm-constant", "udgm", "n2n"])
@pytest.mark.parametrize("lossy", [False, True], ids=["deterministic", "rx-tx-loss"])
def test_reference_media_second_restatement(orc, model, lossy, seed):
    n, t = 120, 14
    x, y, z, ch, en, rxp, txp, packets = random_case(100 * seed + 7, n, t, 140.0, lossy)
    omodel = {"null": orc.MODEL_NULL, "udgm-constant": orc.MODEL_UDGM_CONSTANT, "udgm": orc.MODEL_UDGM, "n2n": orc.MODEL_N2N}[model]
    kw, okw = {}, {}
    if model == "udgm" and seed % 2:
        kw["ratio_rx"] = okw["success_ratio_rx"] = 0.85        # the SUCCESS_RATIO_RX distance law + the tx rule's quirk
    if model == "n2n":
        rng = np.random.default_rng(seed)
        m = rng.choice([0.0, 0.0, 1.0, 0.6], size=(n, n))
        kw["matrix"] = m.tolist(); okw["matrix"] = m
    want = java_transmit(model, x, y, z, ch, en, rxp, txp, packets, JavaRandom(seed), **kw)
    got = oracle_rows(orc, omodel, x, y, z, ch, en, rxp, txp, packets, seed, **okw)
    assert got[0] == want[0]
    if model in ("udgm", "n2n"):
        assert got[1] == want[1]
    assert got[2] == want[2]
    assert len(want[0]) > 20
    if lossy and model in ("udgm", "n2n"):
        assert want[2] > 10 and sum(want[1]) > 0


@pytest.mark.parametrize("seed", range(4))
def test_sinr_medium_second_restatement(orc, seed):
    """The SINR medium has no Java counterpart: the oracle is its specification.  Restated once more
    with libm's log10 / exp (the oracle uses fdlibm, unfused): values within 1e-9 dB, decisions identical wherever the
    margin to a threshold exceeds that."""
    n, t = 150, 12
    x, y, z, ch, en, _, _, packets = random_case(900 + seed, n, t, 120.0, False)
    src = np.array([p[0] for p in packets], dtype=np.int32)
    pw = np.array([p[1] for p in packets]); pc = np.array([p[2] for p in packets], dtype=np.int32)
    r = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x, y, z, channel=ch, enabled=en), orc.make_packets(src, pw, pc))
    got = {(int(a), int(b)): (int(c), float(d), float(e)) for a, b, c, d, e in
           zip(r["pairs"]["tx"], r["pairs"]["rx"], r["pairs"]["outcome"], r["pairs"]["rssi_dbm"], r["pairs"]["sinr_db"])}
    PL0, NEXP, D0, SENS, THR, NOISE = 40.05, 3.0, 1.0, -95.0, 4.0, -100.0
    ln10_10 = math.log(10.0) / 10.0
    seen = 0
    for node in range(n):
        if not en[node]:
            assert not any(k[1] == node for k in got)
            continue
        prx = {}
        for k, (s, p, c) in enumerate(packets):
            if s == node or c != ch[node]:
                continue
            d = max(distance(x, y, z, s, node), D0)
            prx[k] = p - (PL0 + 10.0 * NEXP * math.log10(d / D0))
        total = math.exp(NOISE * ln10_10) + sum(math.exp(v * ln10_10) for v in prx.values())
        for k, v in prx.items():
            if abs(v - SENS) < 1e-9:
                continue
            if v < SENS:
                assert (k, node) not in got
                continue
            o, rssi, sinr = got[(k, node)]
            want_sinr = v - 10.0 * math.log10(total - math.exp(v * ln10_10))
            assert abs(rssi - v) < 1e-9 and abs(sinr - want_sinr) < 1e-7
            if abs(want_sinr - THR) > 1e-6:
                assert o == (DELIVERED if want_sinr >= THR else INTERFERED)
            seen += 1
    assert seen == len(got) and seen > 50
