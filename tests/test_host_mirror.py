"""CPU: the host mirror of the reference simulator (radio_sim_b200/simulator.py, jsonproto.py) -- rows R9-R12 of
SURVEY.md 8(a) and the JSON protocol of 8(b) -- driven by a medium whose decisions come from the oracle
(tests/oracle_medium.py, a test double).  BASELINE configs[0]: the 10-node grid over the JSON protocol."""
import json
import os
import numpy as np
import pytest

from oracle_medium import OracleRadioMedium
from replay import Session, grid10
from radio_sim_b200 import jsonproto
from radio_sim_b200.simulator import Simulator, RadioPacket, RECEIVING, TRANSMITTING, LISTENING, DISABLED

KAT = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "kat_reference_media.json")))
GRID = [c for c in KAT["cases"] if c["name"] == "grid10_udgm_config1"][0]


@pytest.fixture
def oracle_factory(orc):
    return lambda option, params: OracleRadioMedium({"udgm": "udgm", "nullrm": "null", "n2n-link": "n2n",
                                                     "udgm-constant": "udgm-constant", "sinr-b200": "sinr"}[option])


def test_config1_grid_one_broadcast_per_node(oracle_factory):
    s = Session(oracle_factory)
    s.cmd(s.ctrl, "configuration-set", parameters={"propagation-option": "udgm"})
    grid10(s)
    emu = s.emus[0]
    s.step_to(1000, [(emu, i + 1, 0, "0102030405", {}) for i in range(10)])
    got = s.received()
    # hand-derived delivery matrix of the KAT (packet index, receiver index, outcome 2 = delivered)
    want = {(t, rx) for t, rx, o in GRID["expect"] if o == 2}
    assert len(got) == len(want)
    # one `receive` per delivered pair; fields as RadioPacket.toJsonDestination writes them (RadioPacket.java:101-111)
    for m in got:
        assert list(m) == ["command", "node-id", "time-start", "time-end", "rf-power", "wireless-channel", "packet-data"]
        assert m["time-start"] == 0 and m["time-end"] == 320          # 10 hex chars x 32 us (RadioPacket.java:66-73)
        assert m["rf-power"] == 0.0 and m["wireless-channel"] == 26 and m["packet-data"] == "0102030405"
    # every receiver got exactly the frames of its in-range neighbours: count per receiver matches the matrix
    per_rx = {}
    for m in got:
        per_rx[int(m["node-id"])] = per_rx.get(int(m["node-id"]), 0) + 1
    for rx in range(10):
        assert per_rx.get(rx + 1, 0) == sum(1 for (t, r) in want if r == rx)
    # the controller is told when the step is done (JSONClientConnection.java:369-375)
    assert s.ctrl.sent[-1]["reply"] == "OK"


def test_reception_spanning_a_step_boundary_and_node_info(oracle_factory):
    s = Session(oracle_factory)
    s.cmd(s.ctrl, "configuration-set", parameters={"propagation-option": "udgm"})
    grid10(s)
    emu = s.emus[0]
    s.step_to(1000, [(emu, 1, 900, "0102030405", {"rf-power": -3.0})])   # on air 900..1220
    assert s.received() == []                                        # end flank (1220) not before time 1000
    n1, n2 = s.sim.get_node("1"), s.sim.get_node("2")
    assert n1.radio.get_receiving_state() == TRANSMITTING and n2.radio.get_receiving_state() == RECEIVING
    mark = len(emu.sent)
    s.step_to(2000)
    step = [m for m in emu.sent[mark:] if m.get("command") == "time-step"][0]
    info = {i["node-id"]: i for i in step["parameters"]["node-info"]}
    assert info["2"]["receiving"] == RECEIVING and info["2"]["rssi"] == -3.0     # Transciever.getRSSI while receiving
    assert info["1"]["receiving"] == TRANSMITTING and info["1"]["rssi"] == -100.0  # AbstractRadioMedium.baseRSSI
    assert info["5"]["receiving"] == LISTENING and info["5"]["wireless-channel"] == 26
    got = s.received(mark)
    assert sorted(int(m["node-id"]) for m in got) == [2, 6, 7]       # 30 m, 30 m, 42.4 m neighbours of node 1
    assert all(m["time-start"] == 900 and m["time-end"] == 1220 and m["rf-power"] == -3.0 for m in got)
    assert n2.radio.get_receiving_state() == LISTENING


def test_packet_start_time_is_clamped_to_the_current_time(oracle_factory):
    s = Session(oracle_factory)
    s.cmd(s.ctrl, "configuration-set", parameters={"propagation-option": "udgm"})
    grid10(s)
    s.step_to(1000)
    s.step_to(2000, [(s.emus[0], 1, 10, "01", {})])     # stamped in the past: events start at currentTime = 1000
    # start 1000, end 1064 (< 2000): delivered within this step; the message keeps the packet's own times
    got = s.received()
    assert len(got) == 3 and all(m["time-start"] == 10 and m["time-end"] == 74 for m in got)


def test_events_fire_strictly_before_the_step_time(oracle_factory):
    s = Session(oracle_factory)
    s.cmd(s.ctrl, "configuration-set", parameters={"propagation-option": "udgm"})
    grid10(s)
    s.step_to(1000, [(s.emus[0], 1, 680, "0102030405", {})])         # ends exactly at 1000: not processed (strict <)
    assert s.received() == []
    s.step_to(1001)
    assert len(s.received()) == 3


def test_filters_disabled_and_channel_over_the_protocol(oracle_factory):
    s = Session(oracle_factory)
    s.cmd(s.ctrl, "configuration-set", parameters={"propagation-option": "udgm"})
    grid10(s)
    emu = s.emus[0]
    s.cmd(emu, "node-config-set", parameters={"node-id": 2, "radio-state": "disabled"})
    rid = s.cmd(emu, "node-config-set", parameters={"node-id": 6, "wireless-channel": 11})
    reply = [m for m in emu.sent if m.get("id") == rid][0]
    assert reply["reply"] == "OK" and reply["reply-object"]["node-info"]["wireless-channel"] == 11
    assert s.sim.get_node("2").radio.get_receiving_state() == DISABLED
    s.step_to(1000, [(emu, 1, 0, "01", {})])
    assert sorted(int(m["node-id"]) for m in s.received()) == [7]
    mark = len(emu.sent)
    s.step_to(2000, [(emu, 1, 1000, "01", {"wireless-channel": 11})])   # per-packet channel override reaches node 6 only
    assert sorted(int(m["node-id"]) for m in s.received(mark)) == [6]


def test_two_emulators_and_a_single_time_controller(oracle_factory):
    s = Session(oracle_factory, n_emulators=2)
    s.cmd(s.ctrl, "configuration-set", parameters={"propagation-option": "udgm"})
    grid10(s, emulator_of=lambda i: i % 2)
    s.step_to(1000, [(s.emus[0], 1, 0, "01", {})])
    # node 1 (emulator 0) reaches nodes 2 (emulator 1), 6 (emulator 1), 7 (emulator 0)
    assert sorted(int(m["node-id"]) for m in s.emus[1].sent if m.get("command") == "receive") == [2, 6]
    assert sorted(int(m["node-id"]) for m in s.emus[0].sent if m.get("command") == "receive") == [7]
    steps = [m for m in s.emus[1].sent if m.get("command") == "time-step"]
    assert [i["node-id"] for i in steps[0]["parameters"]["node-info"]] == ["2", "4", "6", "8", "10"]
    rid = s.cmd(s.emus[0], "time-set", parameters={"time": 5})
    err = [m for m in s.emus[0].sent if m.get("id") == rid][0]
    assert err["reply"] == "error" and err["reply-object"]["class"] == "command-error"
    rid = s.cmd(s.ctrl, "configuration-set", parameters={"propagation-option": "udgm"})
    assert [m for m in s.ctrl.sent if m.get("id") == rid][0]["reply-object"]["description"] == "already initialized"


def test_error_replies_and_node_id_quirk(oracle_factory):
    s = Session(oracle_factory)
    emu = s.emus[0]
    rid = s.cmd(emu, "transmit", **{"node-id": 1, "time": 0, "packet-data": "01"})
    assert [m for m in emu.sent if m.get("id") == rid][0]["reply-object"]["description"] == "could not find source node"
    s.cmd(emu, "node-config-set", parameters={"node-id": 1})
    rid = s.cmd(emu, "transmit", **{"node-id": 1, "time": 0, "packet-data": "01"})
    assert [m for m in emu.sent if m.get("id") == rid][0]["reply-object"]["description"] == "no radio medium available"
    rid = s.cmd(emu, "frobnicate")
    assert "unsupported command" in [m for m in emu.sent if m.get("id") == rid][0]["reply-object"]["description"]
    s.cmd(emu, "node-config-set", parameters={"node-id": "7"})       # a JSON STRING keeps its quotes (toString())
    assert s.sim.get_node('"7"') is not None and s.sim.get_node("7") is None
    assert s.sim.get_node('"7"').get_id_as_integer() == -1


def test_node_array_order_is_the_neighbour_index_and_remove_compacts():
    sim = Simulator()
    for i in range(5):
        sim.add_node(i + 1, None)
    assert [n.id for n in sim.get_nodes()] == ["1", "2", "3", "4", "5"]
    assert sim.add_node(3, None) is sim.get_node("3")                # already handled (Simulator.java:252-256)
    sim.remove_node(2)
    assert [n.id for n in sim.get_nodes()] == ["1", "3", "4", "5"]   # indices shift (util/ArrayUtils.java)
    p = RadioPacket(sim.get_node("1"), 5, "0102030405")
    assert p.get_packet_air_time() == 320 and p.get_end_time() == 325 and p.channel == 26 and p.txpower == 0.0


def test_framing_round_trip():
    msgs = [{"command": "time-get", "id": 1}, {"command": "log", "parameters": {"message": 'a } " { b'}}]
    raw = b"RSIM 0.1\r\n" + jsonproto.encode(msgs[0]) + b"\r\n\r\n" + jsonproto.encode(msgs[1], use_length=True) + jsonproto.encode(msgs[0])
    assert jsonproto.decode_stream(raw) == [msgs[0], msgs[1], msgs[0]]
    assert jsonproto.WELCOME["radio-simulator"]["api-version"] == "0.6"


# ---- round 2: java.util.Random port, bulk event hand-off, the adapter's host logic without a GPU -------------------
def test_java_random_port_matches_the_jdk_algorithm(orc):
    import ctypes as C
    from radio_sim_b200.javarandom import JavaRandom
    assert [JavaRandom(42).next_double() for _ in range(1)] + list(JavaRandom(42).peek_doubles(3)[1:]) == KAT["jrandom"]["next_double"]
    r = JavaRandom(12345)
    a = r.peek_doubles(5000)
    assert np.array_equal(a, orc.jrandom_doubles(12345, 5000))           # the oracle's C port (oracle/medium_oracle.c)
    r.skip_doubles(1234)
    assert r.next_double() == a[1234] and np.array_equal(r.next_doubles(10), a[1235:1245])
    st = orc.JRandomState()
    orc.lib().orc_jrandom_init(C.byref(st), 7)
    r = JavaRandom(7)
    for b in [10, 16, 1_000_000, 3, (1 << 30) + 5, 1] * 40:
        assert r.next_int(b) == orc.lib().orc_jrandom_next_int_bound(C.byref(st), b)
    picks = JavaRandom(3).sample_distinct(1000, 200)
    assert len(set(picks.tolist())) == 200 and picks.min() >= 0 and picks.max() < 1000


def test_bulk_event_hand_off_equals_one_event_pair_per_heard_pair():
    """Simulator.generate_reception_events_bulk (reference-edits.patch + BulkReceptionEvent.java) against the
    reference's per-pair calls (Simulator.java:321-350): same messages in the same order, same transceiver states."""
    def run(bulk):
        sim = Simulator()
        emu = jsonproto.JSONClientConnection("emu")
        nodes = [sim.add_node(i + 1, emu) for i in range(6)]
        sim.current_time = 50
        packets = [RadioPacket(nodes[0], 0, "0102030405"), RadioPacket(nodes[3], 200, "01"), RadioPacket(nodes[5], 60, "0102")]
        tx_first, rx = [0, 3, 3, 5], [1, 2, 4, 0, 2]
        rssi, outcome = [-1.0, -2.0, -3.0, -4.0, -5.0], [2, 1, 2, 2, 2]
        nl = sim.get_nodes()
        if bulk:
            # the same rows as two shards' lists: nodes 0-2 on the first shard, 3-5 on the second; position == packet
            first = [[0, 2, 2, 4], [0, 1, 1, 1]]
            rxs = [[1, 2, 0, 2], [4]]
            rss = [[-1.0, -2.0, -4.0, -5.0], [-3.0]]
            outs = [[2, 1, 2, 2], [2]]
            sim.generate_reception_events_bulk(packets, 3, [0, 1, 2], 2, first, nl, rxs, rss, outs)
        else:
            for t, p in enumerate(packets):
                sim.generate_transmission_events(p)
                for i in range(tx_first[t], tx_first[t + 1]):
                    sim.generate_reception_events(p, nl[rx[i]], rssi[i], outcome[i] == 2)
        states = []
        for t in (51, 61, 130, 201, 1000):
            sim.process_all_events(t)
            states.append([(n.radio.get_receiving_state(), n.radio.get_rssi()) for n in nl])
        return emu.sent, states
    assert run(True) == run(False)
    sent, states = run(True)
    assert len(sent) == 4 and all(m["command"] == "receive" for m in sent)             # delivered rows only (outcome 2)
    assert states[0][1][0] == RECEIVING and states[0][0][0] == TRANSMITTING            # clamp to current_time = 50


class _FakeMedium:
    """Stands in for radio_sim_b200.medium.Medium so that the adapter's HOST logic runs without a GPU: records the
    calls and answers a tick with 'every other node hears every packet' (plus one draw per pair when asked)."""

    def __init__(self, *a, **kw):
        import types
        self.cfg = types.SimpleNamespace(noise_floor_dbm=-100.0)
        self.calls, self.n, self.need_draws = [], 0, False

    def set_nodes(self, x, y, z=None, **kw):
        self.n = len(x)
        self.calls.append(("set_nodes", self.n))

    def update_positions(self, x, y, z=None, idx=None):
        self.calls.append(("update_positions", idx.tolist(), list(x)))

    def update_radio(self, idx=None, **kw):
        self.calls.append(("update_radio", idx.tolist(), kw["channel"].tolist()))

    def tick(self, src, power, channel, uniforms=None, allow_short_uniforms=False, copy=True):
        import types
        t = len(src)
        rx = np.concatenate([np.delete(np.arange(self.n), s) for s in src]).astype(np.int32)
        first = np.arange(t + 1, dtype=np.int32) * (self.n - 1)
        need = len(rx) if self.need_draws else 0
        short = need and (uniforms is None or len(uniforms) < need)
        self.calls.append(("tick", src.tolist(), None if uniforms is None else len(uniforms)))
        used = need if short else need // 2                        # "consumes" half of the upper bound
        part = types.SimpleNamespace(rx=rx, tx_first=first, rssi_dbm_f64=np.repeat(power, self.n - 1), outcome=np.full(len(rx), 2, np.uint8))
        return types.SimpleNamespace(status=-5 if short else 0, n_uniforms_used=used, n_pairs=len(rx), parts=[part],
                                     packet_of=np.arange(t, dtype=np.int32))

    def node_info_f64(self):
        return np.full(self.n, -90.0), np.zeros(self.n, np.uint8)

    def close(self):
        pass


def test_adapter_host_logic_without_a_gpu(monkeypatch):
    from radio_sim_b200 import simulator as S
    from radio_sim_b200.javarandom import JavaRandom
    monkeypatch.setattr(S, "Group", _FakeMedium)
    sim = Simulator(JavaRandom(5))
    emu = jsonproto.JSONClientConnection("emu")
    handler = jsonproto.SimulatorJSONHandler(sim, lambda o, p: S.CudaRadioMedium("udgm", rssi_in_band=True))
    handler.handle_message(emu, {"command": "configuration-set", "id": 1, "parameters": {"propagation-option": "udgm-b200"}})
    med = sim.radio_medium
    for i in range(4):
        handler.handle_message(emu, {"command": "node-config-set", "parameters": {"node-id": i + 1, "position": [10.0 * i, 0.0]}})
    tx = lambda nid, t: handler.handle_message(emu, {"command": "transmit", "node-id": nid, "time": t, "packet-data": "01"})
    tx(1, 0)
    assert med.medium.calls == []                                  # transmit() only queues
    sim.step_time(1000, 7)
    handler.handle_message(emu, {"reply": "OK", "id": sim.waiting_for_time_id})
    assert med.medium.calls == [("set_nodes", 4), ("tick", [0], None)]
    assert len([m for m in emu.sent if m.get("command") == "receive"]) == 3
    # an in-place change of a registered node: incremental update of exactly that index, no re-upload; and the packet
    # queued BEFORE the change is evaluated before it (node_config_changing flushes)
    tx(2, 1000)
    handler.handle_message(emu, {"command": "node-config-set", "parameters": {"node-id": 3, "wireless-channel": 11, "position": [5.0, 5.0]}})
    tx(4, 1001)
    sim.step_time(2000, 8)
    handler.handle_message(emu, {"reply": "OK", "id": sim.waiting_for_time_id})
    assert med.medium.calls[2:] == [("tick", [1], None), ("update_positions", [2], [5.0]), ("update_radio", [2], [11]), ("tick", [3], None)]
    assert med.uploads == 1 and med.updates == 1
    # draws: first call without uniforms (-5) tells the upper bound, the look-ahead supplies it, the generator advances
    # by what the tick consumed
    med.medium.need_draws = True
    before = sim.get_random().seed
    tx(1, 2000)
    sim.step_time(3000, 9)
    handler.handle_message(emu, {"reply": "OK", "id": sim.waiting_for_time_id})
    assert med.medium.calls[-2:] == [("tick", [0], None), ("tick", [0], 3)]
    probe = JavaRandom(0); probe.seed = before; probe.skip_doubles(1)          # the fake consumed 3 // 2 = 1 draw
    assert sim.get_random().seed == probe.seed
    # a node removed while its packet is queued: the packet is dropped, the others are evaluated
    med.medium.need_draws = False
    tx(2, 3000); tx(1, 3001)
    sim.remove_node("2")
    sim.step_time(4000, 10)
    handler.handle_message(emu, {"reply": "OK", "id": sim.waiting_for_time_id})
    assert med.medium.calls[-2:] == [("set_nodes", 3), ("tick", [0], None)]
    # node-info of the next time-step comes from the device buffers for idle nodes (rssi_in_band)
    step = [m for m in emu.sent if m.get("command") == "time-step"][-1]
    sim.step_time(5000, 11)
    step = [m for m in emu.sent if m.get("command") == "time-step"][-1]
    assert [ni["rssi"] for ni in step["parameters"]["node-info"]] == [-90.0, -90.0, -90.0]


def test_link_quality_is_accepted_and_changes_nothing(oracle_factory):
    """`link-quality` is a logged TODO in the reference (net/SimulatorJSONHandler.java:155-168): the command gets the
    plain OK reply and no medium ever sees it -- the same tick delivers the same frames before and after."""
    def run(with_command):
        s = Session(oracle_factory)
        s.cmd(s.ctrl, "configuration-set", parameters={"propagation-option": "udgm"})
        grid10(s)
        emu = s.emus[0]
        if with_command:
            rid = s.cmd(emu, "link-quality", link={"src": 1, "dst": 2, "quality": 0}, **{"wireless-channel": 26})
            assert [m for m in emu.sent if m.get("id") == rid][0]["reply"] == "OK"
        s.step_to(1000, [(emu, i + 1, 0, "0102030405", {}) for i in range(10)])
        return sorted((int(m["node-id"]), m["time-start"], m["packet-data"]) for m in s.received())
    assert run(True) == run(False) and len(run(False)) > 0
