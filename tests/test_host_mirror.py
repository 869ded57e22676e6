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
