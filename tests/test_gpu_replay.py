"""GPU: BASELINE configs[0] and friends replayed over the JSON radio protocol with the CUDA medium behind the
host mirror (radio_sim_b200.simulator.CudaRadioMedium -> rmb_tick), compared MESSAGE FOR MESSAGE with the same
replay driven by the oracle-backed test double.  For the reference's own media the double is EAGER -- it evaluates
every transmit() at once, on its own, with draws from the simulator's java.util.Random, exactly as the reference's
media do -- so these tests also show that batching the tick changes nothing observable."""
import numpy as np
import pytest

from oracle_medium import OracleRadioMedium
from replay import Session, grid10

pytestmark = pytest.mark.gpu
MODEL_OF = {"udgm": "udgm", "nullrm": "null", "n2n-link": "n2n", "udgm-constant": "udgm-constant", "sinr-b200": "sinr"}


def _run(factory, option, script, keep=None):
    s = Session(factory, n_emulators=2)
    params = {"propagation-option": option}
    if option == "n2n-link":
        rng = np.random.default_rng(9)
        params.update({"number-of-nodes": 10, "matrix-data": [float(v) for v in (rng.uniform(size=100) > 0.4).astype(float)]})
    s.cmd(s.ctrl, "configuration-set", parameters=params)
    grid10(s, emulator_of=lambda i: i % 2)
    script(s)
    out = [e.sent for e in s.emus] + [s.ctrl.sent]
    m = s.sim.radio_medium
    if keep is not None:
        keep.update(uploads=getattr(m, "uploads", None), updates=getattr(m, "updates", None), ticks=getattr(m, "ticks", None),
                    seed=s.sim.get_random().seed)
    if hasattr(m, "close"):
        m.close()
    return out


def _script(s):
    e0, e1 = s.emus
    s.step_to(1000, [((e0, e1)[i % 2], i + 1, 0, "0102030405", {}) for i in range(10)])   # config 1: one broadcast per node
    s.cmd(e1, "node-config-set", parameters={"node-id": 4, "radio-state": "disabled"})
    s.cmd(e0, "node-config-set", parameters={"node-id": 7, "position": [61.0, 29.0, 2.0], "rf-power": -7.0})
    s.step_to(2000, [(e0, 7, 1500, "aabbcc", {}), (e0, 1, 1900, "0102030405", {}), (e1, 2, 1950, "01", {"wireless-channel": 11})])
    s.cmd(e1, "node-config-set", parameters={"node-id": 6, "wireless-channel": 11})
    s.step_to(3000, [(e1, 2, 2100, "0102", {"wireless-channel": 11, "rf-power": 3.0})])
    s.step_to(4000)


def _stochastic_script(s):
    """rx-loss / tx-loss below 1 on several nodes, several ticks: every draw comes out of ONE java.util.Random, so a
    tick that over- or under-consumes shifts every later decision."""
    e0, e1 = s.emus
    for i, (rx, tx) in enumerate([(0.5, 1.0), (0.7, 0.6), (1.0, 0.5), (0.3, 1.0), (0.9, 0.9), (0.5, 0.5), (1.0, 1.0), (0.8, 0.2), (0.6, 1.0), (0.4, 0.7)]):
        s.cmd((e0, e1)[i % 2], "node-config-set", parameters={"node-id": i + 1, "rx-loss": rx, "tx-loss": tx})
    for k in range(8):
        # even ticks: everybody broadcasts (under SINR that is one big collision); odd ticks: two distant nodes
        who = range(10) if k % 2 == 0 else (k % 10, (k + 5) % 10)
        s.step_to(1000 * (k + 1), [((e0, e1)[i % 2], i + 1, 1000 * k + 10 * i, "0102030405", {}) for i in who])
    s.step_to(10000)


def _mid_tick_script(s):
    """A node-config-set BETWEEN two transmissions of one tick: the reference evaluates each transmit() against the
    node state of that moment; a node created mid-tick does not hear what was sent before it existed."""
    e0, e1 = s.emus
    s.cmd(s.ctrl, "time-set", parameters={"time": 1000})
    s.handler.handle_message(e0, {"command": "transmit", "node-id": 1, "time": 0, "packet-data": "0102"})
    s.cmd(e1, "node-config-set", parameters={"node-id": 2, "position": [500.0, 500.0]})          # node 2 walks out of range
    s.cmd(e0, "node-config-set", parameters={"node-id": 11, "position": [15.0, 0.0]})            # a new node next to node 1
    s.handler.handle_message(e0, {"command": "transmit", "node-id": 1, "time": 100, "packet-data": "0304"})
    s.cmd(e0, "node-config-set", parameters={"node-id": 6, "radio-state": "disabled"})
    s.handler.handle_message(e1, {"command": "transmit", "node-id": 2, "time": 200, "packet-data": "0506"})
    for e in s.emus:
        for m in [m for m in e.sent if m.get("command") == "time-step"][-1:]:
            s.handler.handle_message(e, {"reply": "OK", "id": m["id"]})
    s.step_to(2000)


@pytest.mark.parametrize("option", ["udgm", "nullrm", "udgm-constant", "n2n-link", "sinr-b200"])
@pytest.mark.parametrize("bulk", [True, False], ids=["bulk-events", "per-pair-events"])
@pytest.mark.parametrize("devices", [(0,), (0, 0)], ids=["one-device", "two-shards"])
def test_replay_matches_the_oracle_backed_double(orc, option, bulk, devices):
    import radio_sim_b200
    from radio_sim_b200.simulator import CudaRadioMedium
    radio_sim_b200.load()
    if option == "n2n-link" and len(devices) > 1:
        pytest.skip("the matrix medium lives on one device (rmb_create refuses RMB_MODEL_N2N with n_shards > 1)")
    want = _run(lambda o, p: OracleRadioMedium(MODEL_OF[o]), option, _script)
    info = {}
    got = _run(lambda o, p: CudaRadioMedium(MODEL_OF[o], bulk_events=bulk, devices=devices), option, _script, keep=info)
    assert sum(1 for m in want[0] + want[1] if m.get("command") == "receive") > 5
    assert got == want
    # node-config-set on registered nodes went through rmb_update_positions / rmb_update_radio, not a re-upload
    assert info["uploads"] == 1 and info["updates"] >= 2


@pytest.mark.parametrize("option", ["udgm", "n2n-link", "sinr-b200"])
@pytest.mark.parametrize("hook", [True, False], ids=["dirty-hook", "shadow-scan"])
def test_stochastic_replay_draws_from_the_simulators_random(orc, option, hook):
    import radio_sim_b200
    from radio_sim_b200.simulator import CudaRadioMedium
    radio_sim_b200.load()
    winfo, ginfo = {}, {}
    want = _run(lambda o, p: OracleRadioMedium(MODEL_OF[o]), option, _stochastic_script, keep=winfo)

    def factory(o, p):
        m = CudaRadioMedium(MODEL_OF[o], scan_nodes=not hook)
        if not hook:
            m.node_config_changing = lambda node: None          # a reference tree without the handler hook
        return m
    got = _run(factory, option, _stochastic_script, keep=ginfo)
    rx = [m for m in want[0] + want[1] if m.get("command") == "receive"]
    assert len(rx) > (3 if option == "sinr-b200" else 10)
    assert got == want
    assert ginfo["seed"] == winfo["seed"]                        # the generator is left exactly where the reference leaves it


@pytest.mark.parametrize("option", ["sinr-b200", "udgm"])
def test_stochastic_replay_with_the_receivers_on_two_handles(orc, option):
    """The same script on a group of two handles: rmb_group_tick gathers the shards' draw counts, the messages and
    the generator's final state are those of the oracle medium driven by the same java.util.Random."""
    import radio_sim_b200
    from radio_sim_b200.simulator import CudaRadioMedium
    radio_sim_b200.load()
    winfo, ginfo = {}, {}
    want = _run(lambda o, p: OracleRadioMedium(MODEL_OF[o]), option, _stochastic_script, keep=winfo)
    got = _run(lambda o, p: CudaRadioMedium(MODEL_OF[o], devices=(0, 0)), option, _stochastic_script, keep=ginfo)
    assert len([m for m in want[0] + want[1] if m.get("command") == "receive"]) > 3
    assert got == want and ginfo["seed"] == winfo["seed"]


@pytest.mark.parametrize("option", ["udgm", "nullrm", "udgm-constant"])
def test_node_config_set_in_the_middle_of_a_tick(orc, option):
    import radio_sim_b200
    from radio_sim_b200.simulator import CudaRadioMedium
    radio_sim_b200.load()
    want = _run(lambda o, p: OracleRadioMedium(MODEL_OF[o]), option, _mid_tick_script)
    got = _run(lambda o, p: CudaRadioMedium(MODEL_OF[o]), option, _mid_tick_script)
    assert sum(1 for m in want[0] + want[1] if m.get("command") == "receive") >= 3
    assert got == want


def test_time_step_node_info_from_the_device_buffers(orc):
    """8(f)2: with rssi_in_band the idle RSSI in every `time-step` node-info is the node's total in-band power of the
    last evaluated tick, read once from the device (rmb_node_info_f64); a receiving node still reports the RSSI of
    the frame it is receiving (Transciever.java:52-61)."""
    import radio_sim_b200
    from radio_sim_b200.simulator import CudaRadioMedium
    radio_sim_b200.load()
    s = Session(lambda o, p: CudaRadioMedium("sinr", rssi_in_band=True), n_emulators=2)
    s.cmd(s.ctrl, "configuration-set", parameters={"propagation-option": "sinr-b200"})
    grid10(s, emulator_of=lambda i: i % 2)
    e0, e1 = s.emus
    s.step_to(1000, [(e0, 1, 0, "0102030405", {}), (e1, 10, 0, "0102030405", {})])
    s.step_to(2000)                                              # all events of tick 1 are over: every node idle again
    x = np.array([30.0 * (i % 5) for i in range(10)]); y = np.array([30.0 * (i // 5) for i in range(10)])
    ref = orc.tick(orc.params(orc.MODEL_SINR), orc.NodeSet(x, y), orc.make_packets(np.array([0, 9]), 0.0, 26), want_total=True)
    steps = [m for e in s.emus for m in e.sent if m.get("command") == "time-step"]
    last = {}
    for m in steps:
        if m["parameters"]["time"] == 2000:
            for ni in m["parameters"]["node-info"]:
                last[int(ni["node-id"])] = ni
    assert sorted(last) == list(range(1, 11))
    for nid, ni in last.items():
        assert ni["receiving"] == 0
        assert abs(ni["rssi"] - ref["total_dbm"][nid - 1]) < 1e-3, (nid, ni["rssi"], ref["total_dbm"][nid - 1])
    assert any(ni["rssi"] > -99.0 for ni in last.values())
    # the reference's rule without the flag: idle = baseRSSI
    first = [m for m in steps if m["parameters"]["time"] == 1000][0]
    assert all(ni["rssi"] == -100.0 for ni in first["parameters"]["node-info"])
    s.sim.radio_medium.close()
