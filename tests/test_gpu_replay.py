"""GPU: BASELINE configs[0] and friends replayed over the JSON radio protocol with the CUDA medium behind the
host mirror (radio_sim_b200.simulator.CudaRadioMedium -> rmb_tick), compared MESSAGE FOR MESSAGE with the same
replay driven by the oracle-backed test double."""
import numpy as np
import pytest

from oracle_medium import OracleRadioMedium
from replay import Session, grid10

pytestmark = pytest.mark.gpu
MODEL_OF = {"udgm": "udgm", "nullrm": "null", "n2n-link": "n2n", "udgm-constant": "udgm-constant", "sinr-b200": "sinr"}


def _run(factory, option, script):
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


@pytest.mark.parametrize("option", ["udgm", "nullrm", "udgm-constant", "n2n-link", "sinr-b200"])
def test_replay_matches_the_oracle_backed_double(orc, option):
    import radio_sim_b200
    from radio_sim_b200.simulator import CudaRadioMedium
    radio_sim_b200.load()
    want = _run(lambda o, p: OracleRadioMedium(MODEL_OF[o]), option, _script)
    got = _run(lambda o, p: CudaRadioMedium(MODEL_OF[o]), option, _script)
    assert sum(1 for m in want[0] + want[1] if m.get("command") == "receive") > 5
    assert got == want
