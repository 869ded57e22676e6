"""Command streams of the JSON radio protocol for the replay tests (what an emulator and a time controller send;
cf. the reference's test/TestEmulator.java:91-125 and test/TestClient.java)."""
from radio_sim_b200.javarandom import JavaRandom
from radio_sim_b200.simulator import Simulator
from radio_sim_b200.jsonproto import SimulatorJSONHandler, JSONClientConnection


class Session:
    """One simulator + one controller + `n_emulators` emulator connections, all in memory."""

    def __init__(self, medium_factory, n_emulators=1, seed=42):
        self.sim = Simulator(JavaRandom(seed))                     # Simulator(Random r), Simulator.java:87-89
        self.handler = SimulatorJSONHandler(self.sim, medium_factory)
        self.ctrl = JSONClientConnection("controller")
        self.emus = [JSONClientConnection(f"emulator{i}") for i in range(n_emulators)]
        self._id = 100

    def cmd(self, client, command, **fields):
        self._id += 1
        msg = {"command": command, "id": self._id}
        msg.update(fields)
        self.handler.handle_message(client, msg)
        return self._id

    def step_to(self, time, transmissions=()):
        """time-set from the controller; every emulator receives its time-step, sends its frames, then answers OK."""
        first = {id(e): len(e.sent) for e in self.emus}
        self.cmd(self.ctrl, "time-set", parameters={"time": time})
        for emu, node_id, t, data, extra in transmissions:
            msg = {"command": "transmit", "node-id": node_id, "time": t, "packet-data": data}
            msg.update(extra)
            self.handler.handle_message(emu, msg)
        for e in self.emus:
            steps = [m for m in e.sent[first[id(e)]:] if m.get("command") == "time-step"]
            for m in steps:
                self.handler.handle_message(e, {"reply": "OK", "id": m["id"]})

    def received(self, since=0):
        return [m for e in self.emus for m in e.sent[since:] if m.get("command") == "receive"]


def grid10(session, emulator_of=lambda i: 0):
    """BASELINE configs[0]: 10 nodes on a 5 x 2 grid, 30 m pitch, ids 1..10 (sent as JSON numbers)."""
    for i in range(10):
        session.cmd(session.emus[emulator_of(i)], "node-config-set",
                    parameters={"node-id": i + 1, "position": [30.0 * (i % 5), 30.0 * (i // 5)]})
