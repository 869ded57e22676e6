"""TEST DOUBLE: a RadioMedium for the host mirror whose link decisions come from the CPU oracle instead of
librmb200.so -- so that the host logic (event timing, transceiver state, JSON protocol) is testable without a
GPU, and so that the GPU replay can be compared with it message for message.  Same transmit()/flush() shape
as radio_sim_b200.simulator.CudaRadioMedium.  Test infrastructure only."""
import numpy as np
import oracle_py as orc
from radio_sim_b200.simulator import AbstractRadioMedium

MODELS = {"null": orc.MODEL_NULL, "udgm": orc.MODEL_UDGM, "udgm-constant": orc.MODEL_UDGM_CONSTANT, "n2n": orc.MODEL_N2N,
          "sinr": orc.MODEL_SINR}


class OracleRadioMedium(AbstractRadioMedium):
    def __init__(self, model="udgm", **params):
        super().__init__()
        self.model, self.params, self.pending, self.matrix = model, params, [], None

    def set_matrix(self, matrix):
        m = np.asarray(matrix, dtype=np.float64)
        n = int(round(m.size ** 0.5))
        self.matrix = m.reshape(n, n)

    def node_state_changed(self):
        pass

    def transmit(self, packet):
        if self.model != "udgm-constant":
            self.simulator.generate_transmission_events(packet)
        self.pending.append(packet)

    def flush(self):
        packets, self.pending = self.pending, []
        nodes = self.simulator.get_nodes()
        index = {id(nd): i for i, nd in enumerate(nodes)}
        keep = [p for p in packets if id(p.node) in index]
        if not keep:
            return 0
        f = lambda g, dt: np.array([g(nd) for nd in nodes], dtype=dt)
        ns = orc.NodeSet(f(lambda nd: nd.position.x, np.float64), f(lambda nd: nd.position.y, np.float64),
                         f(lambda nd: nd.position.z, np.float64), channel=f(lambda nd: nd.radio.wireless_channel, np.int32),
                         enabled=f(lambda nd: 1 if nd.radio.is_enabled else 0, np.uint8),
                         rxprob=f(lambda nd: nd.radio.rx_probability, np.float64),
                         txprob=f(lambda nd: nd.radio.tx_probability, np.float64),
                         id_int=f(lambda nd: nd.get_id_as_integer(), np.int32))
        kw = dict(self.params)
        if self.matrix is not None:
            kw["matrix"] = self.matrix
        pk = orc.make_packets(np.array([index[id(p.node)] for p in keep], dtype=np.int32),
                              np.array([p.txpower for p in keep]), np.array([p.channel for p in keep], dtype=np.int32))
        pairs = orc.tick(orc.params(MODELS[self.model], **kw), ns, pk)["pairs"]
        for t, rx, rssi, outcome in zip(pairs["tx"], pairs["rx"], pairs["rssi_dbm"], pairs["outcome"]):
            if self.model == "udgm-constant":
                self.simulator.deliver_radio_packet(keep[t], nodes[rx], float(rssi))
            else:
                self.simulator.generate_reception_events(keep[t], nodes[rx], float(rssi), outcome == 2)
        return len(pairs)
