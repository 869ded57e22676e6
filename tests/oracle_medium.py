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
    """eager=True is the REFERENCE's behaviour for its own media: every transmit() is evaluated at once, on its own,
    against the node state of that moment, with the draws taken from simulator.get_random() -- the batched CUDA adapter
    must reproduce it message for message.  eager=False (the only meaningful mode for the SINR medium, which
    evaluates the tick's packets jointly) batches like the adapter does."""

    def __init__(self, model="udgm", eager=None, **params):
        super().__init__()
        self.model, self.params, self.pending, self.matrix = model, params, [], None
        self.eager = (model != "sinr") if eager is None else eager

    def set_matrix(self, matrix):
        m = np.asarray(matrix, dtype=np.float64)
        n = int(round(m.size ** 0.5))
        self.matrix = m.reshape(n, n)

    def node_config_changing(self, node):
        if not self.eager:
            self.flush()

    def transmit(self, packet):
        self.pending.append(packet)
        if self.eager:
            self.flush()

    def flush(self, step_time=0):
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
        rnd = self.simulator.get_random()
        res = orc.tick(orc.params(MODELS[self.model], **kw), ns, pk, uniforms=rnd.peek_doubles(len(keep) * (len(nodes) + 1)))
        rnd.skip_doubles(res["n_uniforms_used"])
        pairs = res["pairs"]
        k = 0
        for t, p in enumerate(keep):                               # the reference's call order (UDGMRadioMedium.java:97-114)
            if self.model != "udgm-constant":
                self.simulator.generate_transmission_events(p)
            while k < len(pairs) and pairs["tx"][k] == t:
                rx, rssi, outcome = pairs["rx"][k], float(pairs["rssi_dbm"][k]), pairs["outcome"][k]
                if self.model == "udgm-constant":
                    self.simulator.deliver_radio_packet(p, nodes[rx], rssi)
                else:
                    self.simulator.generate_reception_events(p, nodes[rx], rssi, outcome == 2)
                k += 1
        return len(pairs)
