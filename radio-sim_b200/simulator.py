"""Host-side mirror of the reference simulator AROUND the medium: the objects `RadioMedium.transmit`
reads and the calls it makes back.  File names below are relative to
radio-medium/java/se/sics/emul8/radiomedium/ in the reference tree.

This image has no JVM, so this Python module stands where the Java `Simulator` stands: same class and
method names, same arithmetic (clamp of the packet start time, air time, strict `<` event processing,
transceiver state machine, node-array order == neighbour index).  It exists so that the drop-in can be
exercised end to end -- JSON commands in (jsonproto.py), `receive` messages out -- through the C-ABI.
It is NOT a second implementation of the medium: every link decision comes from librmb200.so.
"""
import heapq

import numpy as np

from .javarandom import JavaRandom
from .medium import Group

LISTENING, TRANSMITTING, RECEIVING, DISABLED = 0, 1, 2, 3          # Transciever.java:7-10


class Position:
    """Position.java:37-64."""

    def __init__(self):
        self.x = self.y = self.z = 0.0

    def set(self, x, y, z=0.0):                                    # Position.java:44-52: set(x, y) zeroes z
        self.x, self.y, self.z = float(x), float(y), float(z)


class Transciever:
    """Transciever.java:5-113 (the reference's spelling)."""

    def __init__(self, node):
        self.node = node
        self.txpower, self.wireless_channel, self.is_enabled = 0.0, 26, True       # Transciever.java:11-13
        self.receiving_packet = self.sending_packet = None
        self.receiving_rssi = 0.0
        self.rx_probability = self.tx_probability = 1.0                             # Transciever.java:17-18

    def get_rssi(self):                                            # Transciever.java:52-61
        if self.receiving_packet is not None:
            return self.receiving_rssi
        medium = self.node.simulator.radio_medium
        return medium.get_base_rssi(self.node) if medium is not None else -99.99

    def get_receiving_state(self):                                 # Transciever.java:67-78
        if not self.is_enabled:
            return DISABLED
        if self.receiving_packet is not None:
            return RECEIVING
        if self.sending_packet is not None:
            return TRANSMITTING
        return LISTENING

    def set_receiving(self, packet, rssi):                         # Transciever.java:80-84: overwrites an earlier reception
        self.sending_packet = None
        self.receiving_packet, self.receiving_rssi = packet, rssi

    def clear_receiving(self):                                     # Transciever.java:86-88
        self.receiving_packet = None

    def set_sending(self, packet):                                 # Transciever.java:106-109
        self.receiving_packet = None
        self.sending_packet = packet

    def clear_sending(self):
        self.sending_packet = None


class Node:
    """Node.java:39-67."""

    def __init__(self, node_id, simulator):
        self.id, self.simulator = str(node_id), simulator
        self.position, self.radio = Position(), Transciever(self)
        self.client = None

    def get_id_as_integer(self):                                   # Node.java:52-58,65-67: -1 if not numeric
        try:
            return int(self.id)
        except ValueError:
            return -1


class RadioPacket:
    """RadioPacket.java:46-111: power and channel are snapshots of the source radio at creation."""

    def __init__(self, node, time, packet_data):
        self.node, self.time, self.packet_data = node, int(time), packet_data
        self.txpower, self.channel = node.radio.txpower, node.radio.wireless_channel     # RadioPacket.java:49-50

    def get_packet_air_time(self):
        # RadioPacket.java:66-73: 32 us per HEX CHARACTER, i.e. 64 us per byte (twice what its comment says)
        return len(self.packet_data) * 32 if self.packet_data is not None else 0

    def get_end_time(self):
        return self.time + self.get_packet_air_time()

    def to_json_destination(self, destination, rssi):              # RadioPacket.java:101-111, field order kept
        return {"command": "receive", "node-id": destination.id, "time-start": self.time, "time-end": self.get_end_time(),
                "rf-power": rssi, "wireless-channel": self.channel, "packet-data": self.packet_data}


class AbstractRadioMedium:
    """AbstractRadioMedium.java:36-52 / RadioMedium.java:35-45."""

    base_rssi = -100.0                                             # AbstractRadioMedium.java:38

    def __init__(self):
        self.simulator = None

    def get_name(self):
        return type(self).__name__

    def set_simulator(self, simulator):
        self.simulator = simulator

    def get_base_rssi(self, node):
        return self.base_rssi

    def transmit(self, packet):
        raise NotImplementedError


class CudaRadioMedium(AbstractRadioMedium):
    """The adapter of INTEGRATION.md -- java/se/sics/emul8/radiomedium/CudaRadioMedium.java is the same logic in Java,
    statement for statement, and this is the copy that runs (tests/test_gpu_replay.py).

    transmit() only queues.  flush(), the FIRST thing Simulator.emulatorTimeStepDone does (before `currentTime =
    stepTime`, Simulator.java:155-157), evaluates all queued packets with ONE rmb_tick and hands the result over
    packet by packet in arrival order: the packet's two transmission events, then its reception events in
    node-array order -- the reference's own call sequence (UDGMRadioMedium.java:97-114), so equal-time events are
    queued in the reference's insertion order.  'udgm-constant' delivers at once and raises no events
    (UDGMConstantLossRadioMedium.java:16-36).

    Node state reaches the device incrementally: a new node list (copy-on-write registry) is a full rmb_set_nodes;
    in-place changes (node-config-set) go through rmb_update_positions / rmb_update_radio for exactly the changed
    indices, found by node_config_changing() (the hook the JSON handler calls BEFORE it mutates a node: packets
    already queued are evaluated against the state they were sent under, as the reference does inside transmit())
    or by a per-tick scan against the shadow copy (scan_nodes, default on, for callers without the hook).

    Stochastic reception draws from simulator.get_random() (Simulator.java:94) in the reference's order: the tick
    consumes host-supplied uniforms and reports how many; a look-ahead of the generator supplies them and the
    generator is then advanced by exactly n_uniforms_used (`uniforms=` overrides the source for tests)."""

    def __init__(self, model="udgm", uniforms=None, devices=(0,), max_tx_per_tick=16384, rssi_in_band=False, scan_nodes=True,
                 bulk_events=True, **params):
        super().__init__()
        self.model = model
        # rmb_group_*: one or several GPUs behind the same calls (receivers in contiguous node blocks, one handle each)
        self.medium = Group(model, devices=tuple(devices), max_tx_per_tick=max_tx_per_tick, f64_values=True, compact=True, **params)
        self.base_rssi = float(params.get("noise_floor_dbm", self.base_rssi))
        self.pending = []
        self.uniforms = uniforms           # optional JavaRandom-like object (peek_doubles / skip_doubles) instead of simulator.get_random()
        self.rssi_in_band, self.scan_nodes, self.bulk_events = rssi_in_band, scan_nodes, bulk_events
        self._registered = None            # the node list as last uploaded (identity == registry version)
        self._index = {}                   # id(node) -> array index
        self._shadow = None                # what the device holds: dict of numpy arrays
        self._dirty = set()
        self._node_info = None
        self.ticks = self.uploads = self.updates = 0

    def get_name(self):
        return "cuda-" + self.model

    def close(self):
        self.medium.close()

    def set_matrix(self, matrix):                                  # N2NRadioMedium.setMatrix, N2NRadioMedium.java:20-22
        self.medium.set_matrix(matrix)

    # -- node table ------------------------------------------------------------------------------------------------
    _FIELDS = (("x", np.float64, lambda nd: nd.position.x), ("y", np.float64, lambda nd: nd.position.y),
               ("z", np.float64, lambda nd: nd.position.z), ("txpower", np.float64, lambda nd: nd.radio.txpower),
               ("rxprob", np.float64, lambda nd: nd.radio.rx_probability), ("txprob", np.float64, lambda nd: nd.radio.tx_probability),
               ("channel", np.int32, lambda nd: nd.radio.wireless_channel), ("enabled", np.uint8, lambda nd: 1 if nd.radio.is_enabled else 0))

    def _read(self, nodes):
        n = len(nodes)
        return {k: (np.array([g(nd) for nd in nodes], dtype=dt) if n else np.zeros(0, dtype=dt)) for k, dt, g in self._FIELDS}

    def _upload(self, nodes):
        s = self._read(nodes)
        ids = np.array([nd.get_id_as_integer() for nd in nodes], dtype=np.int32)          # Node.java:65-67 (N2N)
        self.medium.set_nodes(s["x"], s["y"], s["z"], txpower_dbm=s["txpower"], channel=s["channel"], enabled=s["enabled"],
                              rxprob=s["rxprob"], txprob=s["txprob"], id_int=ids)
        self._registered, self._shadow = nodes, s
        self._index = {id(nd): i for i, nd in enumerate(nodes)}
        self._dirty.clear()
        self._node_info = None
        self.uploads += 1

    def _push_changes(self, nodes):
        if self.scan_nodes and nodes:
            now = self._read(nodes)
            changed = np.zeros(len(nodes), dtype=bool)
            for k in now:
                changed |= now[k] != self._shadow[k]
            self._dirty.update(int(i) for i in np.nonzero(changed)[0])
        if not self._dirty:
            return
        idx = np.array(sorted(self._dirty), dtype=np.int32)
        sub = self._read([nodes[i] for i in idx])
        for k in sub:
            self._shadow[k][idx] = sub[k]
        self.medium.update_positions(sub["x"], sub["y"], sub["z"], idx=idx)
        self.medium.update_radio(idx=idx, txpower_dbm=sub["txpower"], channel=sub["channel"], enabled=sub["enabled"],
                                 rxprob=sub["rxprob"], txprob=sub["txprob"])
        self._dirty.clear()
        self._node_info = None
        self.updates += 1

    def node_config_changing(self, node):
        """Hook of net/SimulatorJSONHandler.java:105 (node-config-set), BEFORE the node is mutated or created."""
        self.flush()
        if node is not None and id(node) in self._index:
            self._dirty.add(self._index[id(node)])

    def node_state_changed(self):
        """Compatibility with callers that only know `something changed`: the next flush re-reads every node."""
        if self._registered is not None:
            self._dirty.update(range(len(self._registered)))

    # -- the tick --------------------------------------------------------------------------------------------------
    def transmit(self, packet):
        if self.simulator is None:
            return                                                 # NullRadioMedium.java:50-53
        self.pending.append(packet)

    def flush(self, step_time=0):
        packets, self.pending = self.pending, []
        if not packets or self.simulator is None:
            return 0
        nodes = self.simulator.get_nodes()
        if nodes is not self._registered:
            self._upload(nodes)                                    # copy-on-write registry: a new list is a registry change
        else:
            self._push_changes(nodes)
        keep = [p for p in packets if id(p.node) in self._index]   # a source that left the registry cannot be evaluated
        if not keep:
            return 0
        src = np.array([self._index[id(p.node)] for p in keep], dtype=np.int32)
        power = np.array([p.txpower for p in keep], dtype=np.float64)            # RadioPacket.java:49,77: snapshot / override
        channel = np.array([p.channel for p in keep], dtype=np.int32)            # RadioPacket.java:50,85
        # deterministic tick: one call.  With draws: the first call (no uniforms; a missing draw reads as 0.0 = success,
        # which maximises the number of later draws) tells how many uniforms the tick can consume at most; the
        # generator's look-ahead supplies them and the generator is advanced by the number actually consumed.
        # (copies, not views of the handles' pinned slabs: the events raised below outlive this tick)
        r = self.medium.tick(src, power, channel, allow_short_uniforms=True)
        if r.status != 0:
            rnd = self.uniforms if self.uniforms is not None else self.simulator.get_random()
            r = self.medium.tick(src, power, channel, uniforms=rnd.peek_doubles(r.n_uniforms_used))
            rnd.skip_doubles(r.n_uniforms_used)
        self.ticks += 1
        self._node_info = None
        sim = self.simulator
        # packet t sits at gathered record position position[t] of every shard's lists; its rows in the reference's
        # order are shard 0's rows of that position, then shard 1's, ... (shards own increasing node blocks)
        position = np.empty(len(keep), dtype=np.int64)
        filled = np.flatnonzero(r.packet_of >= 0)
        position[r.packet_of[filled]] = filled
        tx_first = [p.tx_first for p in r.parts]
        rx, rssi, outcome = [p.rx for p in r.parts], [p.rssi_dbm_f64 for p in r.parts], [p.outcome for p in r.parts]
        if self.model == "udgm-constant":
            for t, p in enumerate(keep):                           # UDGMConstantLossRadioMedium.java:30-32
                for s in range(len(r.parts)):
                    for i in range(tx_first[s][position[t]], tx_first[s][position[t] + 1]):
                        sim.deliver_radio_packet(p, nodes[rx[s][i]], float(rssi[s][i]))
        elif self.bulk_events:
            sim.generate_reception_events_bulk(keep, len(keep), position, len(r.parts), tx_first, nodes, rx, rssi, outcome)
        else:
            for t, p in enumerate(keep):
                sim.generate_transmission_events(p)
                for s in range(len(r.parts)):
                    for i in range(tx_first[s][position[t]], tx_first[s][position[t] + 1]):
                        sim.generate_reception_events(p, nodes[rx[s][i]], float(rssi[s][i]), outcome[s][i] == 2)
        return int(r.n_pairs)

    # -- node-info from the device buffers (8(f)2) -----------------------------------------------------------------------
    def _refresh_node_info(self):
        if self._node_info is None and self._registered:
            self._node_info = self.medium.node_info_f64()
        return self._node_info

    def get_base_rssi(self, node):                                 # Transciever.getRSSI falls back to this (Transciever.java:52-61)
        if not self.rssi_in_band:
            return self.base_rssi                                  # AbstractRadioMedium.java:45-48
        i = self._index.get(id(node))
        info = self._refresh_node_info()
        return self.base_rssi if i is None or info is None or i >= len(info[0]) else float(info[0][i])

    def node_rssi_for_time_step(self, nodes):
        """Idle RSSI of every node for the per-tick node-info (net/JSONClientConnection.java:326-353), index == node
        array index: total in-band power of the last evaluated tick; None = use Transciever.get_rssi() as the reference."""
        if not self.rssi_in_band or nodes is not self._registered:
            return None
        info = self._refresh_node_info()
        return None if info is None else info[0]

    def node_info(self):
        """Total in-band power and 'heard something' per node for the last tick (rmb_group_node_info_f64)."""
        return self.medium.node_info_f64()


class Simulator:
    """Simulator.java: node registry (:245-310), time stepping (:155-194), event queue and the calls a medium
    makes back (:321-364).

    Events of equal time: the reference's ladder queue (com/botbox/scheduler/EventQueue.java:215-222) pops them
    in an order that depends on its bucket layout; here they pop in insertion order.  Parity with the reference
    is therefore defined on the SET of messages per tick, as SURVEY.md 8(f) notes."""

    TIME_STEP_OK = 0                                               # Simulator.java:61

    def __init__(self, random=None):
        self.random = random if random is not None else JavaRandom(0)          # Simulator.java:64,83-89 (`new Random()` there)
        self.nodes, self.node_table = [], {}
        self.emulators, self.event_listeners = [], []
        self.radio_medium = None
        self.time_controller = None
        self.current_time = 0
        self.step_time_value = 0
        self.waiting_for_time_id = -1
        self.time_controller_last_time_id = -1
        self.emulators_left = 0
        self._message_id = 0
        self._queue, self._seq = [], 0

    # -- registry (Simulator.java:245-310): array index == neighbour index; remove compacts ------------
    def get_nodes(self):
        return self.nodes

    def get_node(self, node_id):
        return self.node_table.get(str(node_id))

    def add_node(self, node_id, client):
        node = self.node_table.get(str(node_id))
        if node is not None:
            return node
        if client is not None and client not in self.emulators:
            self.emulators.append(client)
        node = Node(node_id, self)
        node.client = client
        self.node_table[node.id] = node
        self.nodes = self.nodes + [node]                           # copy-on-write, util/ArrayUtils.java:48-79
        return node

    def remove_node(self, node_id):
        node = self.node_table.pop(str(node_id), None)
        if node is None:
            return None
        self.nodes = [n for n in self.nodes if n is not node]
        if node.client is not None and not any(n.client is node.client for n in self.nodes):
            self.emulators = [e for e in self.emulators if e is not node.client]
        return node

    def get_random(self):                                          # Simulator.java:94
        return self.random

    def set_radio_medium(self, medium):                            # Simulator.java:234-239
        if medium is not None:
            medium.set_simulator(self)
        self.radio_medium = medium

    # -- what a medium calls back (Simulator.java:321-364) -------------------------------------------------
    def _add_event(self, time, fn):
        heapq.heappush(self._queue, (time, self._seq, fn))
        self._seq += 1

    def generate_reception_events(self, packet, destination, rssi, do_deliver):
        t0 = max(packet.time, self.current_time)                   # Simulator.java:323-326
        radio = destination.radio
        self._add_event(t0, lambda: radio.set_receiving(packet, rssi))             # events/ReceptionEvent.java:37-39

        def end():                                                 # events/ReceptionEvent.java:40-45
            radio.clear_receiving()
            if do_deliver:
                self.deliver_radio_packet(packet, destination, rssi)
        self._add_event(t0 + packet.get_packet_air_time(), end)

    def generate_reception_events_bulk(self, packets, n_packets, position, n_shards, tx_first, nodes, rx, rssi, outcome):
        """reference-edits.patch `Simulator.generateReceptionEventsBulk` + events/BulkReceptionEvent.java: per packet, in
        arrival order, the two transmission events, then per shard ONE start and ONE end event covering the packet's
        rows tx_first[s][position[t]] .. tx_first[s][position[t] + 1] of that shard's lists (shards in node order)."""
        for t in range(n_packets):
            packet = packets[t]
            self.generate_transmission_events(packet)
            t0 = max(packet.time, self.current_time)
            for s in range(n_shards):
                lo, hi = int(tx_first[s][position[t]]), int(tx_first[s][position[t] + 1])
                if hi <= lo:
                    continue

                def start(packet=packet, lo=lo, hi=hi, rx=rx[s], rssi=rssi[s]):
                    for i in range(lo, hi):
                        nodes[rx[i]].radio.set_receiving(packet, float(rssi[i]))

                def end(packet=packet, lo=lo, hi=hi, rx=rx[s], rssi=rssi[s], outcome=outcome[s]):
                    for i in range(lo, hi):
                        dest = nodes[rx[i]]
                        dest.radio.clear_receiving()
                        if outcome[i] == 2:
                            self.deliver_radio_packet(packet, dest, float(rssi[i]))
                self._add_event(t0, start)
                self._add_event(t0 + packet.get_packet_air_time(), end)

    def generate_transmission_events(self, packet):
        t0 = max(packet.time, self.current_time)                   # Simulator.java:339-342
        radio = packet.node.radio
        self._add_event(t0, lambda: radio.set_sending(packet))                     # events/TransmissionEvent.java:20-26
        self._add_event(t0 + packet.get_packet_air_time(), radio.clear_sending)

    def deliver_radio_packet(self, packet, destination, rssi):     # Simulator.java:356-364
        cc = destination.client
        if cc is None or not cc.is_connected():
            return
        cc.send_packet(packet, destination, -1, rssi)

    # -- time stepping (Simulator.java:155-194) --------------------------------------------------------------
    def get_next_message_id(self):
        self._message_id += 1
        return self._message_id

    def step_time(self, time, msg_id):
        self.waiting_for_time_id = self.get_next_message_id()
        self.time_controller_last_time_id = msg_id
        self.step_time_value = time
        if not self.emulators:
            self.emulators_left = 0
            self._emulator_time_step_done()
            return
        self.emulators_left = len(self.emulators)
        for em in list(self.emulators):
            em.emulate_to_time(self.get_nodes(), time, self.waiting_for_time_id)

    def emulator_time_stepped(self, client, msg_id, ok):           # Simulator.java:131-153
        if self.waiting_for_time_id < 0 or msg_id != self.waiting_for_time_id:
            return
        if client.set_time(self.step_time_value, msg_id):
            self.emulators_left -= 1
        if self.emulators_left == 0:
            self._emulator_time_step_done()

    def _emulator_time_step_done(self):
        # the additive hook of INTEGRATION.md: the tick's packets are evaluated here, while currentTime is still the
        # time the emulators started from (what the reference's media saw when transmit() ran), before the events
        # of the step are processed
        m = self.radio_medium
        if m is not None and hasattr(m, "flush"):
            m.flush(self.step_time_value)
        self.current_time = self.step_time_value                   # Simulator.java:156-157
        self.waiting_for_time_id = -1
        self.process_all_events(self.current_time)
        if self.time_controller is not None:
            self.time_controller.time_step_done(self.time_controller_last_time_id)

    def process_all_events(self, time):                            # Simulator.java:213-228: strictly before `time`
        while self._queue and self._queue[0][0] < time:
            heapq.heappop(self._queue)[2]()
