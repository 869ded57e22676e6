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

from .medium import Medium

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
    """The adapter of INTEGRATION.md (java/.../CudaRadioMedium.java is the same thing in Java): transmit()
    enqueues the packet and raises the two transmission events at once, as every reference medium does
    (UDGMRadioMedium.java:97); flush(), called when the tick's emulators are done, evaluates all queued
    packets with ONE rmb_tick and replays the heard pairs as generateReceptionEvents calls in the
    reference's own order (packets in arrival order, receivers in node-array order).

    model 'udgm-constant' delivers at once instead (UDGMConstantLossRadioMedium.java:16-36: no events)."""

    def __init__(self, model="udgm", uniforms=None, device=0, **params):
        super().__init__()
        self.model = model
        self.medium = Medium(model, device=device, f64_values=True, **params)
        self.pending = []
        self.uniforms = uniforms           # callable(n) -> n doubles standing in for simulator.getRandom().nextDouble()
        self._registered = None            # node objects as last uploaded (identity == array index)
        self._dirty = True
        self.ticks = 0

    def get_name(self):
        return "cuda-" + self.model

    def close(self):
        self.medium.close()

    def node_state_changed(self):
        """node-config-set / addNode / removeNode happened: the SoA copy is re-uploaded before the next tick."""
        self._dirty = True

    def set_matrix(self, matrix):                                  # N2NRadioMedium.java:20-22
        self.medium.set_matrix(matrix)

    def _upload(self):
        nodes = self.simulator.get_nodes()
        n = len(nodes)
        a = lambda f, dt: np.array([f(nd) for nd in nodes], dtype=dt) if n else np.zeros(0, dtype=dt)
        self.medium.set_nodes(a(lambda nd: nd.position.x, np.float64), a(lambda nd: nd.position.y, np.float64),
                              a(lambda nd: nd.position.z, np.float64),
                              txpower_dbm=a(lambda nd: nd.radio.txpower, np.float64),
                              channel=a(lambda nd: nd.radio.wireless_channel, np.int32),
                              enabled=a(lambda nd: 1 if nd.radio.is_enabled else 0, np.uint8),
                              rxprob=a(lambda nd: nd.radio.rx_probability, np.float64),
                              txprob=a(lambda nd: nd.radio.tx_probability, np.float64),
                              id_int=a(lambda nd: nd.get_id_as_integer(), np.int32))
        self._registered = list(nodes)
        self._dirty = False

    def transmit(self, packet):
        if self.model != "udgm-constant":
            self.simulator.generate_transmission_events(packet)
        self.pending.append(packet)

    def flush(self):
        """End of tick (hook before Simulator.processAllEvents, Simulator.java:155-161)."""
        packets, self.pending = self.pending, []
        if not packets:
            return 0
        if self._dirty or self._registered is None:
            self._upload()
        index = {id(nd): i for i, nd in enumerate(self._registered)}
        keep = [p for p in packets if id(p.node) in index]          # a source removed meanwhile cannot transmit
        if not keep:
            return 0
        src = np.array([index[id(p.node)] for p in keep], dtype=np.int32)
        power = np.array([p.txpower for p in keep], dtype=np.float64)
        channel = np.array([p.channel for p in keep], dtype=np.int32)
        u = None
        if self.uniforms is not None:
            u = np.asarray(self.uniforms(len(keep) * (len(self._registered) + 1)), dtype=np.float64)
        r = self.medium.tick(src, power, channel, uniforms=u)
        self.ticks += 1
        sim = self.simulator
        for t, rx, rssi, outcome in zip(r.tx, r.rx, r.rssi_dbm_f64, r.outcome):
            packet, dest = keep[t], self._registered[rx]
            if self.model == "udgm-constant":
                sim.deliver_radio_packet(packet, dest, float(rssi))
            else:
                sim.generate_reception_events(packet, dest, float(rssi), outcome == 2)
        return int(r.n_pairs)

    def node_info(self):
        """Total in-band power and 'heard something' per node for the last tick (rmb_node_info)."""
        return self.medium.node_info()


class Simulator:
    """Simulator.java: node registry (:245-310), time stepping (:155-194), event queue and the calls a medium
    makes back (:321-364).

    Events of equal time: the reference's ladder queue (com/botbox/scheduler/EventQueue.java:215-222) pops them
    in an order that depends on its bucket layout; here they pop in insertion order.  Parity with the reference
    is therefore defined on the SET of messages per tick, as SURVEY.md 8(f) notes."""

    TIME_STEP_OK = 1

    def __init__(self):
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
        self._state_changed()
        return node

    def remove_node(self, node_id):
        node = self.node_table.pop(str(node_id), None)
        if node is None:
            return None
        self.nodes = [n for n in self.nodes if n is not node]
        if node.client is not None and not any(n.client is node.client for n in self.nodes):
            self.emulators = [e for e in self.emulators if e is not node.client]
        self._state_changed()
        return node

    def _state_changed(self):
        m = self.radio_medium
        if m is not None and hasattr(m, "node_state_changed"):
            m.node_state_changed()

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
            m.flush()
        self.current_time = self.step_time_value                   # Simulator.java:156-157
        self.waiting_for_time_id = -1
        self.process_all_events(self.current_time)
        if self.time_controller is not None:
            self.time_controller.time_step_done(self.time_controller_last_time_id)

    def process_all_events(self, time):                            # Simulator.java:213-228: strictly before `time`
        while self._queue and self._queue[0][0] < time:
            heapq.heappop(self._queue)[2]()
