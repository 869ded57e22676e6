"""The JSON radio protocol of the reference (net/SimulatorJSONHandler.java:28-242,
net/JSONClientConnection.java:134-420, net/Server.java:60-66), in memory: a replay harness that feeds
recorded command streams through the host mirror (simulator.py) and the C-ABI, and collects what the
reference would send back.  No sockets: the TCP server is control plane and out of scope; framing is
handled by `decode_stream` / `encode` so that byte-level fixtures can be replayed too.

File names are relative to radio-medium/java/se/sics/emul8/radiomedium/.
"""
import json

from .simulator import RadioPacket, CudaRadioMedium

# propagation-option -> model.  The reference knows n2n-link / udgm / nullrm (net/SimulatorJSONHandler.java:179-214);
# reference-edits.patch adds the "-b200" names that select the CUDA medium there.  In this harness every option is
# served by the CUDA medium (there is no other), so both spellings are accepted.
OPTIONS = {"n2n-link": "n2n", "udgm": "udgm", "nullrm": "null", "udgm-constant": "udgm-constant",
           "n2n-link-b200": "n2n", "udgm-b200": "udgm", "nullrm-b200": "null", "udgm-constant-b200": "udgm-constant",
           "sinr-b200": "sinr"}
WELCOME = {"radio-simulator": {"name": "RSIM 0.1", "api-version": "0.6"}, "status": "OK"}        # net/Server.java:60-66


def encode(obj, use_length=False):
    """net/JSONClientConnection.java:257-287: bare JSON + CRLF, or `len\\n` + JSON."""
    data = json.dumps(obj, separators=(",", ":")).encode("utf-8")
    return (str(len(data)).encode("ascii") + b"\r\n" + data) if use_length else data + b"\r\n"


def decode_stream(data):
    """net/JSONClientConnection.java:134-250: a stream of bare JSON objects (brace matching outside strings) and/or
    `len;attrs\\n` + payload messages; CR ignored, blank lines skipped, an `RSIM ...` banner line skipped."""
    out, i, n = [], 0, len(data)
    while i < n:
        c = data[i:i + 1]
        if c in (b"\r", b"\n", b" ", b"\t"):
            i += 1
            continue
        if c == b"{":
            depth, j, quoted, stuffed = 0, i, False, False
            while j < n:
                ch = data[j:j + 1]
                if stuffed:
                    stuffed = False
                elif ch == b"\\":
                    stuffed = True
                elif ch == b'"':
                    quoted = not quoted
                elif not quoted and ch == b"{":
                    depth += 1
                elif not quoted and ch == b"}":
                    depth -= 1
                    if depth == 0:
                        break
                j += 1
            out.append(json.loads(data[i:j + 1].decode("utf-8")))
            i = j + 1
            continue
        j = data.index(b"\n", i) if b"\n" in data[i:] else n
        line = data[i:j].replace(b"\r", b"").decode("ascii").strip()
        i = j + 1
        if not line or line.startswith("RSIM "):
            continue
        size = int(line.split(";")[0])
        if size > 0:
            out.append(json.loads(data[i:i + size].decode("utf-8")))
            i += size
    return out


class JSONClientConnection:
    """One emulator / controller connection (net/JSONClientConnection.java), its outgoing messages kept in `sent`."""

    def __init__(self, name="client"):
        self.name, self.sent, self.connected, self.emulation_time = name, [], True, 0

    def is_connected(self):
        return self.connected

    def send(self, obj):
        self.sent.append(obj)
        return True

    def emulate_to_time(self, nodes, time, time_id):               # :326-353: node-info of the client's own nodes
        # reference-edits.patch: with the B200 medium the idle RSSI of all nodes is ONE read of the device buffers
        # (index == node-array index); the reference's rule stays -- the receiving RSSI while receiving, else this
        medium = nodes[0].simulator.radio_medium if nodes else None
        device_rssi = medium.node_rssi_for_time_step(nodes) if hasattr(medium, "node_rssi_for_time_step") else None
        infos = [{"node-id": nd.id,
                  "rssi": float(device_rssi[i]) if device_rssi is not None and nd.radio.receiving_packet is None else nd.radio.get_rssi(),
                  "receiving": nd.radio.get_receiving_state(),
                  "wireless-channel": nd.radio.wireless_channel} for i, nd in enumerate(nodes) if nd.client is self]
        self.send({"command": "time-step", "id": time_id, "parameters": {"time": time, "node-info": infos}})

    def set_time(self, time, time_id):                             # :360-366
        if self.emulation_time <= time:
            self.emulation_time = time
            return True
        return False

    def time_step_done(self, time_id):                             # :369-375
        self.send({"reply": "OK", "id": time_id})

    def send_packet(self, packet, destination, msg_id, rssi):      # :377-387
        obj = packet.to_json_destination(destination, rssi)
        if msg_id >= 0:
            obj["id"] = msg_id
        self.send(obj)


def _node_id(value):
    """`json.get("node-id").toString()` (net/SimulatorJSONHandler.java:70,109): a JSON number prints as itself, a JSON
    string KEEPS ITS QUOTES -- so clients must send numbers."""
    return json.dumps(value) if isinstance(value, str) else str(value)


class SimulatorJSONHandler:
    """net/SimulatorJSONHandler.java:28-242.  `medium_factory(option, params)` builds the medium for a
    `configuration-set` propagation-option (default: the CUDA medium of this repo for every option)."""

    def __init__(self, simulator, medium_factory=None):
        self.simulator = simulator
        self.medium_factory = medium_factory or (lambda option, params: CudaRadioMedium(OPTIONS[option]))

    @staticmethod
    def _reply(msg_id, **extra):                                   # createReplyObject, :256-263
        r = {"id": msg_id} if msg_id >= 0 else {}
        r["reply"] = "OK"
        r.update(extra)
        return r

    @staticmethod
    def _error(msg_id, kind, text):                                # createReplyError, :265-273
        r = {"id": msg_id} if msg_id >= 0 else {}
        r["reply"] = "error"
        r["reply-object"] = {"class": kind, "description": text}
        return r

    def handle_message(self, client, msg):
        sim = self.simulator
        status = msg.get("reply")
        if status is not None:                                     # :222-242 handleReply
            mid = msg.get("id", -1)
            if status == "OK" and mid >= 0 and mid == sim.waiting_for_time_id:
                sim.emulator_time_stepped(client, mid, sim.TIME_STEP_OK)
            return True
        mid = msg.get("id", -1)
        reply, noreply = None, False
        command = msg.get("command")
        if command is None:
            reply = self._error(mid, "command-error", "no command specified")
        elif command == "time-get":
            if mid >= 0:
                reply = self._reply(mid, **{"reply-object": {"time": sim.current_time}})
        elif command == "time-set":                                # :46-68
            if mid < 0:
                reply = self._error(mid, "command-error", "time-set must include reply id")
            else:
                if sim.time_controller is None:
                    sim.time_controller = client
                if sim.time_controller is client:
                    try:
                        sim.step_time(int(msg["parameters"]["time"]), mid)
                        noreply = True
                    except Exception as e:                         # noqa: BLE001 -- the reference swallows and reports
                        reply = self._error(mid, "command-error", "failed to set time:" + str(e))
                else:
                    reply = self._error(mid, "command-error", "only one time controller allowed")
        elif command == "transmit":                                # :69-96
            node = sim.get_node(_node_id(msg["node-id"]))
            medium = sim.radio_medium
            if node is None:
                reply = self._error(mid, "command-error", "could not find source node")
            elif medium is None:
                reply = self._error(mid, "command-error", "no radio medium available")
            else:
                packet = RadioPacket(node, int(msg["time"]), msg.get("packet-data", ""))
                v = msg.get("rf-power")
                if isinstance(v, (int, float)) and not isinstance(v, bool):
                    packet.txpower = float(v)
                v = msg.get("wireless-channel")
                if isinstance(v, (int, float)) and not isinstance(v, bool):
                    packet.channel = int(v)
                medium.transmit(packet)
        elif command == "node-config-set":                         # :105-153
            p = msg["parameters"]
            # reference-edits.patch: packets already transmitted in this tick are evaluated against the node state
            # they were sent under (the reference evaluates inside transmit()); None = the node does not exist yet
            if hasattr(sim.radio_medium, "node_config_changing"):
                sim.radio_medium.node_config_changing(sim.get_node(_node_id(p["node-id"])))
            node = sim.add_node(_node_id(p["node-id"]), client)
            pos = p.get("position")
            if isinstance(pos, list):
                if len(pos) > 2:
                    node.position.set(pos[0], pos[1], pos[2])
                elif len(pos) > 1:
                    node.position.set(pos[0], pos[1])
            num = lambda k: p[k] if isinstance(p.get(k), (int, float)) and not isinstance(p.get(k), bool) else None
            if num("rf-power") is not None:
                node.radio.txpower = float(p["rf-power"])
            if num("wireless-channel") is not None:
                node.radio.wireless_channel = int(p["wireless-channel"])
            if num("rx-loss") is not None:
                node.radio.rx_probability = float(p["rx-loss"])
            if num("tx-loss") is not None:
                node.radio.tx_probability = float(p["tx-loss"])
            if isinstance(p.get("radio-state"), str):
                node.radio.is_enabled = p["radio-state"] != "disabled"
            if mid >= 0:
                info = {"node-id": node.id, "rssi": node.radio.get_rssi(), "receiving": node.radio.get_receiving_state(),
                        "wireless-channel": node.radio.wireless_channel}
                reply = self._reply(mid, **{"reply-object": {"node-info": info}})
        elif command == "link-quality":                            # :155-168: a stub in the reference too
            pass
        elif command == "configuration-set":                       # :170-216
            if sim.time_controller is not None:
                reply = self._error(mid, "command-error", "already initialized")
            else:
                p = msg["parameters"]
                option = p.get("propagation-option")
                if option is not None:
                    if option in ("n2n-link", "n2n-link-b200"):
                        vec = p.get("matrix-data")
                        if not isinstance(vec, list) or not vec:
                            reply = self._error(mid, "command-error", "no matrix specified")
                        elif int(p["number-of-nodes"]) != int(len(vec) ** 0.5):
                            reply = self._error(mid, "command-error", "inconsistent data matrix or nodes")
                        else:
                            medium = self.medium_factory(option, p)
                            medium.set_matrix([float(v) for v in vec])
                            sim.set_radio_medium(medium)
                    elif option in ("udgm", "udgm-constant", "sinr-b200", "udgm-b200", "udgm-constant-b200", "nullrm-b200"):
                        sim.set_radio_medium(self.medium_factory(option, p))
                    elif option == "nullrm":
                        if sim.radio_medium is None:               # "the null radio medium is the default" (Main.java:66)
                            sim.set_radio_medium(self.medium_factory(option, p))
                    # anything else: logged and ignored by the reference
        elif command == "subscribe-event":
            if client not in sim.event_listeners:
                sim.event_listeners.append(client)
        elif command == "unsubscribe-event":
            sim.event_listeners = [c for c in sim.event_listeners if c is not client]
        elif command == "log":
            pass
        else:
            reply = self._error(mid, "command-error", "unsupported command: " + str(command))
        if reply is None and mid >= 0 and not noreply:
            reply = self._reply(mid)
        if reply is not None:
            client.send(reply)
        return True
