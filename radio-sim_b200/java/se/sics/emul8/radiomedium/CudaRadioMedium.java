/*
 * CudaRadioMedium -- the reference-side adapter: a RadioMedium (RadioMedium.java:35-45) whose link computation
 * runs in librmb200.so on a B200.  Goes to radio-medium/java/se/sics/emul8/radiomedium/ together with
 * events/BulkReceptionEvent.java and the edits of java/reference-edits.patch (INTEGRATION.md section 3).
 *
 * NOT COMPILED HERE: this image has no JDK.  What is checked instead (tests/test_java_binding.py): every method
 * called on a reference class exists in the reference source with that name and arity, every struct offset below
 * equals offsetof() of include/radiomedium_b200.h under gcc, every downcall descriptor has the arity of the C
 * prototype, and reference-edits.patch applies to the reference tree.  radio-sim_b200/simulator.py holds the
 * same adapter logic in Python, and that one runs against the GPU in tests/test_gpu_replay.py.
 *
 * Binding: Panama FFM (java.lang.foreign, JDK 22+), no JNI glue.  The adapter binds the rmb_group_* entry points: one
 * or several GPUs behind the same calls (Options.devices; receivers in contiguous node-index blocks, one handle and
 * one host worker thread per device inside the library), so a single JVM drives all GPUs of a box.
 *
 * Behaviour (each point is the reference's, file:line given):
 *  - transmit() only queues the packet.  flush() -- FIRST statement of Simulator.emulatorTimeStepDone, i.e.
 *    before `currentTime = stepTime` (Simulator.java:155-157), so that event start times clamp against the same
 *    currentTime a reference medium saw in transmit() (Simulator.java:323-326,339-342) -- evaluates all queued
 *    packets with ONE rmb_tick and hands the result over per packet in arrival order: the packet's two
 *    transmission events, then its reception events in node-array order (UDGMRadioMedium.java:97-114).
 *  - udgm-constant delivers at once and raises no events (UDGMConstantLossRadioMedium.java:16-36).
 *  - RSSI values are the exact doubles (RMB_FLAG_F64_VALUES): the reference media hand on packet.getTransmitPower()
 *    unchanged (UDGMRadioMedium.java:95), a float column would turn -3.3 into -3.2999999523.
 *  - Stochastic reception draws from simulator.getRandom() (Simulator.java:94) in the reference's order.  The tick
 *    consumes host-supplied uniforms and reports how many (rmb_result.n_uniforms_used); a serialised copy of the
 *    generator supplies them and the real one is then advanced by exactly that many nextDouble() calls.
 *  - Node state: a new Node[] from simulator.getNodes() (copy-on-write registry, Simulator.java:249-310) is a full
 *    re-upload; node-config-set mutates nodes in place (net/SimulatorJSONHandler.java:105-143), which reaches the
 *    device through nodeConfigChanging(node) (the one-line hook of the patch) or, without the hook, through a
 *    per-tick scan of all nodes against a shadow copy (Options.scanNodes, default on).
 *  - getBaseRSSI(node) (Transciever.java:52-61) is the reference's constant unless Options.rssiInBand, then the
 *    node's total in-band power of the last evaluated tick from rmb_node_info_f64 (what CC2420 CCA thresholds).
 */
package se.sics.emul8.radiomedium;

import java.io.ByteArrayInputStream;
import java.io.ByteArrayOutputStream;
import java.io.ObjectInputStream;
import java.io.ObjectOutputStream;
import java.lang.foreign.Arena;
import java.lang.foreign.FunctionDescriptor;
import java.lang.foreign.Linker;
import java.lang.foreign.MemorySegment;
import java.lang.foreign.SymbolLookup;
import java.lang.invoke.MethodHandle;
import java.util.ArrayList;
import java.util.IdentityHashMap;
import java.util.Random;

import static java.lang.foreign.ValueLayout.ADDRESS;
import static java.lang.foreign.ValueLayout.JAVA_BYTE;
import static java.lang.foreign.ValueLayout.JAVA_DOUBLE;
import static java.lang.foreign.ValueLayout.JAVA_INT;
import static java.lang.foreign.ValueLayout.JAVA_LONG;

public class CudaRadioMedium extends AbstractRadioMedium {

    /* enum rmb_model */
    public static final int MODEL_NULL = 0, MODEL_UDGM = 1, MODEL_UDGM_CONSTANT = 2, MODEL_N2N = 3, MODEL_SINR = 4;
    /* enum rmb_outcome, rmb_flags, rmb_status */
    private static final int DELIVERED = 2;
    private static final int FLAG_F64_VALUES = 1, FLAG_COMPACT = 2;
    private static final int RMB_OK = 0, RMB_ERR_UNIFORMS = -5;

    /* struct rmb_config: offsetof() of the field named after the prefix, sizeof() for _SIZE */
    private static final long CFG_SIZE = 120, CFG_MODEL = 4, CFG_DEVICE = 8, CFG_FLAGS = 12,
            CFG_SUCCESS_RATIO_TX = 16, CFG_SUCCESS_RATIO_RX = 24, CFG_TRANSMISSION_RANGE = 32, CFG_CONSTANT_RANGE = 40,
            CFG_PL0_DB = 48, CFG_PL_EXPONENT = 56, CFG_D0_M = 64, CFG_SENSITIVITY_DBM = 72, CFG_SINR_THRESHOLD_DB = 80,
            CFG_NOISE_FLOOR_DBM = 88, CFG_HALF_DUPLEX = 112;
    /* struct rmb_result */
    private static final long RES_SIZE = 128, RES_N_PAIRS = 0, RES_N_UNIFORMS_USED = 32, RES_N_TX = 48, RES_RX = 64,
            RES_OUTCOME = 88, RES_RSSI_DBM_F64 = 96, RES_TX_FAILED = 112, RES_TX_FIRST = 120;

    /** Parameters of the medium (`configuration-set`, net/SimulatorJSONHandler.java:170-216); NaN = library default. */
    public static class Options {
        public int[] devices = { 0 };            /* CUDA device ordinals; several = receivers sharded over them */
        public int maxTxPerTick = 16384;         /* sizes the exchange buffers between the devices */
        public double successRatioTx = Double.NaN, successRatioRx = Double.NaN;       /* UDGMRadioMedium.java:17-18 */
        public double transmissionRange = Double.NaN;                                 /* UDGMRadioMedium.java:23 */
        public double constantRange = Double.NaN;                                     /* UDGMConstantLossRadioMedium.java:8 */
        public double pl0Db = Double.NaN, plExponent = Double.NaN, d0M = Double.NaN;
        public double sensitivityDbm = Double.NaN, sinrThresholdDb = Double.NaN;
        public boolean halfDuplex = false;       /* false = the reference: only node == source excluded (UDGMRadioMedium.java:100) */
        public boolean rssiInBand = false;       /* getBaseRSSI = total in-band power of the last tick instead of baseRSSI */
        public boolean scanNodes = true;         /* diff all nodes against the shadow copy every tick (no handler hook needed) */
        public boolean bulkEvents = true;        /* Simulator.generateReceptionEventsBulk instead of one call per heard pair */
    }

    private static final Linker LINKER = Linker.nativeLinker();
    private static final SymbolLookup LIB = SymbolLookup.libraryLookup(
            System.getProperty("rmb200.library", "librmb200.so"), Arena.global());

    private static MethodHandle fn(String name, FunctionDescriptor d) {
        return LINKER.downcallHandle(LIB.find(name).orElseThrow(), d);
    }

    private static final MethodHandle CONFIG_DEFAULT = fn("rmb_config_default", FunctionDescriptor.ofVoid(ADDRESS, JAVA_INT));
    private static final MethodHandle GROUP_CREATE = fn("rmb_group_create", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, JAVA_INT, ADDRESS));
    private static final MethodHandle GROUP_DESTROY = fn("rmb_group_destroy", FunctionDescriptor.of(JAVA_INT, ADDRESS));
    private static final MethodHandle GROUP_LAST_ERROR = fn("rmb_group_last_error", FunctionDescriptor.of(ADDRESS, ADDRESS));
    private static final MethodHandle GROUP_SHARD = fn("rmb_group_shard", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle GROUP_SET_NODES = fn("rmb_group_set_nodes", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT,
            ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle SET_MATRIX = fn("rmb_set_matrix", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT, ADDRESS));
    private static final MethodHandle GROUP_UPDATE_POSITIONS = fn("rmb_group_update_positions", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT,
            ADDRESS, ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle GROUP_UPDATE_RADIO = fn("rmb_group_update_radio", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT,
            ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle GROUP_TICK = fn("rmb_group_tick", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT,
            ADDRESS, ADDRESS, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle GROUP_NODE_INFO_F64 = fn("rmb_group_node_info_f64", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS, ADDRESS));

    private final int model;
    private final Options options;
    private final Arena arena = Arena.ofShared();
    private final MemorySegment handle;              /* rmb_group* */
    private final int world;                         /* devices in the group */
    private final ArrayList<RadioPacket> pending = new ArrayList<>();

    /* the node table as last uploaded: array identity, index of every node, shadow copy of what the device holds */
    private Node[] registered = new Node[0];
    private final IdentityHashMap<Node, Integer> indexOf = new IdentityHashMap<>();
    private double[] sx = new double[0], sy = sx, sz = sx, sPower = sx, sRxp = sx, sTxp = sx;
    private int[] sChannel = new int[0];
    private byte[] sEnabled = new byte[0];
    private int[] dirty = new int[16];
    private int nDirty = 0;
    private boolean[] isDirty = new boolean[0];
    private boolean mayDraw = false;

    /* node-info of the last evaluated tick (Options.rssiInBand) */
    private double[] nodeRssi = new double[0];
    private byte[] nodeReceiving = new byte[0];
    private boolean nodeInfoValid = false;

    public CudaRadioMedium(int model) {
        this(model, new Options());
    }

    public CudaRadioMedium(int model, Options o) {
        this.model = model;
        this.options = o;
        try {
            MemorySegment cfg = arena.allocate(CFG_SIZE, 8);
            CONFIG_DEFAULT.invoke(cfg, model);
            cfg.set(JAVA_INT, CFG_FLAGS, FLAG_F64_VALUES | FLAG_COMPACT);
            cfg.set(JAVA_INT, CFG_HALF_DUPLEX, o.halfDuplex ? 1 : 0);
            setIfGiven(cfg, CFG_SUCCESS_RATIO_TX, o.successRatioTx);
            setIfGiven(cfg, CFG_SUCCESS_RATIO_RX, o.successRatioRx);
            setIfGiven(cfg, CFG_TRANSMISSION_RANGE, o.transmissionRange);
            setIfGiven(cfg, CFG_CONSTANT_RANGE, o.constantRange);
            setIfGiven(cfg, CFG_PL0_DB, o.pl0Db);
            setIfGiven(cfg, CFG_PL_EXPONENT, o.plExponent);
            setIfGiven(cfg, CFG_D0_M, o.d0M);
            setIfGiven(cfg, CFG_SENSITIVITY_DBM, o.sensitivityDbm);
            setIfGiven(cfg, CFG_SINR_THRESHOLD_DB, o.sinrThresholdDb);
            cfg.set(JAVA_DOUBLE, CFG_NOISE_FLOOR_DBM, baseRSSI);                      /* AbstractRadioMedium.java:38 */
            world = o.devices.length;
            MemorySegment devs = arena.allocate(JAVA_INT, world);
            MemorySegment.copy(o.devices, 0, devs, JAVA_INT, 0, world);
            MemorySegment out = arena.allocate(ADDRESS);
            int rc = (int) GROUP_CREATE.invoke(cfg, world, devs, o.maxTxPerTick, out);
            if (rc != RMB_OK) {
                throw new IllegalStateException("rmb_group_create: " + errorText(MemorySegment.NULL));     /* no CPU fallback */
            }
            handle = out.get(ADDRESS, 0);
        } catch (RuntimeException e) {
            throw e;
        } catch (Throwable t) {
            throw new IllegalStateException(t);
        }
    }

    private static void setIfGiven(MemorySegment cfg, long offset, double v) {
        if (!Double.isNaN(v)) {
            cfg.set(JAVA_DOUBLE, offset, v);
        }
    }

    private static String errorText(MemorySegment h) throws Throwable {
        MemorySegment msg = (MemorySegment) GROUP_LAST_ERROR.invoke(h);
        return msg.reinterpret(512).getString(0);
    }

    private void check(int rc, String what) throws Throwable {
        if (rc != RMB_OK) {
            throw new IllegalStateException(what + " rc=" + rc + ": " + errorText(handle));
        }
    }

    @Override
    public String getName() {
        return "B200 medium";
    }

    /** N2NRadioMedium.setMatrix (N2NRadioMedium.java:20-22): rows = source id - 1, columns = destination id - 1. */
    public synchronized void setMatrix(double[][] m) {
        try (Arena a = Arena.ofConfined()) {
            int n = m.length;
            MemorySegment flat = a.allocate(JAVA_DOUBLE, (long) Math.max(1, n * n));
            for (int i = 0; i < n; i++) {
                MemorySegment.copy(m[i], 0, flat, JAVA_DOUBLE, (long) i * n * 8, n);
            }
            /* the matrix medium lives on ONE device (rmb_create refuses RMB_MODEL_N2N with n_shards > 1) */
            MemorySegment shard = a.allocate(ADDRESS);
            check((int) GROUP_SHARD.invoke(handle, 0, shard, MemorySegment.NULL, MemorySegment.NULL), "rmb_group_shard");
            check((int) SET_MATRIX.invoke(shard.get(ADDRESS, 0), n, flat), "rmb_set_matrix");
        } catch (RuntimeException e) {
            throw e;
        } catch (Throwable t) {
            throw new IllegalStateException(t);
        }
    }

    /* RadioMedium.transmit: called from any client reader thread (net/SimulatorJSONHandler.java:93) */
    @Override
    public synchronized void transmit(RadioPacket packet) {
        if (simulator == null) {
            return;                                                     /* NullRadioMedium.java:50-53 */
        }
        pending.add(packet);
    }

    /**
     * Hook for net/SimulatorJSONHandler.java:105 (node-config-set), called BEFORE the handler mutates the node:
     * packets already transmitted in this tick are evaluated against the state they were sent under (the reference
     * evaluates inside transmit()), then the node is marked for re-reading at the next evaluation.
     */
    public synchronized void nodeConfigChanging(Node node) {
        flush(0L);
        Integer i = indexOf.get(node);
        if (i != null) {
            markDirty(i);
        }
    }

    private void markDirty(int i) {
        if (isDirty[i]) {
            return;
        }
        isDirty[i] = true;
        if (nDirty == dirty.length) {
            dirty = java.util.Arrays.copyOf(dirty, 2 * nDirty);
        }
        dirty[nDirty++] = i;
    }

    /** End-of-tick hook: evaluate every transmission queued since the last evaluation. */
    public synchronized void flush(long stepTime) {
        if (pending.isEmpty() || simulator == null) {
            pending.clear();
            return;
        }
        try (Arena tick = Arena.ofConfined()) {
            Node[] nodes = simulator.getNodes();
            if (nodes != registered) {
                upload(nodes);                    /* copy-on-write registry: a new array is a registry change */
            } else {
                pushChanges(nodes, tick);
            }
            /* packets whose source left the registry meanwhile cannot be evaluated (and would not be heard from) */
            int t = 0;
            RadioPacket[] packets = new RadioPacket[pending.size()];
            int[] src = new int[packets.length], chan = new int[packets.length];
            double[] power = new double[packets.length];
            for (RadioPacket p : pending) {
                Integer i = indexOf.get(p.getSource());
                if (i == null) {
                    continue;
                }
                packets[t] = p;
                src[t] = i;
                power[t] = p.getTransmitPower();                        /* RadioPacket.java:49,77: snapshot / override */
                chan[t] = p.getWirelessChannel();                       /* RadioPacket.java:50,85 */
                t++;
            }
            pending.clear();
            if (t == 0) {
                return;
            }
            MemorySegment mSrc = tick.allocate(JAVA_INT, t), mPower = tick.allocate(JAVA_DOUBLE, t), mChan = tick.allocate(JAVA_INT, t);
            MemorySegment.copy(src, 0, mSrc, JAVA_INT, 0, t);
            MemorySegment.copy(power, 0, mPower, JAVA_DOUBLE, 0, t);
            MemorySegment.copy(chan, 0, mChan, JAVA_INT, 0, t);
            MemorySegment res = tick.allocate(RES_SIZE * world, 8);                 /* rmb_result[world] */
            MemorySegment slotsOut = tick.allocate(JAVA_INT), packetOf = tick.allocate(JAVA_INT, (long) world * Math.max(options.maxTxPerTick, t));

            /* deterministic tick: one call.  With draws (on several devices the library gathers the shards' draw counts and
             * every shard walks the one stream: each rmb_result then carries the global count): the first call (no
             * uniforms; missing draws read as 0.0 = "success", which maximises the number of later draws) tells how
             * many uniforms the tick can consume at most; a copy of the simulator's generator supplies that many, and
             * the real generator is advanced by the number actually consumed -- the sequence the reference's media
             * would have left behind. */
            int rc = (int) GROUP_TICK.invoke(handle, t, mSrc, mPower, mChan, MemorySegment.NULL, 0L, res, slotsOut, packetOf);
            if (rc == RMB_ERR_UNIFORMS) {
                long bound = res.get(JAVA_LONG, RES_N_UNIFORMS_USED);
                Random real = simulator.getRandom();
                Random probe = copyOf(real);
                MemorySegment u = tick.allocate(JAVA_DOUBLE, Math.max(1L, bound));
                for (long k = 0; k < bound; k++) {
                    u.setAtIndex(JAVA_DOUBLE, k, probe.nextDouble());
                }
                rc = (int) GROUP_TICK.invoke(handle, t, mSrc, mPower, mChan, u, bound, res, slotsOut, packetOf);
                check(rc, "rmb_group_tick");
                long used = res.get(JAVA_LONG, RES_N_UNIFORMS_USED);
                for (long k = 0; k < used; k++) {
                    real.nextDouble();
                }
            } else {
                check(rc, "rmb_group_tick");
            }

            /* packet k sits at gathered record position position[k] of every shard's lists */
            int slots = slotsOut.get(JAVA_INT, 0);
            int positions = world * slots;
            int[] position = new int[t];
            for (int p = 0; p < positions; p++) {
                int k = packetOf.getAtIndex(JAVA_INT, p);
                if (k >= 0) {
                    position[k] = p;
                }
            }
            int[][] txFirst = new int[world][], rx = new int[world][];
            double[][] rssi = new double[world][];
            byte[][] outcome = new byte[world][];
            for (int s = 0; s < world; s++) {
                MemorySegment r = res.asSlice(RES_SIZE * s, RES_SIZE);
                int n = (int) r.get(JAVA_LONG, RES_N_PAIRS);
                txFirst[s] = new int[positions + 1];
                rx[s] = new int[n];
                rssi[s] = new double[n];
                outcome[s] = new byte[n];
                MemorySegment.copy(r.get(ADDRESS, RES_TX_FIRST).reinterpret(4L * (positions + 1)), JAVA_INT, 0, txFirst[s], 0, positions + 1);
                if (n > 0) {
                    MemorySegment.copy(r.get(ADDRESS, RES_RX).reinterpret(4L * n), JAVA_INT, 0, rx[s], 0, n);
                    MemorySegment.copy(r.get(ADDRESS, RES_RSSI_DBM_F64).reinterpret(8L * n), JAVA_DOUBLE, 0, rssi[s], 0, n);
                    MemorySegment.copy(r.get(ADDRESS, RES_OUTCOME).reinterpret(n), JAVA_BYTE, 0, outcome[s], 0, n);
                }
            }
            nodeInfoValid = false;
            deliver(packets, t, position, txFirst, nodes, rx, rssi, outcome);
        } catch (RuntimeException e) {
            throw e;
        } catch (Throwable e) {
            throw new IllegalStateException(e);
        } finally {
            pending.clear();
        }
    }

    /* the reference's call sequence, packet by packet (UDGMRadioMedium.java:97-114); the rows of packet k are rows
     * txFirst[s][position[k]] .. txFirst[s][position[k] + 1] of shard s, shards in node order */
    private void deliver(RadioPacket[] packets, int t, int[] position, int[][] txFirst, Node[] nodes, int[][] rx, double[][] rssi,
            byte[][] outcome) {
        if (model == MODEL_UDGM_CONSTANT) {
            for (int k = 0; k < t; k++) {                               /* UDGMConstantLossRadioMedium.java:30-32 */
                for (int s = 0; s < world; s++) {
                    for (int i = txFirst[s][position[k]]; i < txFirst[s][position[k] + 1]; i++) {
                        simulator.deliverRadioPacket(packets[k], nodes[rx[s][i]], rssi[s][i]);
                    }
                }
            }
        } else if (options.bulkEvents) {
            simulator.generateReceptionEventsBulk(packets, t, position, world, txFirst, nodes, rx, rssi, outcome);
        } else {
            for (int k = 0; k < t; k++) {
                simulator.generateTransmissionEvents(packets[k]);
                for (int s = 0; s < world; s++) {
                    for (int i = txFirst[s][position[k]]; i < txFirst[s][position[k] + 1]; i++) {
                        simulator.generateReceptionEvents(packets[k], nodes[rx[s][i]], rssi[s][i], outcome[s][i] == DELIVERED);
                    }
                }
            }
        }
    }

    /* java.util.Random is Serializable: a deserialised copy continues the same sequence */
    private static Random copyOf(Random r) throws Exception {
        ByteArrayOutputStream bytes = new ByteArrayOutputStream();
        try (ObjectOutputStream out = new ObjectOutputStream(bytes)) {
            out.writeObject(r);
        }
        try (ObjectInputStream in = new ObjectInputStream(new ByteArrayInputStream(bytes.toByteArray()))) {
            return (Random) in.readObject();
        }
    }

    /* Node[] (objects) -> the SoA arrays of rmb_group_set_nodes; also the shadow copy and the node -> index map */
    private void upload(Node[] nodes) throws Throwable {
        int n = nodes.length;
        sx = new double[n]; sy = new double[n]; sz = new double[n];
        sPower = new double[n]; sRxp = new double[n]; sTxp = new double[n];
        sChannel = new int[n]; sEnabled = new byte[n];
        int[] id = new int[n];
        isDirty = new boolean[n];
        nDirty = 0;
        indexOf.clear();
        for (int i = 0; i < n; i++) {
            readNode(nodes[i], i);
            id[i] = nodes[i].getIdAsInteger();                          /* Node.java:65-67 (N2N rows/columns) */
            indexOf.put(nodes[i], i);
        }
        try (Arena a = Arena.ofConfined()) {
            long c = Math.max(1, n);
            MemorySegment x = a.allocate(JAVA_DOUBLE, c), y = a.allocate(JAVA_DOUBLE, c), z = a.allocate(JAVA_DOUBLE, c);
            MemorySegment pw = a.allocate(JAVA_DOUBLE, c), rxp = a.allocate(JAVA_DOUBLE, c), txp = a.allocate(JAVA_DOUBLE, c);
            MemorySegment ch = a.allocate(JAVA_INT, c), ids = a.allocate(JAVA_INT, c), en = a.allocate(JAVA_BYTE, c);
            MemorySegment.copy(sx, 0, x, JAVA_DOUBLE, 0, n); MemorySegment.copy(sy, 0, y, JAVA_DOUBLE, 0, n);
            MemorySegment.copy(sz, 0, z, JAVA_DOUBLE, 0, n); MemorySegment.copy(sPower, 0, pw, JAVA_DOUBLE, 0, n);
            MemorySegment.copy(sRxp, 0, rxp, JAVA_DOUBLE, 0, n); MemorySegment.copy(sTxp, 0, txp, JAVA_DOUBLE, 0, n);
            MemorySegment.copy(sChannel, 0, ch, JAVA_INT, 0, n); MemorySegment.copy(id, 0, ids, JAVA_INT, 0, n);
            MemorySegment.copy(sEnabled, 0, en, JAVA_BYTE, 0, n);
            check((int) GROUP_SET_NODES.invoke(handle, n, x, y, z, pw, ch, en, rxp, txp, ids), "rmb_group_set_nodes");
        }
        registered = nodes;
        nodeInfoValid = false;
    }

    /* one node's fields into the shadow arrays; true if anything the device holds changed */
    private boolean readNode(Node node, int i) {
        Position pos = node.getPosition();                              /* Position.java:37-39: public doubles */
        Transciever radio = node.getRadio();
        double x = pos.x, y = pos.y, z = pos.z;
        double pw = radio.getTransmitPower(), rxp = radio.getRxProbability(), txp = radio.getTxProbability();
        int ch = radio.getWirelessChannel();
        byte en = (byte) (radio.isEnabled() ? 1 : 0);
        boolean moved = x != sx[i] || y != sy[i] || z != sz[i];
        boolean retuned = pw != sPower[i] || rxp != sRxp[i] || txp != sTxp[i] || ch != sChannel[i] || en != sEnabled[i];
        sx[i] = x; sy[i] = y; sz[i] = z; sPower[i] = pw; sRxp[i] = rxp; sTxp[i] = txp; sChannel[i] = ch; sEnabled[i] = en;
        return moved || retuned;
    }

    /* in-place changes of registered nodes -> rmb_group_update_positions / _radio for exactly those indices */
    private void pushChanges(Node[] nodes, Arena tick) throws Throwable {
        if (options.scanNodes) {
            for (int i = 0; i < nodes.length; i++) {
                if (!isDirty[i] && readNode(nodes[i], i)) {
                    markDirty(i);
                }
            }
        }
        if (nDirty == 0) {
            return;
        }
        int c = nDirty;
        MemorySegment idx = tick.allocate(JAVA_INT, c), x = tick.allocate(JAVA_DOUBLE, c), y = tick.allocate(JAVA_DOUBLE, c),
                z = tick.allocate(JAVA_DOUBLE, c), pw = tick.allocate(JAVA_DOUBLE, c), rxp = tick.allocate(JAVA_DOUBLE, c),
                txp = tick.allocate(JAVA_DOUBLE, c), ch = tick.allocate(JAVA_INT, c), en = tick.allocate(JAVA_BYTE, c);
        for (int k = 0; k < c; k++) {
            int i = dirty[k];
            readNode(nodes[i], i);
            isDirty[i] = false;
            idx.setAtIndex(JAVA_INT, k, i);
            x.setAtIndex(JAVA_DOUBLE, k, sx[i]); y.setAtIndex(JAVA_DOUBLE, k, sy[i]); z.setAtIndex(JAVA_DOUBLE, k, sz[i]);
            pw.setAtIndex(JAVA_DOUBLE, k, sPower[i]); rxp.setAtIndex(JAVA_DOUBLE, k, sRxp[i]); txp.setAtIndex(JAVA_DOUBLE, k, sTxp[i]);
            ch.setAtIndex(JAVA_INT, k, sChannel[i]); en.set(JAVA_BYTE, k, sEnabled[i]);
        }
        nDirty = 0;
        check((int) GROUP_UPDATE_POSITIONS.invoke(handle, c, idx, x, y, z), "rmb_group_update_positions");
        check((int) GROUP_UPDATE_RADIO.invoke(handle, c, idx, pw, ch, en, rxp, txp), "rmb_group_update_radio");
        nodeInfoValid = false;
    }

    /* Transciever.getRSSI falls back to this while the node is not receiving (Transciever.java:52-61) */
    @Override
    public synchronized double getBaseRSSI(Node node) {
        if (!options.rssiInBand) {
            return baseRSSI;                                            /* AbstractRadioMedium.java:45-48 */
        }
        Integer i = indexOf.get(node);
        if (i == null || !refreshNodeInfo() || i >= nodeRssi.length) {
            return baseRSSI;
        }
        return nodeRssi[i];
    }

    /**
     * The idle RSSI of every node for the per-tick node-info (net/JSONClientConnection.java:326-353), index ==
     * node-array index: total in-band power of the last evaluated tick in dBm (noise floor where nothing was on
     * the air).  null when Options.rssiInBand is off or `nodes` is not the array the device holds -- the caller
     * then uses Transciever.getRSSI() as the reference does.  The array is shared, not a copy: read only.
     */
    public synchronized double[] nodeRssiForTimeStep(Node[] nodes) {
        if (!options.rssiInBand || nodes != registered || !refreshNodeInfo()) {
            return null;
        }
        return nodeRssi;
    }

    /** 1 where the node heard at least one frame in the last evaluated tick (same indexing); null if unknown. */
    public synchronized byte[] nodeHeardForTimeStep(Node[] nodes) {
        if (nodes != registered || !refreshNodeInfo()) {
            return null;
        }
        return nodeReceiving;
    }

    private boolean refreshNodeInfo() {
        if (nodeInfoValid) {
            return true;
        }
        int n = registered.length;
        if (n == 0) {
            return false;
        }
        try (Arena a = Arena.ofConfined()) {
            MemorySegment power = a.allocate(JAVA_DOUBLE, n), heard = a.allocate(JAVA_BYTE, n);
            int rc = (int) GROUP_NODE_INFO_F64.invoke(handle, power, heard);
            if (rc != RMB_OK) {
                return false;
            }
            nodeRssi = new double[n];
            nodeReceiving = new byte[n];
            MemorySegment.copy(power, JAVA_DOUBLE, 0, nodeRssi, 0, n);
            MemorySegment.copy(heard, JAVA_BYTE, 0, nodeReceiving, 0, n);
            nodeInfoValid = true;
            return true;
        } catch (Throwable t) {
            return false;
        }
    }

    public synchronized void close() {
        try {
            GROUP_DESTROY.invoke(handle);
        } catch (Throwable t) {
            throw new IllegalStateException(t);
        } finally {
            arena.close();
        }
    }
}
