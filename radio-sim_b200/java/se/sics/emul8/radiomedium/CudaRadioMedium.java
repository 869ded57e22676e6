/*
 * CudaRadioMedium -- the reference-side adapter: a RadioMedium (RadioMedium.java:35-45) whose
 * link computation runs in librmb200.so on a B200.  SOURCE ONLY: this image has no JDK, so this
 * file is not compiled or tested here; the C-ABI underneath is (tests/, through ctypes).
 *
 * Binding: Panama FFM (java.lang.foreign, JDK 22+), no JNI glue.  transmit() only enqueues (and
 * raises the two transmission events immediately, as every reference medium does:
 * UDGMRadioMedium.java:97); flush(time) -- called from Simulator.emulatorTimeStepDone before
 * processAllEvents (Simulator.java:155-161, see INTEGRATION.md) -- evaluates the whole tick with
 * one rmb_tick and replays the heard pairs as generateReceptionEvents calls in the reference's
 * own order.
 */
package se.sics.emul8.radiomedium;

import java.lang.foreign.*;
import java.lang.invoke.MethodHandle;
import java.util.ArrayList;
import static java.lang.foreign.ValueLayout.*;

public class CudaRadioMedium extends AbstractRadioMedium {

    public static final int MODEL_NULL = 0, MODEL_UDGM = 1, MODEL_UDGM_CONSTANT = 2, MODEL_N2N = 3, MODEL_SINR = 4;
    private static final int DELIVERED = 2;

    private static final Linker LINKER = Linker.nativeLinker();
    private static final SymbolLookup LIB = SymbolLookup.libraryLookup(
            System.getProperty("rmb200.library", "librmb200.so"), Arena.global());

    private static MethodHandle fn(String name, FunctionDescriptor d) {
        return LINKER.downcallHandle(LIB.find(name).orElseThrow(), d);
    }

    private static final MethodHandle CONFIG_DEFAULT = fn("rmb_config_default", FunctionDescriptor.ofVoid(ADDRESS, JAVA_INT));
    private static final MethodHandle CREATE = fn("rmb_create", FunctionDescriptor.of(JAVA_INT, ADDRESS, ADDRESS));
    private static final MethodHandle DESTROY = fn("rmb_destroy", FunctionDescriptor.of(JAVA_INT, ADDRESS));
    private static final MethodHandle LAST_ERROR = fn("rmb_last_error", FunctionDescriptor.of(ADDRESS, ADDRESS));
    private static final MethodHandle SET_NODES = fn("rmb_set_nodes", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT,
            ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS, ADDRESS));
    private static final MethodHandle TICK = fn("rmb_tick", FunctionDescriptor.of(JAVA_INT, ADDRESS, JAVA_INT,
            ADDRESS, ADDRESS, ADDRESS, ADDRESS, JAVA_LONG, ADDRESS));

    /* struct rmb_config (include/radiomedium_b200.h): 4 x int32, 10 x double, int64, 2 x int32 = 112 bytes */
    private static final long CFG_SIZE = 112, CFG_MODEL = 4, CFG_PL0 = 48, CFG_EXPONENT = 56, CFG_SENS = 72, CFG_THR = 80;
    /* struct rmb_result: 6 x int64, 2 x int32, 9 pointers = 128 bytes */
    private static final long RES_SIZE = 128, RES_NPAIRS = 0, RES_TX = 56, RES_RX = 64, RES_RSSI = 72, RES_OUTCOME = 88;

    private final Arena arena = Arena.ofShared();
    private final MemorySegment handle;
    private final ArrayList<RadioPacket> pending = new ArrayList<>();
    private Node[] registered = new Node[0];

    public CudaRadioMedium(int model) {
        try {
            MemorySegment cfg = arena.allocate(CFG_SIZE, 8);
            CONFIG_DEFAULT.invoke(cfg, model);
            MemorySegment out = arena.allocate(ADDRESS);
            int rc = (int) CREATE.invoke(cfg, out);
            if (rc != 0) {
                MemorySegment msg = (MemorySegment) LAST_ERROR.invoke(MemorySegment.NULL);
                throw new IllegalStateException("rmb_create: " + msg.reinterpret(256).getString(0));   // no CPU fallback
            }
            handle = out.get(ADDRESS, 0);
        } catch (Throwable t) {
            throw new IllegalStateException(t);
        }
    }

    @Override
    public String getName() {
        return "B200 medium";
    }

    /* RadioMedium.transmit: called from any client reader thread (net/SimulatorJSONHandler.java:93) */
    @Override
    public synchronized void transmit(RadioPacket packet) {
        simulator.generateTransmissionEvents(packet);
        pending.add(packet);
    }

    /* End-of-tick hook: evaluate every transmission of the tick at once. */
    public synchronized void flush(long time) {
        if (pending.isEmpty()) return;
        try (Arena tick = Arena.ofConfined()) {
            Node[] nodes = simulator.getNodes();
            if (nodes != registered) upload(nodes);           // copy-on-write array: identity change == registry change
            int t = pending.size();
            MemorySegment src = tick.allocate(JAVA_INT, t), power = tick.allocate(JAVA_DOUBLE, t), chan = tick.allocate(JAVA_INT, t);
            for (int i = 0; i < t; i++) {
                RadioPacket p = pending.get(i);
                src.setAtIndex(JAVA_INT, i, indexOf(nodes, p.getSource()));
                power.setAtIndex(JAVA_DOUBLE, i, p.getTxpower());
                chan.setAtIndex(JAVA_INT, i, p.getWirelessChannel());
            }
            /* uniforms: the tick may consume at most t + t*n draws; fill lazily in production */
            MemorySegment res = tick.allocate(RES_SIZE, 8);
            int rc = (int) TICK.invoke(handle, t, src, power, chan, MemorySegment.NULL, 0L, res);
            if (rc != 0) throw new IllegalStateException("rmb_tick rc=" + rc);
            long n = res.get(JAVA_LONG, RES_NPAIRS);
            MemorySegment tx = res.get(ADDRESS, RES_TX).reinterpret(4 * n), rx = res.get(ADDRESS, RES_RX).reinterpret(4 * n);
            MemorySegment rssi = res.get(ADDRESS, RES_RSSI).reinterpret(4 * n), out = res.get(ADDRESS, RES_OUTCOME).reinterpret(n);
            for (long i = 0; i < n; i++) {
                simulator.generateReceptionEvents(pending.get(tx.getAtIndex(JAVA_INT, i)), nodes[rx.getAtIndex(JAVA_INT, i)],
                        rssi.getAtIndex(JAVA_FLOAT, i), out.get(JAVA_BYTE, i) == DELIVERED);
            }
        } catch (Throwable e) {
            throw new IllegalStateException(e);
        } finally {
            pending.clear();
        }
    }

    private static int indexOf(Node[] nodes, Node n) {
        for (int i = 0; i < nodes.length; i++) if (nodes[i] == n) return i;
        return -1;
    }

    /* Node[] (AoS objects) -> the SoA arrays of rmb_set_nodes */
    private void upload(Node[] nodes) throws Throwable {
        try (Arena a = Arena.ofConfined()) {
            int n = nodes.length;
            MemorySegment x = a.allocate(JAVA_DOUBLE, n), y = a.allocate(JAVA_DOUBLE, n), z = a.allocate(JAVA_DOUBLE, n);
            MemorySegment pw = a.allocate(JAVA_DOUBLE, n), rxp = a.allocate(JAVA_DOUBLE, n), txp = a.allocate(JAVA_DOUBLE, n);
            MemorySegment ch = a.allocate(JAVA_INT, n), id = a.allocate(JAVA_INT, n), en = a.allocate(JAVA_BYTE, n);
            for (int i = 0; i < n; i++) {
                Position p = nodes[i].getPosition();
                Transciever r = nodes[i].getRadio();
                x.setAtIndex(JAVA_DOUBLE, i, p.x); y.setAtIndex(JAVA_DOUBLE, i, p.y); z.setAtIndex(JAVA_DOUBLE, i, p.z);
                pw.setAtIndex(JAVA_DOUBLE, i, r.getTxpower());
                rxp.setAtIndex(JAVA_DOUBLE, i, r.getRxProbability()); txp.setAtIndex(JAVA_DOUBLE, i, r.getTxProbability());
                ch.setAtIndex(JAVA_INT, i, r.getWirelessChannel()); id.setAtIndex(JAVA_INT, i, nodes[i].getIdAsInteger());
                en.set(JAVA_BYTE, i, (byte) (r.isEnabled() ? 1 : 0));
            }
            int rc = (int) SET_NODES.invoke(handle, n, x, y, z, pw, ch, en, rxp, txp, id);
            if (rc != 0) throw new IllegalStateException("rmb_set_nodes rc=" + rc);
            registered = nodes;
        }
    }

    public void close() throws Throwable {
        DESTROY.invoke(handle);
        arena.close();
    }
}
