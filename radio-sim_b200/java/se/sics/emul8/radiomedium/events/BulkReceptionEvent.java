/*
 * BulkReceptionEvent -- ONE queue entry for all receivers of a packet (goes to
 * radio-medium/java/se/sics/emul8/radiomedium/events/ next to ReceptionEvent.java).
 *
 * The reference raises two ReceptionEvents per heard (packet, node) pair (Simulator.java:321-335): two
 * allocations, one lock round trip and two ladder-queue inserts per pair -- 2.9 million inserts for a tick of
 * 10 000 packets heard by 144 nodes each, far more time than the B200 needs for the whole tick.  The pairs of a
 * packet all share their start and end time, so one start and one end event per packet do the same work:
 * execute() walks rows [from, to) of the tick's result arrays and does for each row exactly what
 * ReceptionEvent.execute does (ReceptionEvent.java:35-46).  The arrays are shared by all events of the tick.
 *
 * Equal-time ordering: the reference's queue leaves the order of events with equal time unspecified
 * (com/botbox/scheduler/EventQueue.java:215-222); here the receivers of one packet are processed together, in
 * node-array order.  The set of setReceiving / clearReceiving / deliverRadioPacket calls is the same.
 */
package se.sics.emul8.radiomedium.events;

import com.botbox.scheduler.TimeEvent;

import se.sics.emul8.radiomedium.Node;
import se.sics.emul8.radiomedium.RadioPacket;
import se.sics.emul8.radiomedium.Simulator;
import se.sics.emul8.radiomedium.Transciever;

public class BulkReceptionEvent extends TimeEvent {

    private static final byte DELIVERED = 2;           /* enum rmb_outcome, include/radiomedium_b200.h */

    private final Simulator simulator;
    private final RadioPacket packet;
    private final Node[] nodes;
    private final int[] rx;
    private final double[] rssi;
    private final byte[] outcome;
    private final int from, to;
    private final boolean isStart;

    public BulkReceptionEvent(long time, Simulator simulator, RadioPacket packet, Node[] nodes, int[] rx,
            double[] rssi, byte[] outcome, int from, int to, boolean isStart) {
        super(time);
        this.simulator = simulator;
        this.packet = packet;
        this.nodes = nodes;
        this.rx = rx;
        this.rssi = rssi;
        this.outcome = outcome;
        this.from = from;
        this.to = to;
        this.isStart = isStart;
    }

    @Override
    public void execute(long currentTime) {
        for (int i = from; i < to; i++) {
            Node destination = nodes[rx[i]];
            Transciever t = destination.getRadio();
            if (isStart) {
                t.setReceiving(packet, rssi[i]);
            } else {
                t.clearReceiving();
                if (outcome[i] == DELIVERED) {
                    /* Only deliver on the end flank */
                    simulator.deliverRadioPacket(packet, destination, rssi[i]);
                }
            }
        }
    }

    @Override
    public String toString() {
        return "BulkReceptionEvent[" + this.time + "," + (isStart ? "start" : "end") + "," + (to - from) + " receivers," + packet + "]";
    }
}
