/*
 * GenGolden -- turns "parity unpinned" into "pinned" on any machine that has a JDK (this image has none).
 *
 * Drives the UNMODIFIED reference media -- NullRadioMedium, UDGMRadioMedium, UDGMConstantLossRadioMedium,
 * N2NRadioMedium -- through their own public API (setSimulator + transmit) with a Simulator subclass that records
 * the calls a medium makes back (generateReceptionEvents / deliverRadioPacket, Simulator.java:321-364) instead of
 * queueing events, and writes what they did as JSON:
 *   - every case of tests/golden/kat_reference_media.json (the hand-derived known answers), as rows
 *     [packet, node index, outcome] + the RSSI handed on + the number of nextDouble() draws;
 *   - BASELINE configs 1-3 (tools/workloads.py: C1 grid, C2, C3 from java.util.Random) under UDGM, Null and
 *     UDGMConstantLoss, as row counts and SHA-256 digests (int32 tx[] | int32 rx[] | uint8 outcome[], little endian).
 *
 * Build and run from the reference tree (JDK 8 because RadioPacket.java:35 imports javax.xml.bind):
 *   javac -cp 'radio-medium/lib/*' -d /tmp/gg radio-medium/java/se/sics/emul8/radiomedium/*.java \
 *         radio-medium/java/se/sics/emul8/radiomedium/{events,net,util}/*.java radio-medium/java/com/botbox/scheduler/*.java \
 *         <this repo>/radio-sim_b200/java/tools/GenGolden.java
 *   java -cp '/tmp/gg:radio-medium/lib/*' se.sics.emul8.radiomedium.GenGolden \
 *         <this repo>/tests/golden/kat_reference_media.json > <this repo>/tests/golden/reference_jvm_output.json
 * tests/test_oracle_kat.py::test_reference_jvm_output_matches_the_oracle then compares the file with the oracle
 * (skipped while the file does not exist).
 */
package se.sics.emul8.radiomedium;

import java.nio.ByteBuffer;
import java.nio.ByteOrder;
import java.nio.charset.StandardCharsets;
import java.nio.file.Files;
import java.nio.file.Paths;
import java.security.MessageDigest;
import java.util.ArrayList;
import java.util.IdentityHashMap;
import java.util.Random;

import com.eclipsesource.json.JsonArray;
import com.eclipsesource.json.JsonObject;
import com.eclipsesource.json.JsonValue;

public class GenGolden {

    /* counts the draws a medium makes from simulator.getRandom() (Simulator.java:94) */
    static class CountingRandom extends Random {
        private static final long serialVersionUID = 1L;
        long draws = 0;

        CountingRandom(long seed) {
            super(seed);
        }

        @Override
        public double nextDouble() {
            draws++;
            return super.nextDouble();
        }
    }

    /* records the calls a medium makes back; nothing is queued, nothing is sent */
    static class RecordingSimulator extends Simulator {
        final ArrayList<int[]> rows = new ArrayList<int[]>();          /* {packet, node index, outcome} */
        final ArrayList<Double> rssi = new ArrayList<Double>();
        final IdentityHashMap<Node, Integer> indexOf = new IdentityHashMap<Node, Integer>();
        int currentPacket = 0;

        RecordingSimulator(Random r) {
            super(r);
        }

        void indexNodes() {
            indexOf.clear();
            Node[] nodes = getNodes();
            for (int i = 0; i < nodes.length; i++) {
                indexOf.put(nodes[i], i);
            }
        }

        @Override
        public void generateReceptionEvents(RadioPacket packet, Node destination, double rssiValue, boolean doDeliver) {
            rows.add(new int[] { currentPacket, indexOf.get(destination), doDeliver ? 2 : 1 });
            rssi.add(rssiValue);
        }

        @Override
        public void generateTransmissionEvents(RadioPacket packet) {
        }

        @Override
        public void deliverRadioPacket(RadioPacket packet, Node destination, double rssiValue) {
            rows.add(new int[] { currentPacket, indexOf.get(destination), 2 });     /* UDGMConstantLossRadioMedium.java:30-32 */
            rssi.add(rssiValue);
        }
    }

    static double[] doubles(JsonValue v, int n, double dflt) {
        double[] a = new double[n];
        for (int i = 0; i < n; i++) {
            a[i] = dflt;
        }
        if (v != null && v.isArray()) {
            JsonArray arr = v.asArray();
            for (int i = 0; i < n && i < arr.size(); i++) {
                a[i] = arr.get(i).asDouble();
            }
        }
        return a;
    }

    static RadioMedium mediumFor(String model, JsonObject params, JsonValue matrix) {
        if ("null".equals(model)) {
            return new NullRadioMedium();
        }
        if ("udgm-constant".equals(model)) {
            return new UDGMConstantLossRadioMedium();
        }
        if ("n2n".equals(model)) {
            JsonArray rowsOrFlat = matrix.asArray();
            int n = (int) Math.round(Math.sqrt(rowsOrFlat.size()));
            double[][] m = new double[n][n];
            if (rowsOrFlat.size() > 0 && rowsOrFlat.get(0).isArray()) {
                n = rowsOrFlat.size();
                m = new double[n][n];
                for (int i = 0; i < n; i++) {
                    for (int j = 0; j < n; j++) {
                        m[i][j] = rowsOrFlat.get(i).asArray().get(j).asDouble();
                    }
                }
            } else {
                for (int i = 0; i < n; i++) {
                    for (int j = 0; j < n; j++) {
                        m[i][j] = rowsOrFlat.get(j + i * n).asDouble();
                    }
                }
            }
            return new N2NRadioMedium(m);
        }
        UDGMRadioMedium udgm = new UDGMRadioMedium();
        if (params != null) {
            JsonValue v = params.get("success_ratio_tx");
            if (v != null) {
                udgm.setSuccessRatioTx(v.asDouble());
            }
            v = params.get("success_ratio_rx");
            if (v != null) {
                udgm.setSuccessRatioRx(v.asDouble());
            }
            v = params.get("transmission_range");
            if (v != null) {
                udgm.setTransmissionRange(v.asDouble());
            }
        }
        return udgm;
    }

    /* one tick through a reference medium: nodes in registration order, packets in arrival order */
    static RecordingSimulator run(RadioMedium medium, long seed, double[] x, double[] y, double[] power, int[] channel,
            boolean[] enabled, double[] rxprob, double[] txprob, int[] src, int[] pktChannel) {
        CountingRandom random = new CountingRandom(seed);
        RecordingSimulator sim = new RecordingSimulator(random);
        sim.setRadioMedium(medium);
        int n = x.length;
        for (int i = 0; i < n; i++) {
            Node node = sim.addNode(Integer.toString(i + 1), null);
            node.getPosition().set(x[i], y[i]);
            Transciever radio = node.getRadio();
            radio.setTransmitPower(power[i]);
            radio.setWirelessChannel(channel[i]);
            radio.setEnabled(enabled[i]);
            radio.setRxProbability(rxprob[i]);
            radio.setTxProbability(txprob[i]);
        }
        sim.indexNodes();
        Node[] nodes = sim.getNodes();
        for (int t = 0; t < src.length; t++) {
            RadioPacket packet = new RadioPacket(nodes[src[t]], 0L, "0102030405");
            if (pktChannel != null) {
                packet.setWirelessChannel(pktChannel[t]);
            }
            sim.currentPacket = t;
            medium.transmit(packet);
        }
        return sim;
    }

    static String sha256(ArrayList<int[]> rows) throws Exception {
        int n = rows.size();
        ByteBuffer b = ByteBuffer.allocate(9 * n).order(ByteOrder.LITTLE_ENDIAN);
        for (int[] r : rows) {
            b.putInt(r[0]);
        }
        for (int[] r : rows) {
            b.putInt(r[1]);
        }
        for (int[] r : rows) {
            b.put((byte) r[2]);
        }
        byte[] d = MessageDigest.getInstance("SHA-256").digest(b.array());
        StringBuilder s = new StringBuilder();
        for (byte v : d) {
            s.append(String.format("%02x", v & 0xff));
        }
        return s.toString();
    }

    /* tools/workloads.py: random_placement and JavaRandom.sample_distinct */
    static int[] sampleDistinct(Random r, int n, int k) {
        int[] a = new int[n];
        for (int i = 0; i < n; i++) {
            a[i] = i;
        }
        int[] out = new int[k];
        for (int i = 0; i < k; i++) {
            int j = i + r.nextInt(n - i);
            int tmp = a[i];
            a[i] = a[j];
            a[j] = tmp;
            out[i] = a[i];
        }
        return out;
    }

    static void appendRows(StringBuilder out, RecordingSimulator sim) {
        out.append("\"rows\": [");
        for (int i = 0; i < sim.rows.size(); i++) {
            int[] r = sim.rows.get(i);
            out.append(i == 0 ? "" : ", ").append("[").append(r[0]).append(", ").append(r[1]).append(", ").append(r[2]).append("]");
        }
        out.append("], \"rssi\": [");
        for (int i = 0; i < sim.rssi.size(); i++) {
            out.append(i == 0 ? "" : ", ").append(Double.toString(sim.rssi.get(i)));
        }
        out.append("], \"draws\": ").append(((CountingRandom) sim.getRandom()).draws);
    }

    public static void main(String[] args) throws Exception {
        String text = new String(Files.readAllBytes(Paths.get(args[0])), StandardCharsets.UTF_8);
        JsonObject kat = JsonObject.readFrom(text);
        StringBuilder out = new StringBuilder();
        out.append("{\n \"java.version\": \"").append(System.getProperty("java.version")).append("\",\n \"cases\": {\n");
        boolean first = true;
        for (JsonValue cv : kat.get("cases").asArray()) {
            JsonObject c = cv.asObject();
            String model = c.get("model").asString();
            if ("sinr".equals(model)) {
                continue;                                   /* the reference has no such medium */
            }
            JsonArray xs = c.get("x").asArray();
            int n = xs.size();
            double[] x = doubles(c.get("x"), n, 0.0), y = doubles(c.get("y"), n, 0.0);
            double[] power = doubles(c.get("txpower"), n, 0.0);
            double[] rxprob = doubles(c.get("rxprob"), n, 1.0), txprob = doubles(c.get("txprob"), n, 1.0);
            double[] chd = doubles(c.get("channel"), n, 26.0), end = doubles(c.get("enabled"), n, 1.0);
            int[] channel = new int[n];
            boolean[] enabled = new boolean[n];
            for (int i = 0; i < n; i++) {
                channel[i] = (int) chd[i];
                enabled[i] = end[i] != 0.0;
            }
            JsonArray sa = c.get("src").asArray();
            int[] src = new int[sa.size()];
            for (int i = 0; i < src.length; i++) {
                src[i] = sa.get(i).asInt();
            }
            int[] pktChannel = null;
            if (c.get("pkt_channel") != null) {
                JsonArray pa = c.get("pkt_channel").asArray();
                pktChannel = new int[pa.size()];
                for (int i = 0; i < pktChannel.length; i++) {
                    pktChannel[i] = pa.get(i).asInt();
                }
            }
            JsonObject params = c.get("params") != null ? c.get("params").asObject() : null;
            if (params != null && params.get("half_duplex") != null && params.get("half_duplex").asInt() != 0) {
                continue;                                   /* a spec flag of the new medium, not reference behaviour */
            }
            long seed = c.get("seed") != null ? c.get("seed").asLong() : 0L;
            RecordingSimulator sim = run(mediumFor(model, params, c.get("matrix")), seed, x, y, power, channel, enabled,
                    rxprob, txprob, src, pktChannel);
            out.append(first ? "" : ",\n").append("  \"").append(c.get("name").asString()).append("\": {");
            appendRows(out, sim);
            out.append("}");
            first = false;
        }
        out.append("\n },\n \"configs\": {\n");

        /* BASELINE configs 1-3 under the three geometric reference media */
        String[] models = { "udgm", "null", "udgm-constant" };
        String[] names = { "c1", "c2", "c3" };
        long[] seeds = { 0L, 1L, 2L };
        int[] ns = { 10, 1000, 100000 }, ts = { 10, 50, 1000 };
        double[] sides = { 0.0, 1000.0, 3162.0 };
        first = true;
        for (int k = 0; k < names.length; k++) {
            int n = ns[k];
            double[] x = new double[n], y = new double[n];
            int[] src;
            if (k == 0) {
                for (int i = 0; i < n; i++) {
                    x[i] = 30.0 * (i % 5);
                    y[i] = 30.0 * (i / 5);
                }
                src = new int[n];
                for (int i = 0; i < n; i++) {
                    src[i] = i;
                }
            } else {
                Random r = new Random(seeds[k]);
                for (int i = 0; i < n; i++) {
                    x[i] = r.nextDouble() * sides[k];
                    y[i] = r.nextDouble() * sides[k];
                }
                src = sampleDistinct(r, n, ts[k]);
            }
            double[] zero = new double[n], one = new double[n];
            int[] channel = new int[n];
            boolean[] enabled = new boolean[n];
            for (int i = 0; i < n; i++) {
                one[i] = 1.0;
                channel[i] = 26;
                enabled[i] = true;
            }
            for (String model : models) {
                RecordingSimulator sim = run(mediumFor(model, null, null), 0L, x, y, zero, channel, enabled, one, one, src, null);
                int delivered = 0;
                for (int[] r : sim.rows) {
                    delivered += r[2] == 2 ? 1 : 0;
                }
                out.append(first ? "" : ",\n").append("  \"").append(names[k]).append("-").append(model).append("\": {\"n_pairs\": ")
                        .append(sim.rows.size()).append(", \"n_delivered\": ").append(delivered).append(", \"sha256\": \"")
                        .append(sha256(sim.rows)).append("\"}");
                first = false;
            }
        }
        out.append("\n }\n}\n");
        System.out.print(out);
    }
}
