package dke.model;

import org.apache.storm.task.OutputCollector;
import org.apache.storm.task.TopologyContext;
import org.apache.storm.topology.OutputFieldsDeclarer;
import org.apache.storm.topology.base.BaseRichBolt;
import org.apache.storm.tuple.Fields;
import org.apache.storm.tuple.Tuple;
import org.apache.storm.tuple.Values;

import java.io.ByteArrayOutputStream;
import java.io.InputStream;
import java.nio.charset.StandardCharsets;
import java.util.ArrayDeque;
import java.util.Map;

/**
 * Drop-in for the reference's dke.model.InferenceBolt: same package, class name, no-arg constructor, BaseRichBolt
 * methods and output field ("message"), so MainTopology.java:54,62 needs no change. All heavy state is created in
 * prepare(). NOT COMPILED here (no JDK/Storm jars on the boxes); the same logic is exercised through its C++ mirror
 * csrc/sib_bolt.cpp by tests/test_bolt_contract.py.
 */
public class InferenceBolt extends BaseRichBolt {
    private transient OutputCollector collector;
    private transient long engine;
    private transient Thread completion;
    private transient volatile boolean stop;
    private transient long seq;
    private transient boolean shared;   // sib.shared = 1: a view of the worker JVM's shared engine (per-tuple poll path)
    /** arrival order == emit order; a null-message entry is a tuple whose JSON did not parse */
    private transient ArrayDeque<Object[]> fifo;   // {Tuple, Boolean nullMsg}

    private static int conf(Map m, String k, int d) { Object v = m.get(k); return v == null ? d : Integer.parseInt(v.toString()); }

    @Override
    public void prepare(Map map, TopologyContext ctx, OutputCollector outputCollector) {
        this.collector = outputCollector;
        this.fifo = new ArrayDeque<>();
        byte[] blob;
        // model/model.sib if the jar has one, else the reference's own model/saved_model.pb (frozen): the native side takes either
        ClassLoader cl = getClass().getClassLoader();
        String res = cl.getResource("model/model.sib") != null ? "model/model.sib" : "model/saved_model.pb";
        try (InputStream in = cl.getResourceAsStream(res)) {
            ByteArrayOutputStream bo = new ByteArrayOutputStream();
            byte[] buf = new byte[1 << 16];
            for (int n; (n = in.read(buf)) > 0; ) bo.write(buf, 0, n);
            blob = bo.toByteArray();
        } catch (Exception e) { throw new RuntimeException("neither model/model.sib nor model/saved_model.pb in the jar", e); }
        String p = String.valueOf(map.getOrDefault("sib.precision", "fp16"));
        int prec = p.equals("fp32") ? 2 : p.equals("bf16") ? 1 : 0;
        int[] cfg = { conf(map, "sib.device", -1), ctx.getThisTaskIndex(), conf(map, "sib.max.batch", 0),
                      conf(map, "sib.deadline.us", 0), prec, conf(map, "sib.slots", 0), conf(map, "sib.guard", 1),
                      conf(map, "sib.json.gpu", 0), conf(map, "sib.strict.batching", 0), conf(map, "sib.shared", 0) };
        shared = cfg[9] != 0;
        engine = NativeInfer.open(blob, cfg, 0f);
        if (engine == 0) throw new RuntimeException("InferenceBolt.prepare: " + NativeInfer.lastError());
        completion = new Thread(shared ? this::completionLoopPerTuple : this::completionLoop, "sib-completion");
        completion.setDaemon(true);
        completion.start();
    }

    @Override
    public void execute(Tuple tuple) {
        byte[] utf8 = tuple.getString(0).getBytes(StandardCharsets.UTF_8);
        Object[] slot = { tuple, Boolean.FALSE };
        synchronized (fifo) { fifo.addLast(slot); fifo.notifyAll(); }   // reserve the emit position first
        for (;;) {
            int rc = NativeInfer.submitJson(engine, seq, utf8);
            if (rc > 0) { seq++; return; }
            if (rc == -1) { synchronized (fifo) { slot[1] = Boolean.TRUE; fifo.notifyAll(); } return; }  // emit null, ack
            if (rc == -4) { Thread.yield(); continue; }                                               // ring full
            synchronized (fifo) { fifo.removeLast(); }
            throw new IllegalArgumentException(NativeInfer.lastError());   // what Tensor.create / copyTo threw
        }
    }

    private void emitLeadingNulls() {   // caller holds fifo's monitor
        while (!fifo.isEmpty() && (Boolean) fifo.peekFirst()[1]) {
            Tuple t = (Tuple) fifo.pollFirst()[0];
            synchronized (collector) { collector.emit(new Values((String) null)); collector.ack(t); }
        }
    }

    private void completionLoopPerTuple() {   // shared engine: results are copied per client, one String per JNI store
        long[] tags = new long[256];
        String[] msgs = new String[256];
        while (true) {
            synchronized (fifo) {
                emitLeadingNulls();
                if (fifo.isEmpty()) {
                    fifo.notifyAll();
                    if (stop) return;
                    try { fifo.wait(); } catch (InterruptedException e) { return; }
                    continue;
                }
            }
            int n = NativeInfer.pollJson(engine, tags, msgs, 2000);
            synchronized (fifo) {
                for (int i = 0; i < n; i++) {
                    emitLeadingNulls();
                    Tuple t = (Tuple) fifo.pollFirst()[0];
                    synchronized (collector) { collector.emit(new Values(msgs[i])); collector.ack(t); }
                }
            }
        }
    }

    private void completionLoop() {
        // Batched egress (sib_poll_json_batch): one JNI crossing and one native formatting pass per poll; the Strings are
        // cut out of one direct buffer and handed to the collector under ONE monitor hold per batch (the per-tuple form
        // paid a NewStringUTF, an array store and a synchronized block per tuple: at ~1 M tuples/s per GPU that monitor was
        // the next ceiling). KafkaBolt downstream still gets one "message" String per tuple (KafkaBolt.java:123-124).
        final int max = 256;
        long[] tags = new long[max];
        int[] offsets = new int[max + 1];
        int[] status = new int[max];
        java.nio.ByteBuffer arena = java.nio.ByteBuffer.allocateDirect(max * 256);
        byte[] text = new byte[max * 256];
        while (true) {
            synchronized (fifo) {
                emitLeadingNulls();
                if (fifo.isEmpty()) {
                    fifo.notifyAll();
                    if (stop) return;
                    try { fifo.wait(); } catch (InterruptedException e) { return; }
                    continue;
                }
            }
            int n = NativeInfer.pollJsonBatch(engine, arena, tags, offsets, status, 2000);
            if (n <= 0) continue;
            arena.position(0);
            arena.get(text, 0, offsets[n]);
            synchronized (fifo) {
                synchronized (collector) {              // OutputCollector is not thread-safe in Storm 1.x (KafkaBolt.java:133)
                    for (int i = 0; i < n; i++) {
                        while (!fifo.isEmpty() && (Boolean) fifo.peekFirst()[1]) {   // decode failures keep their place
                            collector.emit(new Values((String) null));
                            collector.ack((Tuple) fifo.pollFirst()[0]);
                        }
                        Tuple t = (Tuple) fifo.pollFirst()[0];
                        String msg = status[i] == 0 && offsets[i + 1] > offsets[i]
                                ? new String(text, offsets[i], offsets[i + 1] - offsets[i], StandardCharsets.ISO_8859_1) : null;
                        collector.emit(new Values(msg));
                        collector.ack(t);
                    }
                }
            }
        }
    }

    @Override
    public void cleanup() {
        NativeInfer.flush(engine);
        synchronized (fifo) {
            fifo.notifyAll();
            while (!fifo.isEmpty()) { try { fifo.wait(10); } catch (InterruptedException e) { break; } }
            stop = true;
            fifo.notifyAll();
        }
        NativeInfer.close(engine);
    }

    @Override
    public void declareOutputFields(OutputFieldsDeclarer declarer) {
        declarer.declare(new Fields("message"));
    }
}
