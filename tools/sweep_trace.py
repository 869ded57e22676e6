"""Read a per-block time-stamp trace of the sweep kernel (RMB_SWEEP_TRACE=<file>, a debug knob of librmb200.so) and
print where the launch's time goes: block durations by region, the resolve passes, and how the SMs drain at the end.
usage: sweep_trace.py trace.bin"""
import sys
import numpy as np

raw = open(sys.argv[1], "rb").read()
s1, s2, rb_taper, nb = np.frombuffer(raw[:16], dtype=np.int32)
tr = np.frombuffer(raw[16:], dtype=np.uint64).reshape(-1, 12)
t0, t1, t2 = tr[:, 0].astype(np.int64), tr[:, 1].astype(np.int64), tr[:, 2].astype(np.int64)
smid = (tr[:, 3] & 0xffff).astype(np.int64)
passes = (tr[:, 3] >> 16).astype(np.int64)
t2 = np.where(t2 == 0, t1, t2)
base = t0.min()
t0, t1, t2 = (t0 - base) / 1e3, (t1 - base) / 1e3, (t2 - base) / 1e3
n1 = rb_taper * s1
print(f"splits {s1}/{s2}, rx blocks {nb} (tapered from {rb_taper}), blocks {len(tr)} ({n1} + {len(tr) - n1})")
print(f"launch span {t2.max():.1f} us; first block start spread {np.percentile(t0, 1):.1f}..{np.sort(t0)[min(len(t0) - 1, 148 * 9 - 1)]:.1f} us (first wave)")
for name, sel in (("main", np.arange(len(tr)) < n1), ("taper", np.arange(len(tr)) >= n1)):
    if sel.any():
        d = (t1 - t0)[sel]
        print(f"{name:6s} sweep part: n {sel.sum()}, median {np.median(d):.1f} us, p5 {np.percentile(d, 5):.1f}, p95 {np.percentile(d, 95):.1f}, max {d.max():.1f}")
r = passes > 0
d = (t2 - t1)[r]
print(f"resolve: {r.sum()} blocks ran {passes.sum()} passes; epilogue time median {np.median(d):.1f} us, p95 {np.percentile(d, 95):.1f}, max {d.max():.1f}; "
      f"per pass median {np.median(d / passes[r]):.1f} us; block-time {d.sum() / 1e3:.2f} ms of {(t2 - t0).sum() / 1e3:.2f} ms")
w = (~r) & (t2 - t1 > 1.0)
if w.any():
    print(f"waiting finishers that got no pass: {w.sum()}, median wait {np.median((t2 - t1)[w]):.1f} us")
end = t2.max()
last_sweep = t1.max()
print(f"last sweep part ends at {last_sweep:.1f} us, launch ends at {end:.1f} us (resolve-only tail {end - last_sweep:.1f} us)")
# resident blocks over the last 80 us
for back in (80, 60, 40, 30, 20, 10, 5):
    t = end - back
    live = ((t0 <= t) & (t2 > t)).sum()
    sweeping = ((t0 <= t) & (t1 > t)).sum()
    print(f"  t = end - {back:2d} us: resident blocks {live:5d} (sweeping {sweeping:5d}, in epilogue {live - sweeping:4d})")
per_sm_end = np.array([t2[smid == s].max() for s in np.unique(smid)])
print(f"per-SM last activity: median end-{np.median(end - per_sm_end):.1f} us, earliest end-{(end - per_sm_end).max():.1f} us")
busy = (t2 - t0).sum()
print(f"slot occupancy: {busy / (end * 148 * 9):.3f} of 148 x 9 slots over the span")

# inside the last resolve pass a block ran: far sums + slot sort | pooled link budgets | per-receiver sums + atomic | SINR + records
q = tr[passes > 0].astype(np.int64)
d = np.stack([q[:, 5] - q[:, 4], q[:, 6] - q[:, 5], q[:, 7] - q[:, 6], q[:, 8] - q[:, 7]], axis=1) / 1e3
print("last pass of a block, median us: far+sort %.1f | link budgets %.1f | sums+atomic %.1f | sinr+records %.1f | candidates in warp 0: median %d, p95 %d"
      % (*np.median(d, axis=0), np.median(q[:, 9]), np.percentile(q[:, 9], 95)))
late = q[:, 8] > (q[:, 8].max() - 30000)
print("  ... of the blocks ending in the last 30 us (%d): %.1f | %.1f | %.1f | %.1f" % (late.sum(), *np.median(d[late], axis=0)))
sc = tr[tr[:, 10] != 0].astype(np.int64)
if len(sc):
    print("segment scan: %.1f us, ending %.1f us before the launch's end" % ((sc[0, 11] - sc[0, 10]) / 1e3, end - (sc[0, 11] - base) / 1e3))
