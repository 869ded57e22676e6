// kernels.h -- host-callable launchers of the CUDA kernels (one .cu per pipeline stage).
#pragma once
#include "common.cuh"

namespace rmb {

// ---- node state in HBM (SoA, node-array order == Simulator.getNodes() order) -------------
struct NodeArrays {
    int n;
    double *x, *y, *z;                // Position.java:37-39 (exact master copy)
    double *txpower, *rxprob, *txprob;
    int    *channel;
    uint8_t *enabled;
    int    *id_int;                   // Node.getIdAsInteger (N2N)
};

// ---- receivers in work order (grouped by channel, padded) --------------------------------
struct WorkArrays {
    int n_work;                       // multiple of 1024
    int *perm;                        // work slot -> local node index, -1 = padding
    float *fx, *fy, *fz;              // centred fp32 copy used by the sweep
    double *x64, *y64, *z64, *rxp;    // exact coordinates and rxProbability in work order (resolve / recheck)
    const int *tx_epoch; int hd_epoch; // half duplex: node p is on the air in this tick iff tx_epoch[p] == hd_epoch (nullptr: off)
};

// ---- K4: state maintenance ----------------------------------------------------------------
cudaError_t launch_geom(const NodeArrays &nd, unsigned long long *scratch6, Geom *geom, cudaStream_t st);
// work order = enabled receivers grouped by channel group (gid, -1 = disabled), Morton-sorted by (x, y)
// inside a group (stable: ties keep node order), every group padded to 1024 (perm = -1).
struct ReorderScratch { unsigned long long *key, *key_alt; int *val, *val_alt; void *tmp; size_t tmp_bytes; };
size_t      reorder_tmp_bytes(int n);
cudaError_t launch_reorder(const NodeArrays &nd, const int *gid, const Geom *geom, const Groups &groups, int n_work,
                           ReorderScratch sc, int *perm, cudaStream_t st);
cudaError_t launch_regather(const NodeArrays &nd, const WorkArrays &wk, const Geom *geom, cudaStream_t st);
cudaError_t launch_scatter_positions(const NodeArrays &nd, int count, const int *idx,
                                     const double *x, const double *y, const double *z, cudaStream_t st);
cudaError_t launch_init_params(DevParams *dp, cudaStream_t st);

// ---- per tick: tx records and channel bucketing ---------------------------------------------
// packets: src (global index), power (NaN => node's), channel (INT32_MIN => node's)
cudaError_t launch_gather_records(const NodeArrays &nd, int rx_base, int n_tx, const int *src, const double *power,
                                  const int *channel, rmb_tx_record *rec, cudaStream_t st);
// per-transmission staging in arrival order (before the stable placement into channel buckets)
// multi-GPU exchange: store `slots` local records into every peer's gather buffer, then raise this rank's flag
constexpr size_t kPeerFlagBytes = 256;       // flags[kMaxPeers] at the start of a gather buffer, records follow
struct PeerPush { char *peer[kMaxPeers]; int world, rank; size_t rec_off; unsigned epoch; unsigned *done; };
cudaError_t launch_peer_push(const rmb_tx_record *local, int slots, const PeerPush &pp, cudaStream_t st);
cudaError_t launch_peer_barrier(const PeerPush &pp, Counters *ctr, cudaStream_t st);
cudaError_t launch_peer_wait(const unsigned *flags, int world, unsigned epoch, Counters *ctr, cudaStream_t st);   // all ranks' records of `epoch` have landed   // pp.epoch = barrier epoch, second flag row
struct TxStage { float4 *f; float *theta; int *gid; };
// one cooperative launch: peer wait, counter zeroing, per-tx constants, tx-failure rule, stable bucketing
int prepare_grid(int n_tx);
cudaError_t launch_prepare(const NodeArrays &nd, const DevParams *dp, const Geom *geom, Groups groups, int n_tx,
                           const rmb_tx_record *rec, TxStage stg, TxTiles tx, float *theta, float *theta_min_g,
                           Counters *ctr, PrepareAux aux, int grid, cudaStream_t st);

// ---- K1 ---------------------------------------------------------------------------------------
struct SweepArgs {
    const float *fx, *fy, *fz;
    Groups groups;
    TxTiles tx;
    const float *theta, *theta_min_g;
    float neg_h;                      // -n/2 (general-exponent mode)
    float *far_part;                  // [splits][n_work] partial far-field sums
    int *cnt, *slots;                 // cnt must be zero on entry (atomic append of near candidates)
    int K;
};
int sweep_mode_for_exponent(double n);
int sweep_auto_variant(int mode, bool planar, int n_work, int n_tx_bucket);     // variant for a tick of this shape (shallow launches: shorter rx blocks)
// number of tx splits (grid.y) for a bucket of about n_tx_bucket transmissions; forced_splits > 0 overrides
int sweep_plan(int mode, bool planar, int variant, int n_work, int n_tx_bucket, int forced_splits);
// The sweep's fused epilogue (128-thread variants): the last tx-split block of an rx block resolves its receivers
// (exact decisions, pair records) and the last rx block scans the per-packet counts -- see resolve_body.cuh.
struct SlotArrays; struct PairRecs;
struct SweepSplits { int s1, s2, rb_taper, rb_fine, block_rx; bool can_fuse; };
SweepSplits sweep_splits(int mode, bool planar, int variant, int n_work, int n_tx_bucket, int s1, int taper, int fine8);

// ---- exact resolve of the near candidates + guard-band recheck --------------------------------
struct SlotArrays {
    int K;
    int *slots;                       // [n_work*K] bucket position j of the candidate (resolve sorts them ascending)
    int *cnt;                         // [n_work] candidates found (may exceed K => retry)
    float *far_part;                  // [max splits][n_work] fp32 partial far-field interference sums [mW]
    int splits, splits2, w_taper;     // receivers < w_taper were swept in `splits` tx splits, the others in `splits2`
    double *prx, *lin, *sinr;         // [n_work*K] exact values per candidate
    uint8_t *flags;                   // [n_work*K] kHeard | kDecodable
    int *hcnt;                        // [n_work] heard pairs of the receiver
    double *itot;                     // [n_work] total in-band power [mW]
    int *recheck;                     // receivers inside the SINR guard band
    int *recheck_cnt;                 // [entry of recheck] partial sums stored so far (zeroed when the entry is appended)
    int *recheck2;                    // of those, the ones the tree sum cannot decide (sequential redo)
    int *first;                       // [n_work] first pair record of the receiver (its heard slots follow in slot order)
};
// dense heard-pair records, appended by the resolve kernel (one warp-aggregated atomic per 32 receivers)
struct PairRecs {
    long long cap;                    // records beyond cap are dropped; the host grows and redoes the resolve
    int key_shift;
    unsigned long long *key;          // (t << key_shift) | global rx
    double *prx, *sinr;
    uint8_t *flags;                   // kHeard | kDecodable
    int *rank;                        // position inside the packet's bucket (atomic counter, arbitrary order)
    int *tx_cnt;                      // [n_tx] heard pairs per packet (zero on entry)
};
struct SweepFuse {
    int enabled;
    WorkArrays wk; const DevParams *dp; const rmb_tx_record *rec; SlotArrays sl; PairRecs pr; Counters *ctr;
    int n_tx; int *seg;               // seg == nullptr: the kernel behind the sweep scans the per-packet counts itself
    unsigned long long *trace;        // debug (RMB_SWEEP_TRACE): [block][12] start / sweep done / block done [ns], smid | passes << 16, stamps of the last pass
    int *rx_done, *rx_pass;           // [rx blocks] splits finished / resolve passes claimed (zero on entry)
    int rb_spin;                      // rx blocks >= rb_spin: the last finishers of an rx block wait for its last split and share the passes
    int rb_fine;                      // rx blocks >= rb_fine (>= rb_spin): passes of 32 receivers instead of 128
};
// overlap: programmatic dependent launch behind the preparation kernel (no event may sit between the two launches)
cudaError_t launch_sweep(int mode, bool planar, int variant, int n_work, SweepSplits splits, const SweepArgs &a, const SweepFuse &fuse,
                         bool overlap, cudaStream_t st);
// also turns the per-packet heard counts into segment offsets (seg[t] = first pair of packet t, exclusive scan
// of pr.tx_cnt; largest bucket -> ctr->max_bucket): done by the last block to finish
cudaError_t launch_resolve(const NodeArrays &nd, const WorkArrays &wk, const DevParams *dp, TxTiles tx,
                           const rmb_tx_record *rec, SlotArrays sl, PairRecs pr, Counters *ctr, int n_tx, int *seg,
                           int w_first, cudaStream_t st);   // receivers [w_first, n_work) (a multiple of 128)
// per-node RSSI (total in-band power, dBm) and receiving flag from the work-order results of the last tick
cudaError_t launch_node_info(const WorkArrays &wk, SlotArrays sl, float *node_rssi, double *node_rssi64, uint8_t *node_recv, cudaStream_t st);
// part: >= n_work + grid doubles of scratch; cooperative launch of `grid` (= recheck_grid()) blocks with its
// own barrier counter
int recheck_grid();
cudaError_t launch_recheck(const NodeArrays &nd, const WorkArrays &wk, const DevParams *dp, Groups groups, TxTiles tx,
                           const rmb_tx_record *rec, SlotArrays sl, PairRecs pr, Counters *ctr, double *part, int grid,
                           unsigned *bar, unsigned bar_target, cudaStream_t st);

struct PairArrays;
// everything behind the sweep of a deterministic-reception tick in one cooperative kernel (finish.cu): guard-band
// recheck, bucket placement, per-packet ranking and the output rows.  bar: a counter that is ZERO at launch.
// part: >= n_work + grid doubles; items: >= record capacity
// finish_scans(n_tx): the kernel turns the per-packet heard counts into segment offsets itself (every block, in shared
// memory) -- the sweep's epilogue then leaves the scan out (SweepFuse.seg == nullptr).  overlap: launched with
// programmatic stream serialization, so that its blocks become resident while the sweep's last blocks still run.
bool finish_scans(int n_tx);
int finish_grid(int n_tx);
cudaError_t launch_finish(const WorkArrays &wk, const DevParams *dp, Groups groups, TxTiles tx, const rmb_tx_record *rec,
                          SlotArrays sl, PairRecs pr, PairArrays pa, int n_tx, int *seg, unsigned long long *items,
                          double *part, Counters *ctr, Counters *host_ctr, int grid, unsigned *bar, bool overlap, cudaStream_t st);

// ---- emission: compaction, (tx,rx) ordering, stochastic rule, outputs -------------------------
struct PairArrays {
    long long cap;
    int key_shift;                       // key = (t << key_shift) | global rx
    unsigned long long *key, *key_alt;   // (t << key_shift) | global rx
    unsigned *val, *val_alt;             // slot reference (SINR) / unused
    double *p;                           // reception probability of the pair
    double *rssi, *sinr;                 // exact values (sorted order)
    uint8_t *cand;                       // pair may be delivered (decodable / in range)
    int *draw_scan;                      // exclusive scan of "needs a draw" flags
    int *seg;                            // [n_tx+1] first sorted pair of packet t
    long long *tx_base;                  // [n_tx] uniform index of the packet's first receiver draw
    uint8_t *tx_failed;                  // [n_tx]
    int *scan_tmp;                       // block sums for the scans
    void *cub_tmp; size_t cub_tmp_bytes;
    // outputs
    int *o_tx, *o_rx; float *o_rssi, *o_sinr; uint8_t *o_outcome; double *o_rssi64, *o_sinr64;
    // finish.cu only: all columns in ONE buffer, packed for the tick's actual row count (slab_layout), so that the
    // lists travel to the host in a single copy; slab == nullptr: the separate arrays above
    char *slab; int slab_flags;
};
// Where the columns of a tick with m rows and n_tx packets sit in the slab (16-byte aligned); the same arithmetic on
// the device (finish kernel, m from the counters) and on the host (download).  flags: rmb_flags of the handle.
struct SlabLayout { long long rx, rssi, tx, sinr, rssi64, sinr64, outcome, first, failed, total; };
__host__ __device__ inline SlabLayout slab_layout(long long m, int n_tx, int flags)
{
    auto up = [](long long v) { return (v + 15) / 16 * 16; };
    const bool f64 = (flags & RMB_FLAG_F64_VALUES) != 0, compact = (flags & RMB_FLAG_COMPACT) != 0;
    SlabLayout L{};
    long long o = 0;
    L.rx = o; o += up(4 * m);
    L.rssi = o; o += up(4 * m);
    L.tx = L.sinr = -1;
    if (!compact) { L.tx = o; o += up(4 * m); L.sinr = o; o += up(4 * m); }
    L.rssi64 = L.sinr64 = -1;
    if (f64) { L.rssi64 = o; o += up(8 * m); L.sinr64 = o; o += up(8 * m); }
    L.outcome = o; o += up(m);
    L.first = o; o += up(4ll * (n_tx + 1));
    L.failed = o; o += up(n_tx);
    L.total = o;
    return L;
}
// exclusive scan of int32 (n elements) using tmp (>= n/1024+2 ints); total written to *total_out (device)
cudaError_t launch_exclusive_scan(const int *in, int *out, int n, int *tmp, int *total_out, cudaStream_t st);
// SINR medium: pair records -> (packet arrival, node index) order.
//   seg     : seg[t] = first pair of packet t (exclusive scan of tx_cnt, done by the resolve), largest bucket -> ctr->max_bucket
//   order   : order[pos] = record index of the pos-th pair.  Buckets are filled through the atomic ranks and
//             sorted by rx inside one block each (shared-memory bitonic sort, buckets <= kBucketSortMax);
//             larger buckets take the CUB radix-sort fallback over the whole key.
//   emit    : gathers the records in that order and writes the outputs (or, for stochastic reception,
//             the sorted pair arrays the draw-order stage consumes).
constexpr int kBucketSortMax = 2048;
struct OrderScratch { unsigned long long *items; unsigned long long *fb_key, *fb_key_alt; unsigned *fb_val, *fb_val_alt; void *cub_tmp; size_t cub_tmp_bytes; };
size_t      sort_pairs_tmp_bytes(long long n);
cudaError_t launch_order_buckets(PairRecs pr, const Counters *ctr, int n_tx, const int *seg, unsigned long long *items,
                                 unsigned *order, cudaStream_t st);
cudaError_t launch_order_fallback(PairRecs pr, long long m, int key_bits, OrderScratch sc, unsigned *order, cudaStream_t st);
cudaError_t launch_emit_sinr(const NodeArrays &nd, const DevParams *dp, const rmb_tx_record *rec, int n_tx, PairRecs pr,
                             const unsigned *order, PairArrays pa, long long m, bool stochastic, Counters *ctr,
                             Counters *host_ctr, cudaStream_t st);
// first[t] = first row of packet t in the (sorted) tx column, first[n_tx] = n
cudaError_t launch_tx_first(const int *o_tx, long long n, int n_tx, int *first, cudaStream_t st);
cudaError_t launch_emit(const DevParams *dp, const rmb_tx_record *rec, int n_tx, PairArrays pa, long long n,
                        bool stochastic, const double *uniforms, long long n_uniforms, Counters *ctr,
                        cudaStream_t st);

// the stochastic rule in two halves, for receivers sharded over several handles: every shard reports the draws
// its receivers would make per packet (counts_out[n_tx]), the host side gathers them (all_counts[world][n_tx]),
// and every shard then walks the same global uniform sequence.  all_counts == nullptr: unsharded.
cudaError_t launch_emit_draw_prepare(int n_tx, PairArrays pa, long long n, int *counts_out, cudaStream_t st);
cudaError_t launch_emit_draw_finish(const DevParams *dp, const rmb_tx_record *rec, int n_tx, PairArrays pa, long long n,
                                    const double *uniforms, long long n_uniforms, Counters *ctr,
                                    const int *all_counts, int world, int rank, const int *order, int n_order,
                                    cudaStream_t st);   // order: records in original packet arrival order (null = as stored)

// ---- K2: the reference media (Null / UDGM / UDGMConstantLoss / N2N), exact double --------------
// tx_epoch / hd_epoch: half duplex (nullptr: off) -- launch_refmedia_mark stamps the tick's local sources first
cudaError_t launch_refmedia_mark(const NodeArrays &nd, const DevParams *dp, int n_tx, const rmb_tx_record *rec,
                                 int *tx_epoch, int hd_epoch, cudaStream_t st);
cudaError_t launch_refmedia_count(const NodeArrays &nd, const DevParams *dp, const double *matrix, int matrix_n,
                                  int n_tx, const rmb_tx_record *rec, int *blkcnt, Counters *ctr, bool stochastic,
                                  const int *tx_epoch, int hd_epoch, cudaStream_t st);
cudaError_t launch_refmedia_scatter(const NodeArrays &nd, const DevParams *dp, const double *matrix, int matrix_n,
                                    int n_tx, const rmb_tx_record *rec, const int *blkoff, PairArrays pa,
                                    uint8_t *node_recv, const int *tx_epoch, int hd_epoch, cudaStream_t st);
int refmedia_blocks_per_tx(int n);

}  // namespace rmb
