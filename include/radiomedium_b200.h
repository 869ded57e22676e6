/*
 * radiomedium_b200.h -- C-ABI of the B200-native radio medium (librmb200.so).
 *
 * This is the drop-in boundary for radio-sim's radio-medium hot path.  Every entry point
 * names the reference interface it replaces (paths relative to the reference tree,
 * radio-medium/java/se/sics/emul8/radiomedium/ unless a directory is given).
 * Plain C: opaque handle, plain pointers and sizes, no exceptions, no torch types.
 * The Java side binds these with Panama FFM (INTEGRATION.md shows the stub).
 *
 * Threading: a handle is serialised internally by one mutex (RadioMedium.transmit is called
 * from any client reader thread: net/SimulatorJSONHandler.java:93).
 * Ownership: inputs are caller-owned and copied before return.  Result pointers are
 * handle-owned pinned host memory, valid until the next rmb_tick* / rmb_fetch / rmb_destroy.
 * Errors: 0 or a negative rmb_status; text via rmb_last_error().  There is no CPU fallback:
 * without a CUDA device rmb_create fails with RMB_ERR_NO_DEVICE.
 */
#ifndef RADIOMEDIUM_B200_H
#define RADIOMEDIUM_B200_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RMB_ABI_VERSION 2

typedef struct rmb_handle rmb_handle;

enum rmb_status {
    RMB_OK            =  0,
    RMB_ERR_INVALID   = -1,   /* bad argument / bad index                               */
    RMB_ERR_NO_DEVICE = -2,   /* no CUDA device (no CPU fallback exists)                */
    RMB_ERR_CUDA      = -3,   /* CUDA runtime error, see rmb_last_error                 */
    RMB_ERR_CAPACITY  = -4,   /* more heard pairs than max_pairs                        */
    RMB_ERR_UNIFORMS  = -5,   /* uniform stream too short (missing draws read as 0.0)   */
    RMB_ERR_STATE     = -6,   /* call order (e.g. tick before set_nodes)                */
    RMB_ERR_UNSUPPORTED = -7,
    RMB_ERR_PEER      = -8    /* multi-GPU exchange: a peer's records did not arrive in time */
};

/* which RadioMedium implementation the handle behaves as */
enum rmb_model {
    RMB_MODEL_NULL          = 0,  /* NullRadioMedium.java:48-77                         */
    RMB_MODEL_UDGM          = 1,  /* UDGMRadioMedium.java:63-116                        */
    RMB_MODEL_UDGM_CONSTANT = 2,  /* UDGMConstantLossRadioMedium.java:16-36             */
    RMB_MODEL_N2N           = 3,  /* N2NRadioMedium.java:24-72                          */
    RMB_MODEL_SINR          = 4   /* new: log-distance + co-channel interference + SINR */
};

/* per (tx,rx) outcome == what the Java medium does with the pair:
 * no call / generateReceptionEvents(..,false) / generateReceptionEvents(..,true)
 * (Simulator.java:321-335, events/ReceptionEvent.java:12-16) */
enum rmb_outcome { RMB_NOT_HEARD = 0, RMB_INTERFERED = 1, RMB_DELIVERED = 2 };

enum rmb_flags {
    RMB_FLAG_F64_VALUES = 1,  /* also return rssi/sinr as double (exact-path values)    */
    RMB_FLAG_COMPACT    = 2   /* result lists as the reference's callbacks need them and no more: rx, rssi, outcome
                                 per row plus tx_first per packet; the tx and sinr columns stay on the device
                                 (pointers NULL) -- 9 instead of 17 bytes per heard pair over PCIe              */
};

/* Replaces the fields/setters of the media (UDGMRadioMedium.java:17-61,
 * UDGMConstantLossRadioMedium.java:8, AbstractRadioMedium.java:38) and the
 * "configuration-set" parameters (net/SimulatorJSONHandler.java:170-216). */
typedef struct rmb_config {
    int32_t struct_size;          /* = sizeof(rmb_config), ABI check                     */
    int32_t model;                /* rmb_model                                           */
    int32_t device;               /* CUDA device ordinal                                 */
    int32_t flags;                /* rmb_flags                                           */
    double  success_ratio_tx;     /* 1.0  (never read by transmit(): reference quirk)    */
    double  success_ratio_rx;     /* 1.0                                                 */
    double  transmission_range;   /* 50.0                                                */
    double  constant_range;       /* 100.0                                               */
    double  pl0_db;               /* 40.05  path loss at d0                              */
    double  pl_exponent;          /* 3.0                                                 */
    double  d0_m;                 /* 1.0                                                 */
    double  sensitivity_dbm;      /* -95.0                                               */
    double  sinr_threshold_db;    /* 4.0                                                 */
    double  noise_floor_dbm;      /* -100.0 == AbstractRadioMedium.baseRSSI              */
    int64_t max_pairs;            /* cap on heard pairs per tick, 0 = default (64 Mi)    */
    int32_t rx_base;              /* multi-GPU: global index of this shard's first node  */
    int32_t n_shards;             /* multi-GPU: number of receiver shards (0/1 = unsharded) */
    int32_t half_duplex;          /* 0 = the reference: only node == source is excluded as a receiver
                                     (UDGMRadioMedium.java:100), a node that is itself on the air still receives.
                                     1 = spec flag (all media): a node that is the source of ANY packet of the tick
                                     passes no receiver filter in that tick and reports the noise floor            */
    int32_t reserved;             /* 0                                                   */
} rmb_config;

/* Limits (exceeding one is an error return, never a wrong result):
 *   - at most 64 distinct wireless channels among the ENABLED receivers of a handle (RMB_ERR_UNSUPPORTED from
 *     rmb_set_nodes / rmb_update_radio); packets may use any channel;
 *   - max_pairs <= 2^31 - 1 (row and segment offsets are int32; rmb_create rejects more with RMB_ERR_INVALID);
 *   - reference media (NULL/UDGM/UDGM_CONSTANT/N2N): n_tx * ceil(n / 256) < 2^31 per tick (RMB_ERR_CAPACITY);
 *   - SINR: per-receiver near-candidate slots grow on demand (x2) up to 3e9 slots per handle; a tick in which
 *     every receiver hears thousands of packets fails with RMB_ERR_CAPACITY instead;
 *   - N2N needs the source's integer id, which a tx record does not carry: RMB_MODEL_N2N with n_shards > 1 is
 *     rejected by rmb_create (RMB_ERR_UNSUPPORTED);
 *   - stochastic reception (any rx/tx probability strictly between 0 and 1) on sharded receivers needs the
 *     rmb_tick_draws_begin/_end pair (rmb_group_tick goes through it by itself); any other per-handle
 *     tick entry point returns RMB_ERR_UNSUPPORTED for such a tick, also when the probability arrives in a record
 *     built by another shard;
 *   - at most 16 handles exchange records with rmb_peer_* (RMB_ERR_INVALID). */

/* One active transmission as the device sees it (RadioPacket.java:46-52 plus the
 * source's Position.java:37-39 and Transciever.txProbability).  48 bytes; this is the
 * record that is all-gathered between GPUs. */
typedef struct rmb_tx_record {
    double  x, y, z;
    double  power_dbm;
    double  txprob;
    int32_t channel;
    int32_t src;                  /* global node index, -1 = padding (ignored)           */
} rmb_tx_record;

/* Result of one tick: the heard (tx,rx) pairs in the Java call order of
 * generateReceptionEvents -- packets in arrival order, receivers in node-array order
 * (UDGMRadioMedium.java:99-114). */
typedef struct rmb_result {
    int64_t  n_pairs;             /* heard pairs                                         */
    int64_t  n_delivered;
    int64_t  n_interfered;
    uint64_t n_decisions;         /* filter-passing (tx,rx) pairs evaluated              */
    int64_t  n_uniforms_used;
    int64_t  n_recheck;           /* receivers that took the exact double path           */
    int32_t  n_tx;
    int32_t  near_slots;          /* internal per-receiver candidate capacity used       */
    const int32_t *tx;            /* [n_pairs] index into this tick's packet list        */
    const int32_t *rx;            /* [n_pairs] global node index ("neighbour index")     */
    const float   *rssi_dbm;      /* [n_pairs] value for generateReceptionEvents         */
    const float   *sinr_db;       /* [n_pairs] NaN for the reference media               */
    const uint8_t *outcome;       /* [n_pairs] RMB_INTERFERED / RMB_DELIVERED            */
    const double  *rssi_dbm_f64;  /* NULL unless RMB_FLAG_F64_VALUES                     */
    const double  *sinr_db_f64;
    const uint8_t *tx_failed;     /* [n_tx] UDGMRadioMedium.java:88-92 "interference"    */
    const int32_t *tx_first;      /* [n_tx+1] first row of packet t (rows of t: tx_first[t] .. tx_first[t+1]);
                                     NULL unless RMB_FLAG_COMPACT                          */
} rmb_result;

/* device-side durations of the last tick, CUDA events on the handle's stream */
typedef struct rmb_timing {
    float ms_total;               /* first kernel .. last kernel (no host copies)        */
    float ms_prepare;             /* tx record gather + channel bucketing                */
    float ms_sweep;               /* the link-budget / decision sweep (dominant kernel)  */
    float ms_resolve;             /* exact near-pair evaluation + guard-band recheck     */
    float ms_emit;                /* compaction, ordering, stochastic rule, outputs      */
    int32_t sweep_launches;
    int32_t kernel_launches;      /* all kernels launched by the last tick               */
    int32_t recheck_sequential;   /* guard-band receivers that needed the sequential sum */
} rmb_timing;

int         rmb_version(void);
void        rmb_config_default(rmb_config *cfg, int32_t model);
/* Replaces `new XxxRadioMedium()` + Simulator.setRadioMedium (Simulator.java:234-239). */
int         rmb_create(const rmb_config *cfg, rmb_handle **out);
int         rmb_destroy(rmb_handle *h);
const char *rmb_last_error(const rmb_handle *h);   /* h may be NULL: last create error  */

/* Replaces Simulator.addNode + the node-config-set fields (Simulator.java:249-277,
 * net/SimulatorJSONHandler.java:105-143): the whole node array, registration order.
 * NULL arrays take the reference defaults: z=0, txpower=0, channel=26, enabled=1,
 * rxprob=txprob=1 (Transciever.java:11-18), id_int = index+1 (Node.java:52-58). */
int rmb_set_nodes(rmb_handle *h, int32_t n, const double *x, const double *y, const double *z,
                  const double *txpower_dbm, const int32_t *channel, const uint8_t *enabled,
                  const double *rxprob, const double *txprob, const int32_t *id_int);
/* N2NRadioMedium.setN2NMatrix / matrix-data (N2NRadioMedium.java:20-22,
 * net/SimulatorJSONHandler.java:183-206): flat row-major n x n. */
int rmb_set_matrix(rmb_handle *h, int32_t n, const double *matrix);
/* node-config-set "position" (net/SimulatorJSONHandler.java:111-122; Position.java:44-52):
 * idx==NULL => dense replace of all `count`==n nodes; z==NULL => z=0 (Position.set(x,y)). */
int rmb_update_positions(rmb_handle *h, int32_t count, const int32_t *idx,
                         const double *x, const double *y, const double *z);
/* The dense form without the wait: the next positions of ALL nodes are uploaded on a copy stream into a second set of
 * arrays and become the node table at the next call that settles the handle (a tick, an update, rmb_build_tx_records,
 * rmb_fetch ...; rmb_tick_finish alone does not).
 * It does NOT wait for an enqueued tick (rmb_tick_enqueue / rmb_tick_peers_enqueue): the upload of tick k+1's
 * positions runs beside tick k on the GPU -- per-tick mobility of every node (BASELINE configs[4]) costs one PCIe
 * transfer that hides behind the previous tick.  The caller's arrays must stay untouched until that next settling
 * call returns (pinned memory makes the copy asynchronous; pageable memory is staged by the driver inside the call). */
int rmb_update_positions_async(rmb_handle *h, int32_t count, const double *x, const double *y, const double *z);
/* node-config-set rf-power / wireless-channel / rx-loss / tx-loss / radio-state
 * (net/SimulatorJSONHandler.java:123-143).  NULL field arrays are left unchanged. */
int rmb_update_radio(rmb_handle *h, int32_t count, const int32_t *idx, const double *txpower_dbm,
                     const int32_t *channel, const uint8_t *enabled,
                     const double *rxprob, const double *txprob);

/* Replaces the tick's RadioMedium.transmit(RadioPacket) calls (RadioMedium.java:41),
 * batched: packet t = {source node index, power, channel} in arrival order.
 * power_dbm / channel NULL => snapshot from the source radio (RadioPacket.java:49-50).
 * uniforms stands in for simulator.getRandom().nextDouble() (Simulator.java:94), consumed
 * in the Java draw order.  All sources must be local to this handle. */
int rmb_tick(rmb_handle *h, int32_t n_tx, const int32_t *src, const double *power_dbm,
             const int32_t *channel, const double *uniforms, size_t n_uniforms, rmb_result *out);

/* Multi-GPU form of the same step (receivers sharded, tx records all-gathered):
 * rmb_build_tx_records writes n_tx records to DEVICE memory d_records (src indices are
 * global; entries whose source is not local get src=-1 unless take_all!=0);
 * rmb_tick_records evaluates n_tx records (DEVICE pointer, src=-1 entries ignored but they
 * keep their arrival index). */
int rmb_build_tx_records(rmb_handle *h, int32_t n_tx, const int32_t *src, const double *power_dbm,
                         const int32_t *channel, void *d_records);
int rmb_tick_records(rmb_handle *h, int32_t n_tx, const void *d_records,
                     const double *uniforms, size_t n_uniforms, rmb_result *out);
/* Same computation on records already resident in HBM; results stay on the device
 * (out gets counts only, pointers NULL).  rmb_fetch downloads the lists afterwards. */
int rmb_tick_resident(rmb_handle *h, int32_t n_tx, const void *d_records, rmb_result *out);
int rmb_fetch(rmb_handle *h, rmb_result *out);
/* Stochastic reception (rx-loss / tx-loss < 1) when the receivers are sharded over several handles.  The
 * reference draws from ONE java.util.Random in a global order -- per packet the tx draw, then one draw per
 * receiver that needs it in node-array order (UDGMRadioMedium.java:85-92,104-111) -- so a shard has to know
 * how many draws the shards before it make for every packet:
 *   rmb_tick_draws_begin  evaluates the tick up to the draw flags and writes, to DEVICE memory, the draws this
 *                         shard's receivers would make per record (int32[n_tx]);
 *   (the caller all-gathers these: all_counts[world][n_tx], shards in node-index order)
 *   rmb_tick_draws_end    walks the global sequence -- every shard evaluates every tx draw, identically -- and
 *                         finishes the tick; `uniforms` is the SAME host-supplied stream on every shard, and
 *                         n_uniforms_used is the global count.  d_order (DEVICE int32[n_order], may be NULL =
 *                         as stored): the record positions in the tick's original packet arrival order -- an
 *                         exchanged record list is grouped by owner rank, the reference draws in arrival order.
 * Works on an unsharded handle too (world = 1).  SINR and UDGM media (the other media draw nothing, or -- N2N --
 * cannot be sharded). */
int rmb_tick_draws_begin(rmb_handle *h, int32_t n_tx, const void *d_records, int32_t *d_counts_out);
int rmb_tick_draws_end(rmb_handle *h, const int32_t *d_all_counts, int32_t world, int32_t rank,
                       const int32_t *d_order, int32_t n_order,
                       const double *uniforms, size_t n_uniforms, rmb_result *out);
/* Asynchronous form: rmb_tick_enqueue puts the whole tick on the handle's stream and returns at once
 * (d_records must stay valid until the tick is settled); rmb_tick_finish waits for it and returns the
 * counts (pointers NULL; rmb_fetch downloads the lists).  Any other call on the handle settles an
 * enqueued tick first.  Lets the Java side overlap the next tick's marshalling with the GPU
 * (Simulator.emulatorTimeStepDone, Simulator.java:155-161) and keeps the host round trip out of the
 * device time of a tick.  Stochastic reception (draws) is evaluated synchronously inside the call. */
int rmb_tick_enqueue(rmb_handle *h, int32_t n_tx, const void *d_records);
int rmb_tick_finish(rmb_handle *h, rmb_result *out);

/* Multi-GPU, second form: the exchange of the tick's tx records done by the library itself over
 * NVLink peer memory instead of a collective the caller runs (north_star: "the small per-tick
 * active-transmitter set is all-gathered").  One handle per GPU, in one process or one process per
 * GPU.  Setup: every handle allocates its gather buffer and describes it (rmb_peer_export:
 * RMB_PEER_HANDLE_BYTES bytes holding a CUDA IPC handle plus, for peers living in the same process, the
 * raw pointer); the host passes the blobs around by any means (torch.distributed in this repo) and
 * every handle maps its peers' buffers (rmb_peer_open, blobs in rank order).
 * Per tick, on every rank with the same `slots` (<= max_slots):
 *   rmb_peer_push          builds the records of the locally owned sources -- src[slot] is a global node
 *                          index owned by this handle, or -1 for padding -- STORES them into every
 *                          peer's buffer and then raises this rank's epoch flag there;
 *   rmb_peer_push_records  the same for records already in this GPU's memory (16-byte aligned);
 *   rmb_tick_peers         waits ON THE DEVICE for all ranks' flags (no host synchronisation, no
 *                          collective) and evaluates the world*slots gathered records exactly like
 *                          rmb_tick_records; arrival index of a record = rank*slots + slot.
 *   rmb_peer_tick          rmb_peer_push + rmb_tick_peers in ONE call, host packet list in, host lists out.  On the
 *                          SINR medium without draws, every handle on a device of its own, the exchange is part of
 *                          the tick's first kernel (no separate push launch); otherwise the two steps run back to
 *                          back.  A handle's records leave when its tick starts, so every rank has to be inside
 *                          this call at the same time -- one thread or process per handle; calling the handles one
 *                          after another from one thread runs into the time-out.
 * Pushes and ticks must alternate, in step on all ranks (buffers are double-buffered by tick parity).
 * A rank that never pushes makes the others fail with RMB_ERR_PEER after ~20 s instead of hanging. */
#define RMB_PEER_HANDLE_BYTES 128
int rmb_peer_export(rmb_handle *h, int32_t world, int32_t rank, int32_t max_slots, void *blob_out);
int rmb_peer_open(rmb_handle *h, const void *blobs /* world * RMB_PEER_HANDLE_BYTES */);
int rmb_peer_push(rmb_handle *h, int32_t slots, const int32_t *src, const double *power_dbm, const int32_t *channel);
int rmb_peer_push_records(rmb_handle *h, int32_t slots, const void *d_records);
int rmb_tick_peers(rmb_handle *h, const double *uniforms, size_t n_uniforms, rmb_result *out);
int rmb_peer_tick(rmb_handle *h, int32_t slots, const int32_t *src, const double *power_dbm, const int32_t *channel, rmb_result *out);
int rmb_tick_peers_resident(rmb_handle *h, rmb_result *out);   /* results stay on the device, see rmb_fetch */
int rmb_tick_peers_enqueue(rmb_handle *h);                     /* asynchronous, settled by rmb_tick_finish */
int rmb_tick_peers_draws_begin(rmb_handle *h, int32_t *d_counts_out);   /* rmb_tick_draws_begin on the gathered records */
/* rmb_peer_push_records + rmb_tick_peers_enqueue as ONE step: the tick's first kernel stores this rank's
 * records into the peers' buffers, raises the flags and then waits for the peers' -- exchange and the
 * preparation that consumes it in a single launch.  Asynchronous; every rank must call it for the tick.
 * Needs one GPU per handle: the preparation kernel is a cooperative launch, and two cooperative kernels do not
 * run side by side on one device (handles sharing a GPU use rmb_peer_push + rmb_tick_peers). */
int rmb_peer_exchange_tick_enqueue(rmb_handle *h, int32_t slots, const void *d_records);
/* Device-side rendezvous of all ranks' streams: a one-warp kernel that raises this rank's barrier flag in every
 * peer's buffer and waits for all of theirs (no host synchronisation, no collective).  Every rank must call it the
 * same number of times.  What a caller puts on the stream behind it starts within an NVLink round trip on all
 * GPUs -- bench.py lines the ranks up with it in front of the start event.  Asynchronous. */
int rmb_peer_barrier_enqueue(rmb_handle *h);

/* Replaces Transciever.getRSSI / RadioMedium.getBaseRSSI for the per-tick node-info
 * (Transciever.java:52-61, net/JSONClientConnection.java:331-341): total in-band power
 * of the last tick per local node [dBm] (noise floor when idle), and 1 if the node heard
 * at least one frame. NULL outputs are skipped. */
int rmb_node_info(rmb_handle *h, float *rssi_dbm, uint8_t *receiving);
/* the same in double: what `node-info.rssi` carries (a JSON double, net/JSONClientConnection.java:337) and what
 * the parity tests compare with the oracle's sequential sum */
int rmb_node_info_f64(rmb_handle *h, double *rssi_dbm, uint8_t *receiving);

/* ---- several GPUs behind one object ---------------------------------------------------------------------------
 * For a caller that lives in ONE process and thinks in whole node arrays and whole packet lists -- the reference's
 * Simulator with its one RadioMedium (Simulator.java:155-165, 234-239).  A group owns one handle per device
 * (receivers in contiguous node-index blocks, rmb_config.rx_base / n_shards filled in), maps the handles' exchange
 * buffers onto each other and runs the shards' calls on one host worker thread per device, so that kernels and
 * device-to-host copies of all GPUs overlap.  Every call below is the whole-array form of the handle call of the
 * same name.
 *   rmb_group_create   cfg: as for rmb_create (device / rx_base / n_shards are overwritten); devices NULL = 0..n-1
 *                      (the same device may appear twice: development on a one-GPU box); max_tx_per_tick sizes the
 *                      exchange buffers (a tick with more packets from one shard returns RMB_ERR_CAPACITY).
 *   rmb_group_tick     results[r] = shard r's lists (r = 0..n_devices-1, handle-owned pinned memory as for rmb_tick).
 *                      Row order inside a shard: gathered record position, then node index; record position
 *                      p = r * slots + i is packet packet_of[p] of the caller's list (-1 = padding).  Shards own
 *                      increasing node blocks, so packet t's rows in the reference's order are shard 0's rows of t,
 *                      then shard 1's, ...  (*slots_out, packet_of[n_devices * max_tx_per_tick] may be NULL for
 *                      n_devices == 1, where position == packet).  uniforms: as for rmb_tick (RMB_ERR_UNIFORMS tells
 *                      how many a tick needs).  On several devices stochastic reception goes through the handles'
 *                      rmb_tick_peers_draws_begin / rmb_tick_draws_end with the per-packet draw counts gathered in
 *                      between, in the caller's packet order: the same draws as one device makes; every results[r]
 *                      then carries the GLOBAL n_uniforms_used. */
typedef struct rmb_group rmb_group;
int         rmb_group_create(const rmb_config *cfg, int32_t n_devices, const int32_t *devices, int32_t max_tx_per_tick, rmb_group **out);
int         rmb_group_destroy(rmb_group *g);
const char *rmb_group_last_error(const rmb_group *g);      /* g may be NULL: last create error */
int32_t     rmb_group_size(const rmb_group *g);
int         rmb_group_shard(const rmb_group *g, int32_t r, rmb_handle **handle, int32_t *first_node, int32_t *n_nodes);
int rmb_group_set_nodes(rmb_group *g, int32_t n, const double *x, const double *y, const double *z,
                        const double *txpower_dbm, const int32_t *channel, const uint8_t *enabled,
                        const double *rxprob, const double *txprob, const int32_t *id_int);
int rmb_group_update_positions(rmb_group *g, int32_t count, const int32_t *idx, const double *x, const double *y, const double *z);
int rmb_group_update_radio(rmb_group *g, int32_t count, const int32_t *idx, const double *txpower_dbm, const int32_t *channel,
                           const uint8_t *enabled, const double *rxprob, const double *txprob);
int rmb_group_tick(rmb_group *g, int32_t n_tx, const int32_t *src, const double *power_dbm, const int32_t *channel,
                   const double *uniforms, size_t n_uniforms, rmb_result *results, int32_t *slots_out, int32_t *packet_of);
int rmb_group_node_info_f64(rmb_group *g, double *rssi_dbm, uint8_t *receiving);
/* Moves a shard: the global index of the handle's first node (rmb_config.rx_base) for the node table that the next
 * rmb_set_nodes uploads.  What rmb_group_set_nodes uses when the node count, and with it the partition, changes. */
int rmb_set_shard_base(rmb_handle *h, int32_t rx_base);
/* 1 when a tick on this handle may draw from the uniform stream -- some node of its table has an rx or tx probability
 * below 1, or the model draws by itself (N2N; UDGM with success_ratio_rx < 1: UDGMRadioMedium.java:104-111) -- else 0.
 * When ANY shard of a partition says 1, every shard has to evaluate its ticks through rmb_tick_draws_begin/_end (a
 * packet's source may sit on another shard); rmb_group_tick asks this. */
int32_t rmb_may_draw(rmb_handle *h);

int rmb_last_timing(const rmb_handle *h, rmb_timing *out);
/* CUDA stream the handle's kernels and copies run on (a cudaStream_t passed as void*).
 * rmb_set_stream makes the handle use the caller's stream instead of its own, so that work the
 * caller orders on that stream (e.g. the NCCL all-gather of the tx records) needs no host
 * synchronisation with the tick; rmb_build_tx_records then does not synchronise either.
 * NULL restores the handle's own stream.  No reference counterpart (device plumbing). */
int rmb_get_stream(rmb_handle *h, void **stream);
int rmb_set_stream(rmb_handle *h, void *stream);
/* tuning knob for benchmarks/tests: near-slot capacity, and in `sweep_variant`
 *   bits 0..7   sweep kernel variant (0 = auto)            bits 8..15  forced number of tx splits (0 = planned)
 *   bit 16      stand-alone resolve kernel instead of the sweep's fused epilogue
 *   bits 17..19 tx-split multiplier of the launch's final wave (0 = default 2, 1 = none)
 *   bit 20      separate recheck / ordering / emit kernels instead of the one finish kernel
 *   bits 21..23 eighths of a wave resolved in 32-receiver passes (0 = default 4, 7 = none)
 * Results do not depend on any of these. */
int rmb_set_tuning(rmb_handle *h, int32_t sweep_variant, int32_t near_slots);

/* NO COUNTERPART IN THE REFERENCE.  One tick of the UDGM medium with every
 * probability at 1 (the rule of UDGMRadioMedium.java:67-81,100-102 without its draws: heard <=> node != source,
 * enabled, on the packet's channel, d*d / (R*R) <= 1; heard == delivered; RSSI = the packet's tx power) for callers
 * that cannot allocate -- code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
 * No handle; the call allocates NOTHING (no cudaMalloc / cudaFree / allocating library routine): inputs and outputs
 * are the caller's DEVICE buffers, and the one piece of scratch the computation needs -- a count per (packet, block
 * of 256 nodes), scanned in place -- lives in a file-scope __device__ array of RMB_STATIC_MAX_CELLS ints (2 MiB per
 * GPU that loads the library, held until exit).
 *   limit      n_tx * ceil(n / 256) <= RMB_STATIC_MAX_CELLS (e.g. 100 000 nodes x 1 000 packets); beyond it the call
 *              returns RMB_ERR_CAPACITY and says so in error_text.  rmb_tick has no such limit.
 *   sharing    every call on a device uses the same array, which keeps the last call's contents; nothing in it has to
 *              start at zero (each call writes every cell it reads).  Calls that OVERLAP on one GPU (two streams, two
 *              threads) race on it: order them on one stream.
 *   inputs     device: node-array order as for rmb_set_nodes (d_z, d_enabled may be NULL = 0 / all on); n_tx packets
 *              by source node index (outside 0..n-1, e.g. -1: heard by nobody), power in dBm, channel (d_pkt_channel
 *              NULL = the source's channel).  stream: a cudaStream_t (NULL = the default stream); the call only
 *              enqueues (three kernels) and does not synchronise.
 *   outputs    rows in the reference's call order (packets in arrival order, receivers in node-array order), as
 *              rmb_tick returns them: d_tx (may be NULL), d_rx, d_rssi_dbm, d_outcome hold `capacity` rows each;
 *              d_tx_first[n_tx + 1] = first row of every packet; *d_n_pairs = rows the tick has -- when that exceeds
 *              capacity, rows beyond capacity were not written.
 *   set-up     the first call on a device looks the array's device address up (cudaGetSymbolAddress): make one call
 *              (n_tx = 0 will do) before starting a graph capture on that device.
 *   error_text (may be NULL) receives the reason for a non-zero return. */
#define RMB_STATIC_MAX_CELLS (1 << 19)
int rmb_udgm_tick_static(int32_t device, void *stream, int32_t n, const double *d_x, const double *d_y, const double *d_z,
                         const int32_t *d_channel, const uint8_t *d_enabled, int32_t n_tx, const int32_t *d_src,
                         const double *d_power_dbm, const int32_t *d_pkt_channel, double transmission_range,
                         int64_t capacity, int32_t *d_tx, int32_t *d_rx, float *d_rssi_dbm, uint8_t *d_outcome,
                         int32_t *d_tx_first, int64_t *d_n_pairs, char *error_text, size_t error_text_bytes);

#ifdef __cplusplus
}
#endif
#endif
