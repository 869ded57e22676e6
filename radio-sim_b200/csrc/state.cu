// state.cu -- K4 (per-tick node/transmission state in HBM) and the tx-side preparation.
//
// Replaces the AoS objects Node/Position/Transciever/RadioPacket (Node.java:39-59,
// Position.java:37-52, Transciever.java:5-113, RadioPacket.java:46-52) by SoA device
// buffers: an exact double master copy in node-array order plus a centred fp32 copy in
// "work order" (receivers grouped by channel) for the sweep.
#include "common.cuh"
#include "kernels.h"
#include "exact_math.cuh"
#include <cfloat>
#include <climits>
#include <cub/cub.cuh>

namespace rmb {

// ---- bounding box ------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long enc(double d)
{
    unsigned long long u = (unsigned long long)__double_as_longlong(d);
    return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}
__device__ __forceinline__ double dec(unsigned long long u)
{
    u = (u >> 63) ? (u & 0x7fffffffffffffffull) : ~u;
    return __longlong_as_double((long long)u);
}

__global__ void geom_init_kernel(unsigned long long *s)
{
    if (threadIdx.x < 3) s[threadIdx.x] = ~0ull;          // minima
    else if (threadIdx.x < 6) s[threadIdx.x] = 0ull;      // maxima
}

__global__ void geom_reduce_kernel(NodeArrays nd, unsigned long long *s)
{
    double mn[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, mx[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nd.n; i += gridDim.x * blockDim.x) {
        double v[3] = {nd.x[i], nd.y[i], nd.z[i]};
#pragma unroll
        for (int k = 0; k < 3; ++k) { mn[k] = fmin(mn[k], v[k]); mx[k] = fmax(mx[k], v[k]); }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k)
        for (int o = 16; o > 0; o >>= 1) {
            mn[k] = fmin(mn[k], __shfl_xor_sync(0xffffffffu, mn[k], o));
            mx[k] = fmax(mx[k], __shfl_xor_sync(0xffffffffu, mx[k], o));
        }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) { atomicMin(&s[k], enc(mn[k])); atomicMax(&s[3 + k], enc(mx[k])); }
    }
}

__global__ void geom_final_kernel(int n, const unsigned long long *s, Geom *g)
{
    if (n <= 0) { *g = Geom{0, 0, 0, 0, 0, 0, 1, 0}; return; }
    double mn[3], mx[3];
    for (int k = 0; k < 3; ++k) { mn[k] = dec(s[k]); mx[k] = dec(s[3 + k]); }
    g->cx = 0.5 * (mn[0] + mx[0]); g->cy = 0.5 * (mn[1] + mx[1]); g->cz = 0.5 * (mn[2] + mx[2]);
    double m = fmax(fmax(mx[0] - g->cx, mx[1] - g->cy), mx[2] - g->cz);
    g->maxabs = m;
    g->zmin = mn[2]; g->zmax = mx[2];
    g->planar = (mn[2] == mx[2]) ? 1 : 0;
}

cudaError_t launch_geom(const NodeArrays &nd, unsigned long long *scratch6, Geom *geom, cudaStream_t st)
{
    geom_init_kernel<<<1, 32, 0, st>>>(scratch6);
    if (nd.n > 0) {
        int blocks = min(148 * 4, (nd.n + 255) / 256);
        geom_reduce_kernel<<<blocks, 256, 0, st>>>(nd, scratch6);
    }
    geom_final_kernel<<<1, 1, 0, st>>>(nd.n, scratch6, geom);
    return cudaGetLastError();
}

// ---- work order: channel group, then Morton (Z-curve) order of (x, y) ----------------------
// Spatially close receivers share a warp, so the rare "near" re-scans of the sweep hit the same
// tx tile for all lanes of a warp instead of 32 different ones.  The order only affects speed:
// every result list is re-ordered to (packet, node index) by the emit stage.
__device__ __forceinline__ unsigned spread16(unsigned v)
{
    v &= 0xffffu;
    v = (v | (v << 8)) & 0x00ff00ffu; v = (v | (v << 4)) & 0x0f0f0f0fu;
    v = (v | (v << 2)) & 0x33333333u; v = (v | (v << 1)) & 0x55555555u;
    return v;
}

__global__ void morton_keys_kernel(NodeArrays nd, const int *gid, const Geom *geom, unsigned long long *key, int *val)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nd.n) return;
    const int g = gid[i];
    unsigned long long k = ~0ull;                       // disabled nodes sort to the end
    if (g >= 0) {
        const double ext = geom->maxabs > 0.0 ? 2.0 * geom->maxabs : 1.0;
        const double ux = (nd.x[i] - (geom->cx - geom->maxabs)) / ext, uy = (nd.y[i] - (geom->cy - geom->maxabs)) / ext;
        const unsigned qx = (unsigned)fmin(fmax(ux * 65535.0, 0.0), 65535.0), qy = (unsigned)fmin(fmax(uy * 65535.0, 0.0), 65535.0);
        k = ((unsigned long long)g << 32) | (spread16(qx) | (spread16(qy) << 1));
    }
    key[i] = k; val[i] = i;
}

__global__ void place_perm_kernel(int n_enabled, const unsigned long long *key, const int *val, Groups groups, int *perm)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_enabled) return;
    const int g = (int)(key[i] >> 32);
    perm[groups.w_begin[g] + (i - groups.dense_begin[g])] = val[i];
}

size_t reorder_tmp_bytes(int n)
{
    size_t bytes = 0;
    cub::DoubleBuffer<unsigned long long> k(nullptr, nullptr);
    cub::DoubleBuffer<int> v(nullptr, nullptr);
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, k, v, n, 0, 40, 0);
    return bytes;
}

cudaError_t launch_reorder(const NodeArrays &nd, const int *gid, const Geom *geom, const Groups &groups, int n_work,
                           ReorderScratch sc, int *perm, cudaStream_t st)
{
    RMB_CUDA_OK(cudaMemsetAsync(perm, 0xff, (size_t)(n_work > 0 ? n_work : 1) * sizeof(int), st));    // -1 = padding
    if (nd.n <= 0) return cudaSuccess;
    morton_keys_kernel<<<(nd.n + 255) / 256, 256, 0, st>>>(nd, gid, geom, sc.key, sc.val);
    cub::DoubleBuffer<unsigned long long> k(sc.key, sc.key_alt);
    cub::DoubleBuffer<int> v(sc.val, sc.val_alt);
    size_t bytes = sc.tmp_bytes;
    RMB_CUDA_OK(cub::DeviceRadixSort::SortPairs(sc.tmp, bytes, k, v, nd.n, 0, 40, st));     // 6 bits group + 32 Morton; all-ones = disabled
    const int n_enabled = groups.dense_begin[groups.ngroups];
    if (n_enabled > 0)
        place_perm_kernel<<<(n_enabled + 255) / 256, 256, 0, st>>>(n_enabled, k.Current(), v.Current(), groups, perm);
    return cudaGetLastError();
}

// ---- fp32 work-order copy ----------------------------------------------------------------
__global__ void regather_kernel(NodeArrays nd, WorkArrays wk, const Geom *geom)
{
    int w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= wk.n_work) return;
    int p = wk.perm[w];
    if (p < 0) {                                                                    // padding: infinitely far
        wk.fx[w] = 1e15f; wk.fy[w] = 1e15f; wk.fz[w] = 1e15f;   // (1e15 c)^2 stays finite in fp32 for any sane c
        wk.x64[w] = 0.0; wk.y64[w] = 0.0; wk.z64[w] = 0.0; wk.rxp[w] = 0.0;
        return;
    }
    const double X = nd.x[p], Y = nd.y[p], Z = nd.z[p];
    wk.fx[w] = (float)(X - geom->cx);
    wk.fy[w] = (float)(Y - geom->cy);
    wk.fz[w] = (float)(Z - geom->cz);
    wk.x64[w] = X; wk.y64[w] = Y; wk.z64[w] = Z; wk.rxp[w] = nd.rxprob[p];
}

cudaError_t launch_regather(const NodeArrays &nd, const WorkArrays &wk, const Geom *geom, cudaStream_t st)
{
    if (wk.n_work <= 0) return cudaSuccess;
    regather_kernel<<<(wk.n_work + 255) / 256, 256, 0, st>>>(nd, wk, geom);
    return cudaGetLastError();
}

// node-config-set "position" for a sparse set of nodes (net/SimulatorJSONHandler.java:111-122)
__global__ void scatter_positions_kernel(NodeArrays nd, int count, const int *idx, const double *x, const double *y,
                                         const double *z)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    int p = idx[i];
    if (p < 0 || p >= nd.n) return;
    nd.x[p] = x[i]; nd.y[p] = y[i]; nd.z[p] = z ? z[i] : 0.0;
}

cudaError_t launch_scatter_positions(const NodeArrays &nd, int count, const int *idx, const double *x, const double *y,
                                     const double *z, cudaStream_t st)
{
    if (count <= 0) return cudaSuccess;
    scatter_positions_kernel<<<(count + 255) / 256, 256, 0, st>>>(nd, count, idx, x, y, z);
    return cudaGetLastError();
}

__global__ void init_params_kernel(DevParams *dp)
{
    dp->n0_mw = xm::dbm_to_mw(dp->noise_floor_dbm);
    dp->sinr_guard_db = 1.5e-3;
}

cudaError_t launch_init_params(DevParams *dp, cudaStream_t st)
{
    init_params_kernel<<<1, 1, 0, st>>>(dp);
    return cudaGetLastError();
}

// ---- RadioPacket snapshot (RadioPacket.java:46-52): source position, power, channel ------
__global__ void gather_records_kernel(NodeArrays nd, int rx_base, int n_tx, const int *src, const double *power,
                                      const int *channel, rmb_tx_record *rec)
{
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tx) return;
    int s = src[t] - rx_base;
    rmb_tx_record r;
    if (s < 0 || s >= nd.n) {
        r.x = r.y = r.z = 0.0; r.power_dbm = 0.0; r.txprob = 1.0; r.channel = 0; r.src = -1;
    } else {
        r.x = nd.x[s]; r.y = nd.y[s]; r.z = nd.z[s];
        double pw = power ? power[t] : nan("");
        r.power_dbm = isnan(pw) ? nd.txpower[s] : pw;
        int ch = channel ? channel[t] : INT_MIN;
        r.channel = (ch == INT_MIN) ? nd.channel[s] : ch;
        r.txprob = nd.txprob[s];
        r.src = src[t];
    }
    rec[t] = r;
}

cudaError_t launch_gather_records(const NodeArrays &nd, int rx_base, int n_tx, const int *src, const double *power,
                                  const int *channel, rmb_tx_record *rec, cudaStream_t st)
{
    if (n_tx <= 0) return cudaSuccess;
    gather_records_kernel<<<(n_tx + 255) / 256, 256, 0, st>>>(nd, rx_base, n_tx, src, power, channel, rec);
    return cudaGetLastError();
}

// ---- multi-GPU exchange of the tick's records (replaces an all-gather collective) -------------
// Every thread moves one 48-byte record as three 16-byte stores per peer, straight into the peers'
// HBM over NVLink.  When the last block has arrived, one thread raises this rank's flag in every
// peer's buffer (release, system scope): the consumers' preparation kernels spin on their LOCAL copy.
__global__ void __launch_bounds__(256) peer_push_kernel(const rmb_tx_record *local, int slots, PeerPush pp)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < slots) {
        const uint4 *s = reinterpret_cast<const uint4 *>(local + t);
        const uint4 a = s[0], b = s[1], c = s[2];
        for (int p = 0; p < pp.world; ++p) {
            uint4 *d = reinterpret_cast<uint4 *>(pp.peer[p] + pp.rec_off) + ((size_t)pp.rank * slots + t) * 3;
            d[0] = a; d[1] = b; d[2] = c;
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        if (atomicAdd(pp.done, 1u) == gridDim.x - 1) {
            *pp.done = 0u;
            __threadfence_system();
            for (int p = 0; p < pp.world; ++p) st_relaxed_sys(reinterpret_cast<unsigned *>(pp.peer[p]) + pp.rank, pp.epoch);
        }
    }
}

// Device-side rendezvous of all ranks' streams (rmb_peer_barrier_enqueue): every rank raises its slot of the SECOND
// flag row (bytes 64..127 of a gather buffer) in every peer's buffer to `epoch`, then waits until all its own slots
// show it.  The ranks leave within an NVLink round trip of each other -- a tighter line-up than a collective gives.
__global__ void peer_barrier_kernel(PeerPush pp, Counters *ctr)
{
    const int p = threadIdx.x;
    if (p < pp.world) st_release_sys(reinterpret_cast<unsigned *>(pp.peer[p] + 64) + pp.rank, pp.epoch);
    __syncthreads();
    if (p < pp.world && !peer_wait(reinterpret_cast<const unsigned *>(pp.peer[pp.rank] + 64) + p, pp.epoch)) ctr->peer_timeout = 1;
}

// the reference media's tick has no preparation kernel to wait for the peers' records in: a one-warp kernel does it
__global__ void peer_wait_kernel(const unsigned *flags, int world, unsigned epoch, Counters *ctr)
{
    if ((int)threadIdx.x < world && !peer_wait(flags + threadIdx.x, epoch)) ctr->peer_timeout = 1;
}

cudaError_t launch_peer_wait(const unsigned *flags, int world, unsigned epoch, Counters *ctr, cudaStream_t st)
{
    peer_wait_kernel<<<1, 32, 0, st>>>(flags, world, epoch, ctr);
    return cudaGetLastError();
}

cudaError_t launch_peer_barrier(const PeerPush &pp, Counters *ctr, cudaStream_t st)
{
    peer_barrier_kernel<<<1, 32, 0, st>>>(pp, ctr);
    return cudaGetLastError();
}

cudaError_t launch_peer_push(const rmb_tx_record *local, int slots, const PeerPush &pp, cudaStream_t st)
{
    peer_push_kernel<<<max(1, (slots + 255) / 256), 256, 0, st>>>(local, slots, pp);
    return cudaGetLastError();
}

// ---- channel bucketing + fast-path constants ----------------------------------------------
// For every receiver channel group g (in work order) the tick's records on that channel are
// compacted, ARRIVAL ORDER PRESERVED (warp match/ballot + prefix sums), into bucket g of the sweep's
// tx tiles; buckets are padded to kTile with infinitely weak dummies.
//   c_t     = P_t^(-1/n),  P_t = 10^((Ptx - PL0)/10) * d0^n        (power folded into geometry)
//   theta_t = min(sensitivity (minus 0.05 dB), P_t * Dnear^-n)     (near threshold, mW)
//   Dnear   = max(d0, n * delta_d / 1e-4), delta_d = fp32 geometry error of the centred copy
// ONE cooperative launch does the whole tx side of a tick (it used to be three kernels and three
// memsets): wait for the peers' records (multi-GPU exchange), zero the tick's counters, per-
// transmission constants + the deterministic tx-failure rule, grid barrier, stable placement.
// Block b owns the contiguous arrival range [b*per, (b+1)*per), so "blocks in order, threads in
// order" is arrival order.
__device__ __forceinline__ double tx_success_p(const DevParams *dp, const rmb_tx_record &r)
{
    if (dp->model == RMB_MODEL_UDGM) return dp->success_ratio_rx * r.txprob;   // UDGMRadioMedium.java:63-65 (sic: Rx)
    if (dp->model == RMB_MODEL_N2N || dp->model == RMB_MODEL_SINR) return r.txprob;   // N2NRadioMedium.java:24-26
    return 1.0;
}

__global__ void __launch_bounds__(kPrepThreads)
prepare_all_kernel(NodeArrays nd, const DevParams *dp, const Geom *geom, Groups groups, int n_tx,
                   const rmb_tx_record *rec, TxStage stg, TxTiles tx, float *theta, float *theta_min_g,
                   Counters *ctr, PrepareAux aux)
{
    constexpr int kWarps = kPrepThreads / 32;
    __shared__ double s_const[3];
    __shared__ int s_cnt[kMaxGroups], s_base[kMaxGroups], s_goff[kMaxGroups + 1], s_thmin[kMaxGroups];
    __shared__ int s_wcnt[kWarps][kMaxGroups];
    __shared__ unsigned long long s_dec;
    __shared__ int s_timeout, s_needs;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int ng = groups.ngroups;
    // the sweep behind this kernel is launched with programmatic stream serialization: its blocks may become resident
    // (and fetch their receivers) while this kernel still runs; they wait for its completion before they read a tile
    asm volatile("griddepcontrol.launch_dependents;");

    // fused exchange: this rank's records go out first (16-byte stores over NVLink, a record per thread) ...
    if (aux.push_local) {
        for (int t = blockIdx.x * kPrepThreads + tid; t < aux.push_slots; t += gridDim.x * kPrepThreads) {
            const uint4 *s = reinterpret_cast<const uint4 *>(aux.push_local + t);
            const uint4 a = s[0], b = s[1], c = s[2];
            for (int p = 0; p < aux.world; ++p) {
                uint4 *d = reinterpret_cast<uint4 *>(aux.push_peer[p] + aux.push_off) + ((size_t)aux.push_rank * aux.push_slots + t) * 3;
                d[0] = a; d[1] = b; d[2] = c;
            }
        }
        __syncthreads();
        if (tid == 0) {
            __threadfence_system();
            if (atomicAdd(aux.push_done, 1u) == gridDim.x - 1) {
                *aux.push_done = 0u;
                __threadfence_system();
                for (int p = 0; p < aux.world; ++p) st_relaxed_sys(reinterpret_cast<unsigned *>(aux.push_peer[p]) + aux.push_rank, aux.epoch);
            }
        }
    }
    if (tid == 0) { s_timeout = 0; s_needs = 0; }
    __syncthreads();

    if (tid < 3) {
        const double n = dp->pl_exponent, d0 = dp->d0_m;
        const double delta_d = geom->maxabs * (geom->planar ? 1.7e-7 : 2.1e-7);
        const double dnear = fmax(d0, n * delta_d / 1e-4);
        s_const[tid] = tid == 0 ? pow(d0, n) : tid == 1 ? pow(dnear, -n) : pow(10.0, (dp->sensitivity_dbm - 0.05) / 10.0);
    }
    if (tid < kMaxGroups) { s_cnt[tid] = 0; s_thmin[tid] = __float_as_int(FLT_MAX); }
    if (tid == 0) s_dec = 0ull;
    // counters of the tick (they used to be cudaMemsetAsync nodes on the critical path)
    {
        const int gtid = blockIdx.x * kPrepThreads + tid, gsz = gridDim.x * kPrepThreads;
        int4 *c4 = reinterpret_cast<int4 *>(aux.cnt);                    // n_work is a multiple of 1024
        for (int i = gtid; i < aux.n_work / 4; i += gsz) c4[i] = make_int4(0, 0, 0, 0);
        for (int i = gtid; i <= n_tx; i += gsz) aux.tx_cnt[i] = 0;
        for (int i = gtid; i < aux.n_rx_done; i += gsz) aux.rx_done[i] = 0;
        if (gtid == 0) { Counters z{}; z.t_prepare_begin = global_timer_ns(); *ctr = z; aux.bar[2] = 0u; }      // bar[2]: the barrier counter of finish.cu, zero at its launch
    }
    // ... and the peers' records of this tick must have landed before anybody reads `rec` (everything above is
    // independent of them and runs while the records are on their way)
    if (aux.peer_flags && tid < aux.world && !peer_wait(aux.peer_flags + tid, aux.epoch)) s_timeout = 1;
    __syncthreads();

    // 1. per-transmission constants of this block's arrival range
    const int per = (n_tx + (int)gridDim.x - 1) / (int)gridDim.x;
    const int t_begin = min(n_tx, (int)blockIdx.x * per), t_end = min(n_tx, t_begin + per);
    unsigned long long decisions = 0;
    for (int c0 = t_begin; c0 < t_end; c0 += kPrepThreads) {
        const int t = c0 + tid;
        if (t < t_end) {
            const rmb_tx_record r = rec[t];
            int g = -1;
            if (r.src >= 0)
                for (int k = 0; k < ng; ++k) if (groups.channel[k] == r.channel) { g = k; break; }
            stg.gid[t] = g;
            aux.tx_failed[t] = (r.src >= 0 && tx_success_p(dp, r) <= 0.0) ? 1 : 0;   // no draws: the tx rule is p_tx <= 0
            aux.tx_base[t] = 0;
            if (g >= 0) {
                const double n = dp->pl_exponent;
                const double d0n = s_const[0], dnear_n = s_const[1], theta_s = s_const[2];
                // P^(-1/n) = 10^(-e/n) / d0: two exp10 instead of a pow(10, .) and a general pow (half the latency chain)
                const double e10 = (r.power_dbm - dp->pl0_db) / 10.0;
                const double P = exp10(e10) * d0n;
                const float cf = (float)(exp10(-e10 / n) / dp->d0_m);
                const double c = (double)cf;
                stg.f[t] = make_float4((float)((r.x - geom->cx) * c), (float)((r.y - geom->cy) * c),
                                       (float)((r.z - geom->cz) * c), cf);
                const float th = (float)(fmin(theta_s, P * dnear_n) * (1.0 - 3e-5));   // margin: fp32 / Newton rsqrt error of the sweep
                stg.theta[t] = th;
                atomicAdd(&s_cnt[g], 1);
                atomicMin(&s_thmin[g], __float_as_int(th));           // th > 0: integer order == float order
                // decisions of this packet: every enabled co-channel receiver except its own source
                const int s = r.src - dp->rx_base;
                const int self = (s >= 0 && s < nd.n && nd.enabled[s] && nd.channel[s] == r.channel) ? 1 : 0;
                if (!aux.tx_epoch) decisions += (unsigned long long)(groups.n_rx[g] - self);
            }
            // a record with a tx probability strictly inside (0, 1) needs a draw: the deterministic path must not
            // evaluate it (the host turns this into an error; a shard cannot see it in its own node table)
            if (r.src >= 0 && !aux.stochastic) { const double pt = tx_success_p(dp, r); if (pt > 0.0 && pt < 1.0) s_needs = 1; }
            // half duplex: stamp the local sources of this tick and count them per receiver group (each node once)
            if (aux.tx_epoch && r.src >= 0) {
                const int s = r.src - dp->rx_base;
                if (s >= 0 && s < nd.n && atomicExch(&aux.tx_epoch[s], aux.hd_epoch) != aux.hd_epoch) {
                    const int gg = aux.node_gid[s];
                    if (gg >= 0) atomicAdd(&aux.hd_cnt[(aux.hd_epoch & 1) * kMaxGroups + gg], 1);
                }
            }
        }
    }
    if (aux.tx_epoch && blockIdx.x == 0 && tid < kMaxGroups) aux.hd_cnt[((aux.hd_epoch + 1) & 1) * kMaxGroups + tid] = 0;   // the next tick's
    for (int o = 16; o > 0; o >>= 1) decisions += __shfl_xor_sync(0xffffffffu, decisions, o);
    if (lane == 0 && decisions) atomicAdd(&s_dec, decisions);
    __syncthreads();
    if (tid < ng) {
        aux.blk_cnt[blockIdx.x * kMaxGroups + tid] = s_cnt[tid];
        aux.blk_thmin[blockIdx.x * kMaxGroups + tid] = s_thmin[tid];
    }
    grid_barrier(aux.bar, aux.bar_target);

    // 2. bucket offsets: groups in order, blocks in order inside a group
    for (int g = wid; g < ng; g += kWarps) {
        int before = 0, total = 0, thm = __float_as_int(FLT_MAX);
        for (int b = lane; b < (int)gridDim.x; b += 32) {
            const int v = __ldcg(&aux.blk_cnt[b * kMaxGroups + g]);
            total += v;
            if (b < (int)blockIdx.x) before += v;
            if (blockIdx.x == 0) thm = min(thm, __ldcg(&aux.blk_thmin[b * kMaxGroups + g]));
        }
        for (int o = 16; o > 0; o >>= 1) {
            before += __shfl_xor_sync(0xffffffffu, before, o);
            total += __shfl_xor_sync(0xffffffffu, total, o);
            thm = min(thm, __shfl_xor_sync(0xffffffffu, thm, o));
        }
        if (lane == 0) {
            s_cnt[g] = total; s_base[g] = before;
            if (blockIdx.x == 0) { theta_min_g[g] = __int_as_float(thm); tx.gcount[g] = total; }
        }
    }
    if (aux.tx_epoch) {
        // half duplex: a packet is evaluated by the enabled co-channel receivers that are not on the air themselves
        // (its own source is one of the transmitting nodes, if it sits in the group at all)
        unsigned long long dec = 0;
        for (int t = t_begin + tid; t < t_end; t += kPrepThreads) {
            const int g = stg.gid[t];
            if (g >= 0) dec += (unsigned long long)(groups.n_rx[g] - __ldcg(&aux.hd_cnt[(aux.hd_epoch & 1) * kMaxGroups + g]));
        }
        for (int o = 16; o > 0; o >>= 1) dec += __shfl_xor_sync(0xffffffffu, dec, o);
        if (lane == 0 && dec) atomicAdd(&s_dec, dec);
    }
    __syncthreads();
    if (tid == 0) {
        int cursor = 0;
        for (int g = 0; g < ng; ++g) { s_goff[g] = cursor; s_base[g] += cursor; cursor += (s_cnt[g] + kTile - 1) / kTile * kTile; }
        s_goff[ng] = cursor;
        if (s_dec) atomicAdd(&ctr->decisions, s_dec);
        if (s_timeout) ctr->peer_timeout = 1;
        if (s_needs) ctr->needs_draws = 1;
    }
    __syncthreads();
    if (blockIdx.x == 0) {
        for (int g = tid; g <= ng; g += kPrepThreads) tx.goff[g] = s_goff[g];
        for (int i = tid; i < ng * kTile; i += kPrepThreads) {           // pad every bucket to a whole tile
            const int g = i / kTile, j = s_goff[g] + s_cnt[g] + i % kTile;
            if (j < s_goff[g + 1]) {
                tx.c[j] = 0.f; tx.nx[j] = tx.ny[j] = tx.nz[j] = -1e18f;     // d2' ~ 1e36 -> lin underflows to 0
                theta[j] = FLT_MAX;
                tx.arrival[j] = -1;
            }
        }
    }

    // 3. stable placement of this block's range
    for (int c0 = t_begin; c0 < t_end; c0 += kPrepThreads) {
        const int t = c0 + tid;
        const int g = t < t_end ? stg.gid[t] : -1;
        for (int i = tid; i < kWarps * kMaxGroups; i += kPrepThreads) (&s_wcnt[0][0])[i] = 0;
        __syncthreads();
        const unsigned m = __match_any_sync(0xffffffffu, g);
        const int r = __popc(m & ((1u << lane) - 1));
        if (g >= 0 && r == 0) s_wcnt[wid][g] = __popc(m);
        __syncthreads();
        if (tid < ng) {
            int acc = s_base[tid];
            for (int w = 0; w < kWarps; ++w) { const int v = s_wcnt[w][tid]; s_wcnt[w][tid] = acc; acc += v; }
            s_base[tid] = acc;
        }
        __syncthreads();
        if (g >= 0) {
            const int j = s_wcnt[wid][g] + r;
            const float4 f = stg.f[t];
            tx.c[j] = f.w; tx.nx[j] = -f.x; tx.ny[j] = -f.y; tx.nz[j] = -f.z;
            theta[j] = stg.theta[t]; tx.arrival[j] = t;
        }
        __syncthreads();
    }
}

int prepare_grid(int n_tx)
{
    // every SM takes one block: the tick's counters are zeroed by the whole grid, the per-transmission
    // double-precision pow() chains are latency bound and spread best thin
    (void)n_tx;
    const int sms = device_sms();
    return sms < kPrepMaxGrid ? sms : kPrepMaxGrid;
}

cudaError_t launch_prepare(const NodeArrays &nd, const DevParams *dp, const Geom *geom, Groups groups, int n_tx,
                           const rmb_tx_record *rec, TxStage stg, TxTiles tx, float *theta, float *theta_min_g,
                           Counters *ctr, PrepareAux aux, int grid, cudaStream_t st)
{
    NodeArrays a_nd = nd; Groups a_groups = groups;
    void *args[] = {(void *)&a_nd, (void *)&dp, (void *)&geom, (void *)&a_groups, (void *)&n_tx, (void *)&rec, (void *)&stg,
                    (void *)&tx, (void *)&theta, (void *)&theta_min_g, (void *)&ctr, (void *)&aux};
    return launch_grid_synced((const void *)&prepare_all_kernel, dim3((unsigned)grid), dim3(kPrepThreads), args, st);
}

}  // namespace rmb
