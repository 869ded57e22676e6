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

// ---- channel bucketing + fast-path constants ----------------------------------------------
// For every receiver channel group g (in work order) the tick's records on that channel are
// compacted, ARRIVAL ORDER PRESERVED (warp ballot + prefix sums), into bucket g of the sweep's
// tx tiles; buckets are padded to kTile with infinitely weak dummies.
//   c_t     = P_t^(-1/n),  P_t = 10^((Ptx - PL0)/10) * d0^n        (power folded into geometry)
//   theta_t = min(sensitivity (minus 0.05 dB), P_t * Dnear^-n)     (near threshold, mW)
//   Dnear   = max(d0, n * delta_d / 1e-4), delta_d = fp32 geometry error of the centred copy
// Three launches: init (bucket counters), per-transmission constants (one thread per tx, any
// number of blocks), stable placement (one block per group).
__global__ void prepare_init_kernel(const DevParams *dp, const Geom *geom, int ngroups, int *gcount, float *theta_min_g,
                                    double *pconst)
{
    const int g = threadIdx.x;
    if (g < ngroups) { gcount[g] = 0; theta_min_g[g] = FLT_MAX; }
    if (g == 0) {                          // per-tick constants of the per-transmission kernel
        const double n = dp->pl_exponent, d0 = dp->d0_m;
        const double delta_d = geom->maxabs * (geom->planar ? 1.7e-7 : 2.1e-7);
        const double dnear = fmax(d0, n * delta_d / 1e-4);
        pconst[0] = pow(d0, n);
        pconst[1] = pow(dnear, -n);
        pconst[2] = pow(10.0, (dp->sensitivity_dbm - 0.05) / 10.0);
    }
}

__global__ void __launch_bounds__(256)
prepare_tx_kernel(NodeArrays nd, const DevParams *dp, const Geom *geom, Groups groups, int n_tx,
                  const rmb_tx_record *rec, TxStage stg, int *gcount, float *theta_min_g, Counters *ctr)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long decisions = 0;
    if (t < n_tx) {
        const rmb_tx_record r = rec[t];
        int g = -1;
        if (r.src >= 0)
            for (int k = 0; k < groups.ngroups; ++k) if (groups.channel[k] == r.channel) { g = k; break; }
        stg.gid[t] = g;
        if (g >= 0) {
            const double n = dp->pl_exponent;
            const double d0n = stg.pconst[0], dnear_n = stg.pconst[1], theta_s = stg.pconst[2];
            const double P = pow(10.0, (r.power_dbm - dp->pl0_db) / 10.0) * d0n;
            const float cf = (float)pow(P, -1.0 / n);
            const double c = (double)cf;
            stg.f[t] = make_float4((float)((r.x - geom->cx) * c), (float)((r.y - geom->cy) * c),
                                   (float)((r.z - geom->cz) * c), cf);
            const float th = (float)(fmin(theta_s, P * dnear_n) * (1.0 - 3e-5));   // margin: fp32 / Newton rsqrt error of the sweep
            stg.theta[t] = th;
            atomicAdd(&gcount[g], 1);
            atomicMin((int *)&theta_min_g[g], __float_as_int(th));        // th > 0: integer order == float order
            // decisions of this packet: every enabled co-channel receiver except its own source
            const int s = r.src - dp->rx_base;
            const int self = (s >= 0 && s < nd.n && nd.enabled[s] && nd.channel[s] == r.channel) ? 1 : 0;
            decisions = (unsigned long long)(groups.n_rx[g] - self);
        }
    }
    for (int o = 16; o > 0; o >>= 1) decisions += __shfl_xor_sync(0xffffffffu, decisions, o);
    if ((threadIdx.x & 31) == 0 && decisions) atomicAdd(&ctr->decisions, decisions);
}

__global__ void __launch_bounds__(1024)
prepare_place_kernel(Groups groups, int n_tx, TxStage stg, TxTiles tx, float *theta)
{
    __shared__ int s_warp[32];
    __shared__ int s_total;
    const int g = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    int cursor = 0;
    for (int k = 0; k < g; ++k) cursor += (tx.gcount[k] + kTile - 1) / kTile * kTile;
    int count_g = 0;
    // every thread owns 4 consecutive transmissions of a 4096-chunk (arrival order = thread order)
    for (int c0 = 0; c0 < n_tx; c0 += 4096) {
        const int t0 = c0 + tid * 4;
        bool match[4];
        int mine = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) { match[k] = t0 + k < n_tx && stg.gid[t0 + k] == g; mine += match[k]; }
        int incl = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += y; }
        if (lane == 31) s_warp[wid] = incl;
        __syncthreads();
        if (wid == 0) {
            int v = s_warp[lane], wi = v;
            for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += y; }
            s_warp[lane] = wi - v;
            if (lane == 31) s_total = wi;
        }
        __syncthreads();
        int j = cursor + count_g + s_warp[wid] + incl - mine;
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (match[k]) { tx.f[j] = stg.f[t0 + k]; theta[j] = stg.theta[t0 + k]; tx.arrival[j] = t0 + k; ++j; }
        count_g += s_total;
        __syncthreads();
    }
    const int padded = (count_g + kTile - 1) / kTile * kTile;
    for (int j = cursor + count_g + tid; j < cursor + padded; j += 1024) {
        tx.f[j] = make_float4(1e18f, 1e18f, 1e18f, 0.f);      // d2' ~ 1e36 -> lin underflows to 0
        theta[j] = FLT_MAX;
        tx.arrival[j] = -1;
    }
    if (tid == 0) {
        tx.goff[g] = cursor;
        if (g == groups.ngroups - 1) tx.goff[groups.ngroups] = cursor + padded;
    }
}

cudaError_t launch_prepare(const NodeArrays &nd, const DevParams *dp, const Geom *geom, Groups groups, int n_tx,
                           const rmb_tx_record *rec, TxStage stg, TxTiles tx, float *theta, float *theta_min_g,
                           Counters *ctr, cudaStream_t st)
{
    prepare_init_kernel<<<1, kMaxGroups, 0, st>>>(dp, geom, groups.ngroups, tx.gcount, theta_min_g, stg.pconst);
    if (groups.ngroups <= 0) return cudaMemsetAsync(tx.goff, 0, sizeof(int), st);
    if (n_tx > 0)
        prepare_tx_kernel<<<(n_tx + 255) / 256, 256, 0, st>>>(nd, dp, geom, groups, n_tx, rec, stg, tx.gcount, theta_min_g, ctr);
    prepare_place_kernel<<<groups.ngroups, 1024, 0, st>>>(groups, n_tx, stg, tx, theta);
    return cudaGetLastError();
}

}  // namespace rmb
