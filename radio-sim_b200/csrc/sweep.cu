// sweep.cu -- K1, the link-budget / interference sweep (the hot kernel).
//
// For every enabled receiver (work order: grouped by channel) and every co-channel
// transmission of the tick it evaluates the log-distance link budget
//     lin(t,r) = P_t * d(t,r)^-n   [mW]              
// and accumulates the co-channel interference sum in the linear domain (N2).  Replaces the
// O(N) receiver loop of RadioMedium.transmit (UDGMRadioMedium.java:99-114) for all the
// tick's packets at once.
//
// Formulation (DESIGN.md "K1"): the tx power is folded into the geometry.  With
// c_t = P_t^(-1/n) the kernel works on scaled coordinates x' = c_t*x, so that
//     d2' = c_t^2 * d^2   and   lin = (d2')^(-n/2)
// i.e. per pair 2 FFMA (scaled differences) + FMUL + FFMA (d2') + ONE MUFU (rsqrt / rcp;
// lg2+ex2 for a general exponent) + 1-2 FP32 ops to accumulate.  The MUFU (XU) pipe,
// 16 lanes/SM/clk, is the binding unit.
//
// "Near" pairs -- those that might be heard (lin >= theta_t) or whose fp32 geometry is
// not accurate enough (d <= D_near) -- are NOT summed here: they are detected per tile of
// kTile transmissions through the tile's partial sum, re-scanned, and appended to the
// receiver's candidate slots for the exact double-precision resolve kernel.  So the fp32
// sum only ever holds weak far-field terms, and every decision is taken in double.
#include "common.cuh"
#include "kernels.h"

namespace rmb {

enum { MODE_N2 = 0, MODE_N3 = 1, MODE_N4 = 2, MODE_GEN = 3 };

__device__ __forceinline__ float mufu_rsq(float x) { float y; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_rcp(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_lg2(float x) { float y; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

// scaled squared distance; t = {x', y', z', c}
template <bool PLANAR>
__device__ __forceinline__ float scaled_d2(float rx, float ry, float rz, const float4 &t)
{
    float dx = __fmaf_rn(rx, t.w, -t.x);
    float dy = __fmaf_rn(ry, t.w, -t.y);
    float d2 = __fmaf_rn(dy, dy, __fmul_rn(dx, dx));
    if (!PLANAR) {
        float dz = __fmaf_rn(rz, t.w, -t.z);
        d2 = __fmaf_rn(dz, dz, d2);
    }
    return d2;
}

// acc + (d2)^(-n/2)
template <int MODE>
__device__ __forceinline__ float accumulate(float d2, float acc, float neg_h)
{
    if (MODE == MODE_N3) { float q = mufu_rsq(d2); return __fmaf_rn(__fmul_rn(q, q), q, acc); }
    if (MODE == MODE_N2) { return __fadd_rn(acc, mufu_rcp(d2)); }
    if (MODE == MODE_N4) { float q = mufu_rcp(d2); return __fmaf_rn(q, q, acc); }
    return __fadd_rn(acc, mufu_ex2(__fmul_rn(neg_h, mufu_lg2(d2))));
}

template <int MODE, bool PLANAR, int R, int THREADS>
__global__ void __launch_bounds__(THREADS)
sweep_kernel(const float *__restrict__ fx, const float *__restrict__ fy, const float *__restrict__ fz,
             Groups groups, TxTiles tx, const float *__restrict__ theta, const float *__restrict__ theta_min_g,
             float neg_h, float *__restrict__ far_out, int *__restrict__ cnt_out, int *__restrict__ slots,
             int K, Counters *ctr)
{
    __shared__ float4 s_tx[kStageTx];
    __shared__ float  s_theta[kStageTx];

    const int w0 = (blockIdx.x * THREADS + threadIdx.x) * R;
    // the block's receivers all belong to one channel group (groups are padded to 1024)
    const int wb = blockIdx.x * THREADS * R;
    int g = 0;
    while (g + 1 < groups.ngroups && wb >= groups.w_begin[g + 1]) ++g;

    float rx[R], ry[R], rz[R], sum[R], ps[R];
    int cnt[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        rx[r] = fx[w0 + r]; ry[r] = fy[w0 + r]; rz[r] = PLANAR ? 0.f : fz[w0 + r];
        sum[r] = 0.f; ps[r] = 0.f; cnt[r] = 0;
    }
    const int j_begin = tx.goff[g], j_end = tx.goff[g + 1];
    const float th_min = theta_min_g[g];

    for (int s0 = j_begin; s0 < j_end; s0 += kStageTx) {
        const int ns = min(kStageTx, j_end - s0);
        __syncthreads();
        for (int i = threadIdx.x; i < ns; i += THREADS) { s_tx[i] = tx.f[s0 + i]; s_theta[i] = theta[s0 + i]; }
        __syncthreads();
        for (int j = 0; j < ns; j += kTile) {
#pragma unroll
            for (int u = 0; u < kTile; ++u) {
                const float4 t = s_tx[j + u];
#pragma unroll
                for (int r = 0; r < R; ++r)
                    ps[r] = accumulate<MODE>(scaled_d2<PLANAR>(rx[r], ry[r], rz[r], t), ps[r], neg_h);
            }
            float m = ps[0];
#pragma unroll
            for (int r = 1; r < R; ++r) m = fmaxf(m, ps[r]);
            if (!(m < th_min)) {
                // rare: some pair of this tile may be near -- re-scan the affected receivers
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    if (!(ps[r] < th_min)) {
                        float far = 0.f;
                        for (int u = 0; u < kTile; ++u) {
                            const float4 t = s_tx[j + u];
                            const float lin = accumulate<MODE>(scaled_d2<PLANAR>(rx[r], ry[r], rz[r], t), 0.f, neg_h);
                            if (!(lin < s_theta[j + u])) {
                                if (cnt[r] < K) slots[(size_t)(w0 + r) * K + cnt[r]] = s0 + j + u;
                                ++cnt[r];
                            } else {
                                far = __fadd_rn(far, lin);
                            }
                        }
                        ps[r] = far;
                    }
                }
            }
#pragma unroll
            for (int r = 0; r < R; ++r) { sum[r] = __fadd_rn(sum[r], ps[r]); ps[r] = 0.f; }
        }
    }
    int over = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        far_out[w0 + r] = sum[r];
        cnt_out[w0 + r] = cnt[r];
        over = max(over, cnt[r]);
    }
    if (over > K) atomicMax(&ctr->max_near, over);
}

template <int MODE, bool PLANAR>
static cudaError_t launch_variant(int variant, int n_work, const SweepArgs &a, cudaStream_t st)
{
#define RMB_LAUNCH(RR, TT)                                                                              \
    do {                                                                                                \
        int blocks = n_work / ((RR) * (TT));                                                            \
        sweep_kernel<MODE, PLANAR, RR, TT><<<blocks, TT, 0, st>>>(a.fx, a.fy, a.fz, a.groups, a.tx,     \
            a.theta, a.theta_min_g, a.neg_h, a.far_out, a.cnt_out, a.slots, a.K, a.ctr);                \
    } while (0)
    switch (variant) {
    case 2:  RMB_LAUNCH(8, 128); break;
    case 3:  RMB_LAUNCH(2, 64);  break;
    case 4:  RMB_LAUNCH(4, 64);  break;
    case 5:  RMB_LAUNCH(8, 64);  break;
    case 6:  RMB_LAUNCH(1, 32);  break;
    default: RMB_LAUNCH(4, 128); break;
    }
#undef RMB_LAUNCH
    return cudaGetLastError();
}

int sweep_mode_for_exponent(double n)
{
    if (n == 2.0) return MODE_N2;
    if (n == 3.0) return MODE_N3;
    if (n == 4.0) return MODE_N4;
    return MODE_GEN;
}

cudaError_t launch_sweep(int mode, bool planar, int variant, int n_work, const SweepArgs &a, cudaStream_t st)
{
    if (n_work <= 0) return cudaSuccess;
    switch (mode) {
    case MODE_N2:  return planar ? launch_variant<MODE_N2, true>(variant, n_work, a, st)  : launch_variant<MODE_N2, false>(variant, n_work, a, st);
    case MODE_N3:  return planar ? launch_variant<MODE_N3, true>(variant, n_work, a, st)  : launch_variant<MODE_N3, false>(variant, n_work, a, st);
    case MODE_N4:  return planar ? launch_variant<MODE_N4, true>(variant, n_work, a, st)  : launch_variant<MODE_N4, false>(variant, n_work, a, st);
    default:       return planar ? launch_variant<MODE_GEN, true>(variant, n_work, a, st) : launch_variant<MODE_GEN, false>(variant, n_work, a, st);
    }
}

}  // namespace rmb
