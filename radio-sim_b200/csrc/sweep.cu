// sweep.cu -- K1, the link-budget / interference sweep (the hot kernel).
//
// For every enabled receiver (work order: grouped by channel, spatially sorted inside a group)
// and every co-channel transmission of the tick it evaluates the log-distance link budget
//     lin(t,r) = P_t * d(t,r)^-n   [mW]              
// and accumulates the co-channel interference sum in the linear domain (N2).  Replaces the
// O(N) receiver loop of RadioMedium.transmit (UDGMRadioMedium.java:99-114) for all the
// tick's packets at once.
//
// Formulation (DESIGN.md "K1"): the tx power is folded into the geometry.  With
// c_t = P_t^(-1/n) the kernel works on scaled coordinates x' = c_t*x, so that
//     d2' = c_t^2 * d^2   and   lin = (d2')^(-n/2)
// Blackwell packed fp32 (FFMA2 / FMUL2, `fma.rn.f32x2`): one instruction handles the SAME
// receiver against TWO transmissions (the shared-memory tx stage is SoA, so a 64-bit LDS
// yields {c_j, c_j+1} directly; the receiver coordinate is duplicated once in registers).
// Per pair: 2.5 FMA-pipe issue slots + ONE MUFU (rsqrt / rcp; lg2+ex2 for a general
// exponent).  The MUFU (XU) pipe, 16 lanes/SM/clk, is the binding unit.
//
// Decomposition: grid = (receiver blocks) x (tx splits).  A block owns THREADS*R receivers
// and 1/S of its channel bucket; it writes one fp32 partial far-field sum per receiver
// (far_part[split][w], summed in fixed order by the resolve kernel -> deterministic).  S is
// chosen on the host so that the grid is many waves deep (tail < 5 %) even when the
// receivers are sharded over 8 GPUs.
//
// "Near" pairs -- those that might be heard (lin >= theta_t) or whose fp32 geometry is
// not accurate enough (d <= D_near) -- are NOT summed here: they are detected per tile of
// kTile transmissions through the tile's partial sum, re-scanned, and appended to the
// receiver's candidate slots for the exact double-precision resolve kernel.  So the fp32
// sum only ever holds weak far-field terms, and every decision is taken in double.
#include "common.cuh"
#include "kernels.h"
#include "resolve_body.cuh"

namespace rmb {

enum { MODE_N2 = 0, MODE_N3 = 1, MODE_N4 = 2, MODE_GEN = 3 };

typedef unsigned long long f32x2;     // two fp32 lanes in one 64-bit register pair

__device__ __forceinline__ float mufu_rsq(float x) { float y; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_rcp(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_lg2(float x) { float y; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

__device__ __forceinline__ f32x2 pack2(float lo, float hi) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpack2(f32x2 v, float &lo, float &hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) { f32x2 d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

// (d2)^(-n/2), one pair, scalar (rescan path)
template <int MODE>
__device__ __forceinline__ float lin_of(float d2, float neg_h)
{
    if (MODE == MODE_N3) { float q = mufu_rsq(d2); return __fmul_rn(__fmul_rn(q, q), q); }
    if (MODE == MODE_N2) return mufu_rcp(d2);
    if (MODE == MODE_N4) { float q = mufu_rcp(d2); return __fmul_rn(q, q); }
    return mufu_ex2(__fmul_rn(neg_h, mufu_lg2(d2)));
}

// acc + (d2)^(-n/2) on both lanes
template <int MODE>
__device__ __forceinline__ f32x2 accumulate2(f32x2 d2, f32x2 acc, float neg_h)
{
    float a, b;
    unpack2(d2, a, b);
    if (MODE == MODE_N3) { f32x2 q = pack2(mufu_rsq(a), mufu_rsq(b)); return fma2(mul2(q, q), q, acc); }
    if (MODE == MODE_N2) return add2(acc, pack2(mufu_rcp(a), mufu_rcp(b)));
    if (MODE == MODE_N4) { f32x2 q = pack2(mufu_rcp(a), mufu_rcp(b)); return fma2(q, q, acc); }
    return add2(acc, pack2(mufu_ex2(__fmul_rn(neg_h, mufu_lg2(a))), mufu_ex2(__fmul_rn(neg_h, mufu_lg2(b)))));
}

// (d2)^(-3/2) on both lanes WITHOUT the XU pipe: integer seed + two Newton steps in packed fp32.  The sweep is bound by
// the MUFU (XU) pipe at 16 lanes/SM/clk while the FMA pipe idles a quarter of the time (ncu: XU 90 %, FMA-heavy 68 %),
// so one pair in eight could take this route: 7 packed FMA-pipe instructions + 4 integer ones for two pairs instead of two
// MUFU.RSQ -- on paper both pipes then fill up together (-10 %); on the GPU the launch got 5.5 % LONGER (the extra 13 % of
// issued instructions and the 7-deep dependent chain cost more than the freed XU slots give back), so it is compiled out.  Relative error of q after two steps <= 4.7e-6 (always low), i.e.
// 1.4e-5 on q^3 for an eighth of the far terms: two orders of magnitude inside the guard band.
__device__ __forceinline__ f32x2 accumulate2_fma_n3(f32x2 d2, f32x2 acc)
{
    float a, b;
    unpack2(d2, a, b);
    const f32x2 y0 = pack2(__int_as_float(0x5f3759df - (__float_as_int(a) >> 1)), __int_as_float(0x5f3759df - (__float_as_int(b) >> 1)));
    const f32x2 nh = mul2(d2, pack2(-0.5f, -0.5f)), c15 = pack2(1.5f, 1.5f);
    f32x2 y = mul2(y0, fma2(nh, mul2(y0, y0), c15));
    y = mul2(y, fma2(nh, mul2(y, y), c15));
    return fma2(mul2(y, y), y, acc);
}

#ifndef RMB_FMA_RSQRT
#define RMB_FMA_RSQRT 0      // MEASURED SLOWER (C4, one B200: sweep 2605 us with, 2470 us without): off, kept for the record
#endif
constexpr bool kFmaRsqrt = RMB_FMA_RSQRT != 0;
constexpr int kStage = 1024;          // tx per shared-memory stage (SoA: c, -x', -y', (-z',) theta)

// ---- the tx stage: 1-D bulk copies (TMA, cp.async.bulk) global -> shared, completion on an mbarrier ------------
// The tiles are regular -- five SoA rows, 16-float granules -- so one elected thread moves a whole stage with five
// copy instructions and the other threads never touch the data on its way in (no LDG -> register -> STS round trip,
// no barrier between "loaded" and "stored").  One buffer: a block's share is one stage long in practice (~526
// transmissions), and two half-size buffers measured slower (2544 vs 2465 us at C4: a second wait + barrier per block).
__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, unsigned bytes, unsigned long long *bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned parity)
{
    asm volatile("{\n"
                 ".reg .pred p;\n"
                 "RMB_WAIT:\n"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
                 "@p bra RMB_DONE;\n"
                 "bra RMB_WAIT;\n"
                 "RMB_DONE:\n"
                 "}" :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}

// one stage = NF bulk copies fired by one thread, the barrier armed with their byte count.  Deliberately NOT inlined:
// inlined, its address arithmetic raised the register pressure of the sweep loop (56-register budget) and the loop's
// running sums went to local memory -- 2513 instead of 2465 us per C4 launch (ptxas: 64 vs 16 bytes of spill stores)
template <bool PLANAR>
__device__ __noinline__ void fire_stage(float *s_stage, unsigned long long *bar, const TxTiles &tx, const float *theta, int s0, int ns)
{
    const unsigned bytes = (unsigned)ns * 4u;
    mbar_expect_tx(bar, bytes * (PLANAR ? 4u : 5u));
    bulk_g2s(s_stage, tx.c + s0, bytes, bar);
    bulk_g2s(s_stage + kStage, tx.nx + s0, bytes, bar);
    bulk_g2s(s_stage + 2 * kStage, tx.ny + s0, bytes, bar);
    bulk_g2s(s_stage + 3 * kStage, theta + s0, bytes, bar);
    if (!PLANAR) bulk_g2s(s_stage + 4 * kStage, tx.nz + s0, bytes, bar);
}

template <int MODE, bool PLANAR, int R, int THREADS>
__global__ void __launch_bounds__(THREADS, (THREADS == 128 && R <= 4) ? 9 : 1)
sweep_kernel(const float *__restrict__ fx, const float *__restrict__ fy, const float *__restrict__ fz,
             Groups groups, TxTiles tx, const float *__restrict__ theta, const float *__restrict__ theta_min_g,
             float neg_h, float *__restrict__ far_part, int *__restrict__ cnt, int *__restrict__ slots,
             int K, int n_work, int S1, int S2, int rb_taper, SweepFuse fuse)
{
    // one stage buffer, carved into its SoA rows; one mbarrier
    constexpr int NF = PLANAR ? 4 : 5;
    __shared__ __align__(128) float s_stage[NF * kStage];
    __shared__ __align__(8) unsigned long long s_bar;
    float *const s_c = s_stage, *const s_nx = s_stage + kStage, *const s_ny = s_stage + 2 * kStage,
          *const s_theta = s_stage + 3 * kStage, *const s_nz = s_stage + (PLANAR ? 0 : 4 * kStage);

    if (fuse.trace && threadIdx.x == 0) fuse.trace[12 * (size_t)blockIdx.x] = global_timer_ns();
    // programmatic dependent launch, both ways: the kernel behind this one (finish.cu) may bring its blocks in as soon
    // as every block of this launch has started -- i.e. during the final wave, into slots nobody is waiting for --
    // and this kernel's own blocks may have started before the preparation kernel in front of it was done
    asm volatile("griddepcontrol.launch_dependents;");
    const int n1 = rb_taper * S1;
    const bool tapered = (int)blockIdx.x >= n1;
    const int S = tapered ? S2 : S1;
    const int bb = tapered ? (int)blockIdx.x - n1 : (int)blockIdx.x;
    const int rb = (tapered ? rb_taper : 0) + bb / S, split = bb % S;
    const int w0 = (rb * THREADS + threadIdx.x) * R;
    // the block's receivers all belong to one channel group (groups are padded to 1024)
    const int wb = rb * THREADS * R;
    int g = 0;
    while (g + 1 < groups.ngroups && wb >= groups.w_begin[g + 1]) ++g;

    float rx[R], ry[R], rz[R], sum[R];
    f32x2 X[R], Y[R], Z[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        rx[r] = fx[w0 + r]; ry[r] = fy[w0 + r]; rz[r] = PLANAR ? 0.f : fz[w0 + r];
        X[r] = pack2(rx[r], rx[r]); Y[r] = pack2(ry[r], ry[r]); Z[r] = pack2(rz[r], rz[r]);
        sum[r] = 0.f;
    }
    // (the receivers above do not depend on the preparation kernel; everything below does)
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (threadIdx.x == 0) atomicMax(&fuse.ctr->t_sweep_begin_inv, ~global_timer_ns());

    // this block's share of the bucket: split `split` of S, a multiple of kTile
    const int b0 = tx.goff[g], b1 = tx.goff[g + 1];
    const int per = ((b1 - b0 + S - 1) / S + kTile - 1) / kTile * kTile;
    const int j_begin = min(b1, b0 + split * per), j_end = min(b1, j_begin + per);
    const float th_min = theta_min_g[g];

    if (threadIdx.x == 0) {
        mbar_init(&s_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    unsigned phase = 0;
    for (int s0 = j_begin; s0 < j_end; s0 += kStage) {
        const int ns = min(kStage, j_end - s0);
        __syncthreads();                  // the barrier is initialised / everybody is done with the previous stage
        if (threadIdx.x == 0) fire_stage<PLANAR>(s_stage, &s_bar, tx, theta, s0, ns);
        mbar_wait(&s_bar, phase);
        phase ^= 1u;
        for (int j = 0; j < ns; j += kTile) {
            f32x2 ps[R];
#pragma unroll
            for (int r = 0; r < R; ++r) ps[r] = 0ull;
#pragma unroll
            for (int u = 0; u < kTile; u += 4) {
                // four transmissions = two packed operands per coordinate (LDS.128, warp-uniform address)
                const ulonglong2 c2 = *reinterpret_cast<const ulonglong2 *>(&s_c[j + u]);
                const ulonglong2 x2 = *reinterpret_cast<const ulonglong2 *>(&s_nx[j + u]);
                const ulonglong2 y2 = *reinterpret_cast<const ulonglong2 *>(&s_ny[j + u]);
                ulonglong2 z2 = make_ulonglong2(0ull, 0ull);
                if (!PLANAR) z2 = *reinterpret_cast<const ulonglong2 *>(&s_nz[j + u]);
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    f32x2 dx = fma2(X[r], c2.x, x2.x), dy = fma2(Y[r], c2.x, y2.x);
                    f32x2 d2 = fma2(dy, dy, mul2(dx, dx));
                    if (!PLANAR) { f32x2 dz = fma2(Z[r], c2.x, z2.x); d2 = fma2(dz, dz, d2); }
                    ps[r] = accumulate2<MODE>(d2, ps[r], neg_h);
                    dx = fma2(X[r], c2.y, x2.y); dy = fma2(Y[r], c2.y, y2.y);
                    d2 = fma2(dy, dy, mul2(dx, dx));
                    if (!PLANAR) { f32x2 dz = fma2(Z[r], c2.y, z2.y); d2 = fma2(dz, dz, d2); }
                    if (MODE == MODE_N3 && kFmaRsqrt && u == kTile - 4) ps[r] = accumulate2_fma_n3(d2, ps[r]);   // one pair in eight off the XU pipe
                    else ps[r] = accumulate2<MODE>(d2, ps[r], neg_h);
                }
            }
            float s[R];
#pragma unroll
            for (int r = 0; r < R; ++r) { float a, b; unpack2(ps[r], a, b); s[r] = __fadd_rn(a, b); }
            float m = s[0];
#pragma unroll
            for (int r = 1; r < R; ++r) m = fmaxf(m, s[r]);
            if (__any_sync(0xffffffffu, !(m < th_min))) {
                // rare: some pair of this tile may be near.  The warp re-scans the affected receivers TOGETHER:
                // a flagged (lane, r) is broadcast and the tile's kTile transmissions are spread over a half warp,
                // two flagged receivers per pass.  (A per-thread scalar re-scan costs the whole warp kTile
                // iterations per event: 8 % of the kernel's instructions on a 1/8 shard, whose warps cover 8x the area.)
                const int lane = threadIdx.x & 31, half = lane >> 4, uu = lane & (kTile - 1);
                static_assert(kTile == 16, "the cooperative re-scan maps one tile onto a half warp");
                const float c = s_c[j + uu], nx = s_nx[j + uu], ny = s_ny[j + uu], th = s_theta[j + uu];
                const float nz = PLANAR ? 0.f : s_nz[j + uu];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    unsigned fl = __ballot_sync(0xffffffffu, !(s[r] < th_min));
                    while (fl) {
                        const int a = __ffs(fl) - 1;
                        fl &= fl - 1;
                        const bool two = fl != 0;
                        const int b = two ? __ffs(fl) - 1 : a;
                        fl &= fl - 1;                                  // (0 & anything) stays 0
                        const int src = half ? b : a;
                        const bool active = half == 0 || two;
                        const float qx = __shfl_sync(0xffffffffu, rx[r], src), qy = __shfl_sync(0xffffffffu, ry[r], src);
                        const float dx = __fmaf_rn(qx, c, nx), dy = __fmaf_rn(qy, c, ny);
                        float d2 = __fmaf_rn(dy, dy, __fmul_rn(dx, dx));
                        if (!PLANAR) { const float qz = __shfl_sync(0xffffffffu, rz[r], src); const float dz = __fmaf_rn(qz, c, nz); d2 = __fmaf_rn(dz, dz, d2); }
                        const float lin = lin_of<MODE>(d2, neg_h);
                        const bool near = !(lin < th);
                        if (active && near) {
                            // order of arrival in the slots is arbitrary; resolve sorts them
                            const int w = w0 + (src - lane) * R + r;
                            const int pos = atomicAdd(&cnt[w], 1);
                            if (pos < K) slots[(size_t)w * K + pos] = s0 + j + uu;
                        }
                        float far = (active && !near) ? lin : 0.f;
#pragma unroll
                        for (int o = kTile / 2; o > 0; o >>= 1) far = __fadd_rn(far, __shfl_xor_sync(0xffffffffu, far, o));   // fixed tree: deterministic
                        const float far_a = __shfl_sync(0xffffffffu, far, 0), far_b = __shfl_sync(0xffffffffu, far, 16);
                        if (lane == a) s[r] = far_a;                 // the owners may sit in either half
                        if (two && lane == b) s[r] = far_b;
                    }
                }
            }
#pragma unroll
            for (int r = 0; r < R; ++r) sum[r] = __fadd_rn(sum[r], s[r]);
        }
    }
    float *out = far_part + (size_t)split * n_work + w0;
#pragma unroll
    for (int r = 0; r < R; ++r) out[r] = sum[r];
    if (fuse.trace && threadIdx.x == 0) {
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        fuse.trace[12 * (size_t)blockIdx.x + 1] = global_timer_ns();
        fuse.trace[12 * (size_t)blockIdx.x + 2] = 0ull;
        fuse.trace[12 * (size_t)blockIdx.x + 3] = smid;
        fuse.trace[12 * (size_t)blockIdx.x + 10] = 0ull;
    }

    // ---- fused epilogue: the exact decisions of the rx block's receivers (resolve_body.cuh) are taken by the
    // sweep blocks themselves -- double-precision, latency-bound work on pipes the sweep leaves idle, hidden behind
    // the other resident blocks instead of being a kernel of its own after the sweep.  The LAST split of an rx
    // block to finish does it, in R passes of 128 receivers claimed through a per-rx-block counter.  In the
    // launch's final wave (rb >= rb_spin; nothing is waiting for a block slot there) what counts is how soon the
    // last receiver is resolved, not throughput: the passes are 32 receivers there (8 per warp, all 32 lanes
    // evaluating their pooled candidates), and the 4R - 1 finishers before the last one wait for it and share them.
    if (THREADS == 128 && fuse.enabled) {
        __shared__ int s_role, s_q, s_flag;
        const bool fine = rb >= fuse.rb_fine;
        const int npass = fine ? R * (THREADS / 32) : R, pass_rx = fine ? 32 : THREADS;
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            const int k = atomicAdd(&fuse.rx_done[rb], 1);
            int role = k == S - 1;
            if (!role && rb >= fuse.rb_spin && k >= S - npass) {
                const unsigned *done = reinterpret_cast<const unsigned *>(&fuse.rx_done[rb]);
                const unsigned long long t0 = global_timer_ns();
                while (ld_acquire_gpu(done) < (unsigned)S && global_timer_ns() - t0 < 500000ull) __nanosleep(200);
                role = ld_acquire_gpu(done) >= (unsigned)S;        // gave up waiting: the last finisher does the passes
            }
            s_role = role;
        }
        __syncthreads();
        if (s_role) {
            __threadfence();
            for (;;) {
                if (threadIdx.x == 0) s_q = atomicAdd(&fuse.rx_pass[rb], 1);
                __syncthreads();
                const int q = s_q;
                __syncthreads();
                if (q >= npass) break;
                if (fuse.trace && threadIdx.x == 0) fuse.trace[12 * (size_t)blockIdx.x + 3] += 1ull << 16;
                const int lane = threadIdx.x & 31, wv = threadIdx.x >> 5;
                const int w = fine ? wb + q * 32 + wv * 8 + lane : wb + q * THREADS + (int)threadIdx.x;
                resolve_receiver<true>(w, !fine || lane < 8, fuse.wk, fuse.dp, tx, fuse.rec, fuse.sl, fuse.pr, fuse.ctr,
                                       fuse.trace ? fuse.trace + 12 * (size_t)blockIdx.x + 4 : nullptr);
                if (fuse.trace && threadIdx.x == 0) fuse.trace[12 * (size_t)blockIdx.x + 8] = global_timer_ns();
                // the last group of receivers of the launch turns the per-packet heard counts into segment offsets
                __syncthreads();
                if (threadIdx.x == 0) { __threadfence(); s_flag = atomicAdd(&fuse.ctr->resolve_done, (unsigned)pass_rx) + (unsigned)pass_rx == (unsigned)n_work; }
                __syncthreads();
                if (s_flag && fuse.seg) {
                    __threadfence();
                    if (fuse.trace && threadIdx.x == 0) fuse.trace[12 * (size_t)blockIdx.x + 10] = global_timer_ns();
                    scan_tx_counts(fuse.pr.tx_cnt, fuse.n_tx, fuse.seg, fuse.ctr);
                    if (fuse.trace && threadIdx.x == 0) fuse.trace[12 * (size_t)blockIdx.x + 11] = global_timer_ns();
                }
            }
        }
        if (fuse.trace && threadIdx.x == 0) fuse.trace[12 * (size_t)blockIdx.x + 2] = global_timer_ns();
    }
    if (threadIdx.x == 0) atomicMax(&fuse.ctr->t_sweep_end, global_timer_ns());
}

// ---- host side: variant table, split planning, launch ---------------------------------------
struct SweepVariant { const void *fn; int R, threads; };

template <int MODE, bool PLANAR>
static SweepVariant variant_of(int variant)
{
#define RMB_V(RR, TT) SweepVariant{(const void *)&sweep_kernel<MODE, PLANAR, RR, TT>, RR, TT}
    switch (variant) {
    case 2:  return RMB_V(8, 128);
    case 3:  return RMB_V(2, 64);
    case 4:  return RMB_V(4, 64);
    case 5:  return RMB_V(8, 64);
    case 6:  return RMB_V(1, 32);
    case 7:  return RMB_V(2, 128);
    case 8:  return RMB_V(4, 256);
    case 9:  return RMB_V(1, 128);
    default: return RMB_V(4, 128);
    }
#undef RMB_V
}

static SweepVariant pick(int mode, bool planar, int variant)
{
    switch (mode) {
    case MODE_N2:  return planar ? variant_of<MODE_N2, true>(variant)  : variant_of<MODE_N2, false>(variant);
    case MODE_N3:  return planar ? variant_of<MODE_N3, true>(variant)  : variant_of<MODE_N3, false>(variant);
    case MODE_N4:  return planar ? variant_of<MODE_N4, true>(variant)  : variant_of<MODE_N4, false>(variant);
    default:       return planar ? variant_of<MODE_GEN, true>(variant) : variant_of<MODE_GEN, false>(variant);
    }
}

// resident blocks per SM of a variant (queried once per kernel: this sits on the tick's launch path)
static int variant_occupancy(const SweepVariant &v)
{
    static const void *fn[64]; static int occ[64]; static int n = 0;
    for (int i = 0; i < n; ++i) if (fn[i] == v.fn) return occ[i];
    int o = 1;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, v.fn, v.threads, 0) != cudaSuccess || o < 1) { o = 1; cudaGetLastError(); }
    if (n < 64) { fn[n] = v.fn; occ[n] = o; ++n; }
    return o;
}

int sweep_plan(int mode, bool planar, int variant, int n_work, int n_tx_bucket, int forced_splits);
// Which variant a tick takes when the caller did not force one: 4 receivers per thread (the fewest shared-memory loads
// per pair) unless that launch would be shallow -- under ~3 waves of blocks the tick's time is the latency chain of the
// fused resolve (one rx block = R passes of 128 receivers by its last finisher), not the arithmetic, and blocks of
// 2 x 128 receivers halve it.  Measured (tools/c5_shard_probe.py, tools/shard_probe.py): 16 channels on a 1/8 shard
// 135 -> 119 us, on one GPU 312 -> 308 us; C4 on a 1/8 shard (3.5 waves) 413 us with R = 4, 419 with R = 2.
int sweep_auto_variant(int mode, bool planar, int n_work, int n_tx_bucket)
{
    const SweepVariant v4 = pick(mode, planar, 1);
    const double slots = (double)device_sms() * variant_occupancy(v4);
    const int s = sweep_plan(mode, planar, 1, n_work, n_tx_bucket, 0);
    const double waves = (double)(n_work / (v4.R * v4.threads)) * s / slots;
    return waves < 3.25 ? 7 : 1;
}

int sweep_mode_for_exponent(double n)
{
    if (n == 2.0) return MODE_N2;
    if (n == 3.0) return MODE_N3;
    if (n == 4.0) return MODE_N4;
    return MODE_GEN;
}

// Number of tx splits S.  Two requirements: (1) the grid (rx blocks x S) must fill the resident block
// slots of the GPU over its waves (quantisation tail): the smallest S reaching >= 95 %, else the best one;
// (2) blocks should be SHORT, about kTargetSplitTx transmissions each: the warp scheduler favours the
// oldest warps, so the blocks of an SM finish one after the other and long blocks leave the XU pipe
// under-occupied at the end of every wave (measured on B200: 1 M receivers x 10 k tx, R=4: 2521 us at
// S=4, 2374 us at S=16..20; a single resident wave of equal "stream-K" blocks was the slowest at 2600 us).
// Each split keeps at least kMinSplitTx transmissions.
int sweep_plan(int mode, bool planar, int variant, int n_work, int n_tx_bucket, int forced_splits)
{
    constexpr int kMinSplitTx = 256, kTargetSplitTx = 512, kMaxSplits = 64;
    const SweepVariant v = pick(mode, planar, variant);
    const int max_s = max(1, min(kMaxSplits, n_tx_bucket / kMinSplitTx));
    if (forced_splits > 0) return min(forced_splits, kMaxSplits);
    const int sms = device_sms(), occ = variant_occupancy(v);
    const double slots = (double)sms * occ;
    const int rx_blocks = max(1, n_work / (v.R * v.threads));
    int best = 1; double best_eff = 0.0;
    for (int s = 1; s <= max_s; ++s) {
        const double waves = rx_blocks * (double)s / slots;
        const double eff = waves / ceil(waves);
        if (eff > best_eff + 1e-9) { best_eff = eff; best = s; }
        if (eff >= 0.95) { best = s; break; }
    }
    return min(max_s, max(best, n_tx_bucket / kTargetSplitTx));
}

// The launch's final wave (with a margin): rx blocks >= rb_taper are cut into s2 = taper * s1 splits (shorter blocks
// where the work runs out), and -- when the variant fuses the resolve -- the last finishers of such an rx block
// share its resolve passes; in the last fine8/8 of a wave (rx blocks >= rb_fine) the passes are 32 receivers wide.
// n_tx_bucket bounds s2 (every split keeps >= 128 transmissions).
SweepSplits sweep_splits(int mode, bool planar, int variant, int n_work, int n_tx_bucket, int s1, int taper, int fine8)
{
    const SweepVariant v = pick(mode, planar, variant);
    const int sms = device_sms(), occ = variant_occupancy(v);
    const int nb = n_work / (v.R * v.threads);
    const int tail = (int)ceil(1.25 * sms * occ / max(1, s1));
    SweepSplits sp;
    sp.s1 = max(1, s1);
    // a launch of fewer than three waves is "final wave" almost everywhere (16 channel groups of 625 transmissions
    // each: 1.5 waves of 8 us blocks): cutting its blocks further only multiplies the per-block overhead
    const bool deep = (long long)nb * sp.s1 >= 3ll * sms * occ;
    sp.s2 = deep ? max(sp.s1, min(min(128, sp.s1 * max(1, taper)), max(1, n_tx_bucket / 128))) : sp.s1;
    sp.rb_taper = max(0, nb - tail);
    // the narrow passes trade throughput for latency: at most the last sixteenth of the receivers takes them, and
    // only when an rx block has as many finishers (tx splits) as narrow passes to share -- with two splits per
    // rx block (16 channel groups of 625 transmissions) they would run eight deep instead (C5: +44 us)
    const int fine_blocks = min((int)ceil(0.125 * fine8 * sms * occ / max(1, sp.s2)), nb / 16);
    const bool can_share = sp.s2 >= v.R * (v.threads / 32);
    sp.rb_fine = (fine8 >= 7 || !can_share) ? nb : max(sp.rb_taper, nb - fine_blocks);
    sp.block_rx = v.R * v.threads;
    sp.can_fuse = v.threads == 128;
    return sp;
}

cudaError_t launch_sweep(int mode, bool planar, int variant, int n_work, SweepSplits splits, const SweepArgs &a, const SweepFuse &fuse,
                         bool overlap, cudaStream_t st)
{
    if (n_work <= 0) return cudaSuccess;
    const SweepVariant v = pick(mode, planar, variant);
    SweepArgs b = a;
    SweepFuse f = fuse;
    if (v.threads != 128) f.enabled = 0;
    int nw = n_work, S1 = max(1, splits.s1), S2 = max(1, splits.s2), rbt = splits.rb_taper;
    const int nb = n_work / (v.R * v.threads);
    rbt = max(0, min(rbt, nb));
    void *args[] = {(void *)&b.fx, (void *)&b.fy, (void *)&b.fz, (void *)&b.groups, (void *)&b.tx, (void *)&b.theta,
                    (void *)&b.theta_min_g, (void *)&b.neg_h, (void *)&b.far_part, (void *)&b.cnt, (void *)&b.slots,
                    (void *)&b.K, (void *)&nw, (void *)&S1, (void *)&S2, (void *)&rbt, (void *)&f};
    const long long blocks = (long long)rbt * S1 + (long long)(nb - rbt) * S2;
    if (blocks > 0x7fffffffll) return cudaErrorInvalidConfiguration;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)blocks); cfg.blockDim = dim3((unsigned)v.threads); cfg.dynamicSmemBytes = 0; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization; attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = (overlap && use_pdl()) ? 1u : 0u;
    return cudaLaunchKernelExC(&cfg, v.fn, args);
}

}  // namespace rmb
