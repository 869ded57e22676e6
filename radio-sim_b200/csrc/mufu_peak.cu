// mufu_peak.cu -- the 20-line microbenchmark SURVEY.md 8(d) asks for: what the XU (MUFU) pipe and the FMA pipe
// of THIS GPU sustain, measured, so that roofline.peak's "16 MUFU lanes per SM per clock" is a number and not an
// assumption.  Stand-alone program (make -C radio-sim_b200/csrc ../mufu_peak), prints one JSON line; bench.py runs
// it once per invocation and quotes it beside the roofline.
//
// Every SM holds 2 blocks x 1024 threads (full occupancy); a thread runs 8 independent dependency chains of the
// instruction under test.  Throughput per SM per clock = threads x ops / cycles, cycles from clock64() around the
// loop of one block; the wall-clock rate (CUDA events) gives the sustained clock.
#include <cstdio>
#include <cuda_runtime.h>

constexpr int kIters = 4096, kChains = 8, kThreads = 1024;

__device__ __forceinline__ float mufu_rsq(float x) { float y; asm volatile("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_lg2(float x) { float y; asm volatile("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float mufu_ex2(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

template <int OP>
__global__ void __launch_bounds__(kThreads, 2) probe(float *sink, long long *cycles, float seed)
{
    float v[kChains];
#pragma unroll
    for (int k = 0; k < kChains; ++k) v[k] = seed + (float)(threadIdx.x + k);
    unsigned long long pv[kChains / 2];
#pragma unroll
    for (int k = 0; k < kChains / 2; ++k) pv[k] = ((unsigned long long)__float_as_uint(v[2 * k]) << 32) | __float_as_uint(v[2 * k + 1]);
    const unsigned long long pc = ((unsigned long long)__float_as_uint(0.999f) << 32) | __float_as_uint(0.999f);
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < kIters; ++i) {
#pragma unroll
        for (int k = 0; k < kChains; ++k) {
            if (OP == 0) v[k] = mufu_rsq(v[k]);                        // MUFU.RSQ
            if (OP == 1) v[k] = mufu_ex2(mufu_lg2(v[k]));              // MUFU.LG2 + MUFU.EX2 (the general-exponent pair)
            if (OP == 2) v[k] = __fmaf_rn(v[k], 0.999f, 0.001f);       // FFMA
        }
        if (OP == 3) {
#pragma unroll
            for (int k = 0; k < kChains / 2; ++k) asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(pv[k]) : "l"(pc));   // FFMA2
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < kChains; ++k) s += v[k];
#pragma unroll
    for (int k = 0; k < kChains / 2; ++k) s += __uint_as_float((unsigned)pv[k]);
    if (s == 12345.678f) sink[0] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int OP>
static void run(const char *name, double ops_per_iter_thread, int sms, bool last)
{
    float *sink; long long *cyc;
    const int blocks = 2 * sms;
    cudaMalloc(&sink, 4); cudaMalloc(&cyc, blocks * sizeof(long long));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<OP><<<blocks, kThreads>>>(sink, cyc, 1.5f);                   // warm-up
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0); probe<OP><<<blocks, kThreads>>>(sink, cyc, 1.5f); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    long long h[1024]; cudaMemcpy(h, cyc, blocks * sizeof(long long), cudaMemcpyDeviceToHost);
    double c = 0; for (int b = 0; b < blocks; ++b) c += (double)h[b]; c /= blocks;
    const double per_thread = ops_per_iter_thread * kIters;
    const double per_sm_clk = 2.0 * kThreads * per_thread / c;
    const double gops = (double)blocks * kThreads * per_thread / (best * 1e-3) / 1e9;
    printf("\"%s\": {\"per_sm_per_clk\": %.2f, \"gops\": %.1f, \"implied_clock_mhz\": %.0f}%s", name, per_sm_clk, gops,
           gops * 1e3 / (per_sm_clk * sms), last ? "" : ", ");
    cudaFree(sink); cudaFree(cyc);
}

int main()
{
    int sms = 148;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0) != cudaSuccess) { printf("{\"error\": \"no CUDA device\"}\n"); return 1; }
    printf("{\"sms\": %d, ", sms);
    run<0>("mufu_rsq", kChains, sms, false);
    run<1>("mufu_lg2_ex2_pairs", kChains, sms, false);     // counted as PAIRS (2 MUFU instructions each)
    run<2>("ffma", kChains, sms, false);
    run<3>("ffma2_lanes", kChains, sms, true);             // counted as fp32 lanes (2 per instruction)
    printf("}\n");
    return cudaDeviceSynchronize() == cudaSuccess ? 0 : 1;
}
