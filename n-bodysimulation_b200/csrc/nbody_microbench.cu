// nbody_microbench.cu -- pipe micro-benchmarks that decide the kernel design on
// sm_100a: scalar FFMA vs packed FFMA2/FADD2/FMUL2
// throughput, MUFU.RSQ throughput and co-issue, broadcast LDS.128 cost, and the
// n-body instruction mix (12 FP32 + 1 MUFU per interaction) in isolation.
// Output: one JSON object per line (kind, warps/SM, lane-ops per clock per SM).
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "nbody_ptx.cuh"

using namespace nbody;

enum Kind { K_FFMA = 0, K_FFMA2, K_FADD2, K_FMUL2, K_MUFU, K_MIX_SCALAR, K_MIX_PACKED, K_MIX_PACKED_LDS, K_FFMA_MUFU_PAR, K_COUNT };
static const char* kNames[] = {"ffma", "ffma2", "fadd2", "fmul2", "mufu_rsq", "mix_scalar_12ffma_1mufu", "mix_packed_12x2_2mufu",
                               "mix_packed_lds", "ffma2_plus_mufu_8to1"};
// FP32 lane-operations (one FMA/ADD/MUL on one lane = 1) per inner iteration per thread
static const int kLaneOps[] = {8, 16, 16, 16, 0, 12 * 4, 24 * 4, 24 * 4, 16 * 4};
static const int kMufuOps[] = {0, 0, 0, 0, 8, 4, 8, 8, 4};

template <int KIND>
__global__ void __launch_bounds__(1024) mb_kernel(float* out, long long* cycles, int iters, float seed)
{
    __shared__ __align__(16) float4 sh[64];
    if (threadIdx.x < 64) sh[threadIdx.x] = make_float4(seed, seed * 0.5f, seed * 0.25f, 1.0f);
    __syncthreads();
    float a[8];
    u64 p[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { a[i] = seed + i + threadIdx.x * 1e-3f; p[i] = pack2(a[i], a[i] + 0.5f); }
    const float b = 0.999f, c = 1e-3f;
    const u64 b2 = pack2(b, b), c2 = pack2(c, c);
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
        if (KIND == K_FFMA) {
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], b, c);
        } else if (KIND == K_FFMA2) {
#pragma unroll
            for (int i = 0; i < 8; ++i) p[i] = fma2(p[i], b2, c2);
        } else if (KIND == K_FADD2) {
#pragma unroll
            for (int i = 0; i < 8; ++i) p[i] = add2(p[i], c2);
        } else if (KIND == K_FMUL2) {
#pragma unroll
            for (int i = 0; i < 8; ++i) p[i] = mul2(p[i], b2);
        } else if (KIND == K_MUFU) {
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = rsqrt_mufu(a[i]);
        } else if (KIND == K_MIX_SCALAR) {
            // 4 independent scalar interactions: 3 FADD, 3 FFMA, 2 FMUL, MUFU, FMUL, 3 FFMA
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float dx = b - a[i], dy = c - a[i], dz = seed - a[i];
                const float r2 = fmaf(dz, dz, fmaf(dy, dy, fmaf(dx, dx, c)));
                const float s = b * rsqrt_mufu(r2 * r2 * r2);
                a[i] = fmaf(dx, s, a[i]); a[i + 4] = fmaf(dy, s, a[i + 4]); a[i] = fmaf(dz, s, a[i]);
            }
        } else if (KIND == K_MIX_PACKED || KIND == K_MIX_PACKED_LDS) {
            u64 sx = b2, sy = c2, sz = b2, sm = c2;
            if (KIND == K_MIX_PACKED_LDS) {
                const ulonglong2 l0 = reinterpret_cast<const ulonglong2*>(sh)[(2 * it) & 63];
                const ulonglong2 l1 = reinterpret_cast<const ulonglong2*>(sh)[(2 * it + 1) & 63];
                sx = l0.x; sy = l0.y; sz = l1.x; sm = l1.y;
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const u64 dx = add2(sx, p[i]), dy = add2(sy, p[i]), dz = add2(sz, p[i]);
                u64 r2 = fma2(dx, dx, c2); r2 = fma2(dy, dy, r2); r2 = fma2(dz, dz, r2);
                const u64 d3 = mul2(mul2(r2, r2), r2);
                float lo, hi;
                unpack2(d3, lo, hi);
                const u64 s = mul2(pack2(rsqrt_mufu(lo), rsqrt_mufu(hi)), sm);
                p[i] = fma2(dx, s, p[i]); p[i + 4] = fma2(dy, s, p[i + 4]); p[i] = fma2(dz, s, p[i]);
            }
        } else if (KIND == K_FFMA_MUFU_PAR) {
            // 8 independent FFMA2 chains plus one independent MUFU chain per 2 FFMA2
#pragma unroll
            for (int i = 0; i < 8; ++i) p[i] = fma2(p[i], b2, c2);
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = rsqrt_mufu(a[i]);
#pragma unroll
            for (int i = 0; i < 8; ++i) p[i] = fma2(p[i], b2, c2);
#pragma unroll
            for (int i = 0; i < 8; ++i) p[i] = fma2(p[i], b2, c2);
#pragma unroll
            for (int i = 0; i < 8; ++i) p[i] = fma2(p[i], b2, c2);
        }
    }
    const long long t1 = clock64();
    float acc = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) { float lo, hi; unpack2(p[i], lo, hi); acc += a[i] + lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int KIND>
static void run(int sms, int warps_per_sm, int iters, float* d_out, long long* d_cyc)
{
    // one CTA per SM holding all the warps, so the per-SM clock window is exact
    const int threads = warps_per_sm * 32;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    mb_kernel<KIND><<<sms, threads>>>(d_out, d_cyc, iters / 10 + 1, 1.5f);   // warm-up
    cudaEventRecord(e0);
    mb_kernel<KIND><<<sms, threads>>>(d_out, d_cyc, iters, 1.5f);
    cudaEventRecord(e1);
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) { printf("{\"kind\": \"%s\", \"error\": \"%s\"}\n", kNames[KIND], cudaGetErrorString(err)); return; }
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    std::vector<long long> cyc(sms);
    cudaMemcpy(cyc.data(), d_cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost);
    double mean = 0;
    for (int i = 0; i < sms; ++i) mean += (double)cyc[i];
    mean /= sms;
    const double lane_ops = (double)kLaneOps[KIND] * threads * iters, mufu = (double)kMufuOps[KIND] * threads * iters;
    printf("{\"kind\": \"%s\", \"warps_per_sm\": %d, \"fp32_lane_ops_per_clk_per_sm\": %.2f, \"mufu_per_clk_per_sm\": %.2f, "
           "\"cycles\": %.0f, \"ms\": %.4f, \"implied_sm_mhz\": %.0f}\n",
           kNames[KIND], warps_per_sm, lane_ops / mean, mufu / mean, mean, ms, mean / (ms * 1e3));
    cudaEventDestroy(e0); cudaEventDestroy(e1);
}

int main(int argc, char** argv)
{
    int iters = argc > 1 ? atoi(argv[1]) : 20000;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) { fprintf(stderr, "no CUDA device\n"); return 1; }
    const int sms = prop.multiProcessorCount;
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_khz\": %d}\n", prop.name, sms, prop.clockRate);
    float* d_out; long long* d_cyc;
    cudaMalloc(&d_out, (size_t)sms * 1024 * sizeof(float));
    cudaMalloc(&d_cyc, sms * sizeof(long long));
    const int ws[] = {4, 8, 16, 32};
    for (int w : ws) {
        run<K_FFMA>(sms, w, iters, d_out, d_cyc);
        run<K_FFMA2>(sms, w, iters, d_out, d_cyc);
        run<K_FADD2>(sms, w, iters, d_out, d_cyc);
        run<K_FMUL2>(sms, w, iters, d_out, d_cyc);
        run<K_MUFU>(sms, w, iters, d_out, d_cyc);
        run<K_MIX_SCALAR>(sms, w, iters / 4, d_out, d_cyc);
        run<K_MIX_PACKED>(sms, w, iters / 4, d_out, d_cyc);
        run<K_MIX_PACKED_LDS>(sms, w, iters / 4, d_out, d_cyc);
        run<K_FFMA_MUFU_PAR>(sms, w, iters / 4, d_out, d_cyc);
    }
    return 0;
}
