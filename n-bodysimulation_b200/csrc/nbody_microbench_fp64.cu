// nbody_microbench_fp64.cu -- one bounded experiment (round 2): does the FP64 pipe of a B200 SM run
// beside the packed-FP32 pipe?  The step kernel is FMA-pipe bound at ~53 % of its issue slots.  If one
// extra target per thread could be carried in FP64 (12 D-ops + integer-ALU float<->double bit
// conversions + 1 MUFU per interaction, no FMA-pipe work), 7 targets would cost the time of 6 -- if the
// FP64 pipe is wide enough.  This measures the packed mix alone (6 targets), the FP64 mix alone (1 target)
// and both interleaved, per 2 source bodies, in interactions per clock per SM, plus the raw DFMA rate.
// Output: one JSON object per line.
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "nbody_ptx.cuh"

using namespace nbody;

// double -> float by integer ops (truncating; d > 0, clamped into the normal float range)
__device__ __forceinline__ float d2f_bits(double d)
{
    int hi = __double2hiint(d), lo = __double2loint(d);
    hi = min(hi - (896 << 20), (254 << 20) | 0xFFFFF);
    return __int_as_float(__funnelshift_l(lo, hi, 3));
}
// float -> double by integer ops (exact for normal floats)
__device__ __forceinline__ double f2d_bits(float f)
{
    const unsigned b = __float_as_uint(f);
    return __hiloint2double((int)((b >> 3) + (896u << 20)), (int)(b << 29));
}

enum Kind { K_F6 = 0, K_D1, K_F6D1, K_F6D1_CVT, K_DFMA, K_COUNT };
static const char* kNames[] = {"packed_6_targets", "fp64_1_target", "packed_6_plus_fp64_1", "packed_6_plus_fp64_1_f2f", "dfma"};
static const int kInter[] = {12, 2, 14, 14, 0};   // interactions per iteration per thread

template <int KIND>
__global__ void __launch_bounds__(256) mb_kernel(float* out, long long* cycles, int iters, float seed)
{
    __shared__ __align__(16) float4 sh[128];
    __shared__ __align__(16) double2 shd[128];
    if (threadIdx.x < 128) {
        sh[threadIdx.x] = make_float4(seed, seed * 0.5f, seed * 0.25f, 1.0f);
        shd[threadIdx.x] = make_double2(seed + threadIdx.x, 0.5 * seed);
    }
    __syncthreads();
    u64 tx[6], ty[6], tz[6], ax[6], ay[6], az[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        const float t = seed + i + threadIdx.x * 1e-3f;
        tx[i] = pack2(-t, -t); ty[i] = pack2(-t * 0.5f, -t * 0.5f); tz[i] = pack2(-t * 0.25f, -t * 0.25f);
        ax[i] = ay[i] = az[i] = 0ull;
    }
    double dtx = -(seed + threadIdx.x * 1e-3), dty = 0.5 * dtx, dtz = 0.25 * dtx, dax = 0, day = 0, daz = 0;
    double dchain[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) dchain[i] = seed + i;
    const u64 e2 = pack2(1e-3f, 1e-3f);
    const long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
        if (KIND == K_DFMA) {
#pragma unroll
            for (int i = 0; i < 8; ++i) dchain[i] = fma(dchain[i], 0.999, 1e-3);
            continue;
        }
        const ulonglong2 l0 = reinterpret_cast<const ulonglong2*>(sh)[(2 * it) & 127];
        const ulonglong2 l1 = reinterpret_cast<const ulonglong2*>(sh)[(2 * it + 1) & 127];
        if (KIND != K_D1) {
#pragma unroll
            for (int i = 0; i < 6; ++i) {
                const u64 dx = add2(l0.x, tx[i]), dy = add2(l0.y, ty[i]), dz = add2(l1.x, tz[i]);
                u64 r2 = fma2(dx, dx, e2); r2 = fma2(dy, dy, r2); r2 = fma2(dz, dz, r2);
                const u64 d3 = mul2(mul2(r2, r2), r2);
                float lo, hi;
                unpack2(d3, lo, hi);
                const u64 s = mul2(pack2(rsqrt_mufu(lo), rsqrt_mufu(hi)), l1.y);
                ax[i] = fma2(dx, s, ax[i]); ay[i] = fma2(dy, s, ay[i]); az[i] = fma2(dz, s, az[i]);
            }
        }
        if (KIND != K_F6) {
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const double2 a = shd[(4 * it + 2 * j) & 127], b = shd[(4 * it + 2 * j + 1) & 127];
                const double dx = a.x + dtx, dy = a.y + dty, dz = b.x + dtz;
                const double r2 = fma(dz, dz, fma(dy, dy, fma(dx, dx, 1e-3)));
                const double d3 = r2 * r2 * r2;
                const float rs = rsqrt_mufu(KIND == K_F6D1_CVT ? (float)d3 : d2f_bits(d3));
                const double s = (KIND == K_F6D1_CVT ? (double)rs : f2d_bits(rs)) * b.y;
                dax = fma(dx, s, dax); day = fma(dy, s, day); daz = fma(dz, s, daz);
            }
        }
    }
    const long long t1 = clock64();
    float acc = (float)(dax + day + daz);
#pragma unroll
    for (int i = 0; i < 8; ++i) acc += (float)dchain[i];
#pragma unroll
    for (int i = 0; i < 6; ++i) { float lo, hi; unpack2(ax[i], lo, hi); acc += lo + hi; unpack2(ay[i], lo, hi); acc += lo + hi; unpack2(az[i], lo, hi); acc += lo + hi; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int KIND>
static void run(int sms, int iters, float* d_out, long long* d_cyc)
{
    const int threads = 256;
    mb_kernel<KIND><<<sms, threads>>>(d_out, d_cyc, iters / 10 + 1, 1.5f);
    mb_kernel<KIND><<<sms, threads>>>(d_out, d_cyc, iters, 1.5f);
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) { printf("{\"kind\": \"%s\", \"error\": \"%s\"}\n", kNames[KIND], cudaGetErrorString(err)); return; }
    std::vector<long long> cyc(sms);
    cudaMemcpy(cyc.data(), d_cyc, sms * sizeof(long long), cudaMemcpyDeviceToHost);
    double mean = 0;
    for (int i = 0; i < sms; ++i) mean += (double)cyc[i];
    mean /= sms;
    if (KIND == K_DFMA)
        printf("{\"kind\": \"dfma\", \"dfma_per_clk_per_sm\": %.2f}\n", 8.0 * threads * iters / mean);
    else
        printf("{\"kind\": \"%s\", \"interactions_per_clk_per_sm\": %.3f, \"cycles_per_iteration\": %.1f}\n",
               kNames[KIND], (double)kInter[KIND] * threads * iters / mean, mean / iters);
}

int main(int argc, char** argv)
{
    int iters = argc > 1 ? atoi(argv[1]) : 20000;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, 0) != cudaSuccess) { fprintf(stderr, "no CUDA device\n"); return 1; }
    const int sms = prop.multiProcessorCount;
    float* d_out; long long* d_cyc;
    cudaMalloc(&d_out, (size_t)sms * 256 * sizeof(float));
    cudaMalloc(&d_cyc, sms * sizeof(long long));
    run<K_DFMA>(sms, iters, d_out, d_cyc);
    run<K_F6>(sms, iters, d_out, d_cyc);
    run<K_D1>(sms, iters, d_out, d_cyc);
    run<K_F6D1>(sms, iters, d_out, d_cyc);
    run<K_F6D1_CVT>(sms, iters, d_out, d_cyc);
    return 0;
}
