// kernel_shim.cu -- the reference's C++ launch interface on top of the C ABI.
// Exports the same mangled symbols as TestProject/kernel.cu:628 and
// Sim-Without-OpenGL-Integration/kernel.cu:85, so the reference's unmodified
// main.cpp / validation.cpp objects link against libnbody_b200.so.
#include <stdexcept>
#include <string>

#include "../../include/kernel.cuh"
#include "../../include/nbody_b200.h"

namespace {
float g_eps2 = 0.002f;   // TestProject/constants.h:25
float g_dt = 0.1f;       // TestProject/constants.h:26
}

void simulate_set_parameters(float eps2, float dt)
{
    g_eps2 = eps2;
    g_dt = dt;
}

// The reference's launcher asks for device 0 and nothing else (kernel.cu:630).  simulate_set_devices(n) spreads
// every following simulate() / simulateHeadless*() call over n GPUs of the box: the current device (where the
// caller's arrays live) plus the n-1 lowest-numbered others.  n <= 1: back to one device.
void simulate_set_devices(int n)
{
    int cur = 0, ndev = 0;
    if (cudaGetDevice(&cur) != cudaSuccess || cudaGetDeviceCount(&ndev) != cudaSuccess) throw std::runtime_error("simulate_set_devices: no CUDA device");
    if (n > ndev) throw std::runtime_error("simulate_set_devices: " + std::to_string(n) + " devices asked for, " + std::to_string(ndev) + " present");
    int list[NBODY_MAX_DEVICES];
    int k = 0;
    if (n > NBODY_MAX_DEVICES) n = NBODY_MAX_DEVICES;
    if (n > 1) {
        list[k++] = cur;
        for (int d = 0; d < ndev && k < n; ++d)
            if (d != cur) list[k++] = d;
    }
    if (nbody_set_devices(k > 1 ? list : nullptr, k) != NBODY_OK) throw std::runtime_error(std::string("simulate_set_devices: ") + nbody_last_error());
}

void simulate(float4* d_bodies, float4* d_accelerations, float4* d_velocity, int N)
{
    const int rc = nbody_step_inplace_f4(reinterpret_cast<float*>(d_bodies), reinterpret_cast<float*>(d_accelerations),
                                         reinterpret_cast<float*>(d_velocity), N, g_eps2, g_dt);
    if (rc != NBODY_OK) throw std::runtime_error(std::string("simulate: ") + nbody_last_error());
}

void simulate(float4* bodies, float3* accelerations, float3* velocity, int N)
{
    // constants of the headless directory: EPS2 0.002, DT 0.01 (double literals)
    const int rc = nbody_step_host_f3(reinterpret_cast<float*>(bodies), reinterpret_cast<float*>(accelerations),
                                      reinterpret_cast<float*>(velocity), N, 0.002, 0.01);
    if (rc != NBODY_OK) throw std::runtime_error(std::string("simulate: ") + nbody_last_error());
}

// ---- headless forms of TestProject/kernel.cuh:3-7 --------------------------------------------
namespace {
void step_opts(float4* d_bodies, void* d_acc, void* d_vel, int N, int integrator, int f3)
{
    nbody_step_opts o;
    nbody_step_opts_default(&o);
    o.eps2 = g_eps2; o.dt = g_dt; o.integrator = integrator; o.va_float3 = f3;
    const int rc = nbody_step_inplace_opts(reinterpret_cast<float*>(d_bodies), d_acc, d_vel, N, &o);
    if (rc != NBODY_OK) throw std::runtime_error(std::string("simulate: ") + nbody_last_error());
}
}

// simulateVisual (kernel.cu:599-625) runs the same `kernel` as simulate()
void simulateHeadless(float4* d_bodies, float4* d_accelerations, float4* d_velocity, int N)
{
    step_opts(d_bodies, d_accelerations, d_velocity, N, NBODY_INTEG_F32, 0);
}
// kernel_reduction + inter_block_reduction (kernel.cu:163-246,406-436; launcher :536-565)
void simulateHeadless_embParallel(float4* d_bodies, float4* d_accelerations, float4* d_velocity, float4*, int N)
{
    step_opts(d_bodies, d_accelerations, d_velocity, N, NBODY_INTEG_MIXED, 0);
}
// kernel_reduction_float3 + inter_block_reduction_float3 (kernel.cu:250-318,439-469; launcher :567-595)
void simulateHeadless_embParallel_float3(float4* d_bodies, float3* d_accelerations, float3* d_velocity, float3*, int N)
{
    step_opts(d_bodies, d_accelerations, d_velocity, N, NBODY_INTEG_MIXED, 1);
}
// "first add during load" forms (kernel_reduction_fadl, kernel.cu:321-402; launchers :472-534): same arithmetic, other launch shape
void simulateHeadless_embParallel_fadl(float4* d_bodies, float4* d_accelerations, float4* d_velocity, float4*, int N)
{
    step_opts(d_bodies, d_accelerations, d_velocity, N, NBODY_INTEG_MIXED, 0);
}
void simulateHeadless_embParallel_fadl4(float4* d_bodies, float4* d_accelerations, float4* d_velocity, float4*, int N)
{
    step_opts(d_bodies, d_accelerations, d_velocity, N, NBODY_INTEG_MIXED, 0);
}
