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
