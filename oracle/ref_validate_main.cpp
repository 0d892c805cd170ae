// ref_validate_main.cpp -- TEST INFRASTRUCTURE.  Drop-in proof: the reference's UNMODIFIED
// validation.cpp / utils.cpp objects linked against libnbody_b200.so's `simulate`, driving
// the reference's own compareHostToDevice (TestProject/validation.cpp:55-103: 1000 lock-stepped
// steps at N_BODIES, then verify_still_bodies on X, V, A).  Set-up follows TestProject/main.cpp:244-284,352-354.
// The reference's checker only prints; this main counts its "Problem at body" lines via a pipe-free
// convention: it prints a final summary line and always exits 0 unless simulate() threw.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include "constants.h"
#include "utils.h"
#include "validation.h"

int main()
{
    const size_t size4 = sizeof(float4) * N_BODIES;
    float4 *bodies, *velocity, *accelerations;
    cudaMallocHost(&bodies, size4);
    cudaMallocHost(&velocity, size4);
    cudaMallocHost(&accelerations, size4);
    fill_with_random4(bodies, N_BODIES);
    fill_with_zeroes4(velocity, N_BODIES);
    fill_with_zeroes4(accelerations, N_BODIES);
    float4 *d_bodies, *d_velocity, *d_accelerations;
    cudaMalloc((void**)&d_velocity, size4);
    cudaMalloc((void**)&d_accelerations, size4);
    cudaMalloc((void**)&d_bodies, size4);
    cudaMemcpy(d_velocity, velocity, size4, cudaMemcpyHostToDevice);
    cudaMemcpy(d_accelerations, accelerations, size4, cudaMemcpyHostToDevice);
    cudaMemcpy(d_bodies, bodies, size4, cudaMemcpyHostToDevice);
    const int rc = compareHostToDevice(d_bodies, d_accelerations, d_velocity, bodies, accelerations, velocity);
    printf("ref_validate: compareHostToDevice returned %d (N_BODIES=%d)\n", rc, N_BODIES);
    return rc;
}
