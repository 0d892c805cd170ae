/*
 * kernel.cuh -- replacement for the reference's launch interface
 * (TestProject/kernel.cuh:2-7 and Sim-Without-OpenGL-Integration/kernel.cuh:5).
 *
 * Swap this header in (or keep the reference's own: the symbols below have the
 * same C++ signatures, so objects compiled against either link to
 * libnbody_b200.so) and main.cpp's simulationLoopNoVisual and validation.cpp's
 * compareHostToDevice run on the B200 engine unchanged.
 *
 * Behaviour kept from the reference: one call = one full step, synchronous,
 * on the current device; failures surface as std::runtime_error
 * (kernel.cu:634,640), which the callers catch as std::exception
 * (main.cpp:150-153, validation.cpp:69-72).
 * Behaviour dropped on purpose: the "Shared memory request too large" guard
 * (kernel.cu:639-641), which rejects every N above ~2.16 M on a B200.
 */
#ifndef NBODY_KERNEL_CUH
#define NBODY_KERNEL_CUH

#include <cuda_runtime.h>
#include "nbody_constants.h"

/* device pointers, float4 state, in place -- TestProject/kernel.cuh:2 */
void simulate(float4* d_bodies, float4* d_accelerations, float4* d_velocity, int N);

/* host pointers, float3 velocity/acceleration, copies included --
 * TestProject/Sim-Without-OpenGL-Integration/kernel.cuh:5 */
void simulate(float4* bodies, float3* accelerations, float3* velocity, int N);

/*
 * TestProject/kernel.cuh:3-7 (simulateVisual, simulateVisual_embParallel and its
 * _float3 / _fadl / _fadl4 forms) take a cudaGraphicsResource* that only an
 * OpenGL context can produce: "unavailable on this box" (DESIGN.md).  Their
 * HEADLESS forms below keep every other argument and take the device pointer
 * the reference obtains from cudaGraphicsResourceGetMappedPointer
 * (kernel.cu:476-480) directly.  d_reduceMatrix (the reference's O(N^2/32)
 * scratch, main.cpp:257-269) is accepted and ignored -- it may be null: the
 * engine's split of the source range between CTAs keeps its partial sums in
 * O(N) scratch of its own.  Integrator as the reduction flavours write it
 * (kernel.cu:455-463: half-kick in double).
 */
void simulateHeadless(float4* d_bodies, float4* d_accelerations, float4* d_velocity, int N);
void simulateHeadless_embParallel(float4* d_bodies, float4* d_accelerations, float4* d_velocity, float4* d_reduceMatrix, int N);
void simulateHeadless_embParallel_float3(float4* d_bodies, float3* d_accelerations, float3* d_velocity, float3* d_reduceMatrix, int N);
void simulateHeadless_embParallel_fadl(float4* d_bodies, float4* d_accelerations, float4* d_velocity, float4* d_reduceMatrix, int N);
void simulateHeadless_embParallel_fadl4(float4* d_bodies, float4* d_accelerations, float4* d_velocity, float4* d_reduceMatrix, int N);

/* Run-time replacements for the compile-time EPS2 / DT the reference bakes into
 * its kernel; they affect the float4 overload above.  Defaults: EPS2, DT. */
void simulate_set_parameters(float eps2, float dt);

/* The reference runs on device 0 only (kernel.cu:630).  After simulate_set_devices(n) every device-pointer
 * launcher above spreads its step over n GPUs of the box -- the current device, where the caller's arrays stay,
 * plus the n-1 lowest-numbered others (NVLink peer access) -- so main.cpp's simulationLoopNoVisual scales with
 * one added line and no other source change.  n <= 1 restores the single-device behaviour.  Throws
 * std::runtime_error when n exceeds the devices present. */
void simulate_set_devices(int n);

#endif /* NBODY_KERNEL_CUH */
