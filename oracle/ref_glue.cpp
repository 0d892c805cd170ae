// ref_glue.cpp -- TEST INFRASTRUCTURE.  C-callable doors into the reference's own,
// unmodified CPU object code (TestProject/validation.cpp, TestProject/utils.cpp),
// compiled where it lies under /root/reference by oracle/Makefile into
// oracle/_ref/libref_cpu*.so.  Used to pin oracle/nbody_oracle.c and (kind
// "reference") as the timed CPU baseline.  No reference source is copied here.
#include <cuda_runtime.h>
#include <stdexcept>
#include "validation.h"   // resolved with -I/root/reference/TestProject
#include "utils.h"
#include "constants.h"

extern "C" {

// TestProject/validation.cpp:28-52, in place (literal reference semantics)
void ref_CPU_compute(float* X, float* A, float* V, int N)
{
    CPU_compute(reinterpret_cast<float4*>(X), reinterpret_cast<float4*>(A), reinterpret_cast<float4*>(V), N);
}

// TestProject/validation.cpp:9-24
void ref_bodyInteractions_CPU(const float* bi, const float* bj, float* ai)
{
    float4 r = bodyInteractions_CPU(make_float4(bi[0], bi[1], bi[2], bi[3]),
                                    make_float4(bj[0], bj[1], bj[2], bj[3]),
                                    make_float4(ai[0], ai[1], ai[2], ai[3]));
    ai[0] = r.x; ai[1] = r.y; ai[2] = r.z; ai[3] = r.w;
}

// Snapshot reading of CPU_compute built from the reference's own pair function
// (object code of validation.cpp:9-24) and its integrator expressions (:43-49).
void ref_snapshot_step(float* X, float* A, float* V, int N)
{
    float4* gX = reinterpret_cast<float4*>(X);
    float4* gA = reinterpret_cast<float4*>(A);
    float4* gV = reinterpret_cast<float4*>(V);
    float4* S = new float4[N];
    for (int i = 0; i < N; ++i) S[i] = gX[i];
#pragma omp parallel for
    for (int i = 0; i < N; ++i) {
        float4 a = { 0.0f, 0.0f, 0.0f };
        for (int j = 0; j < N; ++j)
            if (i != j) a = bodyInteractions_CPU(S[i], S[j], a);
        gA[i] = a;
        gV[i].x += 0.5f * DT * a.x; gV[i].y += 0.5f * DT * a.y; gV[i].z += 0.5f * DT * a.z;
        gX[i].x += DT * gV[i].x;    gX[i].y += DT * gV[i].y;    gX[i].z += DT * gV[i].z;
    }
    delete[] S;
}

// TestProject/utils.cpp:30-37 (unseeded rand(): made explicit as srand(1))
void ref_fill_with_random4(float* X, int N)
{
    srand(1);
    fill_with_random4(reinterpret_cast<float4*>(X), N);
}

// TestProject/validation.cpp:143-164 prints its findings; nothing to return.
void ref_verify_still_bodies(float* v, float* x, int N)
{
    verify_still_bodies(reinterpret_cast<float4*>(v), reinterpret_cast<float4*>(x), N);
}

void ref_constants(float* eps2, float* dt, int* n_bodies, float* max_xyz, float* min_w, float* max_w)
{
    *eps2 = EPS2; *dt = DT; *n_bodies = N_BODIES; *max_xyz = MAX_X; *min_w = MIN_W; *max_w = MAX_W;
}

}  // extern "C"

// compareHostToDevice (validation.cpp:67) needs a `simulate`; the CPU-only
// library gets a weak one that refuses, the drop-in proof binary links ours.
__attribute__((weak)) void simulate(float4*, float4*, float4*, int)
{
    throw std::runtime_error("libref_cpu: no GPU engine linked behind simulate()");
}
