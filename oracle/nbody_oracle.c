/*
 * nbody_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the all-pairs gravity step of LienoPC/N-BodySimulation,
 * used only as the checker by tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py.  Nothing under
 * n-bodysimulation_b200/ may link, import or call this file.
 *
 * Parity status: PINNED.  tests/test_oracle_cpu.py checks every function here
 * bit-for-bit against (a) the reference's own validation.cpp / utils.cpp object
 * code compiled unchanged into oracle/_ref/libref_cpu.so (when present) and
 * (b) the golden vectors in tests/golden/ that were generated from that
 * object code by tests/golden/make_golden.py.  The two flavours the reference
 * has as GPU code only (oracle_step_headless_f3, oracle_step_mixed) are checked
 * against the output of the reference's kernels on a B200
 * (tests/golden/headless_ref_gpu.npz), to summation rounding.  The point-splat
 * frame (oracle_render) restates shader + fixed-function arithmetic and stays
 * UNPINNED: no OpenGL implementation exists on these machines.
 *
 * Build:  gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC  (see oracle/Makefile)
 * -ffp-contract=off keeps every multiply and add separately rounded, which is
 * how the reference is built (MSVC /O2 and g++ -O2 on x86-64 emit no FMA).
 *
 * Layout: bodies are float[4*N] = (x, y, z, mass) like the reference's float4
 * (TestProject/main.cpp:232,244-254); velocities / accelerations float[4*N].
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

/* ---- one softened pair interaction ------------------------------------ */
/* follows TestProject/validation.cpp:9-24 (bodyInteractions_CPU); same
 * left-to-right evaluation: ((dx*dx + dy*dy) + dz*dz) + eps2, then
 * 1.0f / sqrtf(d*d*d), then mass * that, then three multiply-adds. */
static inline void pair_f32(const float *bi, const float *bj, float *ai, float eps2)
{
    float rx = bj[0] - bi[0];
    float ry = bj[1] - bi[1];
    float rz = bj[2] - bi[2];
    float d = rx * rx + ry * ry + rz * rz + eps2;
    float inv = 1.0f / sqrtf(d * d * d);
    float s = bj[3] * inv;
    ai[0] += rx * s;
    ai[1] += ry * s;
    ai[2] += rz * s;
}

void oracle_pair_f32(const float *bi, const float *bj, float *ai, float eps2)
{
    pair_f32(bi, bj, ai, eps2);
}

/* integrator lines, TestProject/validation.cpp:43-49: v += 0.5f*DT*a ; x += DT*v.
 * `0.5f * DT` is one float constant (C left-to-right: (0.5f*DT)*a). */
static inline void kick_drift_f32(float *x, float *v, const float *a, float dt)
{
    float hdt = 0.5f * dt;
    v[0] += hdt * a[0];
    v[1] += hdt * a[1];
    v[2] += hdt * a[2];
    x[0] += dt * v[0];
    x[1] += dt * v[1];
    x[2] += dt * v[2];
}

/* O1: the reference's CPU_compute exactly as written and as built (serial):
 * TestProject/validation.cpp:28-52.  Body i is advanced IN PLACE inside the
 * i loop, so bodies j<i are seen at their new positions. */
void oracle_step_literal_f32(float *X, float *A, float *V, int N, float eps2, float dt)
{
    for (int i = 0; i < N; ++i) {
        float body[4] = { X[4*i], X[4*i+1], X[4*i+2], X[4*i+3] };
        float acc[4] = { 0.0f, 0.0f, 0.0f, 0.0f };
        for (int j = 0; j < N; ++j)
            if (i != j) pair_f32(body, X + 4*j, acc, eps2);
        A[4*i] = acc[0]; A[4*i+1] = acc[1]; A[4*i+2] = acc[2]; A[4*i+3] = acc[3];
        kick_drift_f32(X + 4*i, V + 4*i, acc, dt);
    }
}

/* O2 (parity target): same pair math, same j order, same integrator lines,
 * but every body reads the positions of the START of the step -- what the
 * reference's GPU kernel intends (TestProject/kernel.cu:94-129).
 * Rows are independent, so the OpenMP loop is bit-identical to the serial one. */
void oracle_step_snapshot_f32(float *X, float *A, float *V, int N, float eps2, float dt)
{
    float *S = (float *)malloc((size_t)N * 4 * sizeof(float));
    memcpy(S, X, (size_t)N * 4 * sizeof(float));
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; ++i) {
        float acc[4] = { 0.0f, 0.0f, 0.0f, 0.0f };
        for (int j = 0; j < N; ++j)
            if (i != j) pair_f32(S + 4*i, S + 4*j, acc, eps2);
        A[4*i] = acc[0]; A[4*i+1] = acc[1]; A[4*i+2] = acc[2]; A[4*i+3] = acc[3];
        kick_drift_f32(X + 4*i, V + 4*i, acc, dt);
    }
    free(S);
}

/* O4a: fp32 accelerations (reference pair math and j order) of K sampled
 * targets idx[0..K) against all N sources; out is float[3*K]. */
void oracle_accel_f32_sampled(const float *X, int N, float eps2,
                              const int *idx, int K, float *out)
{
#pragma omp parallel for schedule(static)
    for (int k = 0; k < K; ++k) {
        int i = idx[k];
        float acc[4] = { 0.0f, 0.0f, 0.0f, 0.0f };
        for (int j = 0; j < N; ++j)
            if (i != j) pair_f32(X + 4*i, X + 4*j, acc, eps2);
        out[3*k] = acc[0]; out[3*k+1] = acc[1]; out[3*k+2] = acc[2];
    }
}

/* O3/O4b: the same formula in double precision (the "truth" the fp32 paths
 * are measured against); out is double[3*K].  idx == NULL means targets 0..K). */
void oracle_accel_f64_sampled(const float *X, int N, double eps2,
                              const int *idx, int K, double *out)
{
#pragma omp parallel for schedule(static)
    for (int k = 0; k < K; ++k) {
        int i = idx ? idx[k] : k;
        double xi = X[4*i], yi = X[4*i+1], zi = X[4*i+2];
        double ax = 0.0, ay = 0.0, az = 0.0;
        for (int j = 0; j < N; ++j) {
            if (i == j) continue;
            double rx = (double)X[4*j] - xi, ry = (double)X[4*j+1] - yi, rz = (double)X[4*j+2] - zi;
            double d = rx*rx + ry*ry + rz*rz + eps2;
            double s = (double)X[4*j+3] / (d * sqrt(d));
            ax += rx * s; ay += ry * s; az += rz * s;
        }
        out[3*k] = ax; out[3*k+1] = ay; out[3*k+2] = az;
    }
}

/* Conserved quantity of the reference's half-kick-only scheme:
 * E_eff = sum 1/2 m v^2 - 1/2 * sum_{i<j} m_i m_j / sqrt(r^2 + eps2).
 * out[0] = kinetic, out[1] = full pair potential (sum_{i<j}), E_eff = out[0] + 0.5*out[1]. */
void oracle_energy_f64(const float *X, const float *V, int N, double eps2, double *out)
{
    double ke = 0.0, pe = 0.0;
#pragma omp parallel for schedule(dynamic, 64) reduction(+:ke,pe)
    for (int i = 0; i < N; ++i) {
        double m = X[4*i+3];
        double vx = V[4*i], vy = V[4*i+1], vz = V[4*i+2];
        ke += 0.5 * m * (vx*vx + vy*vy + vz*vz);
        double xi = X[4*i], yi = X[4*i+1], zi = X[4*i+2], p = 0.0;
        for (int j = i + 1; j < N; ++j) {
            double rx = (double)X[4*j] - xi, ry = (double)X[4*j+1] - yi, rz = (double)X[4*j+2] - zi;
            p += (double)X[4*j+3] / sqrt(rx*rx + ry*ry + rz*rz + eps2);
        }
        pe -= m * p;
    }
    out[0] = ke; out[1] = pe;
}

/* Headless-directory flavour of the step, snapshot semantics:
 * TestProject/Sim-Without-OpenGL-Integration/kernel.cu:5-26 (pair math; EPS2
 * is a DOUBLE literal there, so the float sum is widened, eps2 added in double
 * and the result rounded back to float) and :71-77 (integrator with double
 * 0.5*DT: every update is evaluated in double and rounded once).
 * V is float[3*N] (float3), A is not returned by that path (it zeroes gA). */
void oracle_step_headless_f3(float *X, float *V3, int N, double eps2, double dt)
{
    float *S = (float *)malloc((size_t)N * 4 * sizeof(float));
    memcpy(S, X, (size_t)N * 4 * sizeof(float));
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; ++i) {
        const float *bi = S + 4*i;
        float ax = 0.0f, ay = 0.0f, az = 0.0f;
        for (int j = 0; j < N; ++j) {
            /* the GPU kernel has no i != j test; the self term is exactly +0 */
            const float *bj = S + 4*j;
            float rx = bj[0] - bi[0], ry = bj[1] - bi[1], rz = bj[2] - bi[2];
            float d = (float)((double)(rx*rx + ry*ry + rz*rz) + eps2);
            float inv = 1.0f / sqrtf(d * d * d);
            float s = bj[3] * inv;
            ax += rx * s; ay += ry * s; az += rz * s;
        }
        float a[3] = { ax, ay, az };
        for (int c = 0; c < 3; ++c) {
            V3[3*i+c] = (float)((double)V3[3*i+c] + 0.5 * dt * (double)a[c]);
            X[4*i+c]  = (float)((double)X[4*i+c] + dt * (double)V3[3*i+c]);
        }
    }
    free(S);
}

/* Reference initial conditions: TestProject/utils.cpp:6,30-37 with the
 * constants of TestProject/constants.h:15-19.  The reference never calls
 * srand(), which glibc defines as srand(1); we make that explicit so the
 * result does not depend on earlier rand() users in the process. */
void oracle_ic_reference(float *X, int N, float max_xyz, float min_w, float max_w)
{
    srand(1);
    for (int i = 0; i < N; ++i) {
        X[4*i]   = ((float)rand() / RAND_MAX) * (max_xyz - -max_xyz) + -max_xyz;
        X[4*i+1] = ((float)rand() / RAND_MAX) * (max_xyz - -max_xyz) + -max_xyz;
        X[4*i+2] = ((float)rand() / RAND_MAX) * (max_xyz - -max_xyz) + -max_xyz;
        X[4*i+3] = ((float)rand() / RAND_MAX) * (max_w - min_w) + min_w;
    }
}

/* The reference's own acceptance test, TestProject/validation.cpp:143-164
 * (verify_still_bodies): per component |v-x| <= 1% * min(|v|,|x|).
 * Returns the number of bodies the reference would print as "Problem at body". */
int oracle_verify_still_bodies(const float *v, const float *x, int N)
{
    float tolFactor = 1.0 / 100;
    int bad = 0;
    for (int i = 0; i < N; ++i) {
        int fail = 0;
        for (int c = 0; c < 3; ++c) {
            float a = v[4*i+c], b = x[4*i+c];
            float tol = fminf(fabsf(a * tolFactor), fabsf(b * tolFactor));
            if (fabsf(a - b) > tol) fail = 1;
        }
        bad += fail;
    }
    return bad;
}

/* Thread count of every OpenMP region that follows in this process (ours and, sharing the runtime, the
 * reference's own CPU_compute in oracle/_ref/libref_cpu_omp*.so).  bench.py sets it to all host cores
 * explicitly: torchrun exports OMP_NUM_THREADS=1 to its workers. */
void oracle_set_num_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int oracle_num_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* Reduction-flavour step (TestProject/kernel.cu:163-246 kernel_reduction +
 * :406-436 inter_block_reduction; float3 twin :250-318 + :439-469), snapshot
 * semantics.  Pair math as validation.cpp:9-24; the integrator is the one of
 * inter_block_reduction: `v += 0.5 * DT * a` evaluates in DOUBLE (0.5 is a
 * double literal, DT a float macro) and rounds once; `x += DT * v` is float
 * and -- that code exists only as GPU code, built by nvcc with its default
 * -fmad=true -- one fused multiply-add.  PINNED on the reference kernels'
 * own output on a B200 (tests/golden/headless_ref_gpu.npz, made by
 * tests/golden/make_golden_headless_gpu.py through oracle/ref_reduction_glue.cu,
 * which replaces only the OpenGL-bound launch sequence): agreement to summation
 * rounding, <= 4.4e-6 per-body norm-relative at N <= 2048.
 * V, A: float[stride*N], stride 4 (float4, A.w = 0) or 3 (float3). */
void oracle_step_mixed(float *X, float *A, float *V, int N, float eps2, float dt, int stride)
{
    float *S = (float *)malloc((size_t)N * 4 * sizeof(float));
    memcpy(S, X, (size_t)N * 4 * sizeof(float));
#pragma omp parallel for schedule(static)
    for (int i = 0; i < N; ++i) {
        float a[3] = { 0.0f, 0.0f, 0.0f };
        for (int j = 0; j < N; ++j)
            if (i != j) pair_f32(S + 4*i, S + 4*j, a, eps2);
        for (int c = 0; c < 3; ++c) {
            float v = (float)((double)V[stride*i+c] + 0.5 * (double)dt * (double)a[c]);
            V[stride*i+c] = v;
            X[4*i+c] = fmaf(dt, v, X[4*i+c]);
            A[stride*i+c] = a[c];
        }
        if (stride == 4) A[4*i+3] = 0.0f;
    }
    free(S);
}

/* ---- point-splat frame --------------------------------------------------
 * CPU restatement of what the reference draws each step:
 * TestProject/simulation_visualization.cpp:15-37 (vertex shader: weight and
 * NDC position), :40-56 (fragment shader: mix(green, red, weight), alpha 1),
 * :132-135 (depth test on, 1-pixel points), :165-169 (glDrawArrays GL_POINTS
 * in body order), main.cpp:48 (clear colour and depth every frame), uniforms
 * :203-207, window 800x600 (simulation_visualization.h:8-9).
 * Fixed-function steps are restated per the OpenGL 3.3 spec: a point whose
 * clip-space position lies outside [-1,1]^3 is clipped; window coordinates
 * xw = (ndc.x+1)*W/2, yw = (ndc.y+1)*H/2, the 1-pixel point covers pixel
 * (floor(xw), floor(yw)); depth = (ndc.z+1)/2 compared with GL_LESS against a
 * buffer cleared to 1 (kept here as float); colour written as
 * round(c*255).  Rows are stored top-down (image order), GL's y up.
 * depth_test = 0 restates the frame with the test off (last body drawn wins),
 * which is what a driver does if glEnable ran before a context existed
 * (setupOpenGL() is called before glfwInit(), :183-184).
 * PARITY UNPINNED: no GL stack exists on this box to produce a reference frame.
 * rgba: uint8[4*W*H], cleared to (0,0,0,0). */
void oracle_render(const float *X, int N, int W, int H, float max_x, float max_y, float max_z,
                   float min_w, float max_w, int depth_test, unsigned char *rgba)
{
    float *zbuf = (float *)malloc((size_t)W * H * sizeof(float));
    for (int p = 0; p < W * H; ++p) zbuf[p] = 1.0f;
    memset(rgba, 0, (size_t)W * H * 4);
    const float mx[3] = { max_x, max_y, max_z };
    for (int i = 0; i < N; ++i) {
        float ndc[3];
        int inside = 1;
        for (int c = 0; c < 3; ++c) {
            ndc[c] = 2.0f * (X[4*i+c] - (-mx[c])) / (mx[c] - (-mx[c])) - 1.0f;
            if (!(ndc[c] >= -1.0f && ndc[c] <= 1.0f)) inside = 0;
        }
        if (!inside) continue;
        const float xw = (ndc[0] + 1.0f) * ((float)W * 0.5f), yw = (ndc[1] + 1.0f) * ((float)H * 0.5f);
        const int px = (int)floorf(xw), py = (int)floorf(yw);
        if (px < 0 || px >= W || py < 0 || py >= H) continue;   /* ndc == +1 exactly */
        const float depth = (ndc[2] + 1.0f) * 0.5f;
        const int p = (H - 1 - py) * W + px;
        if (depth_test && !(depth < zbuf[p])) continue;
        zbuf[p] = depth;
        float wgt = (X[4*i+3] - min_w) / (max_w - min_w);
        /* mix(a, b, t) = a*(1-t) + b*t per the GLSL spec; written to an 8-bit UNORM target */
        float r = 0.0f * (1.0f - wgt) + 1.0f * wgt, g = 1.0f * (1.0f - wgt) + 0.0f * wgt;
        r = r < 0.0f ? 0.0f : (r > 1.0f ? 1.0f : r);
        g = g < 0.0f ? 0.0f : (g > 1.0f ? 1.0f : g);
        rgba[4*p+0] = (unsigned char)(int)floorf(r * 255.0f + 0.5f);
        rgba[4*p+1] = (unsigned char)(int)floorf(g * 255.0f + 0.5f);
        rgba[4*p+2] = 0;
        rgba[4*p+3] = 255;
    }
    free(zbuf);
}
