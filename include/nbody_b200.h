/*
 * nbody_b200.h -- C ABI of the B200-native all-pairs N-body step.
 *
 * This is the drop-in boundary for the ONE hot path of LienoPC/N-BodySimulation:
 * the O(N^2) softened-gravity step with its fused integrator.  Every entry
 * point names the reference interface it replaces (paths relative to the
 * reference repository root).  Plain pointers and sizes only; no CUDA, C++ or
 * torch types.  "float4 array" below means 4 consecutive floats per body,
 * 16-byte aligned: (x, y, z, mass) for bodies, (vx, vy, vz, pad) for
 * velocities, (ax, ay, az, 0) for accelerations -- the reference's data model
 * (TestProject/main.cpp:232,238-239,250-252).
 *
 * Error convention: every int-returning call gives 0 on success and a negative
 * NBODY_E_* code on failure; nbody_last_error() then holds a message for the
 * calling thread.  The reference instead throws std::runtime_error from its
 * launcher (TestProject/kernel.cu:634,640); include/kernel.cuh converts.
 * There is no CPU fallback: without a CUDA device every compute call fails
 * with NBODY_E_CUDA.
 */
#ifndef NBODY_B200_H
#define NBODY_B200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NBODY_OK          0
#define NBODY_E_INVALID  -1   /* bad argument (N <= 0, eps2 <= 0, null pointer, ...) */
#define NBODY_E_CUDA     -2   /* a CUDA runtime call or kernel launch failed        */
#define NBODY_E_STATE    -3   /* call not valid for this context (e.g. sharded ctx) */

/* integrator arithmetic */
#define NBODY_INTEG_F32   0   /* TestProject/kernel.cu:116-124: float, 0.5f*DT            */
#define NBODY_INTEG_F64   1   /* Sim-Without-OpenGL-Integration/kernel.cu:71-77: double   */
#define NBODY_INTEG_MIXED 2   /* TestProject/kernel.cu:455-463 (reduction flavours):      */
                              /* `0.5 * DT * a` in double, drift `DT * v` in float (FMA)  */

/* summation of a target's N pair terms in the large-N kernels */
#define NBODY_SUM_TWO_LEVEL 0 /* groups of 8 source units (4 096 bodies) summed from zero, then added to a total: */
                              /* 7.7e-7 from the fp64 sum at N = 1 M where one running fp32 sum is at 6.6e-5  */
                              /* and the reference's own sequential sum at 2.6e-4       */
#define NBODY_SUM_SINGLE    1 /* one running sum per target and lane, as the reference's kernel.cu:55-65 has: */
                              /* 2 % faster (ptxas schedules the inner loop better), ~100 x less accurate     */

/* kernel variant: 0 = choose from N; others are listed by nbody_variant_info() */
#define NBODY_VARIANT_AUTO 0

/* position exchange of a multi-device context (nbody_params.n_devices > 1) */
#define NBODY_EXCHANGE_P2P  0 /* fused: the step kernel's epilogue stores the new positions into every    */
                              /* device's next source buffer (NVLink peer stores) and waits for its peers */
                              /* on flags in peer memory; one launch per device per step, no library call */
#define NBODY_EXCHANGE_NCCL 1 /* ncclAllGather of the packed positions on a second stream per device,     */
                              /* overlapped with the own-slice interactions (libnccl.so.2 opened at run   */
                              /* time; NBODY_E_STATE when it is not installed)                            */
#define NBODY_EXCHANGE_P2P_EVENTS 2 /* the fused stores of P2P, the devices ordered by CUDA events between their  */
                              /* streams instead of the in-kernel wait (what P2P itself falls back to for the   */
                              /* small-N kernels); the only form that accepts one device listed more than once  */
#define NBODY_MAX_DEVICES 16

typedef struct nbody_ctx nbody_ctx;

/* Runtime form of TestProject/constants.h:11-13,25-26 (compile-time macros in
 * the reference) plus the sharding a multi-GPU run needs. */
typedef struct nbody_params {
    int   n_bodies;      /* N: number of source bodies (all GPUs)            [N_BODIES 8192]  */
    int   rank;          /* shard index of this context (multi-GPU)          [0]              */
    int   world;         /* number of shards; rank g owns bodies             [1]              */
                         /*   [g*per, min(N,(g+1)*per)), per = even ceil(N/world)             */
    float eps2;          /* softening, added to r^2 directly; must be > 0    [EPS2 0.002f]    */
    float dt;            /* time step                                        [DT 0.1f]        */
    int   integrator;    /* NBODY_INTEG_F32 | NBODY_INTEG_F64                [F32]            */
    int   device;        /* CUDA device ordinal                              [0, as kernel.cu:630] */
    int   variant;       /* kernel variant id, NBODY_VARIANT_AUTO = pick     [0]              */
    int   grid_ctas;     /* persistent CTAs to launch; 0 = SMs x occupancy   [0]              */
    double eps2_f64;     /* used when integrator == F64 (headless constants.h:14)  [0.002]    */
    double dt_f64;       /* used when integrator == F64 (headless constants.h:15)  [0.01]     */
    int   deterministic; /* 1: no timing-based kernel choice (first of the family)    [0]     */
                         /*    -- results are bit-identical either way, see nbody_autotune_report */
    int   n_devices;     /* > 1: ONE context spanning devices[0..n_devices): device g owns    */
                         /*    the g-th slice of the bodies; upload / step / download / energy */
                         /*    / snapshot work on the whole system from one host thread.      */
                         /*    Needs rank 0, world 1.  0 or 1: the single `device` above [0]  */
    int   exchange;      /* NBODY_EXCHANGE_P2P | _NCCL | _P2P_EVENTS (n_devices > 1)  [P2P]   */
    int   devices[NBODY_MAX_DEVICES];   /* CUDA ordinals, all different (see _P2P_EVENTS)     */
    int   summation;     /* NBODY_SUM_TWO_LEVEL | NBODY_SUM_SINGLE                    [TWO_LEVEL] */
} nbody_params;

/* Fills *p with the defaults listed above (TestProject/constants.h). */
int nbody_params_default(nbody_params* p);

/* ---- context API: state stays resident in HBM across steps ------------- */

/* Replaces the allocation half of main(): TestProject/main.cpp:275-283,352-354. */
int nbody_create(nbody_ctx** out, const nbody_params* p);
void nbody_destroy(nbody_ctx* ctx);

/* Host -> device.  x4, v4: float4[N] for ALL bodies (every shard needs all
 * positions as sources and picks the velocities of the bodies it owns).
 * Replaces the cudaMemcpy calls at TestProject/main.cpp:282-283,354. */
int nbody_upload(nbody_ctx* ctx, const float* x4, const float* v4);

/* Bodies owned by this context: [begin, begin+count). */
int nbody_owned_range(nbody_ctx* ctx, int* begin, int* count);

/* Device -> host of the OWNED bodies, float4[count] each (any pointer may be
 * NULL).  The reference has no such call on its headless path (results are
 * dropped, main.cpp:356-366); validation.cpp:79-81 does these copies by hand. */
int nbody_download(nbody_ctx* ctx, float* x4, float* v4, float* a4);

/* nsteps full steps on an unsharded context; returns after the device has
 * finished.  Replaces simulationLoopNoVisual's loop body,
 * TestProject/main.cpp:146-155 -> simulate(), TestProject/kernel.cu:628-645. */
int nbody_step(nbody_ctx* ctx, int nsteps);

/* Same, but only enqueued on the context's stream (no host sync). */
int nbody_step_async(nbody_ctx* ctx, int nsteps);
int nbody_sync(nbody_ctx* ctx);

/* Device time in ms of the last nbody_step / nbody_step_async call (CUDA
 * events on the context's stream).  The reference has no timing. */
int nbody_elapsed_ms(nbody_ctx* ctx, float* ms);

/* SM clock the last stream-K force kernel of this context actually ran at: cycles
 * (clock64) over wall time (globaltimer) between entry and exit of its first CTA.
 * NVML reports the requested clock; under a sustained FP32 load the delivered one
 * is a few per cent lower on some boards (SURVEY section 8d asks for both). */
int nbody_kernel_clock_mhz(nbody_ctx* ctx, float* mhz);

/* Kinetic energy and pair potential sum_{i<j} -m_i m_j/sqrt(r^2+eps2) of the
 * current state, accumulated in double on the device (unsharded contexts).
 * The scheme's conserved quantity is ke + 0.5*pe. */
int nbody_energy(nbody_ctx* ctx, double* ke, double* pe);

/* ---- sharded stepping: one context per GPU, caller owns the exchange ----
 * Each step of rank g:
 *   nbody_shard_local_async   : owned targets x owned sources  -> partial sums
 *   (caller all-gathers the source buffer returned by nbody_shard_sources)
 *   nbody_shard_remote_async  : owned targets x all other sources, integrate,
 *                               write new owned positions into the NEXT source
 *                               buffer, flip buffers.
 * `stream` is a cudaStream_t passed as void* (NULL = the context's stream).  */
int nbody_shard_local_async(nbody_ctx* ctx, void* stream);
int nbody_shard_remote_async(nbody_ctx* ctx, void* stream);
/* Device pointer / geometry of the packed source buffer: which = 0 current
 * (read by the next local/remote calls), 1 = next.  slot_bytes is the size of
 * one rank's slice (all slices equal; own slice starts at rank*slot_bytes). */
int nbody_shard_sources(nbody_ctx* ctx, int which, void** dptr, size_t* total_bytes, size_t* slot_bytes);
/* Fused exchange.  Instead of all-gathering the new positions after the step, the
 * step's epilogue stores each new position into its slot of EVERY rank's next
 * source buffer (16-byte P2P stores over NVLink / NVSwitch): one kernel per step
 * over all sources, no gather, no partial-sum round trip.  The caller needs one
 * cross-rank barrier between steps (any tiny collective on the same stream).
 *   - other processes: exchange nbody_shard_ipc_handles() (2 x 64 bytes, the CUDA IPC
 *     handles of the two source buffers) and attach each peer with nbody_shard_attach_ipc;
 *   - same process (several contexts, e.g. tests): nbody_shard_buffers + nbody_shard_attach_ptr.
 * All ranks must step in lock step from the same upload (buffer parity is shared). */
int nbody_shard_ipc_handles(nbody_ctx* ctx, void* handles128);
int nbody_shard_attach_ipc(nbody_ctx* ctx, int peer_rank, const void* handles128);
int nbody_shard_buffers(nbody_ctx* ctx, void** buf0, void** buf1);
int nbody_shard_attach_ptr(nbody_ctx* ctx, int peer_rank, void* buf0, void* buf1);
int nbody_shard_step_fused_async(nbody_ctx* ctx, void* stream);
/* In-kernel barrier for the fused step (choose before the first fused step, the same on every
 * rank): the step kernel itself waits until every peer has finished the previous step (flags in
 * this GPU's memory, written by the peers' last CTA with st.release.sys over NVLink) and publishes
 * its own completion the same way, so a step is ONE launch with no collective call at all.  Only
 * for one rank per GPU (ranks sharing a GPU would wait on each other).  A wait that sees no
 * progress for 20 s gives up; nbody_shard_barrier_timeouts() != 0 then means the state is invalid. */
int nbody_shard_set_device_barrier(nbody_ctx* ctx, int enable);
int nbody_shard_barrier_timeouts(nbody_ctx* ctx, int* timed_out);
/* Time this GPU's step kernels stood in that flag wait since the last reset (measured by thread 0 of the
 * kernel's first CTA with %globaltimer): total ns, longest single wait, number of waits.  Any pointer may be
 * NULL; reset != 0 zeroes the counters afterwards.  On a multi-device context: the device with the largest total.
 * This is where lock-step on the slowest GPU of a box shows up (the peers' stores themselves are 0.03 % of a step). */
int nbody_shard_barrier_wait_ns(nbody_ctx* ctx, unsigned long long* total_ns, unsigned long long* max_ns,
                                unsigned long long* waits, int reset);
/* nbody_upload / nbody_snapshot_load on a context whose buffers peers are attached to rewrite BOTH source
 * buffers: every rank must have finished its last step (and no rank may start the next) around that call --
 * put a cross-rank barrier on each side.  A multi-device context (n_devices > 1) does this itself. */
/* Geometry helper: bytes of one rank's padded slice for (N, world). */
size_t nbody_shard_slot_bytes(int n_bodies, int world);

/* ---- the reference's launch interface, as C ---------------------------- */

/* What `void simulate(float4* d_bodies, float4* d_accelerations, float4*
 * d_velocity, int N)` (TestProject/kernel.cuh:2, kernel.cu:628-645) forwards
 * to: DEVICE pointers, one step, in place, synchronous, current device. */
int nbody_step_inplace_f4(float* d_bodies4, float* d_accel4, float* d_vel4, int n, float eps2, float dt);

/* What `void simulate(float4* bodies, float3* accelerations, float3* velocity,
 * int N)` (Sim-Without-OpenGL-Integration/kernel.cuh:5, kernel.cu:85-125)
 * forwards to: HOST pointers, float3 velocity, one step including the copies.
 * As in the reference, accelerations are not returned (it zeroes them). */
int nbody_step_host_f3(float* h_bodies4, float* h_accel3, float* h_vel3, int n, double eps2, double dt);

/* One in-place step on DEVICE arrays with every compile-time switch of the
 * reference (TestProject/constants.h:4-7 REDUCTION / FLOAT_3 / FADL) as a
 * run-time option.  This is what the headless forms of the reduction-variant
 * launchers (TestProject/kernel.cuh:4-7, see include/kernel.cuh) forward to. */
typedef struct nbody_step_opts {
    float eps2, dt;      /* [EPS2, DT]                                                     */
    int   integrator;    /* NBODY_INTEG_F32 | NBODY_INTEG_MIXED                            */
    int   va_float3;     /* 1: d_accel / d_vel are float3[N] (FLOAT_3 1, main.cpp:233-269,  */
                         /*    kernel.cu:439-469); 0: float4[N]                            */
    int   variant;       /* kernel variant, NBODY_VARIANT_AUTO = pick from N               */
} nbody_step_opts;
int nbody_step_opts_default(nbody_step_opts* o);
int nbody_step_inplace_opts(float* d_bodies4, void* d_accel, void* d_vel, int n, const nbody_step_opts* o);

/* Devices the in-place forms above (nbody_step_inplace_f4 / _opts, i.e. the reference's simulate()) spread a
 * step over: the arrays stay where the caller put them (the current device, as kernel.cu:630 hard-codes device
 * 0); device g of the list computes the g-th slice of the targets, reading X / V and writing X / V / A of that
 * slice through NVLink peer access, against its own copy of the packed positions (one peer copy per step).
 * So the reference's unmodified caller loop (main.cpp:142-160) scales over the GPUs of a box.  n <= 1 or
 * devices == NULL: current device only (the default).  The first entry must be the device holding the arrays. */
int nbody_set_devices(const int* devices, int n);

/* Releases the per-(device, N) scratch the calls above keep between calls. */
void nbody_release_cached(void);

/* One in-place step on DEVICE float4 arrays like nbody_step_inplace_f4, but ALLOCATING NOTHING: no cudaMalloc,
 * cudaFree or allocating library routine on any call, first or later.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid it).  The step kernel's scratch (packed snapshot of the sources,
 * partial sums of split target tiles, tile counters) is a static 8 MiB device array inside the library
 * (csrc/nbody_noalloc.inl), which every GPU that loads the library holds until the process exits.
 *   - n <= nbody_noalloc_max_bodies() = 16 384; larger n is refused with NBODY_E_INVALID (use the call above).
 *   - the work is ENQUEUED on `stream` (a cudaStream_t passed as void*, NULL = the legacy default stream) on the
 *     current device; the call does not synchronise.
 *   - the first call on a device asks the runtime for the array's address and sets the kernels' attributes (no
 *     allocation, but not stream work either): make that call outside a stream capture.
 *   - all calls on one GPU share the array: calls that can overlap on the GPU (different streams, threads,
 *     graphs replayed side by side) race on it -- serialise them.  Nothing is assumed about the array's contents
 *     between calls (the counters are cleared on `stream` before each step).
 *   - always the stream-K kernel (the one that needs the scratch); nbody_step_inplace_f4 picks its kernel family by
 *     size, so the two agree to summation rounding (accelerations <= 1e-5, positions <= 1e-6), not bit for bit. */
int nbody_step_inplace_noalloc_f4(float* d_bodies4, float* d_accel4, float* d_vel4, int n, float eps2, float dt, void* stream);
int nbody_noalloc_max_bodies(void);

/* ---- headless frame: the reference's per-step OpenGL draw as CUDA --------
 * Shader arithmetic of TestProject/simulation_visualization.cpp:15-56 (weight,
 * NDC position, mix(green, red, weight)), 1-pixel points with the depth test
 * of :132-135, window 800x600 (simulation_visualization.h:8-9), cleared every
 * frame (main.cpp:48).  The GL/GLFW path itself is unavailable on this box;
 * the frame is a device RGBA8 buffer, rows top-down, 4*width*height bytes. */
typedef struct nbody_view {
    int   width, height;         /* [800, 600]  WINDOW_WIDTH / WINDOW_HEIGHT                  */
    float max_x, max_y, max_z;   /* [200000]    MAX_VIEWX/Y/Z, constants.h:21-23 (uniforms)   */
    float min_w, max_w;          /* [1e5, 1e9]  MIN_W / MAX_W, constants.h:18-19 (uniforms)   */
    int   depth_test;            /* [1] GL_LESS; 0 = off (the last body drawn wins a pixel)   */
} nbody_view;
int nbody_view_default(nbody_view* v);
/* DEVICE float4 positions in, DEVICE RGBA8 frame out; synchronous, current device.
 * Replaces renderBodies (simulation_visualization.cpp:165-169) + both shaders. */
int nbody_render_device(const float* d_bodies4, int n, const nbody_view* v, unsigned char* d_rgba);
/* Frame of the context's current positions into a HOST buffer (unsharded contexts). */
int nbody_render(nbody_ctx* ctx, const nbody_view* v, unsigned char* h_rgba);
/* Writes a host RGBA8 frame as a binary PPM (alpha dropped). */
int nbody_write_ppm(const char* path, const unsigned char* h_rgba, int width, int height);

/* ---- run records: snapshot / resume ---------------------------------------
 * The reference drops its results (main.cpp:356-366).  A snapshot is the raw
 * float4 X and V arrays (main.cpp:250-251) after steps_done steps behind a
 * 64-byte header (csrc/nbody_records.inl).  load works on sharded contexts too
 * (every rank reads the file and keeps its slice); save needs an unsharded one. */
int nbody_snapshot_save(nbody_ctx* ctx, const char* path);
int nbody_snapshot_load(nbody_ctx* ctx, const char* path);
int nbody_snapshot_info(const char* path, int* n_bodies, long long* steps_done, float* eps2, float* dt);
/* Full steps taken since the last upload / snapshot load. */
int nbody_steps_done(nbody_ctx* ctx, long long* steps);

/* ---- introspection ------------------------------------------------------ */
const char* nbody_last_error(void);
/* Kernel launches issued by this library in this process (all contexts). */
long long nbody_launch_count(void);
/* Describes variant `id` into buf (empty string for an unknown id); returns the
 * number of variants.  nbody_variant_id(i), i = 0..count-1, lists the ids. */
int nbody_variant_info(int id, char* buf, size_t buflen);
int nbody_variant_id(int index);
/* "id:ms id:ms ..." of the large-N candidates timed when the first large context of
 * this process resolved NBODY_VARIANT_AUTO (empty before that).  The candidates run the
 * same instructions per interaction in different source order; which register
 * allocation ptxas gave each is worth +-3.5 %, so the engine measures instead of
 * guessing.  Environment NBODY_LARGE_VARIANT=<id> forces the choice. */
int nbody_autotune_report(char* buf, size_t buflen);
/* The kernel variant NBODY_VARIANT_AUTO resolves to for a context owning n_targets of n_sources
 * bodies on a device with sm_count SMs, by the static rules alone (small-N family below 8 192
 * sources, tile shape x unit size by unit balance up to 65 536, one of the measured candidates
 * above -- for that last case the id returned is the first candidate and *measured is set to 1:
 * the real choice is timed on the device).  Host-only; needs no GPU. */
int nbody_choose_variant(int n_targets, int n_sources, int sm_count, int* measured);
/* Variant / grid actually used by ctx (after AUTO resolution). */
int nbody_ctx_info(nbody_ctx* ctx, int* variant, int* grid_ctas, int* ctas_per_sm, int* sm_count);
/* *spread = 1 when the step kernel of ctx finishes a target tile whose sources were cut between CTAs in ALL of those
 * CTAs (each sums and integrates its share of the tile's targets; cooperative launch), 0 when the last CTA to arrive
 * does it alone.  Chosen for mid-size N (8 192 .. 20 480 bodies: 14 or more cuts per tile); the sums and their order are
 * the same either way, so the results are bit-identical.  NBODY_NO_SPREAD_FIXUP=1 in the environment turns it off. */
int nbody_ctx_fixup(nbody_ctx* ctx, int* spread);
/* Multi-device context: number of devices (1 for a plain context), NBODY_EXCHANGE_* in use, and whether
 * the cross-device barrier lives in the step kernel (1) or is expressed with events between streams (0). */
int nbody_group_info(nbody_ctx* ctx, int* n_devices, int* exchange, int* in_kernel_barrier);

/* ---- initial conditions (host side; TestProject/utils.cpp:30-37) -------- */
/* Reference distribution, seeded counter-based generator: positions
 * U[-max_xyz,max_xyz]^3, mass U[min_w,max_w] (constants.h:15-19). */
int nbody_ic_uniform(float* x4, int n, unsigned long long seed, float max_xyz, float min_w, float max_w);
/* Plummer sphere in N-body units (M=1, a=3pi/16), velocities scaled by
 * sqrt(1/2) for the scheme's effective G = 1/2. */
int nbody_ic_plummer(float* x4, float* v4, int n, unsigned long long seed);

#ifdef __cplusplus
}
#endif
#endif /* NBODY_B200_H */
