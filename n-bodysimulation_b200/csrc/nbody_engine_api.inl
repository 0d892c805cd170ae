// nbody_engine_api.inl -- second half of nbody_engine.cu: data movement, stepping,
// the reference-shaped one-call entry points and the initial conditions.

namespace {

// multi-device contexts (nbody_group.inl)
int group_upload(nbody_ctx* c, const float* x4, const float* v4);
int group_download(nbody_ctx* c, float* x4, float* v4, float* a4);
int group_step_async(nbody_ctx* c, int nsteps);
int group_sync(nbody_ctx* c);
int group_elapsed_ms(nbody_ctx* c, float* ms);
int group_energy(nbody_ctx* c, double* ke, double* pe);
int group_barrier_timeouts(nbody_ctx* c, int* timed_out);
int group_barrier_wait_ns(nbody_ctx* c, unsigned long long* total_ns, unsigned long long* max_ns, unsigned long long* waits, int reset);
inline bool is_group(const nbody_ctx* c) { return !c->kids.empty(); }
#define NOT_FOR_GROUP(c) do { if (is_group(c)) return fail(NBODY_E_STATE, "%s: not valid on a multi-device context (it drives its own shards)", __func__); } while (0)

// one unsharded step: all sources, fused integrate, new positions into the other buffer
int enqueue_full_step(nbody_ctx* c, cudaStream_t st)
{
    ForceArgs a = base_args(c);
    a.ra_begin = 0; a.ra_count = units_in_use(c);
    a.rb_begin = 0; a.rb_count = 0;
    a.last_unit_bodies = last_unit_fill(c);
    a.next_src = c->packed[c->cur ^ 1];
    a.next_off = c->p.rank * c->slot_bodies;
    a.maxseg = c->plan_full.maxseg;
    int rc = launch_force(c->var, c->plan_full, a, st);
    if (rc) return rc;
    c->cur ^= 1;
    ++c->steps_done;
    return NBODY_OK;
}

// A step of a few microseconds is bound by launch latency, not by the kernel: replay kGraphSteps
// steps as one CUDA graph.  Captured lazily, once per context, with buffer parity 0 at entry
// (kGraphSteps is even, so parity is 0 again at exit).  Returns false when capture is not possible;
// the caller then launches step by step.
constexpr int kGraphSteps = 32;
constexpr int kGraphMaxBodies = 16384;   // above this a step is > 100 us and launches hide behind it

bool ensure_step_graph(nbody_ctx* c)
{
    if (c->step_graph) return true;
    if (c->graph_failed || c->p.world != 1 || c->n > kGraphMaxBodies || c->cur != 0) return false;
    cudaGraph_t g = nullptr;
    if (cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) { c->graph_failed = true; cudaGetLastError(); return false; }
    bool ok = true;
    const long long steps0 = c->steps_done;
    for (int s = 0; s < kGraphSteps && ok; ++s) ok = enqueue_full_step(c, c->stream) == NBODY_OK;
    const cudaError_t e = cudaStreamEndCapture(c->stream, &g);
    c->steps_done = steps0;          // nothing ran yet
    c->cur = 0;                      // kGraphSteps is even
    g_launches.fetch_sub(ok ? kGraphSteps : 0, std::memory_order_relaxed);
    if (ok && e == cudaSuccess && cudaGraphInstantiate(&c->step_graph, g, 0) == cudaSuccess) {
        cudaGraphDestroy(g);
        return true;
    }
    if (g) cudaGraphDestroy(g);
    cudaGetLastError();
    c->step_graph = nullptr;
    c->graph_failed = true;
    return false;
}

// host -> device of one context, enqueued on its stream (nbody_upload synchronises; a multi-device
// context enqueues on every device first, so the copies of the G devices overlap)
int upload_enqueue(nbody_ctx* c, const float* x4, const float* v4)
{
    int rc = set_device(c);
    if (rc) return rc;
    // stage all N positions once, pack each rank's slice into its padded slot
    float4* stage = c->stage;   // persistent: a cudaMalloc/cudaFree per call costs up to 100 ms (profiles/r1 notes)
    cudaError_t e = cudaMemcpyAsync(stage, x4, (size_t)c->n * sizeof(float4), cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(c->packed[0], 0, (size_t)c->n_src_pad * sizeof(float4), c->stream);
    if (e == cudaSuccess) e = cudaMemsetAsync(c->packed[1], 0, (size_t)c->n_src_pad * sizeof(float4), c->stream);
    c->cur = 0;
    c->steps_done = 0;
    for (int r = 0; r < c->p.world && e == cudaSuccess && rc == 0; ++r) {
        const int b = r * c->per;
        const int cnt = b >= c->n ? 0 : (c->n - b < c->per ? c->n - b : c->per);
        if (cnt > 0) rc = launch_pack(stage + b, cnt, c->packed[0], r * c->slot_bodies, round_up(cnt, 2), c->stream);
    }
    if (e == cudaSuccess && rc == 0)
        e = cudaMemcpyAsync(c->X, stage + c->t_begin, (size_t)c->n_targets * sizeof(float4), cudaMemcpyDeviceToDevice, c->stream);
    if (e == cudaSuccess && rc == 0) {
        if (v4) e = cudaMemcpyAsync(c->V, v4 + 4 * (size_t)c->t_begin, (size_t)c->n_targets * sizeof(float4), cudaMemcpyHostToDevice, c->stream);
        else e = cudaMemsetAsync(c->V, 0, (size_t)c->n_targets * sizeof(float4), c->stream);
    }
    if (rc) return rc;
    if (e != cudaSuccess) return fail(NBODY_E_CUDA, "upload failed: %s", cudaGetErrorString(e));
    return NBODY_OK;
}

// kinetic energy and pair potential of n bodies in device arrays X, V on c's device and stream
int energy_of(nbody_ctx* c, const float4* X, const float4* V, int n, double eps2, double* ke, double* pe)
{
    int rc = set_device(c);
    if (rc) return rc;
    const int blocks = (n + 255) / 256;
    if (!c->energy_part) CU(cudaMalloc(&c->energy_part, (size_t)blocks * 2 * sizeof(double)));
    energy_kernel<<<blocks, 256, 0, c->stream>>>(X, V, n, eps2, c->energy_part);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CU(cudaGetLastError());
    std::vector<double> h((size_t)blocks * 2);
    CU(cudaMemcpyAsync(h.data(), c->energy_part, h.size() * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    double k = 0.0, p = 0.0;
    for (int b = 0; b < blocks; ++b) { k += h[2 * b]; p += h[2 * b + 1]; }
    *ke = k;
    *pe = 0.5 * p;   // ordered pairs -> i<j
    return NBODY_OK;
}

// device -> host of the owned bodies, enqueued on the context's stream (ordered after its steps)
int download_enqueue(nbody_ctx* c, float* x4, float* v4, float* a4)
{
    int rc = set_device(c);
    if (rc) return rc;
    const size_t tb = (size_t)c->n_targets * sizeof(float4);
    if (x4) CU(cudaMemcpyAsync(x4, c->X, tb, cudaMemcpyDeviceToHost, c->stream));
    if (v4) CU(cudaMemcpyAsync(v4, c->V, tb, cudaMemcpyDeviceToHost, c->stream));
    if (a4) CU(cudaMemcpyAsync(a4, c->A, tb, cudaMemcpyDeviceToHost, c->stream));
    return NBODY_OK;
}

}  // namespace

extern "C" {

int nbody_upload(nbody_ctx* c, const float* x4, const float* v4)
{
    TRACE_CALL();
    if (!c || !x4) return fail(NBODY_E_INVALID, "null argument");
    if (is_group(c)) return group_upload(c, x4, v4);
    int rc = upload_enqueue(c, x4, v4);
    if (rc) return rc;
    CU(cudaStreamSynchronize(c->stream));
    return NBODY_OK;
}

int nbody_download(nbody_ctx* c, float* x4, float* v4, float* a4)
{
    TRACE_CALL();
    if (!c) return fail(NBODY_E_INVALID, "null context");
    if (is_group(c)) return group_download(c, x4, v4, a4);
    int rc = download_enqueue(c, x4, v4, a4);
    if (rc) return rc;
    CU(cudaStreamSynchronize(c->stream));
    return NBODY_OK;
}

int nbody_step_async(nbody_ctx* c, int nsteps)
{
    TRACE_CALL();
    if (!c) return fail(NBODY_E_INVALID, "null context");
    if (nsteps < 0) return fail(NBODY_E_INVALID, "nsteps must be >= 0");
    if (is_group(c)) return group_step_async(c, nsteps);
    if (c->p.world != 1) return fail(NBODY_E_STATE, "nbody_step on a sharded context; use nbody_shard_*");
    int rc = set_device(c);
    if (rc) return rc;
    CU(cudaEventRecord(c->ev0, c->stream));
    int s = 0;
    if (nsteps >= kGraphSteps + 1 && c->n <= kGraphMaxBodies && !c->graph_failed) {
        if (c->cur != 0) {           // graphs start from buffer parity 0
            rc = enqueue_full_step(c, c->stream);
            if (rc) return rc;
            ++s;
        }
        if (ensure_step_graph(c)) {
            for (; s + kGraphSteps <= nsteps; s += kGraphSteps) {
                CU(cudaGraphLaunch(c->step_graph, c->stream));
                g_launches.fetch_add(kGraphSteps, std::memory_order_relaxed);   // kernels run, not API calls
                c->steps_done += kGraphSteps;
            }
        }
    }
    for (; s < nsteps; ++s) {
        rc = enqueue_full_step(c, c->stream);
        if (rc) return rc;
    }
    CU(cudaEventRecord(c->ev1, c->stream));
    c->timed = true;
    return NBODY_OK;
}

int nbody_sync(nbody_ctx* c)
{
    TRACE_CALL();
    if (!c) return fail(NBODY_E_INVALID, "null context");
    if (is_group(c)) return group_sync(c);
    int rc = set_device(c);
    if (rc) return rc;
    CU(cudaStreamSynchronize(c->stream));
    return NBODY_OK;
}

int nbody_step(nbody_ctx* c, int nsteps)
{
    int rc = nbody_step_async(c, nsteps);
    if (rc) return rc;
    return nbody_sync(c);
}

int nbody_elapsed_ms(nbody_ctx* c, float* ms)
{
    if (!c || !ms) return fail(NBODY_E_INVALID, "null argument");
    if (is_group(c)) return group_elapsed_ms(c, ms);
    if (!c->timed) return fail(NBODY_E_STATE, "no timed call yet");
    CU(cudaEventSynchronize(c->ev1));
    CU(cudaEventElapsedTime(ms, c->ev0, c->ev1));
    return NBODY_OK;
}

int nbody_kernel_clock_mhz(nbody_ctx* c, float* mhz)
{
    if (!c || !mhz) return fail(NBODY_E_INVALID, "null argument");
    if (is_group(c)) c = c->kids[0];
    int rc = set_device(c);
    if (rc) return rc;
    unsigned long long h[4];
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaMemcpy(h, c->clk, sizeof h, cudaMemcpyDeviceToHost));
    if (h[3] <= h[1] || h[2] <= h[0]) return fail(NBODY_E_STATE, "no stream-K force kernel has run on this context yet");
    *mhz = (float)((double)(h[2] - h[0]) / (double)(h[3] - h[1]) * 1e3);   // cycles per ns -> MHz
    return NBODY_OK;
}

int nbody_energy(nbody_ctx* c, double* ke, double* pe)
{
    TRACE_CALL();
    if (!c || !ke || !pe) return fail(NBODY_E_INVALID, "null argument");
    if (is_group(c)) return group_energy(c, ke, pe);
    if (c->p.world != 1) return fail(NBODY_E_STATE, "nbody_energy needs an unsharded context");
    const double eps2 = c->p.integrator == NBODY_INTEG_F64 ? c->p.eps2_f64 : (double)c->p.eps2;
    return energy_of(c, c->X, c->V, c->n, eps2, ke, pe);
}

// ---- sharded stepping ---------------------------------------------------------------------------------
int nbody_shard_sources(nbody_ctx* c, int which, void** dptr, size_t* total_bytes, size_t* slot_bytes)
{
    if (!c || (which != 0 && which != 1)) return fail(NBODY_E_INVALID, "bad argument");
    NOT_FOR_GROUP(c);
    if (dptr) *dptr = c->packed[c->cur ^ which];
    if (total_bytes) *total_bytes = (size_t)c->n_src_pad * sizeof(float4);
    if (slot_bytes) *slot_bytes = (size_t)c->slot_bodies * sizeof(float4);
    return NBODY_OK;
}

int nbody_shard_local_async(nbody_ctx* c, void* stream)
{
    TRACE_CALL();
    if (!c) return fail(NBODY_E_INVALID, "null context");
    NOT_FOR_GROUP(c);
    int rc = set_device(c);
    if (rc) return rc;
    cudaStream_t st = stream ? (cudaStream_t)stream : c->stream;
    if (c->p.world == 1) return enqueue_full_step(c, st);   // nothing remote: one fused launch
    const int UJ = c->var->UJ, own = c->slot_bodies / UJ;
    ForceArgs a = base_args(c);
    a.ra_begin = c->p.rank * own; a.ra_count = units_own_in_use(c);
    a.accum_out = c->accum;
    a.maxseg = c->plan_local.maxseg;
    return launch_force(c->var, c->plan_local, a, st);
}

int nbody_shard_remote_async(nbody_ctx* c, void* stream)
{
    TRACE_CALL();
    if (!c) return fail(NBODY_E_INVALID, "null context");
    NOT_FOR_GROUP(c);
    int rc = set_device(c);
    if (rc) return rc;
    if (c->p.world == 1) return NBODY_OK;
    cudaStream_t st = stream ? (cudaStream_t)stream : c->stream;
    const int UJ = c->var->UJ, own = c->slot_bodies / UJ, total = units_in_use(c);
    ForceArgs a = base_args(c);
    a.ra_begin = 0; a.ra_count = c->p.rank * own;
    a.rb_begin = (c->p.rank + 1) * own; a.rb_count = total > a.rb_begin ? total - a.rb_begin : 0;   // the last shard has nothing behind it
    a.accum_in = c->accum;
    a.next_src = c->packed[c->cur ^ 1];
    a.next_off = c->p.rank * c->slot_bodies;
    a.maxseg = c->plan_remote.maxseg;
    rc = launch_force(c->var, c->plan_remote, a, st);
    if (rc) return rc;
    c->cur ^= 1;
    ++c->steps_done;
    return NBODY_OK;
}

// ---- fused exchange: epilogue stores into the peers' next buffers -----------------------------------
int nbody_shard_ipc_handles(nbody_ctx* c, void* handles128)
{
    if (!c || !handles128) return fail(NBODY_E_INVALID, "null argument");
    NOT_FOR_GROUP(c);
    int rc = set_device(c);
    if (rc) return rc;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handles are 64 bytes");
    cudaIpcMemHandle_t h[2];
    CU(cudaIpcGetMemHandle(&h[0], c->packed[0]));
    CU(cudaIpcGetMemHandle(&h[1], c->packed[1]));
    memcpy(handles128, h, sizeof h);
    return NBODY_OK;
}

static int attach_checks(nbody_ctx* c, int peer)
{
    if (!c) return fail(NBODY_E_INVALID, "null context");
    NOT_FOR_GROUP(c);
    if (c->p.world < 2) return fail(NBODY_E_STATE, "an unsharded context has no peers");
    if (c->p.world > kMaxPeers + 1) return fail(NBODY_E_STATE, "fused exchange supports up to %d ranks", kMaxPeers + 1);
    if (peer < 0 || peer >= c->p.world || peer == c->p.rank) return fail(NBODY_E_INVALID, "bad peer rank %d (this is rank %d of %d)", peer, c->p.rank, c->p.world);
    if (c->peer_packed[peer][0]) return fail(NBODY_E_STATE, "peer %d is already attached", peer);
    return set_device(c);
}

int nbody_shard_attach_ipc(nbody_ctx* c, int peer, const void* handles128)
{
    int rc = attach_checks(c, peer);
    if (rc) return rc;
    if (!handles128) return fail(NBODY_E_INVALID, "null handles");
    cudaIpcMemHandle_t h[2];
    memcpy(h, handles128, sizeof h);
    void *p0 = nullptr, *p1 = nullptr;
    CU(cudaIpcOpenMemHandle(&p0, h[0], cudaIpcMemLazyEnablePeerAccess));
    cudaError_t e = cudaIpcOpenMemHandle(&p1, h[1], cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { cudaIpcCloseMemHandle(p0); return fail(NBODY_E_CUDA, "cudaIpcOpenMemHandle: %s", cudaGetErrorString(e)); }
    c->peer_packed[peer][0] = (float4*)p0; c->peer_packed[peer][1] = (float4*)p1;
    c->peer_ipc[peer] = true;
    ++c->peers_attached;
    return NBODY_OK;
}

int nbody_shard_attach_ptr(nbody_ctx* c, int peer, void* buf0, void* buf1)
{
    int rc = attach_checks(c, peer);
    if (rc) return rc;
    if (!buf0 || !buf1) return fail(NBODY_E_INVALID, "null buffer");
    // same process, other device: the stores need peer access from this context's device
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, buf0) == cudaSuccess && at.type == cudaMemoryTypeDevice && at.device != c->p.device) {
        int can = 0;
        CU(cudaDeviceCanAccessPeer(&can, c->p.device, at.device));
        if (!can) return fail(NBODY_E_CUDA, "device %d cannot access device %d (no P2P path)", c->p.device, at.device);
        const cudaError_t e = cudaDeviceEnablePeerAccess(at.device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail(NBODY_E_CUDA, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
        cudaGetLastError();
    }
    c->peer_packed[peer][0] = (float4*)buf0; c->peer_packed[peer][1] = (float4*)buf1;
    ++c->peers_attached;
    return NBODY_OK;
}

int nbody_shard_buffers(nbody_ctx* c, void** buf0, void** buf1)
{
    if (!c || !buf0 || !buf1) return fail(NBODY_E_INVALID, "null argument");
    NOT_FOR_GROUP(c);
    *buf0 = c->packed[0]; *buf1 = c->packed[1];
    return NBODY_OK;
}

int nbody_shard_step_fused_async(nbody_ctx* c, void* stream)
{
    TRACE_CALL();
    if (!c) return fail(NBODY_E_INVALID, "null context");
    NOT_FOR_GROUP(c);
    int rc = set_device(c);
    if (rc) return rc;
    cudaStream_t st = stream ? (cudaStream_t)stream : c->stream;
    if (c->p.world == 1) return enqueue_full_step(c, st);
    if (c->peers_attached != c->p.world - 1) return fail(NBODY_E_STATE, "%d of %d peers attached", c->peers_attached, c->p.world - 1);
    ForceArgs a = base_args(c);
    a.ra_begin = 0; a.ra_count = units_in_use(c);
    a.last_unit_bodies = last_unit_fill(c);
    a.next_src = c->packed[c->cur ^ 1];
    a.next_off = c->p.rank * c->slot_bodies;
    a.maxseg = c->plan_full.maxseg;
    const size_t sb = (size_t)c->n_src_pad * sizeof(float4);
    for (int r = 0; r < c->p.world; ++r) {
        if (r == c->p.rank) continue;
        if (c->device_barrier) a.bar_peer_flags[a.n_peer] = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(c->peer_packed[r][0]) + sb);
        a.peer_next[a.n_peer++] = c->peer_packed[r][c->cur ^ 1];
    }
    if (c->device_barrier) {
        a.bar_flags = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(c->packed[0]) + sb);
        a.bar_step = c->fused_steps;
        a.bar_done = c->bar_done;
        a.bar_err = reinterpret_cast<int*>(c->bar_done) + 16;
        a.rank = c->p.rank; a.world = c->p.world;
    }
    rc = launch_force(c->var, c->plan_full, a, st);
    if (rc) return rc;
    c->cur ^= 1;
    ++c->steps_done;
    ++c->fused_steps;
    return NBODY_OK;
}

int nbody_shard_set_device_barrier(nbody_ctx* c, int enable)
{
    if (!c) return fail(NBODY_E_INVALID, "null context");
    NOT_FOR_GROUP(c);
    if (enable && c->p.world > 1 && c->var->small)
        return fail(NBODY_E_STATE, "variant %s (small-N family) has no in-kernel barrier; synchronise the ranks from the caller", c->var->name);
    if (c->fused_steps != 0 && (enable != 0) != c->device_barrier)
        return fail(NBODY_E_STATE, "the barrier mode must be chosen before the first fused step (every rank the same)");
    c->device_barrier = enable != 0;
    return NBODY_OK;
}

int nbody_shard_barrier_timeouts(nbody_ctx* c, int* timed_out)
{
    if (!c || !timed_out) return fail(NBODY_E_INVALID, "null argument");
    if (is_group(c)) return group_barrier_timeouts(c, timed_out);
    int rc = set_device(c);
    if (rc) return rc;
    CU(cudaMemcpy(timed_out, reinterpret_cast<int*>(c->bar_done) + 16, sizeof(int), cudaMemcpyDeviceToHost));
    return NBODY_OK;
}

int nbody_shard_barrier_wait_ns(nbody_ctx* c, unsigned long long* total_ns, unsigned long long* max_ns, unsigned long long* waits, int reset)
{
    if (!c) return fail(NBODY_E_INVALID, "null context");
    if (is_group(c)) return group_barrier_wait_ns(c, total_ns, max_ns, waits, reset);
    int rc = set_device(c);
    if (rc) return rc;
    unsigned long long h[3] = {0, 0, 0};
    CU(cudaDeviceSynchronize());   // the steps may run on a caller's stream: wait for all of this device
    CU(cudaMemcpy(h, c->clk + 4, sizeof h, cudaMemcpyDeviceToHost));
    if (total_ns) *total_ns = h[0];
    if (max_ns) *max_ns = h[1];
    if (waits) *waits = h[2];
    if (reset) CU(cudaMemset(c->clk + 4, 0, sizeof h));
    return NBODY_OK;
}

}  // extern "C"

// ---- reference-shaped one-call entry points ------------------------------------------------------------
namespace {

struct Cached {
    const Variant* var = nullptr;
    Plan plan;
    float4* packed = nullptr;
    float4* partial = nullptr;
    unsigned* counters = nullptr;
    float4 *X = nullptr, *V = nullptr, *A = nullptr;   // host-pointer flavour only
    int n_pad = 0;
    // launch-bound sizes: pack + step kernel of ONE call as a two-node CUDA graph, re-used while the caller passes the
    // same arrays and parameters (the reference's loop does, main.cpp:146-155): one graph launch instead of two
    // kernel launches per simulate()
    cudaGraphExec_t call_graph = nullptr;
    cudaStream_t capture_stream = nullptr;
    bool graph_failed = false;
    struct CallKey { void *x, *a, *v; float eps2, dt; int integ, f3; double dt64; } key{};
};
void release_render_scratch();   // nbody_render.inl
std::mutex g_cache_mu;
std::map<std::pair<std::pair<int, int>, int>, Cached> g_cache;   // ((device, N), variant)

void free_cached(Cached& k)
{
    if (k.call_graph) cudaGraphExecDestroy(k.call_graph);
    if (k.capture_stream) cudaStreamDestroy(k.capture_stream);
    cudaFree(k.packed); cudaFree(k.partial); cudaFree(k.counters); cudaFree(k.X); cudaFree(k.V); cudaFree(k.A);
    k = Cached();
}

int get_cached(int n, bool host_buffers, Cached** out, int variant = NBODY_VARIANT_AUTO)
{
    int dev = 0;
    CU(cudaGetDevice(&dev));
    Cached& k = g_cache[std::make_pair(std::make_pair(dev, n), variant)];
    if (!k.var) {
        cudaDeviceProp prop;
        CU(cudaGetDeviceProperties(&prop, dev));
        if (prop.major < 10) return fail(NBODY_E_CUDA, "device %d is sm_%d%d; this library is built for sm_100a only", dev, prop.major, prop.minor);
        k.var = find_variant(variant == NBODY_VARIANT_AUTO ? auto_variant(n, n) : variant);
        if (!k.var) { g_cache.erase(std::make_pair(std::make_pair(dev, n), variant)); return fail(NBODY_E_INVALID, "unknown kernel variant %d", variant); }
        k.n_pad = round_up(n, kSlotAlign);
        int rc = make_plan(k.var, prop.multiProcessorCount, 0, n, round_up(n, k.var->UJ) / k.var->UJ, &k.plan);   // units up to the last body (units_in_use)
        if (rc) { k = Cached(); return rc; }
        cudaError_t e = cudaMalloc(&k.packed, (size_t)k.n_pad * sizeof(float4));
        if (e == cudaSuccess) e = cudaMalloc(&k.partial, k.plan.partial_bytes);
        if (e == cudaSuccess) e = cudaMalloc(&k.counters, (size_t)k.plan.num_counters * sizeof(unsigned));
        if (e == cudaSuccess) e = cudaMemset(k.counters, 0, (size_t)k.plan.num_counters * sizeof(unsigned));
        if (e != cudaSuccess) { free_cached(k); return fail(NBODY_E_CUDA, "scratch allocation failed: %s", cudaGetErrorString(e)); }
    }
    if (host_buffers && !k.X) {
        CU(cudaMalloc(&k.X, (size_t)n * sizeof(float4)));
        CU(cudaMalloc(&k.V, (size_t)n * sizeof(float4)));
        CU(cudaMalloc(&k.A, (size_t)n * sizeof(float4)));
    }
    *out = &k;
    return NBODY_OK;
}

int enqueue_one_step(Cached* k, float4* dX, float4* dA, float4* dV, int n, float eps2, float dt, int f64, double dt64, int va_f3, cudaStream_t st)
{
    // snapshot of the step's start in packed form; X is then advanced in place
    int rc = launch_pack(dX, n, k->packed, 0, k->n_pad, st);
    if (rc) return rc;
    ForceArgs a{};
    a.X = dX; a.V = dV; a.A = dA;
    a.n_targets = n;
    a.src = k->packed;
    a.ra_begin = 0; a.ra_count = round_up(n, k->var->UJ) / k->var->UJ;
    a.last_unit_bodies = n % k->var->UJ;
    a.partial = k->partial; a.counters = k->counters; a.maxseg = k->plan.maxseg;
    a.eps2 = eps2; a.dt = dt; a.integ_f64 = f64; a.dt64 = dt64; a.va_f3 = va_f3;
    return launch_force(k->var, k->plan, a, st);
}

int one_step_on_device(Cached* k, float4* dX, float4* dA, float4* dV, int n, float eps2, float dt, int f64, double dt64, int va_f3 = 0)
{
    if (n <= kGraphMaxBodies && !k->graph_failed) {
        const Cached::CallKey key{dX, dA, dV, eps2, dt, f64, va_f3, dt64};
        if (!k->call_graph || memcmp(&key, &k->key, sizeof key) != 0) {
            if (k->call_graph) { cudaGraphExecDestroy(k->call_graph); k->call_graph = nullptr; }
            cudaGraph_t g = nullptr;
            bool ok = (k->capture_stream || cudaStreamCreateWithFlags(&k->capture_stream, cudaStreamNonBlocking) == cudaSuccess) &&
                      cudaStreamBeginCapture(k->capture_stream, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
            if (ok) {
                const bool enq = enqueue_one_step(k, dX, dA, dV, n, eps2, dt, f64, dt64, va_f3, k->capture_stream) == NBODY_OK;
                ok = cudaStreamEndCapture(k->capture_stream, &g) == cudaSuccess && enq;
                g_launches.fetch_sub(enq ? 2 : 0, std::memory_order_relaxed);   // nothing ran yet
                ok = ok && cudaGraphInstantiate(&k->call_graph, g, 0) == cudaSuccess;
                if (g) cudaGraphDestroy(g);
            }
            if (!ok) { cudaGetLastError(); k->call_graph = nullptr; k->graph_failed = true; }
            else memcpy(&k->key, &key, sizeof key);
        }
        if (k->call_graph) {
            if (cudaGraphLaunch(k->call_graph, 0) == cudaSuccess) {   // the caller's (legacy default) stream, like the reference's launch
                g_launches.fetch_add(2, std::memory_order_relaxed);   // kernels run, not API calls
                return NBODY_OK;
            }
            cudaGetLastError();   // not launchable here: plain launches from now on
            cudaGraphExecDestroy(k->call_graph);
            k->call_graph = nullptr;
            k->graph_failed = true;
        }
    }
    return enqueue_one_step(k, dX, dA, dV, n, eps2, dt, f64, dt64, va_f3, 0);
}

// ---- the in-place step spread over several GPUs (nbody_set_devices) --------------------------------------
// The caller's arrays stay on the first device (the reference's launcher knows only device 0, kernel.cu:630).
// Device g computes the g-th slice of the targets: its kernel reads X / V and writes X / V / A of that slice
// straight in the first device's memory (NVLink peer loads / stores, 80 B per body and step) and streams the
// sources from its own copy of the packed snapshot (one peer copy of 16 N bytes per step).
struct MultiDev {
    int device = 0, t_begin = 0, n_targets = 0;
    Plan plan;
    float4 *packed = nullptr, *partial = nullptr;
    unsigned* counters = nullptr;
    cudaStream_t st = nullptr;   // null on the first device: it works on the caller's (legacy default) stream
};
struct CachedMulti {
    const Variant* var = nullptr;
    int n_pad = 0;
    cudaEvent_t packed_ready = nullptr;
    std::vector<MultiDev> d;
};
std::vector<int> g_inplace_devices;                                           // guarded by g_cache_mu
std::map<std::pair<std::vector<int>, std::pair<int, int>>, CachedMulti> g_cache_multi;   // (devices, (N, variant))

void free_cached_multi(CachedMulti& k)
{
    for (MultiDev& m : k.d) {
        cudaSetDevice(m.device);
        cudaFree(m.packed); cudaFree(m.partial); cudaFree(m.counters);
        if (m.st) cudaStreamDestroy(m.st);
    }
    if (!k.d.empty()) { cudaSetDevice(k.d[0].device); if (k.packed_ready) cudaEventDestroy(k.packed_ready); }
    k = CachedMulti();
}

int get_cached_multi(int n, int variant, CachedMulti** out)
{
    const std::vector<int>& devs = g_inplace_devices;
    const int G = (int)devs.size();
    int dev0 = 0;
    CU(cudaGetDevice(&dev0));
    if (dev0 != devs[0]) return fail(NBODY_E_STATE, "nbody_set_devices: the first device of the list (%d) must be the current one (%d), where the arrays live", devs[0], dev0);
    CachedMulti& k = g_cache_multi[std::make_pair(devs, std::make_pair(n, variant))];
    if (k.var) { *out = &k; return NBODY_OK; }
    const int per = per_rank(n, G);
    if ((long long)per * (G - 1) >= n) { g_cache_multi.erase(std::make_pair(devs, std::make_pair(n, variant))); return fail(NBODY_E_INVALID, "N=%d is too small to spread over %d devices", n, G); }
    const Variant* var = find_variant(variant == NBODY_VARIANT_AUTO ? auto_variant(per, n) : variant);
    if (!var) { g_cache_multi.erase(std::make_pair(devs, std::make_pair(n, variant))); return fail(NBODY_E_INVALID, "unknown kernel variant %d", variant); }
    k.n_pad = round_up(n, kSlotAlign);
    k.d.assign(G, MultiDev());
    int rc = NBODY_OK;
    cudaError_t e = cudaSuccess;
    for (int g = 0; g < G && !rc && e == cudaSuccess; ++g) {
        MultiDev& m = k.d[g];
        m.device = devs[g];
        m.t_begin = g * per;
        m.n_targets = n - m.t_begin < per ? n - m.t_begin : per;
        e = cudaSetDevice(m.device);
        cudaDeviceProp prop;
        if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, m.device);
        if (e != cudaSuccess) break;
        if (prop.major < 10) { rc = fail(NBODY_E_CUDA, "device %d is sm_%d%d; this library is built for sm_100a only", m.device, prop.major, prop.minor); break; }
        if (g > 0) {   // its kernel works on the first device's arrays
            int can = 0;
            e = cudaDeviceCanAccessPeer(&can, m.device, dev0);
            if (e == cudaSuccess && !can) { rc = fail(NBODY_E_CUDA, "device %d cannot access device %d (no P2P path)", m.device, dev0); break; }
            if (e == cudaSuccess) {
                const cudaError_t pe = cudaDeviceEnablePeerAccess(dev0, 0);
                if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) e = pe;
                cudaGetLastError();
            }
            if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&m.st, cudaStreamNonBlocking);
        }
        if (e == cudaSuccess) rc = make_plan(var, prop.multiProcessorCount, 0, m.n_targets, round_up(n, var->UJ) / var->UJ, &m.plan);
        if (!rc && e == cudaSuccess) e = cudaMalloc(&m.packed, (size_t)k.n_pad * sizeof(float4));
        if (!rc && e == cudaSuccess) e = cudaMalloc(&m.partial, m.plan.partial_bytes);
        if (!rc && e == cudaSuccess) e = cudaMalloc(&m.counters, (size_t)m.plan.num_counters * sizeof(unsigned));
        if (!rc && e == cudaSuccess) e = cudaMemset(m.counters, 0, (size_t)m.plan.num_counters * sizeof(unsigned));
    }
    if (!rc && e == cudaSuccess) e = cudaSetDevice(dev0);
    if (!rc && e == cudaSuccess) e = cudaEventCreateWithFlags(&k.packed_ready, cudaEventDisableTiming);
    if (rc || e != cudaSuccess) {
        const std::string keep = g_err;
        free_cached_multi(k);
        g_cache_multi.erase(std::make_pair(devs, std::make_pair(n, variant)));
        cudaSetDevice(dev0);
        if (rc) { g_err = keep; return rc; }
        return fail(NBODY_E_CUDA, "multi-device scratch: %s", cudaGetErrorString(e));
    }
    k.var = var;
    *out = &k;
    return NBODY_OK;
}

int one_step_multi(CachedMulti* k, float4* dX, void* dA, void* dV, int n, float eps2, float dt, int integ, int va_f3)
{
    const int dev0 = k->d[0].device;
    // snapshot of the step's start in packed form on the first device, then one copy per peer
    int rc = launch_pack(dX, n, k->d[0].packed, 0, k->n_pad, 0);
    if (rc) return rc;
    CU(cudaEventRecord(k->packed_ready, 0));
    const size_t va_stride = va_f3 ? 12 : 16;
    for (size_t g = 0; g < k->d.size(); ++g) {
        const size_t i = (g + 1) % k->d.size();   // the peers first, the first device last (its launch is on the caller's stream)
        MultiDev& m = k->d[i];
        CU(cudaSetDevice(m.device));
        if (i != 0) {
            CU(cudaStreamWaitEvent(m.st, k->packed_ready, 0));
            CU(cudaMemcpyPeerAsync(m.packed, m.device, k->d[0].packed, dev0, (size_t)k->n_pad * sizeof(float4), m.st));
        }
        ForceArgs a{};
        a.X = dX + m.t_begin;
        a.V = reinterpret_cast<float4*>(static_cast<char*>(dV) + va_stride * (size_t)m.t_begin);
        a.A = reinterpret_cast<float4*>(static_cast<char*>(dA) + va_stride * (size_t)m.t_begin);
        a.n_targets = m.n_targets;
        a.src = m.packed;
        a.ra_begin = 0; a.ra_count = round_up(n, k->var->UJ) / k->var->UJ;
        a.last_unit_bodies = n % k->var->UJ;
        a.partial = m.partial; a.counters = m.counters; a.maxseg = m.plan.maxseg;
        a.eps2 = eps2; a.dt = dt; a.integ_f64 = integ; a.va_f3 = va_f3;
        rc = launch_force(k->var, m.plan, a, m.st);
        if (rc) { cudaSetDevice(dev0); return rc; }
    }
    cudaError_t e = cudaSuccess;
    for (size_t i = 1; i < k->d.size() && e == cudaSuccess; ++i) {
        e = cudaSetDevice(k->d[i].device);
        if (e == cudaSuccess) e = cudaStreamSynchronize(k->d[i].st);
    }
    cudaSetDevice(dev0);
    if (e != cudaSuccess) return fail(NBODY_E_CUDA, "multi-device step: %s", cudaGetErrorString(e));
    return NBODY_OK;
}

// one in-place step on device arrays, on the current device alone or spread over nbody_set_devices' list
int inplace_step(float4* dX, void* dA, void* dV, int n, float eps2, float dt, int integ, int va_f3, int variant)
{
    std::lock_guard<std::mutex> lk(g_cache_mu);
    if (g_inplace_devices.size() > 1) {
        CachedMulti* k = nullptr;
        int rc = get_cached_multi(n, variant, &k);
        if (rc) return rc;
        rc = one_step_multi(k, dX, dA, dV, n, eps2, dt, integ, va_f3);
        if (rc) return rc;
    } else {
        Cached* k = nullptr;
        int rc = get_cached(n, false, &k, variant);
        if (rc) return rc;
        rc = one_step_on_device(k, dX, static_cast<float4*>(dA), static_cast<float4*>(dV), n, eps2, dt, integ, 0.0, va_f3);
        if (rc) return rc;
    }
    CU(cudaDeviceSynchronize());   // the reference's simulate() is synchronous (kernel.cu:644)
    return NBODY_OK;
}

}  // namespace

extern "C" {

int nbody_step_inplace_f4(float* d_bodies4, float* d_accel4, float* d_vel4, int n, float eps2, float dt)
{
    TRACE_CALL();
    if (!d_bodies4 || !d_accel4 || !d_vel4) return fail(NBODY_E_INVALID, "null device pointer");
    if (n <= 0) return fail(NBODY_E_INVALID, "N must be > 0 (got %d)", n);
    if (!(eps2 > 0.f)) return fail(NBODY_E_INVALID, "eps2 must be > 0");
    if (!eps2_representable(eps2)) return fail(NBODY_E_INVALID, "eps2 = %g is too small: eps2^3 must be a normal float (eps2 >= %g)", (double)eps2, (double)kMinEps2);
    return inplace_step((float4*)d_bodies4, d_accel4, d_vel4, n, eps2, dt, NBODY_INTEG_F32, 0, NBODY_VARIANT_AUTO);
}

int nbody_set_devices(const int* devices, int n)
{
    std::lock_guard<std::mutex> lk(g_cache_mu);
    if (!devices || n <= 1) { g_inplace_devices.clear(); return NBODY_OK; }
    if (n > NBODY_MAX_DEVICES) return fail(NBODY_E_INVALID, "at most %d devices", NBODY_MAX_DEVICES);
    int ndev = 0;
    CU(cudaGetDeviceCount(&ndev));
    for (int i = 0; i < n; ++i) {
        if (devices[i] < 0 || devices[i] >= ndev) return fail(NBODY_E_INVALID, "device %d not present (%d devices)", devices[i], ndev);
        for (int j = 0; j < i; ++j)
            if (devices[j] == devices[i]) return fail(NBODY_E_INVALID, "device %d listed twice", devices[i]);
    }
    g_inplace_devices.assign(devices, devices + n);
    return NBODY_OK;
}

int nbody_step_host_f3(float* h_bodies4, float* h_accel3, float* h_vel3, int n, double eps2, double dt)
{
    TRACE_CALL();
    (void)h_accel3;   // never copied back by the reference either (headless kernel.cu:115-124)
    if (!h_bodies4 || !h_vel3) return fail(NBODY_E_INVALID, "null host pointer");
    if (n <= 0) return fail(NBODY_E_INVALID, "N must be > 0 (got %d)", n);
    if (!(eps2 > 0.0)) return fail(NBODY_E_INVALID, "eps2 must be > 0");
    if (!eps2_representable((float)eps2)) return fail(NBODY_E_INVALID, "eps2 = %g is too small: eps2^3 must be a normal float (eps2 >= %g)", eps2, (double)kMinEps2);
    std::lock_guard<std::mutex> lk(g_cache_mu);
    Cached* k = nullptr;
    int rc = get_cached(n, true, &k);
    if (rc) return rc;
    std::vector<float> v4((size_t)n * 4, 0.f);
    for (int i = 0; i < n; ++i) { v4[4 * (size_t)i] = h_vel3[3 * (size_t)i]; v4[4 * (size_t)i + 1] = h_vel3[3 * (size_t)i + 1]; v4[4 * (size_t)i + 2] = h_vel3[3 * (size_t)i + 2]; }
    CU(cudaMemcpy(k->X, h_bodies4, (size_t)n * sizeof(float4), cudaMemcpyHostToDevice));
    CU(cudaMemcpy(k->V, v4.data(), (size_t)n * sizeof(float4), cudaMemcpyHostToDevice));
    float4* dA = k->A;
    rc = one_step_on_device(k, k->X, dA, k->V, n, (float)eps2, (float)dt, 1, dt);
    cudaError_t e = rc ? cudaSuccess : cudaDeviceSynchronize();
    if (!rc && e == cudaSuccess) e = cudaMemcpy(h_bodies4, k->X, (size_t)n * sizeof(float4), cudaMemcpyDeviceToHost);
    if (!rc && e == cudaSuccess) e = cudaMemcpy(v4.data(), k->V, (size_t)n * sizeof(float4), cudaMemcpyDeviceToHost);
    if (rc) return rc;
    if (e != cudaSuccess) return fail(NBODY_E_CUDA, "host-pointer step failed: %s", cudaGetErrorString(e));
    for (int i = 0; i < n; ++i) { h_vel3[3 * (size_t)i] = v4[4 * (size_t)i]; h_vel3[3 * (size_t)i + 1] = v4[4 * (size_t)i + 1]; h_vel3[3 * (size_t)i + 2] = v4[4 * (size_t)i + 2]; }
    return NBODY_OK;
}

int nbody_step_opts_default(nbody_step_opts* o)
{
    if (!o) return fail(NBODY_E_INVALID, "null options");
    o->eps2 = 0.002f; o->dt = 0.1f;   // TestProject/constants.h:25-26
    o->integrator = NBODY_INTEG_F32; o->va_float3 = 0; o->variant = NBODY_VARIANT_AUTO;
    return NBODY_OK;
}

int nbody_step_inplace_opts(float* d_bodies4, void* d_accel, void* d_vel, int n, const nbody_step_opts* o)
{
    TRACE_CALL();
    if (!d_bodies4 || !d_accel || !d_vel || !o) return fail(NBODY_E_INVALID, "null pointer");
    if (n <= 0) return fail(NBODY_E_INVALID, "N must be > 0 (got %d)", n);
    if (!(o->eps2 > 0.f)) return fail(NBODY_E_INVALID, "eps2 must be > 0");
    if (o->integrator != NBODY_INTEG_F32 && o->integrator != NBODY_INTEG_MIXED)
        return fail(NBODY_E_INVALID, "integrator %d not valid for the in-place step", o->integrator);
    if (!eps2_representable(o->eps2)) return fail(NBODY_E_INVALID, "eps2 = %g is too small: eps2^3 must be a normal float (eps2 >= %g)", (double)o->eps2, (double)kMinEps2);
    return inplace_step((float4*)d_bodies4, d_accel, d_vel, n, o->eps2, o->dt, o->integrator, o->va_float3 ? 1 : 0, o->variant);
}

void nbody_release_cached(void)
{
    std::lock_guard<std::mutex> lk(g_cache_mu);
    int cur = 0;
    cudaGetDevice(&cur);
    for (auto& kv : g_cache) {
        cudaSetDevice(kv.first.first.first);
        free_cached(kv.second);
    }
    g_cache.clear();
    for (auto& kv : g_cache_multi) free_cached_multi(kv.second);
    g_cache_multi.clear();
    release_render_scratch();
    cudaSetDevice(cur);
}

// ---- initial conditions (host) -----------------------------------------------------------------------------
static inline unsigned long long splitmix64(unsigned long long z)
{
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static inline double u01(unsigned long long r) { return (double)(r >> 11) * (1.0 / 9007199254740992.0); }

int nbody_ic_uniform(float* x4, int n, unsigned long long seed, float max_xyz, float min_w, float max_w)
{
    if (!x4 || n <= 0) return fail(NBODY_E_INVALID, "bad argument");
    for (int i = 0; i < n; ++i) {
        const unsigned long long base = seed * 0x100000001B3ull + 4ull * (unsigned long long)i;
        for (int c = 0; c < 3; ++c)
            x4[4 * (size_t)i + c] = (float)((2.0 * u01(splitmix64(base + c)) - 1.0) * (double)max_xyz);
        x4[4 * (size_t)i + 3] = (float)((double)min_w + u01(splitmix64(base + 3)) * ((double)max_w - (double)min_w));
    }
    return NBODY_OK;
}

int nbody_ic_plummer(float* x4, float* v4, int n, unsigned long long seed)
{
    if (!x4 || !v4 || n <= 0) return fail(NBODY_E_INVALID, "bad argument");
    const double PI = 3.14159265358979323846;
    const double rscale = 3.0 * PI / 16.0, vscale = std::sqrt(1.0 / rscale) * std::sqrt(0.5);
    double com[3] = {0, 0, 0}, cov[3] = {0, 0, 0};
    for (int i = 0; i < n; ++i) {
        unsigned long long st = splitmix64(seed ^ (0xD1B54A32D192ED03ull * (unsigned long long)(i + 1)));
        auto next = [&]() { st = splitmix64(st); return u01(st); };
        double r;
        do {   // Aarseth, Henon & Wielen 1974; radius cut at 10 scale lengths
            double x1 = next();
            if (x1 < 1e-12) x1 = 1e-12;
            r = 1.0 / std::sqrt(std::pow(x1, -2.0 / 3.0) - 1.0);
        } while (!(r <= 10.0));
        auto iso = [&](double len, double* o) {
            const double z = (1.0 - 2.0 * next()) * len, ph = 2.0 * PI * next(), s = std::sqrt(len * len - z * z);
            o[0] = s * std::cos(ph); o[1] = s * std::sin(ph); o[2] = z;
        };
        double pos[3], vel[3], q, g;
        iso(r, pos);
        do { q = next(); g = 0.1 * next(); } while (g > q * q * std::pow(1.0 - q * q, 3.5));
        iso(q * std::sqrt(2.0) * std::pow(1.0 + r * r, -0.25), vel);
        for (int c = 0; c < 3; ++c) {
            x4[4 * (size_t)i + c] = (float)(pos[c] * rscale);
            v4[4 * (size_t)i + c] = (float)(vel[c] * vscale);
            com[c] += x4[4 * (size_t)i + c]; cov[c] += v4[4 * (size_t)i + c];
        }
        x4[4 * (size_t)i + 3] = (float)(1.0 / n);
        v4[4 * (size_t)i + 3] = 0.f;
    }
    for (int i = 0; i < n; ++i)
        for (int c = 0; c < 3; ++c) {
            x4[4 * (size_t)i + c] -= (float)(com[c] / n);
            v4[4 * (size_t)i + c] -= (float)(cov[c] / n);
        }
    return NBODY_OK;
}

}  // extern "C"
