// nbody_group.inl -- fifth part of nbody_engine.cu: ONE context spanning several GPUs of a box
// (nbody_params.n_devices > 1).
//
// The reference is single-GPU: its launcher asks for device 0 (TestProject/kernel.cu:630) and its caller loop
// (main.cpp:142-160) knows one set of device arrays.  Here the same calls -- nbody_create / nbody_upload /
// nbody_step / nbody_download / nbody_energy / nbody_snapshot_* -- drive G sharded contexts (device g owns the
// g-th slice of the bodies) from one host thread, so a caller written against the
// single-GPU API scales without new host code:
//   NBODY_EXCHANGE_P2P   one fused kernel per device per step; its epilogue stores the new positions into every
//                        device's next source buffer over NVLink and the kernels wait for each other on flags in
//                        peer memory (nbody_kernels.cuh) -- no library call, no host synchronisation in the loop.
//                        Kernel families without the in-kernel wait (small-N) order the devices with events.
//   NBODY_EXCHANGE_NCCL  per device: own x own interactions on the compute stream while ncclAllGather of the
//                        previous step's positions runs on a second stream, then own x others + integrator
//                        (the north-star's formulation); libnccl.so.2 is opened with dlopen, so the library
//                        itself has no link-time dependency on NCCL.
#include <dlfcn.h>
#include <nccl.h>

struct Group {
    int G = 0;
    int exchange = NBODY_EXCHANGE_P2P;
    bool in_kernel_barrier = false;
    bool timed = false;
    std::vector<cudaEvent_t> t0, t1;         // device-timed span of the last step call, per device
    std::vector<cudaEvent_t> ev_step;        // event barrier (P2P without the in-kernel wait)
    // NCCL exchange
    std::vector<ncclComm_t> comms;
    std::vector<cudaStream_t> comm_st;
    std::vector<cudaEvent_t> ev_comp, ev_gather;
    // whole-system copies on device 0 for nbody_energy / nbody_render
    float4 *Xfull = nullptr, *Vfull = nullptr;
};

namespace {

// ---- NCCL, resolved at run time ----------------------------------------------------------------------------
struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;
std::mutex g_nccl_mu;

int load_nccl()
{
    std::lock_guard<std::mutex> lk(g_nccl_mu);
    if (g_nccl.handle) return NBODY_OK;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);   // an already loaded copy (e.g. torch's) is reused
    if (!h) return fail(NBODY_E_STATE, "NBODY_EXCHANGE_NCCL: cannot open libnccl.so.2 (%s)", dlerror());
    NcclApi a;
    a.handle = h;
    *(void**)&a.CommInitAll = dlsym(h, "ncclCommInitAll");
    *(void**)&a.CommDestroy = dlsym(h, "ncclCommDestroy");
    *(void**)&a.AllGather = dlsym(h, "ncclAllGather");
    *(void**)&a.GroupStart = dlsym(h, "ncclGroupStart");
    *(void**)&a.GroupEnd = dlsym(h, "ncclGroupEnd");
    *(void**)&a.GetErrorString = dlsym(h, "ncclGetErrorString");
    if (!a.CommInitAll || !a.CommDestroy || !a.AllGather || !a.GroupStart || !a.GroupEnd || !a.GetErrorString)
        return fail(NBODY_E_STATE, "libnccl.so.2 lacks an expected symbol");
    g_nccl = a;
    return NBODY_OK;
}

// the group calls visit every device; the caller's current device is put back when they return
struct DeviceRestore {
    int dev = -1;
    DeviceRestore() { if (cudaGetDevice(&dev) != cudaSuccess) dev = -1; }
    ~DeviceRestore() { if (dev >= 0) cudaSetDevice(dev); }
    DeviceRestore(const DeviceRestore&) = delete;
    DeviceRestore& operator=(const DeviceRestore&) = delete;
};

#define NC(call)                                                                                   \
    do {                                                                                           \
        ncclResult_t r_ = (call);                                                                  \
        if (r_ != ncclSuccess) return fail(NBODY_E_CUDA, "%s failed: %s", #call, g_nccl.GetErrorString(r_)); \
    } while (0)

void group_destroy(nbody_ctx* c)
{
    DeviceRestore restore_;
    Group* g = c->group;
    if (g) {
        for (size_t i = 0; i < c->kids.size(); ++i) {
            if (c->kids[i]) cudaSetDevice(c->kids[i]->p.device);
            if (i < g->comm_st.size() && g->comm_st[i]) cudaStreamSynchronize(g->comm_st[i]);
            if (c->kids[i]) cudaStreamSynchronize(c->kids[i]->stream);
        }
        for (size_t i = 0; i < c->kids.size(); ++i) {
            if (c->kids[i]) cudaSetDevice(c->kids[i]->p.device);
            if (i < g->comms.size() && g->comms[i]) g_nccl.CommDestroy(g->comms[i]);
            if (i < g->comm_st.size() && g->comm_st[i]) cudaStreamDestroy(g->comm_st[i]);
            for (auto* v : {&g->t0, &g->t1, &g->ev_step, &g->ev_comp, &g->ev_gather})
                if (i < v->size() && (*v)[i]) cudaEventDestroy((*v)[i]);
        }
        if (!c->kids.empty() && c->kids[0]) {
            cudaSetDevice(c->kids[0]->p.device);
            cudaFree(g->Xfull); cudaFree(g->Vfull);
        }
    }
    for (nbody_ctx* k : c->kids)
        if (k) nbody_destroy(k);
    delete g;
    delete c;
}

int group_create(nbody_ctx** out, const nbody_params* p)
{
    const int G = p->n_devices;
    if (G > NBODY_MAX_DEVICES) return fail(NBODY_E_INVALID, "n_devices %d exceeds NBODY_MAX_DEVICES (%d)", G, NBODY_MAX_DEVICES);
    if (p->rank != 0 || p->world != 1) return fail(NBODY_E_INVALID, "a multi-device context is the whole system: rank 0, world 1 (got %d of %d)", p->rank, p->world);
    if (p->exchange != NBODY_EXCHANGE_P2P && p->exchange != NBODY_EXCHANGE_NCCL && p->exchange != NBODY_EXCHANGE_P2P_EVENTS)
        return fail(NBODY_E_INVALID, "unknown exchange %d", p->exchange);
    int ndev = 0;
    CU(cudaGetDeviceCount(&ndev));
    for (int g = 0; g < G; ++g) {
        if (p->devices[g] < 0 || p->devices[g] >= ndev) return fail(NBODY_E_INVALID, "device %d not present (%d devices)", p->devices[g], ndev);
        for (int h = 0; h < g; ++h)
            if (p->devices[h] == p->devices[g] && p->exchange != NBODY_EXCHANGE_P2P_EVENTS) return fail(NBODY_E_INVALID, "device %d listed twice (the step kernels of one device would wait for each other)", p->devices[g]);
    }
    if (p->exchange == NBODY_EXCHANGE_NCCL) {
        const int rc = load_nccl();
        if (rc) return rc;
    }
    int cur = 0;
    cudaGetDevice(&cur);
    nbody_ctx* c = new nbody_ctx;
    c->p = *p;
    c->n = p->n_bodies;
    c->group = new Group;
    Group* g = c->group;
    g->G = G;
    g->exchange = p->exchange;
    c->kids.assign(G, nullptr);
    int rc = NBODY_OK, variant = p->variant;
#define GTRY(x) do { rc = (x); if (rc) { const std::string keep = g_err; group_destroy(c); cudaSetDevice(cur); g_err = keep; return rc; } } while (0)
#define GTRYCU(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { group_destroy(c); cudaSetDevice(cur); return fail(NBODY_E_CUDA, "%s: %s", #x, cudaGetErrorString(e_)); } } while (0)
    for (int d = 0; d < G; ++d) {
        nbody_params kp = *p;
        kp.n_devices = 0; kp.rank = d; kp.world = G; kp.device = p->devices[d]; kp.variant = variant;
        GTRY(nbody_create(&c->kids[d], &kp));
        if (d == 0) variant = c->kids[0]->var->id;   // every device runs the kernel the first one resolved
    }
    g->t0.assign(G, nullptr); g->t1.assign(G, nullptr); g->ev_step.assign(G, nullptr);
    for (int d = 0; d < G; ++d) {
        GTRYCU(cudaSetDevice(p->devices[d]));
        GTRYCU(cudaEventCreate(&g->t0[d]));
        GTRYCU(cudaEventCreate(&g->t1[d]));
        GTRYCU(cudaEventCreateWithFlags(&g->ev_step[d], cudaEventDisableTiming));
    }
    if (p->exchange != NBODY_EXCHANGE_NCCL) {
        for (int d = 0; d < G; ++d)
            for (int h = 0; h < G; ++h) {
                if (h == d) continue;
                void *b0 = nullptr, *b1 = nullptr;
                GTRY(nbody_shard_buffers(c->kids[h], &b0, &b1));
                GTRY(nbody_shard_attach_ptr(c->kids[d], h, b0, b1));   // enables peer access d -> h
            }
        g->in_kernel_barrier = p->exchange == NBODY_EXCHANGE_P2P && !c->kids[0]->var->small;
        if (g->in_kernel_barrier)
            for (int d = 0; d < G; ++d) GTRY(nbody_shard_set_device_barrier(c->kids[d], 1));
    } else {
        g->comms.assign(G, nullptr); g->comm_st.assign(G, nullptr); g->ev_comp.assign(G, nullptr); g->ev_gather.assign(G, nullptr);
        ncclResult_t r = g_nccl.CommInitAll(g->comms.data(), G, p->devices);
        if (r != ncclSuccess) { group_destroy(c); cudaSetDevice(cur); return fail(NBODY_E_CUDA, "ncclCommInitAll: %s", g_nccl.GetErrorString(r)); }
        for (int d = 0; d < G; ++d) {
            GTRYCU(cudaSetDevice(p->devices[d]));
            GTRYCU(cudaStreamCreateWithFlags(&g->comm_st[d], cudaStreamNonBlocking));
            GTRYCU(cudaEventCreateWithFlags(&g->ev_comp[d], cudaEventDisableTiming));
            GTRYCU(cudaEventCreateWithFlags(&g->ev_gather[d], cudaEventDisableTiming));
        }
    }
#undef GTRY
#undef GTRYCU
    cudaSetDevice(cur);
    *out = c;
    return NBODY_OK;
}

int group_sync(nbody_ctx* c)
{
    DeviceRestore restore_;
    Group* g = c->group;
    for (int d = 0; d < g->G; ++d) {
        int rc = set_device(c->kids[d]);
        if (rc) return rc;
        CU(cudaStreamSynchronize(c->kids[d]->stream));
        if (!g->comm_st.empty()) CU(cudaStreamSynchronize(g->comm_st[d]));
    }
    return NBODY_OK;
}

int group_upload(nbody_ctx* c, const float* x4, const float* v4)
{
    DeviceRestore restore_;
    // both source buffers of every device are rewritten: nobody may still be reading or storing into them
    int rc = group_sync(c);
    for (int d = 0; d < c->group->G && !rc; ++d) rc = upload_enqueue(c->kids[d], x4, v4);
    if (rc) return rc;
    c->steps_done = 0;
    return group_sync(c);
}

int group_download(nbody_ctx* c, float* x4, float* v4, float* a4)
{
    DeviceRestore restore_;
    int rc = NBODY_OK;
    for (int d = 0; d < c->group->G && !rc; ++d) {
        nbody_ctx* k = c->kids[d];
        const size_t off = 4 * (size_t)k->t_begin;
        rc = download_enqueue(k, x4 ? x4 + off : nullptr, v4 ? v4 + off : nullptr, a4 ? a4 + off : nullptr);
    }
    if (rc) return rc;
    return group_sync(c);
}

int group_step_async(nbody_ctx* c, int nsteps)
{
    DeviceRestore restore_;
    Group* g = c->group;
    const int G = g->G;
    int rc = NBODY_OK;
    for (int d = 0; d < G; ++d) {
        rc = set_device(c->kids[d]);
        if (rc) return rc;
        CU(cudaEventRecord(g->t0[d], c->kids[d]->stream));
    }
    for (int s = 0; s < nsteps; ++s) {
        if (g->exchange != NBODY_EXCHANGE_NCCL) {
            for (int d = 0; d < G; ++d) {
                rc = nbody_shard_step_fused_async(c->kids[d], nullptr);
                if (rc) return rc;
                if (!g->in_kernel_barrier) CU(cudaEventRecord(g->ev_step[d], c->kids[d]->stream));
            }
            if (!g->in_kernel_barrier)   // nobody starts step s+1 before everybody has finished step s
                for (int d = 0; d < G; ++d) {
                    rc = set_device(c->kids[d]);
                    if (rc) return rc;
                    for (int h = 0; h < G; ++h)
                        if (h != d) CU(cudaStreamWaitEvent(c->kids[d]->stream, g->ev_step[h], 0));
                }
        } else {
            // own x own while the gather of the previous step's positions is still in flight ...
            for (int d = 0; d < G; ++d) {
                rc = nbody_shard_local_async(c->kids[d], nullptr);
                if (rc) return rc;
            }
            // ... own x others (+ integrator, new own slice into the next buffer) once it has landed
            for (int d = 0; d < G; ++d) {
                nbody_ctx* k = c->kids[d];
                rc = set_device(k);
                if (rc) return rc;
                CU(cudaStreamWaitEvent(k->stream, g->ev_gather[d], 0));
                rc = nbody_shard_remote_async(k, nullptr);
                if (rc) return rc;
                CU(cudaEventRecord(g->ev_comp[d], k->stream));
            }
            // publish: in-place all-gather of the (now current) buffer, every device's slice -> everyone
            NC(g_nccl.GroupStart());
            for (int d = 0; d < G; ++d) {
                nbody_ctx* k = c->kids[d];
                rc = set_device(k);
                if (rc) { g_nccl.GroupEnd(); return rc; }
                cudaStreamWaitEvent(g->comm_st[d], g->ev_comp[d], 0);
                float4* buf = k->packed[k->cur];
                const ncclResult_t r = g_nccl.AllGather(buf + (size_t)d * k->slot_bodies, buf, (size_t)k->slot_bodies * 4, ncclFloat, g->comms[d], g->comm_st[d]);
                if (r != ncclSuccess) { g_nccl.GroupEnd(); return fail(NBODY_E_CUDA, "ncclAllGather: %s", g_nccl.GetErrorString(r)); }
            }
            NC(g_nccl.GroupEnd());
            for (int d = 0; d < G; ++d) {
                rc = set_device(c->kids[d]);
                if (rc) return rc;
                CU(cudaEventRecord(g->ev_gather[d], g->comm_st[d]));
            }
        }
        ++c->steps_done;
    }
    for (int d = 0; d < G; ++d) {
        rc = set_device(c->kids[d]);
        if (rc) return rc;
        // a step ends when its positions are everywhere: after the gather on the NCCL path
        CU(cudaEventRecord(g->t1[d], g->exchange == NBODY_EXCHANGE_NCCL ? g->comm_st[d] : c->kids[d]->stream));
    }
    g->timed = true;
    return NBODY_OK;
}

int group_elapsed_ms(nbody_ctx* c, float* ms)
{
    DeviceRestore restore_;
    Group* g = c->group;
    if (!g->timed) return fail(NBODY_E_STATE, "no timed call yet");
    float worst = 0.f;
    for (int d = 0; d < g->G; ++d) {
        int rc = set_device(c->kids[d]);
        if (rc) return rc;
        float t = 0.f;
        CU(cudaEventSynchronize(g->t1[d]));
        CU(cudaEventElapsedTime(&t, g->t0[d], g->t1[d]));
        if (t > worst) worst = t;
    }
    *ms = worst;   // the devices run in lock step: the slowest one is the step
    return NBODY_OK;
}

// whole-system X (and V) on the first device, in body order
int group_gather_state(nbody_ctx* c, bool with_v)
{
    Group* g = c->group;
    int rc = group_sync(c);
    if (rc) return rc;
    nbody_ctx* k0 = c->kids[0];
    rc = set_device(k0);
    if (rc) return rc;
    const size_t nb = (size_t)c->n * sizeof(float4);
    if (!g->Xfull) CU(cudaMalloc(&g->Xfull, nb));
    if (with_v && !g->Vfull) CU(cudaMalloc(&g->Vfull, nb));
    for (int d = 0; d < g->G; ++d) {
        nbody_ctx* k = c->kids[d];
        const size_t tb = (size_t)k->n_targets * sizeof(float4);
        CU(cudaMemcpyPeerAsync(g->Xfull + k->t_begin, k0->p.device, k->X, k->p.device, tb, k0->stream));
        if (with_v) CU(cudaMemcpyPeerAsync(g->Vfull + k->t_begin, k0->p.device, k->V, k->p.device, tb, k0->stream));
    }
    return NBODY_OK;
}

int group_energy(nbody_ctx* c, double* ke, double* pe)
{
    DeviceRestore restore_;
    int rc = group_gather_state(c, true);
    if (rc) return rc;
    const double eps2 = c->p.integrator == NBODY_INTEG_F64 ? c->p.eps2_f64 : (double)c->p.eps2;
    return energy_of(c->kids[0], c->group->Xfull, c->group->Vfull, c->n, eps2, ke, pe);
}

int group_render(nbody_ctx* c, const nbody_view* v, unsigned char* h_rgba)
{
    DeviceRestore restore_;
    int rc = group_gather_state(c, false);
    if (rc) return rc;
    return render_arrays(c->kids[0], c->group->Xfull, c->n, v, h_rgba);
}

int group_barrier_timeouts(nbody_ctx* c, int* timed_out)
{
    DeviceRestore restore_;
    int total = 0;
    for (nbody_ctx* k : c->kids) {
        int t = 0;
        const int rc = nbody_shard_barrier_timeouts(k, &t);
        if (rc) return rc;
        total += t;
    }
    *timed_out = total;
    return NBODY_OK;
}

int group_barrier_wait_ns(nbody_ctx* c, unsigned long long* total_ns, unsigned long long* max_ns, unsigned long long* waits, int reset)
{
    DeviceRestore restore_;
    unsigned long long bt = 0, bm = 0, bw = 0;
    for (nbody_ctx* k : c->kids) {
        unsigned long long t = 0, m = 0, w = 0;
        const int rc = nbody_shard_barrier_wait_ns(k, &t, &m, &w, reset);
        if (rc) return rc;
        if (t >= bt) { bt = t; bw = w; }
        if (m > bm) bm = m;
    }
    if (total_ns) *total_ns = bt;
    if (max_ns) *max_ns = bm;
    if (waits) *waits = bw;
    return NBODY_OK;
}

}  // namespace

extern "C" {

int nbody_group_info(nbody_ctx* c, int* n_devices, int* exchange, int* in_kernel_barrier)
{
    if (!c) return fail(NBODY_E_INVALID, "null context");
    if (n_devices) *n_devices = c->group ? c->group->G : 1;
    if (exchange) *exchange = c->group ? c->group->exchange : NBODY_EXCHANGE_P2P;
    if (in_kernel_barrier) *in_kernel_barrier = c->group ? (c->group->in_kernel_barrier ? 1 : 0) : (c->device_barrier ? 1 : 0);
    return NBODY_OK;
}

}  // extern "C"
