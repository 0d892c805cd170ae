// nbody_noalloc.inl -- sixth part of nbody_engine.cu (included at its end, so this IS that .cu file: the
// array below, the kernels that use it and the host code that asks for its address are one translation unit).
//
// nbody_step_inplace_noalloc_f4: one in-place step on the caller's DEVICE arrays that allocates nothing.
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// The step kernel (force_kernel, the stream-K form) needs scratch memory: the packed snapshot of the sources
// (nbody_kernels.cuh: 32 B per two bodies), the partial sums of the target tiles whose source range is cut between
// CTAs, and one arrival counter per tile.  The other entry points cudaMalloc that per context / per (device, N);
// here it lives in ONE file-scope device array of a fixed, modest size.  Every GPU that loads the library holds
// those 8 MiB until the process exits.
//
// g_noalloc_scratch is a DEVICE symbol.  In host code its name is only a host-side stand-in, not a device address:
// handing `g_noalloc_scratch` itself to a kernel would fault (or, where the GPU can reach host memory, silently
// use the wrong memory).  So the address is asked for once per device, with that device current
// (cudaGetSymbolAddress), and that pointer goes into the kernel arguments.
//
// All calls share the array and it keeps the last call's contents: the counters are cleared on the caller's stream
// before every step, the packed snapshot and the partial sums are fully rewritten before they are read.  Calls that
// overlap on the GPU (two streams, two host threads, two graphs replayed side by side) RACE on it: serialise them.

constexpr int kNoAllocMaxBodies = 16384;                                         // BASELINE config 2
constexpr size_t kNoAllocCounterBytes = 4096;                                    // 1 024 tile counters
constexpr size_t kNoAllocPackedBytes = (size_t)kNoAllocMaxBodies * sizeof(float4);   // 256 KiB
constexpr size_t kNoAllocScratchBytes = (size_t)8 << 20;                         // counters | packed | partial sums
constexpr size_t kNoAllocPartialBytes = kNoAllocScratchBytes - kNoAllocCounterBytes - kNoAllocPackedBytes;
static_assert(kNoAllocMaxBodies % kSlotAlign == 0, "the packed snapshot is padded to whole slots");
static_assert((kNoAllocCounterBytes + kNoAllocPackedBytes) % 256 == 0, "the partial sums start 256-byte aligned");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

namespace {

std::mutex g_noalloc_mu;
unsigned char* g_noalloc_ptr[64] = {};   // device address of g_noalloc_scratch per CUDA device, asked for once
int g_noalloc_sms[64] = {};

// device address of the scratch array on the CURRENT device (cudaGetSymbolAddress answers for the current device)
int noalloc_scratch(unsigned char** out, int* sms)
{
    int dev = 0;
    CU(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail(NBODY_E_INVALID, "device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_noalloc_ptr[dev]) {
        int major = 0;
        CU(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
        if (major < 10) return fail(NBODY_E_CUDA, "device %d is sm_%dx; this library is built for sm_100a only", dev, major);
        CU(cudaDeviceGetAttribute(&g_noalloc_sms[dev], cudaDevAttrMultiProcessorCount, dev));
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            (void)cudaGetLastError();
            return fail(NBODY_E_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch) on device %d: %s", dev, cudaGetErrorString(e));
        }
        g_noalloc_ptr[dev] = static_cast<unsigned char*>(p);
    }
    *out = g_noalloc_ptr[dev];
    *sms = g_noalloc_sms[dev];
    return NBODY_OK;
}

}  // namespace

extern "C" {

int nbody_noalloc_max_bodies(void) { return kNoAllocMaxBodies; }

int nbody_step_inplace_noalloc_f4(float* d_bodies4, float* d_accel4, float* d_vel4, int n, float eps2, float dt, void* stream)
{
    TRACE_CALL();
    if (!d_bodies4 || !d_accel4 || !d_vel4) return fail(NBODY_E_INVALID, "null device pointer");
    if (n <= 0) return fail(NBODY_E_INVALID, "N must be > 0 (got %d)", n);
    if (n > kNoAllocMaxBodies)
        return fail(NBODY_E_INVALID, "N = %d is beyond the %d bodies the static scratch of the no-allocation step holds; "
                    "use nbody_step_inplace_f4 or a context", n, kNoAllocMaxBodies);
    if (!(eps2 > 0.f)) return fail(NBODY_E_INVALID, "eps2 must be > 0");
    if (!eps2_representable(eps2)) return fail(NBODY_E_INVALID, "eps2 = %g is too small: eps2^3 must be a normal float (eps2 >= %g)", (double)eps2, (double)kMinEps2);
    unsigned char* base = nullptr;
    int sms = 0;
    int rc = noalloc_scratch(&base, &sms);
    if (rc) return rc;
    // always the stream-K force kernel (the one with scratch): the tile shape the mid-size rule picks by unit balance
    // from 8 192 bodies up, the static rules' T2 / T1 kernels below
    const int vid = n >= g_small_below ? mid_variant(n, n, sms) : n >= 2048 ? 3 : 12;
    const Variant* var = find_variant(vid);
    if (!var) return fail(NBODY_E_STATE, "kernel variant %d is not in this build", vid);
    Plan pl;
    rc = make_plan(var, sms, 0, n, round_up(n, var->UJ) / var->UJ, &pl);   // occupancy queries only: allocates nothing
    if (rc) return rc;
    pl.spread = false;   // the last-arriver fix-up: an ordinary launch, the same bits as the cooperative form
    if (pl.partial_bytes > kNoAllocPartialBytes || (size_t)pl.num_counters * sizeof(unsigned) > kNoAllocCounterBytes)
        return fail(NBODY_E_INVALID, "N = %d needs %zu bytes of partial sums and %d counters on this device; the static scratch holds %zu and %zu",
                    n, pl.partial_bytes, pl.num_counters, kNoAllocPartialBytes, kNoAllocCounterBytes / sizeof(unsigned));
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    unsigned* counters = reinterpret_cast<unsigned*>(base);
    float4* packed = reinterpret_cast<float4*>(base + kNoAllocCounterBytes);
    float4* partial = reinterpret_cast<float4*>(base + kNoAllocCounterBytes + kNoAllocPackedBytes);
    // the array keeps the last call's contents (or an aborted call's): the counters must start at zero
    CU(cudaMemsetAsync(counters, 0, (size_t)pl.num_counters * sizeof(unsigned), st));
    // snapshot of the step's start in packed form (zero-mass padding up to whole slots), then the step
    rc = launch_pack(reinterpret_cast<const float4*>(d_bodies4), n, packed, 0, round_up(n, kSlotAlign), st);
    if (rc) return rc;
    ForceArgs a{};
    a.X = reinterpret_cast<float4*>(d_bodies4); a.V = reinterpret_cast<float4*>(d_vel4); a.A = reinterpret_cast<float4*>(d_accel4);
    a.n_targets = n;
    a.src = packed;
    a.ra_begin = 0; a.ra_count = round_up(n, var->UJ) / var->UJ;
    a.partial = partial; a.counters = counters; a.maxseg = pl.maxseg;
    a.eps2 = eps2; a.dt = dt; a.integ_f64 = NBODY_INTEG_F32;
    return launch_force(var, pl, a, st);
}

}  // extern "C"
