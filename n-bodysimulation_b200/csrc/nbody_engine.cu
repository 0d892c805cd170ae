// nbody_engine.cu -- host side of the B200 all-pairs step and its C ABI
// (include/nbody_b200.h).  Replaces the reference's launcher
// TestProject/kernel.cu:628-645 (simulate) and the headless-directory launcher
// Sim-Without-OpenGL-Integration/kernel.cu:85-125; the allocation / copy
// sequence mirrors TestProject/main.cpp:244-284,350-359.
//
// No CPU fallback lives here: every compute entry point needs a CUDA device.
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

#include <nvtx3/nvToolsExt.h>   // header-only NVTX v3: a no-op unless a tool (nsys, ncu --nvtx) is attached

#include "../../include/nbody_b200.h"
#include "nbody_kernels.cuh"

using namespace nbody;

namespace {

thread_local std::string g_err;
std::atomic<long long> g_launches{0};

int fail(int code, const char* fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

// named range around an API call, for timelines (the reference has no tracing at all)
struct TraceRange {
    explicit TraceRange(const char* name) { nvtxRangePushA(name); }
    ~TraceRange() { nvtxRangePop(); }
    TraceRange(const TraceRange&) = delete;
    TraceRange& operator=(const TraceRange&) = delete;
};
#define TRACE_CALL() TraceRange trace_range_(__func__)

#define CU(call)                                                                          \
    do {                                                                                  \
        cudaError_t e_ = (call);                                                          \
        if (e_ != cudaSuccess)                                                            \
            return fail(NBODY_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// ---- kernel variants -----------------------------------------------------------------------
struct Variant {
    int id;
    const char* name;
    int T, NT, UJ, unroll, packed, minb;
    int cs, cmode;   // cluster size (1 = none); 1 = TMA multicast of source units, 2 = j-split + DSMEM reduction
    void (*kernel)(const ForceArgs);
    void (*kernel_single)(const ForceArgs);   // the same kernel with single-level summation (null: small-N family)
    void (*kernel_spread)(const ForceArgs) = nullptr;          // mid-size N: split tiles finished by all their CTAs
    void (*kernel_spread_single)(const ForceArgs) = nullptr;   // (force_kernel's SPREAD; cooperative launch)
    int small = 0;   // 1: small_kernel (T = targets per CTA, sources split over the CTA's threads, shuffle reduction)
                     // 2: rowsplit_kernel (T x H targets per CTA, sources split over the warps and their sub-groups)
    int H = 0;       // rowsplit_kernel: lanes of a warp that hold different targets (4, 8, 16 or 32)
    void (*kernel_trim)(const ForceArgs) = nullptr;   // rowsplit_kernel that skips the padding of a ragged last unit
    size_t smem(bool single = false) const { return small == 1 ? 0 : small == 2 ? (size_t)kStages * UJ * 16 + 16 * kStages + 16 : (size_t)kStages * UJ * 16 + 16 * kStages + 16 + (cmode == 2 ? (size_t)TI() * 16 : 0) + fold_bytes(T, NT, packed != 0, !single); }
    void (*entry(bool single, bool spread = false) const)(const ForceArgs)
    {
        if (spread && kernel_spread) return single && kernel_spread_single ? kernel_spread_single : kernel_spread;
        return single && kernel_single ? kernel_single : kernel;
    }
    int TI() const { return small == 1 ? T : small == 2 ? T * H : T * NT; }
    int TS() const { return cmode == 1 ? TI() * cs : TI(); }          // targets per worker tile
};

#define VARIANT(id, T, NT, UJ, UN, PK, MB) \
    { id, "T" #T "_NT" #NT "_UJ" #UJ "_U" #UN "_" #PK "_B" #MB, T, NT, UJ, UN, PK, MB, 1, 0, force_kernel<T, NT, UJ, UN, PK, MB>, force_kernel<T, NT, UJ, UN, PK, MB, 1, 0, 0> }
// mid-size N: the same kernel also with the spread fix-up of split tiles
#define MVARIANT(id, T, NT, UJ, UN, PK, MB) \
    { id, "T" #T "_NT" #NT "_UJ" #UJ "_U" #UN "_" #PK "_B" #MB, T, NT, UJ, UN, PK, MB, 1, 0, force_kernel<T, NT, UJ, UN, PK, MB>, force_kernel<T, NT, UJ, UN, PK, MB, 1, 0, 0>, \
      force_kernel<T, NT, UJ, UN, PK, MB, 1, 0, kFoldEvery, 1>, force_kernel<T, NT, UJ, UN, PK, MB, 1, 0, 0, 1> }
#define CVARIANT(id, T, NT, UJ, UN, PK, MB, CS, CM) \
    { id, "T" #T "_NT" #NT "_UJ" #UJ "_U" #UN "_" #PK "_B" #MB "_C" #CS "m" #CM, T, NT, UJ, UN, PK, MB, CS, CM, force_kernel<T, NT, UJ, UN, PK, MB, CS, CM>, force_kernel<T, NT, UJ, UN, PK, MB, CS, CM, 0> }

#define SVARIANT(id, TT, NT, UN) \
    { id, "S_TT" #TT "_NT" #NT "_U" #UN, TT, NT, 64, UN, 1, 1, 1, 0, small_kernel<TT, NT, UN>, nullptr, nullptr, nullptr, 1 }
// row-split family (mid-size N that divides evenly over the SMs): T x H targets per CTA, no scratch
#define RVARIANT(id, T, H, UJ, UN, ORD) \
    { id, "R_T" #T "_H" #H "_NT256_UJ" #UJ "_U" #UN "_" #ORD, T, 256, UJ, UN, 1, 1, 1, 0, rowsplit_kernel<T, H, 256, UJ, UN, ORD>, nullptr, nullptr, nullptr, 2, H, rowsplit_kernel<T, H, 256, UJ, UN, ORD, 1> }
#define RVARIANT_L(id, T, H, UJ, UN, ORD, LOOK) \
    { id, "R_T" #T "_H" #H "_NT256_UJ" #UJ "_U" #UN "_" #ORD "_L" #LOOK, T, 256, UJ, UN, 1, 1, 1, 0, rowsplit_kernel<T, H, 256, UJ, UN, ORD, 0, LOOK>, nullptr, nullptr, nullptr, 2, H, rowsplit_kernel<T, H, 256, UJ, UN, ORD, 1, LOOK> }
// packed: 0 = scalar FFMA path (kept as the with/without evidence of the f32x2 choice), 1 = FADD2/FMUL2/FFMA2,
// 5 / 6 / 7 / 8 = the same packed arithmetic in another source order (bit-identical results; ptxas' register
// allocation differs by a few per cent, see tuned_large_variant).
// The default table holds exactly what NBODY_VARIANT_AUTO can resolve to (static_variant, mid_variant, the tuner's
// candidates) plus one scalar and one exemplar of each cluster mode.  The tile-shape sweep of round 1
// (profiles/r1_variant_sweep_*.jsonl) is rebuilt with `make sweep` (-DNBODY_SWEEP_VARIANTS, libnbody_b200_sweep.so).
const Variant kVariants[] = {
    VARIANT(3, 2, 256, 256, 8, 1, 3),        // few targets of many sources (a shard of a big system): 3 CTAs/SM
    VARIANT(9, 4, 256, 256, 4, 0, 2),        // scalar FFMA exemplar
    VARIANT(12, 1, 128, 64, 8, 1, 8),        // fewer than 2 048 targets of many sources
    // large N, 256-body units (12 k .. 64 k sources): one arithmetic, three source orders
    MVARIANT(90, 6, 256, 256, 2, 18, 1),
    VARIANT(97, 6, 256, 256, 2, 7, 1),
    VARIANT(98, 6, 256, 256, 2, 8, 1),
    // large N, 512-body units (>= 64 k sources): one arithmetic, three source orders
    VARIANT(101, 6, 256, 512, 2, 8, 1),
    VARIANT(102, 6, 256, 512, 2, 7, 1),
    VARIANT(105, 6, 256, 512, 2, 18, 1),
    // mid-size N (cost model in mid_variant): T8 tiles and 64- / 128-body units
    MVARIANT(93, 8, 256, 256, 2, 6, 1),
    MVARIANT(110, 8, 256, 128, 2, 6, 1),
    MVARIANT(111, 8, 256, 64, 2, 6, 1),
    MVARIANT(112, 6, 256, 128, 2, 18, 1),
    MVARIANT(113, 6, 256, 64, 2, 18, 1),
    MVARIANT(116, 8, 256, 32, 2, 6, 1),      // 32-body units: 8 k .. 11 k bodies, where a CTA has only 3 - 5 larger units
    MVARIANT(117, 6, 256, 32, 2, 18, 1),
    // thread-block-cluster exemplars (opt-in; measured neutral / slower, profiles/r1_ncu_cluster_evidence.md):
    // m1 = TMA multicast of source units, m2 = j-split with DSMEM reduction
    CVARIANT(60, 6, 256, 256, 2, 1, 1, 2, 1),
    CVARIANT(64, 6, 256, 256, 2, 1, 1, 2, 2),
    // small-N family: j-split inside the CTA, warp-shuffle reduction, no scratch
    SVARIANT(82, 4, 128, 2),
    SVARIANT(85, 2, 128, 4),
    // row-split family: whole target tiles per CTA, every CTA streams all sources (mid_variant picks them when
    // ceil(N / tile) fills the SMs evenly)
    RVARIANT(150, 5, 16, 1024, 4, 0),
    RVARIANT(151, 6, 16, 1024, 4, 0),
    RVARIANT(152, 7, 16, 1024, 4, 0),        // N = 16 384 (BASELINE config 2): 147 CTAs x 112 targets
    RVARIANT(153, 8, 16, 1024, 4, 0),
    RVARIANT(160, 5, 8, 1024, 4, 0),
    RVARIANT(161, 6, 8, 1024, 4, 0),
    RVARIANT(162, 7, 8, 1024, 4, 0),
    RVARIANT(163, 8, 8, 1024, 4, 0),
    RVARIANT(165, 6, 4, 1024, 4, 0),
    RVARIANT(166, 7, 4, 1024, 4, 0),
    RVARIANT(167, 8, 4, 1024, 4, 0),
    RVARIANT(154, 9, 16, 1024, 4, 0),        // T = 9 (242 registers) closes the gaps between the lane classes: tiles of 36, 72, 144
    RVARIANT(164, 9, 8, 1024, 4, 0),
    RVARIANT(168, 9, 4, 1024, 4, 0),
    RVARIANT(170, 5, 32, 1024, 4, 0),
    RVARIANT(171, 6, 32, 1024, 4, 0),
    RVARIANT(172, 7, 32, 1024, 4, 0),
    RVARIANT(173, 8, 32, 1024, 4, 0),
#ifdef NBODY_SWEEP_VARIANTS
    VARIANT(45, 6, 256, 256, 2, 1, 1),       // x,y,z source order: 2 % slower than the orders above once the sums are folded
    VARIANT(46, 6, 256, 256, 2, 5, 1),
    VARIANT(106, 6, 256, 512, 2, 1, 1),
    VARIANT(107, 6, 256, 512, 2, 5, 1),
    VARIANT(129, 6, 256, 1024, 2, 7, 1),
    VARIANT(140, 6, 256, 512, 2, 19, 1),     // scalar FMUL behind each MUFU.RSQ (experiment, profiles/r2_scalar_tail_experiment.txt)
    VARIANT(1, 4, 256, 256, 4, 1, 2),
    VARIANT(2, 2, 256, 256, 4, 1, 4),
    VARIANT(4, 4, 128, 256, 4, 1, 4),
    VARIANT(5, 8, 128, 256, 2, 1, 2),
    VARIANT(7, 4, 512, 256, 4, 1, 1),
    VARIANT(8, 1, 256, 256, 8, 1, 6),
    VARIANT(10, 8, 256, 256, 2, 0, 1),
    VARIANT(11, 2, 256, 256, 4, 0, 4),
    VARIANT(13, 2, 128, 128, 8, 1, 6),
    VARIANT(14, 3, 256, 256, 4, 1, 2),
    VARIANT(15, 6, 128, 256, 2, 1, 3),
    VARIANT(21, 8, 256, 256, 2, 1, 1),
    VARIANT(24, 4, 256, 512, 4, 1, 2),
    VARIANT(25, 4, 256, 1024, 4, 1, 2),
    VARIANT(34, 6, 256, 256, 2, 1, 2),
    VARIANT(35, 6, 512, 256, 2, 1, 1),
    VARIANT(36, 5, 256, 256, 2, 1, 2),
    VARIANT(38, 8, 384, 256, 2, 1, 1),
    VARIANT(40, 12, 256, 256, 1, 1, 1),
    VARIANT(44, 8, 256, 512, 2, 1, 1),
    VARIANT(48, 6, 256, 256, 4, 1, 1),
    VARIANT(53, 7, 256, 256, 2, 1, 1),
    VARIANT(55, 5, 384, 256, 2, 1, 1),
    VARIANT(56, 6, 384, 256, 2, 1, 1),
    VARIANT(91, 6, 256, 256, 2, 6, 1),
    VARIANT(92, 8, 256, 256, 2, 5, 1),
    VARIANT(96, 8, 256, 512, 2, 6, 1),
    VARIANT(103, 6, 256, 1024, 2, 18, 1),
    VARIANT(114, 4, 256, 128, 4, 1, 2),
    VARIANT(115, 8, 256, 128, 2, 1, 1),
    CVARIANT(61, 6, 256, 256, 2, 1, 1, 4, 1),
    CVARIANT(62, 6, 256, 512, 2, 1, 1, 2, 1),
    CVARIANT(65, 6, 256, 256, 2, 1, 1, 4, 2),
    CVARIANT(66, 2, 256, 256, 8, 1, 3, 2, 2),
    CVARIANT(67, 2, 256, 256, 8, 1, 3, 4, 2),
    CVARIANT(70, 6, 256, 256, 2, 1, 1, 8, 2),
    RVARIANT(180, 7, 16, 512, 4, 0),         // unit size / unroll evidence of the row-split family (profiles/r2_rowsplit.txt)
    RVARIANT(181, 7, 16, 512, 2, 0),
    RVARIANT(182, 7, 16, 256, 2, 0),
    RVARIANT(183, 7, 16, 512, 8, 0),
    RVARIANT(185, 7, 16, 1024, 4, 1),
    RVARIANT(187, 7, 16, 1024, 4, 8),
    RVARIANT(188, 7, 16, 1024, 4, 9),
    RVARIANT(189, 7, 16, 1024, 4, 13),
    RVARIANT(190, 7, 16, 1024, 2, 0),
    RVARIANT_L(191, 7, 16, 1024, 4, 0, 3),
    RVARIANT_L(192, 7, 16, 1024, 4, 0, 1),
    RVARIANT(186, 7, 16, 1024, 4, 5),         // source orders, unroll 2 and ring look-ahead 3 / 1 at UJ 1024: none beats 152
    SVARIANT(80, 8, 128, 2),
    SVARIANT(81, 8, 256, 2),
    SVARIANT(83, 16, 256, 1),
    SVARIANT(84, 8, 64, 2),
    SVARIANT(86, 8, 512, 1),
#endif
};
const int kNumVariants = (int)(sizeof kVariants / sizeof kVariants[0]);
constexpr size_t kFlagBytes = 4096;   // barrier flags behind packed[0]
constexpr int kSlotAlign = 1024;   // bodies; multiple of every variant's UJ

const Variant* find_variant(int id)
{
    for (int i = 0; i < kNumVariants; ++i)
        if (kVariants[i].id == id) return &kVariants[i];
    return nullptr;
}

int tuned_large_variant(int n_sources, bool deterministic);
int mid_variant(int n_targets, int n_sources, int sms);

// The static part of the choice (best of the r1 sweeps, profiles/r1_variant_sweep_*.jsonl; the cluster variants
// 60-70 measured neutral (C2 multicast) or slower (C4/C8, DSMEM j-split) and are opt-in only).
// *measured = true: one of the equivalent large-N kernels, to be picked by tuned_large_variant().
int env_int(const char* name, int dflt)
{
    const char* e = getenv(name);
    return e && *e ? atoi(e) : dflt;
}
// Bodies below which the small-N family runs.  Measured with the spread fix-up (tools/experiment_small_below.sh,
// profiles/r2_small_below.txt, best of three 3 200-step runs per size): from 8 192 bodies up the stream-K kernels tie
// or win at every size tried (0 .. 15 %); between 5 632 and 7 680 the two alternate with the rounding of units per
// CTA / CTAs per wave (+8 % at 6 144, -7 % at 7 680), so the rule stays simple.
// NBODY_SMALL_BELOW in the environment moves it (tuning switch).
const int g_small_below = env_int("NBODY_SMALL_BELOW", 8192);
// Row-split family (rowsplit_kernel): whole target tiles of T x H bodies per CTA, one wave.  Its time is one CTA's:
// tile x sources at the shape's measured rate + a fixed 3.5 us (launch, prologue, reduction); it wins where
// ceil(N / tile) fills the SMs evenly, because nothing crosses CTAs (no partial sums through global memory).
// Rates: G interactions / s on 148 SMs at N = 148 x tile (tools/experiment_rowsplit.sh, profiles/r2_rowsplit.txt).
int round_up(int v, int m) { return (v + m - 1) / m * m; }
struct RowCand { int id, T, H; double rate; };
const RowCand kRowCands[] = {
    {165, 6, 4, 2500.}, {166, 7, 4, 2500.}, {167, 8, 4, 2500.}, {168, 9, 4, 2450.},
    {160, 5, 8, 2500.}, {161, 6, 8, 2640.}, {162, 7, 8, 2620.}, {163, 8, 8, 2645.}, {164, 9, 8, 2600.},
    {150, 5, 16, 2550.}, {151, 6, 16, 2600.}, {152, 7, 16, 2690.}, {153, 8, 16, 2600.}, {154, 9, 16, 2600.},
    {170, 5, 32, 2625.}, {171, 6, 32, 2640.}, {172, 7, 32, 2680.}, {173, 8, 32, 2645.},
};
const bool g_rowsplit = getenv("NBODY_NO_ROWSPLIT") == nullptr;   // with / without switch of the experiments
int row_variant(int n, int sms, double* t_us)
{
    int best = 0;
    *t_us = 1e300;
    for (const RowCand& c : kRowCands) {
        const int tile = c.T * c.H;
        if ((n + tile - 1) / tile > sms) continue;   // one wave only: a second wave of whole tiles loses to stream-K
        const int quantum = 2 * (32 / c.H) * 8 * 4;   // bodies per round of the CTA's source split (unroll 4)
        const double t = (double)tile * (double)round_up(n, quantum) * 148.0 / (c.rate * 1e3) + 3.5;
        if (t < *t_us) { *t_us = t; best = c.id; }
    }
    return best;
}

int static_variant(int n_targets, int n_sources, int sms, bool* measured)
{
    *measured = false;
    if (g_rowsplit && n_targets == n_sources && n_sources >= 2048 && n_sources < 65536) {
        // unsharded mid-size system: row-split where its whole tiles fill the SMs well enough to beat the others'
        // measured times (small-N family: N^2 / 0.95e6 + 0.7 us below 3 072 bodies, N^2 / 1.9e6 + 3.2 us above; stream-K
        // kernels: 1.04 N^2 / 2.69e6 + 12 us on 148 SMs;
        // both a little optimistic for the others, so a tie goes to them)
        if (sms <= 0) {
            int dev = 0;
            sms = 148;
            if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        }
        double t_row = 0.0;
        const int id = row_variant(n_sources, sms, &t_row);
        const double n2 = (double)n_sources * (double)n_sources;
        const double t_other = n_sources < 3072 ? n2 / 0.95e6 + 0.7                       // small-N family, 2 targets per CTA
                               : n_sources < g_small_below ? n2 / 1.9e6 + 3.2            // ... 4 targets per CTA
                                                           : 1.04 * n2 / (2.69e6 * sms / 148.0) + 12.0;
        if (id && t_row < 0.99 * t_other) return id;
    }
    if (n_sources < g_small_below) {     // a step is < 50 us: latency decides -> small-N family
        if (n_sources < 3072) return 85; // 2 targets / CTA, sources split over 128 threads, shuffle reduction
        return 82;                       // 4 targets / CTA
    }
    if (n_targets >= g_small_below && n_sources < 65536) return mid_variant(n_targets, n_sources, sms);
    if (n_targets >= 12288) { *measured = true; return n_sources < 65536 ? 90 : 101; }
    if (n_targets >= 2048) return 3;     // T2 x 256 threads, 3 CTAs/SM
    return 12;                           // T1 x 128 threads, 64-body units
}

int auto_variant(int n_targets, int n_sources, bool deterministic = false)
{
    bool measured = false;
    const int id = static_variant(n_targets, n_sources, 0, &measured);
    return measured ? tuned_large_variant(n_sources, deterministic) : id;
}

// Mid-size N (12 k ... 64 k bodies, a step of 0.1 - 1.5 ms): the kernel's time is (units of the busiest CTA) x
// (time per unit), and with only a few units per CTA the rounding of units / CTAs and the padding of the last
// target tile are worth several per cent (N = 16 384, T6 x 256-body units: 704 units on 148 CTAs -> 5 where 4.76
// are needed, and the 11th tile is 2/3 empty).  Pick the tile shape and unit size that minimise that product.
int mid_variant(int n_targets, int n_sources, int sms)
{
    struct Cand { int id, T, UJ; };
    static const Cand cands[] = {{90, 6, 256}, {112, 6, 128}, {113, 6, 64}, {117, 6, 32}, {93, 8, 256}, {110, 8, 128}, {111, 8, 64}, {116, 8, 32}};
    if (sms <= 0) {
        int dev = 0;
        sms = 148;
        if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    // an unsharded context visits the units up to its last body; a shard sees the other shards' slots, padded to kSlotAlign
    const bool whole = n_targets == n_sources;
    double best_cost = 1e300;
    int best = 90;
    for (const Cand& c : cands) {
        const long long src_pad = whole ? (long long)(n_sources + c.UJ - 1) / c.UJ * c.UJ : (long long)(n_sources + kSlotAlign - 1) / kSlotAlign * kSlotAlign;
        const long long tiles = (n_targets + c.T * 256 - 1) / (c.T * 256), J = src_pad / c.UJ, U = tiles * J;
        const long long maxu = (U + sms - 1) / sms;
        // cycles per unit: UJ/4 loop iterations of (T x 24 packed ops x 2 + ~35) cycles, plus ~150 of barrier / producer work
        // (the T8 kernels measure ~2.5 % slower per interaction than T6 at equal balance: 250 vs 192 registers)
        const double cost = (double)maxu * ((c.UJ / 4) * (c.T * 48.0 + 35.0) + 150.0) * (c.T == 8 ? 1.025 : 1.0);
        if (cost < best_cost) { best_cost = cost; best = c.id; }
    }
    return best;
}


// The kernels take rsqrt.approx.ftz of (r^2 + eps2)^3: for the self term and the zero-mass padding r = 0, so a
// subnormal eps2^3 is flushed to 0, rsqrt gives +inf and 0 * inf = NaN in every acceleration.  (The reference
// divides by sqrtf of the subnormal, which is finite.)  Such softenings are rejected instead of answered wrongly.
constexpr float kMinEps2 = 2.2737368e-13f;   // 2^-42: its cube 2^-126 is the smallest normal float
bool eps2_representable(float eps2)
{
    const float e3 = (eps2 * eps2) * eps2;   // as the kernel forms it (two rounded multiplies)
    return e3 >= 1.17549435e-38f;
}

// ---- launch plan for one (variant, targets, J) ---------------------------------------------------
struct Plan {
    int grid = 0, occ = 0, maxseg = 0, num_itiles = 0, num_counters = 0;
    size_t partial_bytes = 0;
    bool single = false;   // single-level summation kernel (nbody_params.summation)
    bool spread = false;   // split tiles finished by all their CTAs (force_kernel's SPREAD): cooperative launch
};

constexpr long long kSpreadBelowSources = 65536 + 1024;
bool g_spread_fixup = getenv("NBODY_NO_SPREAD_FIXUP") == nullptr;   // debugging switch: the last-arriver fix-up everywhere
bool cooperative_launch_ok()
{
    int dev = 0, ok = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return false;
    return cudaDeviceGetAttribute(&ok, cudaDevAttrCooperativeLaunch, dev) == cudaSuccess && ok != 0;
}

int make_plan(const Variant* v, int sm_count, int grid_override, int n_targets, int J, Plan* out, bool single = false)
{
    Plan pl;
    pl.single = single && v->kernel_single;
    if (v->small) {   // one CTA per tile of targets, every CTA visits all J units: nothing to cut, no scratch
        (void)sm_count; (void)grid_override; (void)J;
        pl.occ = 1;
        if (v->small == 2) {
            CU(cudaFuncSetAttribute(v->kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)v->smem()));
            CU(cudaFuncSetAttribute(v->kernel_trim, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)v->smem()));
        }
        pl.num_itiles = (n_targets + v->TI() - 1) / v->TI();
        pl.grid = pl.num_itiles;
        pl.maxseg = 1; pl.num_counters = 1; pl.partial_bytes = sizeof(float4);
        *out = pl;
        return NBODY_OK;
    }
    CU(cudaFuncSetAttribute(v->entry(pl.single), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)v->smem(pl.single)));
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&pl.occ, v->entry(pl.single), v->NT, v->smem(pl.single)));
    if (pl.occ < 1) return fail(NBODY_E_CUDA, "variant %s does not fit on an SM", v->name);
    long long workers = (long long)sm_count * pl.occ;   // CTAs, or clusters for the cluster variants
    if (v->cs > 1) {
        cudaLaunchConfig_t cfg = {};
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = v->cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.gridDim = dim3((unsigned)(sm_count * pl.occ / v->cs * v->cs)); cfg.blockDim = dim3(v->NT);
        cfg.dynamicSmemBytes = v->smem(pl.single); cfg.attrs = at; cfg.numAttrs = 1;
        int ncl = 0;
        CU(cudaOccupancyMaxActiveClusters(&ncl, v->entry(pl.single), &cfg));
        if (ncl < 1) return fail(NBODY_E_CUDA, "no cluster of %d CTAs of variant %s can be resident", v->cs, v->name);
        workers = ncl;
    }
    pl.num_itiles = (n_targets + v->TS() - 1) / v->TS();
    const unsigned long long U = (unsigned long long)pl.num_itiles * (unsigned long long)J;
    if (grid_override > 0) workers = v->cs > 1 ? (grid_override / v->cs > 0 ? grid_override / v->cs : 1) : grid_override;
    if ((unsigned long long)workers > U) workers = (long long)U;
    if (workers < 1) workers = 1;
    pl.grid = (int)workers * v->cs;
    // Mid-size N: few tiles, each cut into many segments, and the last CTA to arrive would sum them all alone.
    // Spread that over the tile's CTAs when the launch can be cooperative (the CTAs wait for each other).  Measured
    // (tools/experiment_spread.sh, profiles/r2_spread_fixup.txt), spread | last arriver: 73.1 | 76.1 us at N = 12 288,
    // 114.0 | 116.9 us at 16 384, 174.8 | 176.1 us at 20 480; with 12 or fewer cuts per tile (24 576 bodies and up) the
    // fix-up is too small to matter and the other instantiation's register allocation costs more than it saves
    // (252 | 242 us at 24 576), so the rule asks for 14 cuts per tile.
    if (v->kernel_spread && g_spread_fixup && (long long)J * v->UJ < kSpreadBelowSources && workers >= 14LL * pl.num_itiles &&
        cooperative_launch_ok()) {
        int occ = 0;
        CU(cudaFuncSetAttribute(v->entry(pl.single, true), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)v->smem(pl.single)));
        CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, v->entry(pl.single, true), v->NT, v->smem(pl.single)));
        pl.spread = workers <= (long long)sm_count * occ;
    }
    const unsigned long long q = U / (unsigned long long)workers;   // smallest range, >= 1 when U > 0
    pl.maxseg = q > 0 ? (int)((unsigned long long)(J - 1) / q) + 2 : 2;
    pl.num_counters = pl.num_itiles * v->cs;
    pl.partial_bytes = (size_t)pl.num_itiles * (v->cmode == 1 ? v->cs : 1) * pl.maxseg * v->TI() * sizeof(float4);
    *out = pl;
    return NBODY_OK;
}

int launch_force(const Variant* v, const Plan& pl, const ForceArgs& args_in, cudaStream_t st)
{
    const ForceArgs& args = args_in;
    if (pl.spread) {
        cudaLaunchConfig_t cfg = {};
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeCooperative;
        at[0].val.cooperative = 1;
        cfg.gridDim = dim3(pl.grid); cfg.blockDim = dim3(v->NT); cfg.dynamicSmemBytes = v->smem(pl.single); cfg.stream = st;
        cfg.attrs = at; cfg.numAttrs = 1;
        const cudaError_t e = cudaLaunchKernelEx(&cfg, v->entry(pl.single, true), args);
        if (e == cudaErrorCooperativeLaunchTooLarge) {
            // fewer SMs than the plan saw (MPS with an SM limit, a green context): the last-arriver kernel needs no
            // co-residency and gives the same bits
            (void)cudaGetLastError();
            v->entry(pl.single)<<<pl.grid, v->NT, v->smem(pl.single), st>>>(args);
        } else {
            CU(e);
        }
    } else if (v->cs > 1) {
        cudaLaunchConfig_t cfg = {};
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = v->cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.gridDim = dim3(pl.grid); cfg.blockDim = dim3(v->NT); cfg.dynamicSmemBytes = v->smem(pl.single); cfg.stream = st;
        cfg.attrs = at; cfg.numAttrs = 1;
        CU(cudaLaunchKernelEx(&cfg, v->entry(pl.single), args));
    } else if (v->small == 2 && args.last_unit_bodies > 0 && args.rb_count == 0) {
        v->kernel_trim<<<pl.grid, v->NT, v->smem(), st>>>(args);   // ragged last unit: its padding is not visited
    } else {
        v->entry(pl.single)<<<pl.grid, v->NT, v->smem(pl.single), st>>>(args);
    }
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CU(cudaGetLastError());
    return NBODY_OK;
}

int launch_pack(const float4* X, int n, float4* packed, int off, int n_cover, cudaStream_t st)
{
    if (n_cover <= 0) return NBODY_OK;
    const int pairs = (n_cover + 1) / 2;
    pack_kernel<<<(pairs + 255) / 256, 256, 0, st>>>(X, n, packed, off, n_cover);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    CU(cudaGetLastError());
    return NBODY_OK;
}

// ---- which of the equivalent large-N kernels is fastest on this device ---------------------------------
// The candidates of one family (T6 x 256 threads with 256-body units below 64 k sources, with 512-body units from
// there on) issue the same instructions per interaction and form every sum in the same order: they differ only
// in the order the source code names the targets / axes, i.e. in the register allocation ptxas happens to give
// them; measured, that is worth +-3.5 % (register-bank conflicts on the 3-operand FFMA2; profiles/r1_variant_sweep_*).
// Rather than trust one allocation, the first large context of a process times each candidate for ~1 ms on
// synthetic bodies and keeps the fastest.  Because tile shape, unit size and grid are the same inside a family,
// the RESULT does not depend on which one wins: bit-identical across processes, boxes and ranks
// (tests/test_parity_gpu.py::test_tuner_candidates_are_bit_identical).  NBODY_LARGE_VARIANT=<id> in the
// environment overrides; nbody_params.deterministic = 1 skips the timing and takes the first of the family.
const int kTuneCandidates[] = {90, 97, 98, 101, 102, 105};
constexpr int kTuneBodies = 49152;   // whole T6 tiles (1536): 0.9 ms per launch
std::mutex g_tune_mu;
std::map<int, std::pair<int, int>> g_tuned;   // device -> (best with 256-body units, best overall)
std::string g_tune_report;

int tuned_large_variant(int n_sources, bool deterministic)
{
    // 512-body units win at N >= 64 k (fewer barrier round trips) but cut a 16 k-body step into too few pieces
    const bool small_units = n_sources < 65536;
    int dev = 0;
    if (deterministic || cudaGetDevice(&dev) != cudaSuccess) return small_units ? 90 : 101;
    std::lock_guard<std::mutex> lk(g_tune_mu);
    auto it = g_tuned.find(dev);
    if (it != g_tuned.end()) return small_units ? it->second.first : it->second.second;
    int best = 101, best256 = 90;
    if (const char* env = getenv("NBODY_LARGE_VARIANT")) {
        const int id = atoi(env);
        if (find_variant(id) && !find_variant(id)->small) { g_tuned[dev] = std::make_pair(id, id); g_tune_report = "forced by NBODY_LARGE_VARIANT"; return id; }
    }
    cudaDeviceProp prop;
    float4 *X = nullptr, *V = nullptr, *A = nullptr, *src = nullptr, *partial = nullptr;
    unsigned* counters = nullptr;
    cudaStream_t st = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    const size_t nb = (size_t)kTuneBodies * sizeof(float4);
    bool ok = cudaGetDeviceProperties(&prop, dev) == cudaSuccess && cudaMalloc(&X, nb) == cudaSuccess && cudaMalloc(&V, nb) == cudaSuccess &&
              cudaMalloc(&A, nb) == cudaSuccess && cudaMalloc(&src, nb) == cudaSuccess && cudaMalloc(&counters, 4096) == cudaSuccess &&
              cudaMemset(X, 0, nb) == cudaSuccess && cudaMemset(V, 0, nb) == cudaSuccess && cudaMemset(src, 0, nb) == cudaSuccess &&
              cudaMemset(counters, 0, 4096) == cudaSuccess && cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) == cudaSuccess &&
              cudaEventCreate(&e0) == cudaSuccess && cudaEventCreate(&e1) == cudaSuccess;
    size_t partial_bytes = 0;
    float best_ms = 1e30f, best256_ms = 1e30f;
    char line[64];
    g_tune_report.clear();
    for (int id : kTuneCandidates) {
        const Variant* v = find_variant(id);
        Plan pl;
        if (!ok || !v || make_plan(v, prop.multiProcessorCount, 0, kTuneBodies, kTuneBodies / v->UJ, &pl) != NBODY_OK) continue;
        if (pl.partial_bytes > partial_bytes) {
            cudaFree(partial);
            partial = nullptr;
            if (cudaMalloc(&partial, pl.partial_bytes) != cudaSuccess) { ok = false; continue; }
            partial_bytes = pl.partial_bytes;
        }
        ForceArgs a{};
        a.X = X; a.V = V; a.A = A; a.n_targets = kTuneBodies; a.src = src;
        a.ra_count = kTuneBodies / v->UJ;
        a.partial = partial; a.counters = counters; a.maxseg = pl.maxseg;
        a.eps2 = 0.002f; a.dt = 0.f;   // dt = 0: the zeroed state stays put, every launch does identical work
        float ms = 1e30f;
        for (int rep = 0; rep < 3; ++rep) {   // first launch also loads the kernel image
            cudaEventRecord(e0, st);
            if (launch_force(v, pl, a, st) != NBODY_OK) { ms = 1e30f; break; }
            cudaEventRecord(e1, st);
            float t = 0.f;
            if (cudaEventSynchronize(e1) != cudaSuccess || cudaEventElapsedTime(&t, e0, e1) != cudaSuccess) { ms = 1e30f; break; }
            if (rep > 0 && t < ms) ms = t;
        }
        snprintf(line, sizeof line, "%s%d:%.3fms", g_tune_report.empty() ? "" : " ", id, ms);
        g_tune_report += line;
        if (v->UJ == 512 && ms < best_ms) { best_ms = ms; best = id; }
        if (v->UJ == 256 && ms < best256_ms) { best256_ms = ms; best256 = id; }
    }
    cudaGetLastError();
    cudaFree(X); cudaFree(V); cudaFree(A); cudaFree(src); cudaFree(partial); cudaFree(counters);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    if (st) cudaStreamDestroy(st);
    g_tuned[dev] = std::make_pair(best256, best);
    return small_units ? best256 : best;
}

}  // namespace

// ---- context ---------------------------------------------------------------------------------------
struct nbody_ctx {
    nbody_params p{};
    int sm_count = 0;
    const Variant* var = nullptr;
    int n = 0, per = 0, t_begin = 0, n_targets = 0;
    int slot_bodies = 0, n_src_pad = 0;          // packed geometry (bodies)
    float4 *X = nullptr, *V = nullptr, *A = nullptr, *accum = nullptr;
    float4* packed[2] = {nullptr, nullptr};
    int cur = 0;
    float4* partial = nullptr;
    unsigned* counters = nullptr;
    double* energy_part = nullptr;
    unsigned long long* clk = nullptr;          // [0..3] {clock64, globaltimer} x {entry, exit} of CTA 0 of the last force kernel; [4..6] flag-wait ns: sum, max, count
    float4* stage = nullptr;                     // upload staging, float4[n]
    Plan plan_full, plan_local, plan_remote;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool timed = false;
    unsigned char* frame = nullptr;              // nbody_render's device RGBA8 frame
    unsigned long long* frame_keys = nullptr;    // per-pixel depth|index keys of that frame
    size_t frame_bytes = 0;
    long long steps_done = 0;                    // full steps since the last upload / snapshot load
    // launch-bound step loops (small N): kGraphSteps consecutive steps captured once as a CUDA graph
    // (buffer parity 0 at entry); replayed by nbody_step_async
    cudaGraphExec_t step_graph = nullptr;
    bool graph_failed = false;
    // fused exchange: peers' packed buffers [peer rank][0/1]; ipc = opened with cudaIpcOpenMemHandle
    float4* peer_packed[kMaxPeers + 1][2] = {};
    bool peer_ipc[kMaxPeers + 1] = {};
    int peers_attached = 0;
    // in-kernel barrier of the fused exchange: flags live behind packed[0] (same allocation, so the peers' IPC
    // mapping of that buffer reaches them); fused_steps only ever grows, so re-uploads need no flag reset
    bool device_barrier = false;
    unsigned long long fused_steps = 0;
    unsigned* bar_done = nullptr;                // 256 B: [0] CTA arrival counter, [16] time-out mark
    // multi-device context (nbody_params.n_devices > 1, nbody_group.inl): this object only holds the per-device
    // sharded contexts and what orders them; every other member above stays empty
    std::vector<nbody_ctx*> kids;
    struct Group* group = nullptr;
};

namespace {

int group_create(nbody_ctx** out, const nbody_params* p);   // nbody_group.inl
void group_destroy(nbody_ctx* c);

int per_rank(int n, int world) { return round_up((n + world - 1) / world, 2); }

ForceArgs base_args(nbody_ctx* c)
{
    ForceArgs a{};
    a.X = c->X; a.V = c->V; a.A = c->A;
    a.n_targets = c->n_targets;
    a.src = c->packed[c->cur];
    a.partial = c->partial;
    a.counters = c->counters;
    a.eps2 = c->p.integrator == NBODY_INTEG_F64 ? (float)c->p.eps2_f64 : c->p.eps2;
    a.dt = c->p.dt;
    a.integ_f64 = c->p.integrator;   // 0 float, 1 double (headless flavour), 2 mixed (reduction flavours)
    a.dt64 = c->p.dt_f64;
    a.clk = c->clk;
    return a;
}

int set_device(nbody_ctx* c)
{
    CU(cudaSetDevice(c->p.device));
    return NBODY_OK;
}

// Source units a step visits: the packed buffer is padded to kSlotAlign bodies per shard (zero-mass bodies), but
// the units behind the last real body hold nothing and are not visited (N = 7 680: 120 units of 64, not 128).
int units_in_use(const nbody_ctx* c)
{
    const int UJ = c->var->UJ, last = c->n - (c->p.world - 1) * c->per;   // bodies of the last shard
    return ((c->p.world - 1) * c->slot_bodies + round_up(last, UJ)) / UJ;
}

// real bodies in the last of those units (0 = the unit is full)
int last_unit_fill(const nbody_ctx* c)
{
    const int last = c->n - (c->p.world - 1) * c->per;
    return last % c->var->UJ;
}

// ... and of this shard's own slot (the last shard's slot ends early)
int units_own_in_use(const nbody_ctx* c)
{
    const int own = c->slot_bodies / c->var->UJ, left = units_in_use(c) - c->p.rank * own;
    return left < own ? left : own;
}

// launch plans of c->var for this context's geometry + the split-tile scratch they need
int plan_scratch(nbody_ctx* c)
{
    const int units_total = units_in_use(c), units_own = units_own_in_use(c);
    const bool single = c->p.summation == NBODY_SUM_SINGLE;
    int rc = make_plan(c->var, c->sm_count, c->p.grid_ctas, c->n_targets, units_total, &c->plan_full, single);
    if (rc) return rc;
    size_t partial_bytes = c->plan_full.partial_bytes;
    if (c->p.world > 1) {
        rc = make_plan(c->var, c->sm_count, c->p.grid_ctas, c->n_targets, units_own, &c->plan_local, single);
        if (rc) return rc;
        rc = make_plan(c->var, c->sm_count, c->p.grid_ctas, c->n_targets, units_total - units_own, &c->plan_remote, single);
        if (rc) return rc;
        if (c->plan_local.partial_bytes > partial_bytes) partial_bytes = c->plan_local.partial_bytes;
        if (c->plan_remote.partial_bytes > partial_bytes) partial_bytes = c->plan_remote.partial_bytes;
    }
    cudaFree(c->partial); cudaFree(c->counters);
    c->partial = nullptr; c->counters = nullptr;
    CU(cudaMalloc(&c->partial, partial_bytes));
    CU(cudaMalloc(&c->counters, (size_t)c->plan_full.num_counters * sizeof(unsigned)));
    CU(cudaMemset(c->counters, 0, (size_t)c->plan_full.num_counters * sizeof(unsigned)));
    return NBODY_OK;
}

// run another kernel variant from now on (nbody_snapshot_load re-pins the saver's)
int repin_variant(nbody_ctx* c, const Variant* v)
{
    int rc = set_device(c);
    if (rc) return rc;
    CU(cudaStreamSynchronize(c->stream));
    if (c->step_graph) { cudaGraphExecDestroy(c->step_graph); c->step_graph = nullptr; }
    c->graph_failed = false;
    const Variant* old = c->var;
    c->var = v;
    rc = plan_scratch(c);
    if (rc) { c->var = old; plan_scratch(c); }
    return rc;
}

}  // namespace

extern "C" {

const char* nbody_last_error(void) { return g_err.c_str(); }
long long nbody_launch_count(void) { return g_launches.load(); }

int nbody_params_default(nbody_params* p)
{
    if (!p) return fail(NBODY_E_INVALID, "null params");
    memset(p, 0, sizeof *p);
    p->n_bodies = 8192;      // TestProject/constants.h:13
    p->rank = 0; p->world = 1;
    p->eps2 = 0.002f;        // constants.h:25
    p->dt = 0.1f;            // constants.h:26
    p->integrator = NBODY_INTEG_F32;
    p->eps2_f64 = 0.002;     // Sim-Without-OpenGL-Integration/constants.h:14
    p->dt_f64 = 0.01;        // Sim-Without-OpenGL-Integration/constants.h:15
    return NBODY_OK;
}

int nbody_variant_info(int id, char* buf, size_t buflen)
{
    const Variant* v = find_variant(id);
    if (buf && buflen) buf[0] = 0;
    if (v && buf && buflen) {
        char cl[192] = "";
        if (v->cs > 1) snprintf(cl, sizeof cl, ", clusters of %d CTAs (%s)", v->cs, v->cmode == 1 ? "TMA multicast of source units" : "j-split, DSMEM reduction");
        if (v->small == 1) snprintf(cl, sizeof cl, ", small-N family: %d targets/CTA, sources split over the CTA, warp-shuffle reduction", v->T);
        if (v->small == 2) snprintf(cl, sizeof cl, ", row-split family: %d targets/CTA (%d lanes x T), every CTA streams all sources, split %d ways over its warps", v->TI(), v->H, (32 / v->H) * (v->NT / 32));
        snprintf(buf, buflen, "%s: %d targets/thread, %d threads/CTA, %d sources/tile, unroll %d pairs, %s FP32, min %d CTAs/SM, %zu B smem%s",
                 v->name, v->T, v->NT, v->UJ, v->unroll,
                 v->packed == 0 ? "scalar" : v->packed == 1 ? "packed f32x2" : v->packed == 5 ? "packed f32x2, targets reversed" : v->packed == 6 ? "packed f32x2, target-outer order" : v->packed == 18 ? "packed f32x2, x,z,y order" : "packed f32x2, z-first order",
                 v->minb, v->smem(), cl);
    }
    return kNumVariants;
}

int nbody_autotune_report(char* buf, size_t buflen)
{
    if (!buf || !buflen) return fail(NBODY_E_INVALID, "null buffer");
    std::lock_guard<std::mutex> lk(g_tune_mu);
    snprintf(buf, buflen, "%s", g_tune_report.c_str());
    return NBODY_OK;
}

int nbody_choose_variant(int n_targets, int n_sources, int sm_count, int* measured)
{
    if (n_targets <= 0 || n_sources < n_targets) {
        fail(NBODY_E_INVALID, "need 0 < n_targets <= n_sources");
        return 0;   // no variant has id 0
    }
    bool m = false;
    const int id = static_variant(n_targets, n_sources, sm_count > 0 ? sm_count : 148, &m);
    if (measured) *measured = m ? 1 : 0;
    return id;
}

int nbody_variant_id(int index)
{
    return index >= 0 && index < kNumVariants ? kVariants[index].id : 0;
}

size_t nbody_shard_slot_bytes(int n_bodies, int world)
{
    if (n_bodies <= 0 || world <= 0) return 0;
    return (size_t)round_up(per_rank(n_bodies, world), kSlotAlign) * sizeof(float4);
}

void nbody_destroy(nbody_ctx* c)
{
    if (!c) return;
    if (c->group || !c->kids.empty()) { group_destroy(c); return; }
    cudaSetDevice(c->p.device);
    for (int r = 0; r <= kMaxPeers; ++r)
        if (c->peer_ipc[r]) { cudaIpcCloseMemHandle(c->peer_packed[r][0]); cudaIpcCloseMemHandle(c->peer_packed[r][1]); }
    if (c->step_graph) cudaGraphExecDestroy(c->step_graph);
    cudaFree(c->X); cudaFree(c->V); cudaFree(c->A); cudaFree(c->accum);
    cudaFree(c->packed[0]); cudaFree(c->packed[1]);
    cudaFree(c->partial); cudaFree(c->counters); cudaFree(c->energy_part); cudaFree(c->clk); cudaFree(c->bar_done); cudaFree(c->stage); cudaFree(c->frame); cudaFree(c->frame_keys);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

int nbody_create(nbody_ctx** out, const nbody_params* p)
{
    TRACE_CALL();
    if (!out || !p) return fail(NBODY_E_INVALID, "null argument");
    *out = nullptr;
    if (p->n_bodies <= 0) return fail(NBODY_E_INVALID, "n_bodies must be > 0 (got %d)", p->n_bodies);
    if (p->world < 1 || p->rank < 0 || p->rank >= p->world) return fail(NBODY_E_INVALID, "bad shard %d of %d", p->rank, p->world);
    const bool f64 = p->integrator == NBODY_INTEG_F64;
    if (!f64 && p->integrator != NBODY_INTEG_F32 && p->integrator != NBODY_INTEG_MIXED) return fail(NBODY_E_INVALID, "unknown integrator %d", p->integrator);
    if (p->summation != NBODY_SUM_TWO_LEVEL && p->summation != NBODY_SUM_SINGLE) return fail(NBODY_E_INVALID, "unknown summation %d", p->summation);
    // the self term is 0 * m / eps2^1.5 and the zero-mass padding sits at the
    // origin: both need eps2 > 0 to stay finite
    if (!((f64 ? p->eps2_f64 : (double)p->eps2) > 0.0)) return fail(NBODY_E_INVALID, "eps2 must be > 0");
    if (!eps2_representable(f64 ? (float)p->eps2_f64 : p->eps2))
        return fail(NBODY_E_INVALID, "eps2 = %g is too small: eps2^3 must be a normal float (eps2 >= %g), or the self term 0 * rsqrt(eps2^3) is NaN", f64 ? p->eps2_f64 : (double)p->eps2, (double)kMinEps2);
    if (p->n_devices > 1) return group_create(out, p);
    int ndev = 0;
    CU(cudaGetDeviceCount(&ndev));
    if (p->device < 0 || p->device >= ndev) return fail(NBODY_E_INVALID, "device %d not present (%d devices)", p->device, ndev);

    nbody_ctx* c = new nbody_ctx;
    c->p = *p;
    int rc = set_device(c);
    if (rc) { delete c; return rc; }
    cudaDeviceProp prop;
    cudaError_t e = cudaGetDeviceProperties(&prop, p->device);
    if (e != cudaSuccess) { delete c; return fail(NBODY_E_CUDA, "cudaGetDeviceProperties: %s", cudaGetErrorString(e)); }
    if (prop.major < 10) { delete c; return fail(NBODY_E_CUDA, "device %d is sm_%d%d; this library is built for sm_100a only", p->device, prop.major, prop.minor); }
    c->sm_count = prop.multiProcessorCount;

    c->n = p->n_bodies;
    c->per = per_rank(c->n, p->world);
    c->t_begin = p->rank * c->per;
    c->n_targets = c->t_begin >= c->n ? 0 : (c->n - c->t_begin < c->per ? c->n - c->t_begin : c->per);
    if (c->n_targets <= 0) { delete c; return fail(NBODY_E_INVALID, "shard %d of %d owns no bodies at N=%d", p->rank, p->world, p->n_bodies); }
    c->slot_bodies = round_up(c->per, kSlotAlign);
    c->n_src_pad = c->slot_bodies * p->world;

    const int vid = p->variant == NBODY_VARIANT_AUTO ? auto_variant(c->n_targets, c->n, p->deterministic != 0) : p->variant;
    c->var = find_variant(vid);
    if (!c->var) { delete c; return fail(NBODY_E_INVALID, "unknown kernel variant %d", vid); }

#define TRY(x) do { rc = (x); if (rc) { nbody_destroy(c); return rc; } } while (0)
#define TRYCU(x) do { cudaError_t e2 = (x); if (e2 != cudaSuccess) { nbody_destroy(c); return fail(NBODY_E_CUDA, "%s: %s", #x, cudaGetErrorString(e2)); } } while (0)
    TRY(plan_scratch(c));
    const size_t tb = (size_t)c->n_targets * sizeof(float4);
    const size_t sb = (size_t)c->n_src_pad * sizeof(float4);
    TRYCU(cudaMalloc(&c->X, tb));
    TRYCU(cudaMalloc(&c->V, tb));
    TRYCU(cudaMalloc(&c->A, tb));
    TRYCU(cudaMalloc(&c->accum, tb));
    TRYCU(cudaMalloc(&c->stage, (size_t)c->n * sizeof(float4)));
    TRYCU(cudaMalloc(&c->packed[0], sb + kFlagBytes));
    TRYCU(cudaMemset(c->packed[0], 0, sb + kFlagBytes));
    TRYCU(cudaMalloc(&c->bar_done, 256));
    TRYCU(cudaMemset(c->bar_done, 0, 256));
    TRYCU(cudaMalloc(&c->packed[1], sb));
    TRYCU(cudaMalloc(&c->clk, 8 * sizeof(unsigned long long)));
    TRYCU(cudaMemset(c->clk, 0, 8 * sizeof(unsigned long long)));
    TRYCU(cudaMemset(c->V, 0, tb));
    TRYCU(cudaMemset(c->A, 0, tb));
    TRYCU(cudaMemset(c->packed[0], 0, sb));
    TRYCU(cudaMemset(c->packed[1], 0, sb));
    TRYCU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    TRYCU(cudaEventCreate(&c->ev0));
    TRYCU(cudaEventCreate(&c->ev1));
    TRYCU(cudaDeviceSynchronize());
#undef TRY
#undef TRYCU
    *out = c;
    return NBODY_OK;
}

int nbody_ctx_info(nbody_ctx* c, int* variant, int* grid_ctas, int* ctas_per_sm, int* sm_count)
{
    if (!c) return fail(NBODY_E_INVALID, "null context");
    if (!c->kids.empty()) c = c->kids[0];   // every device of a multi-device context runs the same plan
    if (variant) *variant = c->var->id;
    if (grid_ctas) *grid_ctas = c->plan_full.grid;
    if (ctas_per_sm) *ctas_per_sm = c->plan_full.occ;
    if (sm_count) *sm_count = c->sm_count;
    return NBODY_OK;
}

int nbody_ctx_fixup(nbody_ctx* c, int* spread)
{
    if (!c || !spread) return fail(NBODY_E_INVALID, "null argument");
    if (!c->kids.empty()) c = c->kids[0];
    *spread = c->plan_full.spread ? 1 : 0;
    return NBODY_OK;
}

int nbody_owned_range(nbody_ctx* c, int* begin, int* count)
{
    if (!c) return fail(NBODY_E_INVALID, "null context");
    if (!c->kids.empty()) {   // the whole system
        if (begin) *begin = 0;
        if (count) *count = c->n;
        return NBODY_OK;
    }
    if (begin) *begin = c->t_begin;
    if (count) *count = c->n_targets;
    return NBODY_OK;
}

}  // extern "C"

#include "nbody_engine_api.inl"
#include "nbody_render.inl"
#include "nbody_records.inl"
#include "nbody_group.inl"
#include "nbody_noalloc.inl"
