// nbody_kernels.cuh -- device code of the all-pairs step for sm_100a.
//
// Replaces TestProject/kernel.cu:9-29 (bodyInteractions), :55-65
// (tile_interaction) and :80-130 (kernel) of the reference -- same physics and
// integrator, different machine mapping:
//   * sources live in HBM in a PACKED PAIR layout (32 B per two bodies:
//     x0 x1 y0 y1 | z0 z1 m0 m1) so that one LDS.128 broadcast feeds the packed
//     FP32 pipe (FADD2/FMUL2/FFMA2) two interactions per lane per instruction;
//   * source tiles are staged by 1-D TMA bulk copies (UBLKCP) into a 4-stage
//     shared-memory ring guarded by full/empty mbarriers;
//   * each thread keeps T target bodies in registers;
//   * work = (target tile) x (source unit) pairs, linearised and cut into equal
//     contiguous ranges, one per persistent CTA ("stream-K" style), so that all
//     148 SMs finish together at any N.  A target tile whose source range is
//     split between CTAs is completed by the LAST CTA to arrive, which adds the
//     partial sums in fixed segment order (deterministic) and integrates -- or, at
//     mid-size N where a tile has many cuts, by ALL of its CTAs, 1 / nseg of the
//     targets each (SPREAD, cooperative launch; same sums, same order);
//   * snapshot-pure: sources are read from the packed buffer of the step's
//     start, new positions go to the other packed buffer.
#pragma once
#include "nbody_ptx.cuh"

namespace nbody {

constexpr int kStages = 4;           // shared-memory ring depth
constexpr int kLookahead = kStages - 2;  // tiles in flight ahead of the consumers
constexpr int kMaxPeers = 15;        // other GPUs a context can store to (one NVSwitch domain)
#ifndef NBODY_FOLD_UNITS
#define NBODY_FOLD_UNITS 8           // 0: single-level summation (the round-1 kernel; `make sweep` evidence builds)
#endif
constexpr int kFoldEvery = NBODY_FOLD_UNITS;   // source units per first-level sum of force_kernel (see FOLD there)
constexpr bool kFoldUnits = kFoldEvery != 0;
// bytes of the per-thread totals behind the ring (+ barriers, flag, j-split scratch) of a force_kernel CTA
__host__ __device__ constexpr size_t fold_bytes(int T, int NT, bool packed, bool two_level = true) { return packed && kFoldUnits && two_level ? (size_t)3 * T * NT * sizeof(float) : 0; }

struct ForceArgs {
    // owned targets, AoS float4, indexed 0..n_targets (caller-visible layout)
    float4* X;            // (x,y,z,m)   read, then advanced in place by the owner thread
    float4* V;            // (vx,vy,vz,w) in place, .w untouched
    float4* A;            // (ax,ay,az,0) out
    int n_targets;
    // sources: packed pair layout, addressed in units of UJ bodies
    const float4* src;    // current packed buffer
    int ra_begin, ra_count;   // unit range A  [ra_begin, ra_begin+ra_count)
    int rb_begin, rb_count;   // unit range B, visited after A
    // new positions of the owned targets in packed layout (may be null)
    float4* next_src;
    int next_off;         // packed body index of target 0 (even)
    // fused exchange (multi-GPU): the same 16-byte stores also go to every peer's NEXT buffer
    // (peer memory mapped over NVLink), replacing the all-gather of the new positions
    float4* peer_next[kMaxPeers];
    int n_peer;
    // in-kernel cross-GPU barrier of the fused exchange (optional; bar_flags == nullptr: the caller
    // synchronises the ranks between steps).  flags[r] in THIS GPU's memory = number of fused steps rank r
    // has completed (written by rank r's last CTA with st.release.sys over NVLink).  Step k may read the
    // peers' positions and overwrite the buffer they read in step k-1 once every flags[r] >= k.
    unsigned long long* bar_flags;              // local, [world]
    unsigned long long* bar_peer_flags[kMaxPeers];   // the peers' flag arrays (peer memory)
    unsigned long long bar_step;                // k: fused steps this rank has completed before this launch
    unsigned* bar_done;                         // local counter of finished CTAs, zero between launches
    int* bar_err;                               // set to 1 when a wait timed out (results are then invalid)
    int rank, world;
    // two-phase accumulation (sharded runs)
    const float4* accum_in;   // added first when non-null
    float4* accum_out;        // when non-null: store sums here, do not integrate
    // split-tile fix-up scratch
    float4* partial;      // [num_itiles][maxseg][TI]
    unsigned* counters;   // [num_itiles], zero between launches
    int maxseg;
    // parameters (TestProject/constants.h:25-26 at run time)
    float eps2, dt;
    int integ_f64;        // 0: float integrator (kernel.cu:116-124 / validation.cpp:43-49), 1: double (headless
                          // flavour), 2: mixed (reduction flavours, kernel.cu:455-463: double half-kick, float drift)
    int va_f3;            // 1: V and A are packed float3 arrays (FLOAT_3 build of the reference, main.cpp:233-269)
    int last_unit_bodies; // real bodies in the last unit of range A (0 = all of it): the zero-mass padding behind them
                          // need not be visited (rowsplit_kernel, whose units are 1 024 bodies; others ignore it)
    double dt64;
    // optional: CTA 0 records {clock64, globaltimer} at entry and exit -> SM clock actually run at;
    // clk[4] += ns its thread 0 spent in the cross-GPU flag wait, clk[5] = the longest single wait, clk[6] = waits
    unsigned long long* clk;
};

// ---- integrator + stores for one target (TestProject/kernel.cu:116-129) -------------
// Separately rounded multiply and add, exactly like validation.cpp:43-49 as
// built for x86 (no FMA contraction), so that equal accelerations give
// bit-equal positions and velocities.
__device__ __forceinline__ void integrate_one(const ForceArgs& a, int gi, bool valid, float fx, float fy, float fz, int lane)
{
    float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
    if (valid) {
        if (a.accum_in) {
            const float4 p = a.accum_in[gi];
            fx = __fadd_rn(p.x, fx); fy = __fadd_rn(p.y, fy); fz = __fadd_rn(p.z, fz);
        }
        if (a.accum_out) {
            a.accum_out[gi] = make_float4(fx, fy, fz, 0.f);
            return;
        }
        x = a.X[gi];
        float4 v;
        if (a.va_f3) {
            const float* v3 = reinterpret_cast<const float*>(a.V) + 3 * (size_t)gi;
            v = make_float4(v3[0], v3[1], v3[2], 0.f);
        } else {
            v = a.V[gi];
        }
        if (a.integ_f64 == 0) {
            const float hdt = 0.5f * a.dt;
            v.x = __fadd_rn(v.x, __fmul_rn(hdt, fx));
            v.y = __fadd_rn(v.y, __fmul_rn(hdt, fy));
            v.z = __fadd_rn(v.z, __fmul_rn(hdt, fz));
            x.x = __fadd_rn(x.x, __fmul_rn(a.dt, v.x));
            x.y = __fadd_rn(x.y, __fmul_rn(a.dt, v.y));
            x.z = __fadd_rn(x.z, __fmul_rn(a.dt, v.z));
        } else if (a.integ_f64 == 1) {
            // Sim-Without-OpenGL-Integration/kernel.cu:71-77: 0.5 * DT * a in double
            const double hdt = 0.5 * a.dt64;
            v.x = (float)((double)v.x + hdt * (double)fx);
            v.y = (float)((double)v.y + hdt * (double)fy);
            v.z = (float)((double)v.z + hdt * (double)fz);
            x.x = (float)((double)x.x + a.dt64 * (double)v.x);
            x.y = (float)((double)x.y + a.dt64 * (double)v.y);
            x.z = (float)((double)x.z + a.dt64 * (double)v.z);
        } else {
            // TestProject/kernel.cu:455-463 (inter_block_reduction*): `0.5 * DT * a` promotes the
            // half-kick to double (DT is a float literal there); the drift `DT * v` stays float and is
            // an FMA in the reference's GPU build (nvcc -fmad default).
            const double hdt = 0.5 * (double)a.dt;
            v.x = (float)((double)v.x + hdt * (double)fx);
            v.y = (float)((double)v.y + hdt * (double)fy);
            v.z = (float)((double)v.z + hdt * (double)fz);
            x.x = fmaf(a.dt, v.x, x.x);
            x.y = fmaf(a.dt, v.y, x.y);
            x.z = fmaf(a.dt, v.z, x.z);
        }
        a.X[gi] = x;
        if (a.va_f3) {
            float* v3 = reinterpret_cast<float*>(a.V) + 3 * (size_t)gi;
            float* a3 = reinterpret_cast<float*>(a.A) + 3 * (size_t)gi;
            v3[0] = v.x; v3[1] = v.y; v3[2] = v.z;
            a3[0] = fx; a3[1] = fy; a3[2] = fz;
        } else {
            a.V[gi] = v;
            a.A[gi] = make_float4(fx, fy, fz, 0.f);
        }
    } else if (a.accum_out) {
        return;
    }
    if (a.next_src) {
        // even lane holds body 2p, odd lane body 2p+1 (tile bases are even):
        // exchange halves so each lane writes one 16-byte half of the pair.
        const bool odd = lane & 1;
        const float s0 = odd ? x.x : x.z, s1 = odd ? x.y : x.w;
        const float r0 = __shfl_xor_sync(0xffffffffu, s0, 1);
        const float r1 = __shfl_xor_sync(0xffffffffu, s1, 1);
        const float4 o = odd ? make_float4(r0, x.z, r1, x.w) : make_float4(x.x, r0, x.y, r1);
        const bool pair_valid = __shfl_xor_sync(0xffffffffu, (int)valid, 1) || valid;
        if (pair_valid) {
            const size_t at = (size_t)(a.next_off + gi - (int)odd) + (odd ? 1 : 0);
            a.next_src[at] = o;
            for (int p = 0; p < a.n_peer; ++p) a.peer_next[p][at] = o;   // P2P stores (NVLink)
        }
    }
}

// worker that owns linear unit u when U units are cut into W equal ranges
__device__ __forceinline__ int unit_owner(u64 u, u64 U, u64 W) { return (int)(((u + 1) * W - 1) / U); }

// MODE: 0 scalar FFMA | 1 packed FADD2/FMUL2/FFMA2 | 5, 6, 7, 8, 18 packed, the same arithmetic in another source
//       order (5 targets reversed, 6 target-outer, 7 z first, 8 both 5 and 7, 18 x,z,y): every target's sums are formed by
//       the same operations in the same order, so the results are bit-identical; only ptxas' register
//       allocation differs (worth a few per cent, see tuned_large_variant in nbody_engine.cu).
// CS / CMODE: thread-block clusters (launch attribute clusterDim = CS).
//   CMODE 1 "multicast": the CS CTAs of a cluster own CS neighbouring target tiles and stream
//           the SAME source units; CTA r fetches slice r of each unit and TMA-multicasts it
//           into every CTA's ring, so one L2/HBM read feeds CS SMs.  The stream-K cut is made
//           between clusters.
//   CMODE 2 "j-split": the CS CTAs share ONE target tile and take every CS-th source unit of
//           the segment; the partial sums are combined through distributed shared memory
//           (ld.shared::cluster, fixed rank order), CTA r finishing slice r of the tile.
// FOLDK: source units per first-level sum (0 = single-level summation, nbody_params.summation = NBODY_SUM_SINGLE)
// SPREAD 1: the fix-up of a split tile is spread over the CTAs that worked on it (each sums and integrates 1 / nseg of
//        the tile's targets once all partial sums are there) instead of falling to the last one to arrive.  The CTAs
//        wait for each other, so the launch must be cooperative (whole grid resident).  Same sums in the same order.
//        A separate instantiation (mid-size N only): the large-N kernels keep their code and register allocation.
template <int T, int NT, int UJ, int UNROLL, int MODE, int MINB, int CS = 1, int CMODE = 0, int FOLDK = kFoldEvery, int SPREAD = 0>
__global__ void __launch_bounds__(NT, MINB) force_kernel(const ForceArgs a)
{
    constexpr bool PACKED = MODE != 0;
    // Two-level summation: the packed accumulators cover kFoldEvery source units (4 096 bodies with 512-body units)
    // and are then folded into per-thread scalar totals.  A target with one dominant neighbour otherwise adds ~N small
    // terms to a large accumulator one by one, each rounded at the big term's ulp: measured at N = 1 M, uniform cube,
    // that cost the worst bodies 1.7e-3 of their acceleration (the reference's own sequential fp32 sum: 3.3e-3) and
    // 2e-4 on 64 sampled targets; folded, the big accumulator sees N / 4 096 additions and the same samples sit 1.3e-6
    // from the fp64 sum.  The totals live in shared memory, one private column per thread ([3 T][NT] floats behind
    // the ring): in registers they are 18 more live values across the inner loop.  Cost (profiles/r2_fold_sweep.txt):
    // the fold itself is < 0.1 % of a unit's work, but ptxas schedules the inner loop around it differently --
    // 2 738 G/s without, 2 645 .. 2 710 with, depending only on the source order of the three accumulations
    // (x,z,y and z,y,x orders: 2 700 - 2 710; x,y,z: 2 645), hence the orders the dispatch table keeps.
    constexpr bool FOLD = PACKED && FOLDK != 0;
    constexpr bool MCAST = CS > 1 && CMODE == 1;
    constexpr bool JSPLIT = CS > 1 && CMODE == 2;
    constexpr int TI = T * NT;
    constexpr int TS = MCAST ? TI * CS : TI;   // targets of one worker's tile
    constexpr uint32_t kStageBytes = UJ * 16;
    static_assert((kStages & (kStages - 1)) == 0, "ring depth must be a power of two");
    static_assert((FOLDK & (FOLDK - 1)) == 0, "fold group must be a power of two (or 0 = off)");
    static_assert((UJ / 2) % UNROLL == 0, "unroll must divide the pairs per unit");
    static_assert(CS == 1 || CMODE == 1 || CMODE == 2, "cluster variants need a mode");
    static_assert(!SPREAD || CS == 1, "the spread fix-up is for the non-cluster kernels");
    static_assert(!MCAST || (kStageBytes / CS) % 16 == 0, "multicast slices must be 16-byte multiples");
    static_assert(!JSPLIT || (TI / CS) % 32 == 0, "j-split slices must be whole warps");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const uint32_t smem_base = smem_u32(smem_raw);
    const uint32_t bar_full = smem_base + kStages * kStageBytes;
    const uint32_t bar_empty = bar_full + 8 * kStages;
    volatile int* s_flag = reinterpret_cast<volatile int*>(smem_raw + kStages * kStageBytes + 16 * kStages);
    float4* s_red = reinterpret_cast<float4*>(smem_raw + kStages * kStageBytes + 16 * kStages + 16);   // JSPLIT only: [TI]
    float* s_tot = reinterpret_cast<float*>(smem_raw + kStages * kStageBytes + 16 * kStages + 16 + (JSPLIT ? TI * 16 : 0)) + threadIdx.x;   // FOLD only: [3 T][NT]

    const int tid = threadIdx.x, lane = tid & 31;
    if (a.clk && blockIdx.x == 0 && tid == 0) { a.clk[0] = clock64(); a.clk[1] = globaltimer_ns(); }
    const uint32_t crank = CS > 1 ? cluster_ctarank() : 0u;
    const int J = a.ra_count + a.rb_count;
    const int num_itiles = (a.n_targets + TS - 1) / TS;
    const u64 U = (u64)num_itiles * (u64)J;
    const u64 W = gridDim.x / CS, w = blockIdx.x / CS;   // workers = clusters
    const u64 u_begin = w * U / W, u_end = (w + 1) * U / W;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) {
            mbar_init(bar_full + 8 * s, 1);
            mbar_init(bar_empty + 8 * s, MCAST ? CS * (NT / 32) : NT / 32);
        }
        mbar_fence_init();
    }
    if (CS > 1) cluster_sync_all();   // peers' barriers exist before any multicast / remote arrive
    else __syncthreads();

    // ---- producer state (thread 0 only) ------------------------------------------------
    // JSPLIT: within each segment (part of one tile's unit row inside [u_begin, u_end)) this
    // CTA takes the units at offset crank, crank+CS, ... from the segment start.
    // Producer bookkeeping is 32-bit and division-free outside JSPLIT: it runs in warp 0 once per
    // unit, and a 64-bit modulo there cost 3 % of the whole kernel (r1: 2 730 -> 2 630 G/s).
    u64 pu = u_begin;                   // next linear unit to consider (JSPLIT walks it unit by unit)
    int pj = (int)(u_begin % (u64)J);   // its source unit, index into A|B
    int pleft = (int)(u_end - u_begin); // units still to fetch (non-JSPLIT)
    int kf = 0;                         // fetches issued so far
    auto issue = [&]() -> bool {
        if (JSPLIT) {
            while (pu < u_end) {
                u64 seg0 = (pu / (u64)J) * (u64)J;
                if (seg0 < u_begin) seg0 = u_begin;
                if ((pu - seg0) % CS == crank) break;
                ++pu;
            }
            if (pu >= u_end) return false;
            pj = (int)(pu % (u64)J);
            ++pu;
        } else {
            if (pleft <= 0) return false;
            --pleft;
        }
        const int s = kf & (kStages - 1), n = kf / kStages;
        if (n > 0) mbar_wait(bar_empty + 8 * s, (n - 1) & 1);
        const int ju = pj < a.ra_count ? a.ra_begin + pj : a.rb_begin + (pj - a.ra_count);
        mbar_arrive_expect_tx(bar_full + 8 * s, kStageBytes);
        if (MCAST) {
            constexpr uint32_t kSlice = kStageBytes / CS;
            tma_bulk_g2s_mcast(smem_base + s * kStageBytes + crank * kSlice, reinterpret_cast<const unsigned char*>(a.src + (size_t)ju * UJ) + crank * kSlice,
                               kSlice, bar_full + 8 * s, (uint16_t)((1u << CS) - 1u));
        } else {
            tma_bulk_g2s(smem_base + s * kStageBytes, a.src + (size_t)ju * UJ, kStageBytes, bar_full + 8 * s);
        }
        if (!JSPLIT && ++pj == J) pj = 0;
        ++kf;
        return true;
    };
    if (tid == 0) {
        if (a.bar_flags) {
            // every source fetch below and every peer store of the epilogue is ordered after this wait (thread 0's
            // TMA issues -> full barriers -> all consumers).  Polling is local; 20 s without progress = a dead peer.
            const unsigned long long t0 = globaltimer_ns();
            for (int r = 0; r < a.world; ++r) {
                if (r == a.rank) continue;
                while (ld_acquire_sys_u64(a.bar_flags + r) < a.bar_step) {
                    if (globaltimer_ns() - t0 > 20000000000ull) { *a.bar_err = 1; break; }
                    __nanosleep(200);
                }
            }
            __threadfence_system();
            if (a.clk && blockIdx.x == 0) {   // how long this GPU stood waiting for its peers (all CTAs start together)
                const unsigned long long waited = globaltimer_ns() - t0;
                a.clk[4] += waited;
                if (waited > a.clk[5]) a.clk[5] = waited;
                a.clk[6] += 1ull;
            }
        }
        while (kf < kLookahead && issue()) {}
    }

    const u64 eps2p = pack2(a.eps2, a.eps2);
    int k = 0;   // units consumed so far
    // SPREAD: the split segments of this CTA (a contiguous unit range has at most two: the tail of its first tile
    // and the head of its last), to be finished after the loop: {tile or -1, segments of the tile << 16 | this CTA's
    // index} twice, in the 16 bytes of s_flag (unused in this mode) so that nothing more stays live across the loop
    volatile int* s_pend = s_flag;
    if (SPREAD && tid == 0) { s_pend[0] = -1; s_pend[2] = -1; }
    u64 u = u_begin;
    while (u < u_end) {
        const int it = (int)(u / (u64)J);
        const int j0 = (int)(u - (u64)it * (u64)J);
        const u64 left = u_end - u;
        const int jn = (u64)(J - j0) < left ? (J - j0) : (int)left;

        // ---- targets into registers (negated: dx = xj + (-xi) is one FADD2) ----------
        u64 nx[T], ny[T], nz[T], ax[T], ay[T], az[T];
        float sx[T], sy[T], sz[T], qx[T], qy[T], qz[T];   // scalar variant

#pragma unroll
        for (int t = 0; t < T; ++t) {
            const int gi = (MCAST ? it * CS + (int)crank : it) * TI + tid + t * NT;
            float4 xi = make_float4(0.f, 0.f, 0.f, 0.f);
            if (gi < a.n_targets) xi = a.X[gi];
            if (PACKED) {
                nx[t] = pack2(-xi.x, -xi.x); ny[t] = pack2(-xi.y, -xi.y); nz[t] = pack2(-xi.z, -xi.z);
                ax[t] = 0ull; ay[t] = 0ull; az[t] = 0ull;
            } else {
                sx[t] = xi.x; sy[t] = xi.y; sz[t] = xi.z;
                qx[t] = 0.f; qy[t] = 0.f; qz[t] = 0.f;
            }
            if (FOLD) { s_tot[(3 * t) * NT] = 0.f; s_tot[(3 * t + 1) * NT] = 0.f; s_tot[(3 * t + 2) * NT] = 0.f; }
        }

        // ---- stream the source units of this segment ------------------------------------
        // FOLD: groups of kFoldEvery units, each summed from zero in the packed accumulators and then folded into the
        // totals -- a loop nest (group { zero; unit { pairs } ; fold }) rather than a conditional inside the unit
        // loop, so that the unit loop keeps the shape ptxas schedules best (the same as without folding).
        constexpr int JSTEP = JSPLIT ? CS : 1;
        int jj = JSPLIT ? (int)crank : 0;
#pragma unroll 1
        while (jj < jn) {
        const int jg_end = FOLD ? (jj + FOLDK * JSTEP < jn ? jj + FOLDK * JSTEP : jn) : jn;
        if (FOLD) {
#pragma unroll
            for (int t = 0; t < T; ++t) { ax[t] = 0ull; ay[t] = 0ull; az[t] = 0ull; }
        }
#pragma unroll 1
        for (; jj < jg_end; jj += JSTEP, ++k) {
            if (tid == 0) issue();
            const int s = k & (kStages - 1);
            mbar_wait(bar_full + 8 * s, (k / kStages) & 1);
            const ulonglong2* sp = reinterpret_cast<const ulonglong2*>(smem_raw + s * kStageBytes);
            // two interactions (one source pair) against every target held by this thread
            auto interact2 = [&](const ulonglong2 xy, const ulonglong2 zm) {
#pragma unroll
                for (int tt = 0; tt < T; ++tt) {
                    // source order only (same operations, same order inside every sum): REV targets back to front;
                    // order of the three differences (DORD) and of the three accumulations (AORD): 0 xyz, 1 zyx, 5 xzy
                    constexpr bool REV = MODE == 5 || MODE == 8;
                    constexpr int DORD = (MODE == 7 || MODE == 8 || MODE == 19) ? 1 : MODE == 18 ? 5 : 0;
                    constexpr int AORD = DORD;
                    const int t = REV ? T - 1 - tt : tt;
                    u64 dx, dy, dz;
                    if (DORD == 1) { dz = add2(zm.x, nz[t]); dy = add2(xy.y, ny[t]); dx = add2(xy.x, nx[t]); }
                    else if (DORD == 5) { dx = add2(xy.x, nx[t]); dz = add2(zm.x, nz[t]); dy = add2(xy.y, ny[t]); }
                    else { dx = add2(xy.x, nx[t]); dy = add2(xy.y, ny[t]); dz = add2(zm.x, nz[t]); }
                    u64 r2 = fma2(dx, dx, eps2p);
                    r2 = fma2(dy, dy, r2);
                    r2 = fma2(dz, dz, r2);
                    const u64 d3 = mul2(mul2(r2, r2), r2);      // (r^2+eps2)^3, as kernel.cu:20
                    float lo, hi;
                    unpack2(d3, lo, hi);
                    u64 sc;
                    if (MODE == 19) {
                        // experiment (DESIGN section 8, round 1 "untried"): the two m * r^-3 products as SCALAR FMULs, one
                        // behind each MUFU.RSQ, so that a 1-cycle FP32 op rather than a 2-cycle packed one sits next to
                        // the MUFU issue (which costs the FMA pipe 1.17 issue cycles, profiles/r1_microbench_pipes.jsonl)
                        float mlo, mhi;
                        unpack2(zm.y, mlo, mhi);
                        sc = pack2(__fmul_rn(rsqrt_mufu(lo), mlo), __fmul_rn(rsqrt_mufu(hi), mhi));
                    } else {
                        sc = mul2(pack2(rsqrt_mufu(lo), rsqrt_mufu(hi)), zm.y);
                    }
                    if (AORD == 1) { az[t] = fma2(dz, sc, az[t]); ay[t] = fma2(dy, sc, ay[t]); ax[t] = fma2(dx, sc, ax[t]); }
                    else if (AORD == 5) { ax[t] = fma2(dx, sc, ax[t]); az[t] = fma2(dz, sc, az[t]); ay[t] = fma2(dy, sc, ay[t]); }
                    else { ax[t] = fma2(dx, sc, ax[t]); ay[t] = fma2(dy, sc, ay[t]); az[t] = fma2(dz, sc, az[t]); }
                }
            };
            if (MODE == 6) {
                // MODE 6: same arithmetic, target-outer / pair-inner order inside a group of UNROLL pairs
#pragma unroll 1
                for (int p = 0; p < UJ / 2; p += UNROLL) {
                    ulonglong2 gxy[UNROLL], gzm[UNROLL];
#pragma unroll
                    for (int q = 0; q < UNROLL; ++q) { gxy[q] = sp[2 * (p + q)]; gzm[q] = sp[2 * (p + q) + 1]; }
#pragma unroll
                    for (int t = 0; t < T; ++t) {
#pragma unroll
                        for (int q = 0; q < UNROLL; ++q) {
                            const u64 dx = add2(gxy[q].x, nx[t]), dy = add2(gxy[q].y, ny[t]), dz = add2(gzm[q].x, nz[t]);
                            u64 r2 = fma2(dx, dx, eps2p);
                            r2 = fma2(dy, dy, r2);
                            r2 = fma2(dz, dz, r2);
                            const u64 d3 = mul2(mul2(r2, r2), r2);
                            float lo, hi;
                            unpack2(d3, lo, hi);
                            const u64 sc = mul2(pack2(rsqrt_mufu(lo), rsqrt_mufu(hi)), gzm[q].y);
                            ax[t] = fma2(dx, sc, ax[t]);
                            ay[t] = fma2(dy, sc, ay[t]);
                            az[t] = fma2(dz, sc, az[t]);
                        }
                    }
                }
            } else {
#pragma unroll 1
            for (int p = 0; p < UJ / 2; p += UNROLL) {
#pragma unroll
                for (int q = 0; q < UNROLL; ++q) {
                    if (PACKED) {
                        interact2(sp[2 * (p + q)], sp[2 * (p + q) + 1]);   // (x0,x1)(y0,y1) | (z0,z1)(m0,m1)
                    } else {
                        const float4* sf = reinterpret_cast<const float4*>(sp);
                        const float4 xy = sf[2 * (p + q)];
                        const float4 zm = sf[2 * (p + q) + 1];
#pragma unroll
                        for (int t = 0; t < T; ++t) {
                            {
                                const float dx = xy.x - sx[t], dy = xy.z - sy[t], dz = zm.x - sz[t];
                                const float r2 = fmaf(dz, dz, fmaf(dy, dy, fmaf(dx, dx, a.eps2)));
                                const float sc = zm.z * rsqrt_mufu(r2 * r2 * r2);
                                qx[t] = fmaf(dx, sc, qx[t]); qy[t] = fmaf(dy, sc, qy[t]); qz[t] = fmaf(dz, sc, qz[t]);
                            }
                            {
                                const float dx = xy.y - sx[t], dy = xy.w - sy[t], dz = zm.y - sz[t];
                                const float r2 = fmaf(dz, dz, fmaf(dy, dy, fmaf(dx, dx, a.eps2)));
                                const float sc = zm.w * rsqrt_mufu(r2 * r2 * r2);
                                qx[t] = fmaf(dx, sc, qx[t]); qy[t] = fmaf(dy, sc, qy[t]); qz[t] = fmaf(dz, sc, qz[t]);
                            }
                        }
                    }
                }
            }
            }
            __syncwarp();
            if (MCAST) {   // every producer of the cluster writes this stage: release it in all CTAs
                if (lane < CS) mbar_arrive_cluster(mapa_shared(bar_empty + 8 * s, (uint32_t)lane));
            } else if (lane == 0) {
                mbar_arrive(bar_empty + 8 * s);
            }
        }
        if (FOLD) {   // this group's sums into the totals
#pragma unroll
            for (int t = 0; t < T; ++t) {
                float lo, hi;
                unpack2(ax[t], lo, hi); s_tot[(3 * t) * NT] = __fadd_rn(s_tot[(3 * t) * NT], __fadd_rn(lo, hi));
                unpack2(ay[t], lo, hi); s_tot[(3 * t + 1) * NT] = __fadd_rn(s_tot[(3 * t + 1) * NT], __fadd_rn(lo, hi));
                unpack2(az[t], lo, hi); s_tot[(3 * t + 2) * NT] = __fadd_rn(s_tot[(3 * t + 2) * NT], __fadd_rn(lo, hi));
            }
        }
        }

        // ---- segment done: fused epilogue, or split-tile fix-up ------------------------
        float fx[T], fy[T], fz[T];
#pragma unroll
        for (int t = 0; t < T; ++t) {
            if (FOLD) {
                fx[t] = s_tot[(3 * t) * NT]; fy[t] = s_tot[(3 * t + 1) * NT]; fz[t] = s_tot[(3 * t + 2) * NT];
            } else if (PACKED) {
                float lo, hi;
                unpack2(ax[t], lo, hi); fx[t] = lo + hi;
                unpack2(ay[t], lo, hi); fy[t] = lo + hi;
                unpack2(az[t], lo, hi); fz[t] = lo + hi;
            } else {
                fx[t] = qx[t]; fy[t] = qy[t]; fz[t] = qz[t];
            }
        }
        const bool split = !(j0 == 0 && jn == J);
        int w_first = 0, nseg = 1;
        if (split) {
            const u64 first_u = (u64)it * (u64)J;
            w_first = unit_owner(first_u, U, W);
            nseg = unit_owner(first_u + (u64)J - 1, U, W) - w_first + 1;
        }
        if (!JSPLIT) {
            const int vt = MCAST ? it * CS + (int)crank : it;   // tile whose T*NT targets this CTA holds
            bool finish = true;
            if (split) {
                float4* tile_part = a.partial + (size_t)vt * a.maxseg * TI;
                float4* slot = tile_part + (size_t)((int)w - w_first) * TI;
#pragma unroll
                for (int t = 0; t < T; ++t) slot[tid + t * NT] = make_float4(fx[t], fy[t], fz[t], 0.f);
                __threadfence();
                __syncthreads();
                if (SPREAD) {   // counted now, finished by all of the tile's CTAs after the loop
                    if (tid == 0) {
                        atomicAdd(&a.counters[vt], 1u);
                        const int at = s_pend[0] < 0 ? 0 : 2;
                        s_pend[at] = vt; s_pend[at + 1] = (nseg << 16) | ((int)w - w_first);
                    }
                    u += (u64)jn;
                    continue;
                }
                if (tid == 0) {
                    const unsigned old = atomicAdd(&a.counters[vt], 1u);
                    *s_flag = (old == (unsigned)(nseg - 1));
                }
                __syncthreads();
                finish = (*s_flag != 0);
                if (finish) {
                    __threadfence();
#pragma unroll
                    for (int t = 0; t < T; ++t) { fx[t] = 0.f; fy[t] = 0.f; fz[t] = 0.f; }
                    for (int sg = 0; sg < nseg; ++sg) {
#pragma unroll
                        for (int t = 0; t < T; ++t) {
                            const float4 p = __ldcg(tile_part + (size_t)sg * TI + tid + t * NT);
                            fx[t] = __fadd_rn(fx[t], p.x); fy[t] = __fadd_rn(fy[t], p.y); fz[t] = __fadd_rn(fz[t], p.z);
                        }
                    }
                    if (tid == 0) a.counters[vt] = 0u;
                }
            }
            if (finish) {
#pragma unroll
                for (int t = 0; t < T; ++t) {
                    const int gi = vt * TI + tid + t * NT;
                    integrate_one(a, gi, gi < a.n_targets, fx[t], fy[t], fz[t], lane);
                }
            }
        } else {
            // ---- j-split: combine the CS partial sums through distributed shared memory ----
            constexpr int SL = TI / CS;                 // targets finished by each CTA
            constexpr int RT = (SL + NT - 1) / NT;
#pragma unroll
            for (int t = 0; t < T; ++t) s_red[tid + t * NT] = make_float4(fx[t], fy[t], fz[t], 0.f);
            cluster_sync_all();
            float rx[RT], ry[RT], rz[RT];
            const uint32_t red_mine = smem_u32(s_red) + (uint32_t)(crank * SL) * 16u;
#pragma unroll
            for (int i = 0; i < RT; ++i) {
                rx[i] = 0.f; ry[i] = 0.f; rz[i] = 0.f;
                const int kk = tid + i * NT;
                if (kk < SL) {
#pragma unroll
                    for (int q = 0; q < CS; ++q) {   // fixed rank order: deterministic
                        const float4 p = ld_dsmem_f4(mapa_shared(red_mine + (uint32_t)kk * 16u, (uint32_t)q));
                        rx[i] = __fadd_rn(rx[i], p.x); ry[i] = __fadd_rn(ry[i], p.y); rz[i] = __fadd_rn(rz[i], p.z);
                    }
                }
            }
            cluster_sync_all();   // every peer has read this CTA's sums: s_red may be reused
            bool finish = true;
            if (split) {
                float4* tile_part = a.partial + (size_t)it * a.maxseg * TI + (size_t)crank * SL;
                float4* slot = tile_part + (size_t)((int)w - w_first) * TI;
#pragma unroll
                for (int i = 0; i < RT; ++i)
                    if (tid + i * NT < SL) slot[tid + i * NT] = make_float4(rx[i], ry[i], rz[i], 0.f);
                __threadfence();
                __syncthreads();
                if (tid == 0) {
                    const unsigned old = atomicAdd(&a.counters[it * CS + (int)crank], 1u);
                    *s_flag = (old == (unsigned)(nseg - 1));
                }
                __syncthreads();
                finish = (*s_flag != 0);
                if (finish) {
                    __threadfence();
#pragma unroll
                    for (int i = 0; i < RT; ++i) { rx[i] = 0.f; ry[i] = 0.f; rz[i] = 0.f; }
                    for (int sg = 0; sg < nseg; ++sg) {
#pragma unroll
                        for (int i = 0; i < RT; ++i) {
                            if (tid + i * NT < SL) {
                                const float4 p = __ldcg(tile_part + (size_t)sg * TI + tid + i * NT);
                                rx[i] = __fadd_rn(rx[i], p.x); ry[i] = __fadd_rn(ry[i], p.y); rz[i] = __fadd_rn(rz[i], p.z);
                            }
                        }
                    }
                    if (tid == 0) a.counters[it * CS + (int)crank] = 0u;
                }
            }
            if (finish) {
#pragma unroll
                for (int i = 0; i < RT; ++i) {
                    if (tid + i * NT < SL) {   // warp-uniform: SL and NT are multiples of 32
                        const int gi = it * TI + (int)crank * SL + tid + i * NT;
                        integrate_one(a, gi, gi < a.n_targets, rx[i], ry[i], rz[i], lane);
                    }
                }
            }
        }
        u += (u64)jn;
    }
    if (SPREAD) {
        // ---- spread fix-up: CTA number ix of the nseg that worked on a split tile finishes slice ix of its targets
        // (whole warps), summing the nseg partial sums in segment order exactly as the last-arriver path does.  With
        // a tile cut 18 ways (N = 16 384) the last arriver alone reads 18 x TI x 16 B = 590 KB through one SM while
        // 140 SMs idle (13 us of a 126 us kernel, profiles/r2_ncu_full_16k_raw.csv: SM active min / avg / max 177k /
        // 205k / 231k cycles); spread, every SM reads 32 KB.
        __syncthreads();
        for (int q = 0; q < 2; ++q) {
            const int vt = s_pend[2 * q];
            if (vt < 0) break;
            const int nseg = s_pend[2 * q + 1] >> 16, ix = s_pend[2 * q + 1] & 0xffff;
            if (tid == 0) {
                const unsigned long long t0 = globaltimer_ns();
                while (ld_acquire_gpu_u32(a.counters + vt) < (unsigned)nseg) {
                    if (globaltimer_ns() - t0 > 4000000000ull) __trap();   // cannot happen on a cooperative launch
                }
            }
            __syncthreads();
            __threadfence();
            const int chunk = (((TI + nseg - 1) / nseg) + 31) & ~31;
            const int i_end = (ix + 1) * chunk < TI ? (ix + 1) * chunk : TI;
            const float4* tile_part = a.partial + (size_t)vt * a.maxseg * TI;
            for (int i = ix * chunk + tid; i < i_end; i += NT) {
                float rx = 0.f, ry = 0.f, rz = 0.f;
#pragma unroll 4
                for (int sg = 0; sg < nseg; ++sg) {
                    const float4 p = __ldcg(tile_part + (size_t)sg * TI + i);
                    rx = __fadd_rn(rx, p.x); ry = __fadd_rn(ry, p.y); rz = __fadd_rn(rz, p.z);
                }
                const int gi = vt * TI + i;
                integrate_one(a, gi, gi < a.n_targets, rx, ry, rz, lane);
            }
            __syncthreads();
            if (tid == 0) {   // the last CTA to finish its slice leaves the counter at zero for the next launch
                const unsigned old = atomicAdd(&a.counters[vt], 1u);
                if (old == 2u * (unsigned)nseg - 1u) a.counters[vt] = 0u;
            }
        }
    }
    if (a.clk && blockIdx.x == 0 && tid == 0) { a.clk[2] = clock64(); a.clk[3] = globaltimer_ns(); }
    if (a.bar_flags) {
        // the last CTA of this GPU to get here publishes "step k done" to every peer: all P2P stores of all CTAs
        // are fenced (system scope) before their CTA counts itself, the publisher fences again before the flags
        __threadfence_system();
        __syncthreads();
        if (tid == 0) {
            const unsigned old = atomicAdd(a.bar_done, 1u);
            if (old == gridDim.x - 1) {
                *a.bar_done = 0u;
                __threadfence_system();
                for (int p = 0; p < a.n_peer; ++p) st_release_sys_u64(a.bar_peer_flags[p] + a.rank, a.bar_step + 1);
            }
        }
    }
    if (CS > 1) cluster_sync_all();   // no CTA leaves while peers may still touch its shared memory
}

// ---- small-N kernel: j-split inside the CTA, warp-shuffle reduction ---------------------------------
// Below a few thousand bodies a step is a handful of microseconds and the stream-K machinery
// (TMA ring, partial sums through global memory, last-arriver fix-up) is mostly latency.  Here a
// CTA owns TT targets, every thread keeps ALL of them in registers and takes every NT-th source
// pair straight from the packed buffer (two coalesced LDG.128 per pair, L2-resident: the whole
// system is <= 128 KB), so the source range is split across the threads of the CTA; the NT
// partial sums per target are combined by shfl.xor inside each warp and through shared memory
// across the warps, in fixed order (deterministic), and warp 0 runs the same fused integrator.
// One launch, no scratch, no atomics.  Same ForceArgs as force_kernel (both source ranges, the
// two-phase accumulators and the peer stores of the fused exchange work unchanged); UJ = 64.
template <int TT, int NT, int UNROLL>
__global__ void __launch_bounds__(NT) small_kernel(const ForceArgs a)
{
    constexpr int UJ = 64, NW = NT / 32;
    static_assert(TT % 2 == 0 && TT <= 32 && NT % 32 == 0, "targets per CTA: even, at most one warp");
    __shared__ float s_part[NW][3 * TT];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int g0 = blockIdx.x * TT;

    u64 nx[TT], ny[TT], nz[TT], ax[TT], ay[TT], az[TT];
#pragma unroll
    for (int t = 0; t < TT; ++t) {
        float4 xi = make_float4(0.f, 0.f, 0.f, 0.f);
        if (g0 + t < a.n_targets) xi = __ldg(a.X + g0 + t);
        nx[t] = pack2(-xi.x, -xi.x); ny[t] = pack2(-xi.y, -xi.y); nz[t] = pack2(-xi.z, -xi.z);
        ax[t] = 0ull; ay[t] = 0ull; az[t] = 0ull;
    }
    const u64 eps2p = pack2(a.eps2, a.eps2);
    const int pairs_a = a.ra_count * (UJ / 2), pairs = pairs_a + a.rb_count * (UJ / 2);
    const ulonglong2* src = reinterpret_cast<const ulonglong2*>(a.src);
    auto pair_at = [&](int p) -> size_t {   // packed index (in float4 units / 2) of pair p of range A|B
        return p < pairs_a ? (size_t)a.ra_begin * (UJ / 2) + p : (size_t)a.rb_begin * (UJ / 2) + (p - pairs_a);
    };
#pragma unroll 1
    for (int p0 = tid; p0 < pairs; p0 += NT * UNROLL) {
        ulonglong2 xy[UNROLL], zm[UNROLL];
#pragma unroll
        for (int q = 0; q < UNROLL; ++q) {
            const int p = p0 + q * NT;
            if (p < pairs) {
                const size_t at = pair_at(p);
                xy[q] = __ldg(src + 2 * at); zm[q] = __ldg(src + 2 * at + 1);
            } else {   // zero-mass pair at the origin contributes exactly +0 (eps2 > 0)
                xy[q] = make_ulonglong2(0ull, 0ull); zm[q] = make_ulonglong2(0ull, 0ull);
            }
        }
#pragma unroll
        for (int q = 0; q < UNROLL; ++q) {
#pragma unroll
            for (int t = 0; t < TT; ++t) {
                const u64 dx = add2(xy[q].x, nx[t]), dy = add2(xy[q].y, ny[t]), dz = add2(zm[q].x, nz[t]);
                u64 r2 = fma2(dx, dx, eps2p);
                r2 = fma2(dy, dy, r2);
                r2 = fma2(dz, dz, r2);
                const u64 d3 = mul2(mul2(r2, r2), r2);
                float lo, hi;
                unpack2(d3, lo, hi);
                const u64 sc = mul2(pack2(rsqrt_mufu(lo), rsqrt_mufu(hi)), zm[q].y);
                ax[t] = fma2(dx, sc, ax[t]);
                ay[t] = fma2(dy, sc, ay[t]);
                az[t] = fma2(dz, sc, az[t]);
            }
        }
    }
    // ---- reduce the NT partial sums of every target: lanes (shfl.xor), then warps (smem) ----------
#pragma unroll
    for (int t = 0; t < TT; ++t) {
        float lo, hi, f[3];
        unpack2(ax[t], lo, hi); f[0] = lo + hi;
        unpack2(ay[t], lo, hi); f[1] = lo + hi;
        unpack2(az[t], lo, hi); f[2] = lo + hi;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
#pragma unroll
            for (int m = 16; m > 0; m >>= 1) f[c] = __fadd_rn(f[c], __shfl_xor_sync(0xffffffffu, f[c], m));
            if (lane == 0) s_part[wid][3 * t + c] = f[c];
        }
    }
    __syncthreads();
    if (wid == 0) {
        float fx = 0.f, fy = 0.f, fz = 0.f;
        if (lane < TT) {
#pragma unroll
            for (int w = 0; w < NW; ++w) {
                fx = __fadd_rn(fx, s_part[w][3 * lane]); fy = __fadd_rn(fy, s_part[w][3 * lane + 1]); fz = __fadd_rn(fz, s_part[w][3 * lane + 2]);
            }
        }
        const int gi = g0 + lane;
        integrate_one(a, gi, lane < TT && gi < a.n_targets, fx, fy, fz, lane);
    }
}

// ---- row-split kernel: mid-size N whose targets divide evenly over the SMs -----------------------------
// At N = 16 384 the stream-K cut leaves every target tile shared by ~18 CTAs: each writes its partial sums to
// global memory, waits for the others and reads 18 of them back (profiles/r2_ncu_full_16k_spread_raw.csv; about
// a tenth of a 112 us step is prologue + that round trip).  When ceil(N / 148) is close to a shape a CTA can hold
// -- T x H targets: H = 4, 8, 16 or 32 lanes of every warp hold T targets each -- nothing needs to cross CTAs at all:
// CTA b owns targets [b TI, (b + 1) TI) and streams ALL source units through the same TMA ring as force_kernel;
// the source pairs of a unit are split 32 / H x NT / 32 ways over the sub-groups of a warp and the warps (the
// north-star's j-split), every thread keeping its T targets in registers for the whole kernel, and the partial
// sums meet once at the end: shfl.xor across the sub-groups of a warp, then shared memory across the warps in
// fixed order (deterministic).  One launch, no scratch, no atomics, no cooperative launch.  Each accumulator sees
// N / (2 x split) additions (512 at N = 16 384), so no second summation level is needed below 64 k sources.
// Same ForceArgs as force_kernel (both source ranges, two-phase accumulators, peer stores); no in-kernel barrier.
// TRIM 1: the zero-mass padding behind the last real body of the last unit is skipped (ForceArgs.last_unit_bodies); a
// separate instantiation, launched only when N is not a multiple of the unit: with the variable trip count in the
// code ptxas allocates the common loop differently (N = 16 384: 104.8 us instead of 103.6).  Same sums either way.
template <int T, int H, int NT, int UJ, int UNROLL, int ORD = 0, int TRIM = 0, int LOOK = kLookahead>
__global__ void __launch_bounds__(NT, 1) rowsplit_kernel(const ForceArgs a)
{
    constexpr int SUB = 32 / H, NW = NT / 32, SPLIT = SUB * NW, TI = T * H;
    constexpr int PER = (UJ / 2) / SPLIT;                 // pairs of one unit per thread
    constexpr uint32_t kStageBytes = UJ * 16;
    static_assert(H == 4 || H == 8 || H == 16 || H == 32, "lanes per target group");
    static_assert(PER >= UNROLL && PER % UNROLL == 0, "the unit's pairs must divide over the split and the unroll");
    static_assert(TI % 2 == 0 && TI <= NT, "one finishing thread per target, pairs stay inside the tile");
    static_assert((size_t)NW * 3 * TI * sizeof(float) <= (size_t)kStages * kStageBytes, "the reduction area aliases the ring");
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const uint32_t smem_base = smem_u32(smem_raw);
    const uint32_t bar_full = smem_base + kStages * kStageBytes;
    const uint32_t bar_empty = bar_full + 8 * kStages;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int hl = lane % H, myq = wid * SUB + lane / H;   // target column, index in the source split
    if (a.clk && blockIdx.x == 0 && tid == 0) { a.clk[0] = clock64(); a.clk[1] = globaltimer_ns(); }
    const int J = a.ra_count + a.rb_count;
    // pairs per thread of the last unit: the zero-mass padding behind the last real body is skipped (whole unrolls)
    const int last_per = (TRIM && a.last_unit_bodies > 0 && a.rb_count == 0)
                             ? min(PER, (a.last_unit_bodies + 2 * SPLIT * UNROLL - 1) / (2 * SPLIT * UNROLL) * UNROLL) : PER;
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, NW); }
        mbar_fence_init();
    }
    __syncthreads();
    int pj = 0, kf = 0;   // producer (thread 0): next source unit, fetches issued
    auto issue = [&]() {
        if (pj >= J) return;
        const int s = kf & (kStages - 1), n = kf / kStages;
        if (n > 0) mbar_wait(bar_empty + 8 * s, (n - 1) & 1);
        const int ju = pj < a.ra_count ? a.ra_begin + pj : a.rb_begin + (pj - a.ra_count);
        mbar_arrive_expect_tx(bar_full + 8 * s, kStageBytes);
        tma_bulk_g2s(smem_base + s * kStageBytes, a.src + (size_t)ju * UJ, kStageBytes, bar_full + 8 * s);
        ++pj; ++kf;
    };
    if (tid == 0)
        for (int i = 0; i < LOOK; ++i) issue();

    const int g0 = blockIdx.x * TI;
    u64 nx[T], ny[T], nz[T], ax[T], ay[T], az[T];
#pragma unroll
    for (int t = 0; t < T; ++t) {
        const int gi = g0 + hl + t * H;
        float4 xi = make_float4(0.f, 0.f, 0.f, 0.f);
        if (gi < a.n_targets) xi = a.X[gi];
        nx[t] = pack2(-xi.x, -xi.x); ny[t] = pack2(-xi.y, -xi.y); nz[t] = pack2(-xi.z, -xi.z);
        ax[t] = 0ull; ay[t] = 0ull; az[t] = 0ull;
    }
    const u64 eps2p = pack2(a.eps2, a.eps2);
    // one source unit against this thread's targets; `per` = pairs per thread (a constant for all units but the last,
    // so the body is instantiated twice and the common case keeps its compile-time trip count)
    auto consume = [&](const int k, const int per) {
        if (tid == 0) issue();
        const int s = k & (kStages - 1);
        mbar_wait(bar_full + 8 * s, (k / kStages) & 1);
        const ulonglong2* sp = reinterpret_cast<const ulonglong2*>(smem_raw + s * kStageBytes) + 2 * myq;
#pragma unroll 1
        for (int i = 0; i < per; i += UNROLL) {
#pragma unroll
            for (int q = 0; q < UNROLL; ++q) {
                const ulonglong2 xy = sp[2 * SPLIT * (i + q)], zm = sp[2 * SPLIT * (i + q) + 1];   // (x0,x1)(y0,y1) | (z0,z1)(m0,m1)
#pragma unroll
                for (int tt = 0; tt < T; ++tt) {
                    const int t = ORD >= 8 ? T - 1 - tt : tt;   // source order only: the same sums either way
                    u64 dx, dy, dz;
                    if ((ORD & 7) == 1) { dz = add2(zm.x, nz[t]); dy = add2(xy.y, ny[t]); dx = add2(xy.x, nx[t]); }
                    else if ((ORD & 7) == 5) { dx = add2(xy.x, nx[t]); dz = add2(zm.x, nz[t]); dy = add2(xy.y, ny[t]); }
                    else { dx = add2(xy.x, nx[t]); dy = add2(xy.y, ny[t]); dz = add2(zm.x, nz[t]); }
                    u64 r2 = fma2(dx, dx, eps2p);
                    r2 = fma2(dy, dy, r2);
                    r2 = fma2(dz, dz, r2);
                    const u64 d3 = mul2(mul2(r2, r2), r2);      // (r^2+eps2)^3, as kernel.cu:20
                    float lo, hi;
                    unpack2(d3, lo, hi);
                    const u64 sc = mul2(pack2(rsqrt_mufu(lo), rsqrt_mufu(hi)), zm.y);
                    if ((ORD & 7) == 1) { az[t] = fma2(dz, sc, az[t]); ay[t] = fma2(dy, sc, ay[t]); ax[t] = fma2(dx, sc, ax[t]); }
                    else if ((ORD & 7) == 5) { ax[t] = fma2(dx, sc, ax[t]); az[t] = fma2(dz, sc, az[t]); ay[t] = fma2(dy, sc, ay[t]); }
                    else { ax[t] = fma2(dx, sc, ax[t]); ay[t] = fma2(dy, sc, ay[t]); az[t] = fma2(dz, sc, az[t]); }
                }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_empty + 8 * s);
    };
    if (TRIM) {
#pragma unroll 1
        for (int k = 0; k < J - 1; ++k) consume(k, PER);
        consume(J - 1, last_per);
    } else {
#pragma unroll 1
        for (int k = 0; k < J; ++k) consume(k, PER);
    }
    // ---- the SPLIT partial sums of every target: sub-groups of a warp (shfl.xor), then warps (shared memory) ----
    float f[T][3];
#pragma unroll
    for (int t = 0; t < T; ++t) {
        float lo, hi;
        unpack2(ax[t], lo, hi); f[t][0] = __fadd_rn(lo, hi);
        unpack2(ay[t], lo, hi); f[t][1] = __fadd_rn(lo, hi);
        unpack2(az[t], lo, hi); f[t][2] = __fadd_rn(lo, hi);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
#pragma unroll
            for (int m = H; m < 32; m <<= 1) f[t][c] = __fadd_rn(f[t][c], __shfl_xor_sync(0xffffffffu, f[t][c], m));
        }
    }
    __syncthreads();   // every warp has consumed every unit: the ring is free
    float* s_red = reinterpret_cast<float*>(smem_raw);   // [NW][3][TI]
    if (lane < H) {
#pragma unroll
        for (int t = 0; t < T; ++t) {
#pragma unroll
            for (int c = 0; c < 3; ++c) s_red[(wid * 3 + c) * TI + hl + t * H] = f[t][c];
        }
    }
    __syncthreads();
    if (tid < ((TI + 31) & ~31)) {   // whole warps: integrate_one exchanges pair halves by shuffle
        float fx = 0.f, fy = 0.f, fz = 0.f;
        if (tid < TI) {
#pragma unroll
            for (int w = 0; w < NW; ++w) {
                fx = __fadd_rn(fx, s_red[(w * 3 + 0) * TI + tid]);
                fy = __fadd_rn(fy, s_red[(w * 3 + 1) * TI + tid]);
                fz = __fadd_rn(fz, s_red[(w * 3 + 2) * TI + tid]);
            }
        }
        const int gi = g0 + tid;
        integrate_one(a, gi, tid < TI && gi < a.n_targets, fx, fy, fz, lane);
    }
    if (a.clk && blockIdx.x == 0 && tid == 0) { a.clk[2] = clock64(); a.clk[3] = globaltimer_ns(); }
}

// ---- AoS float4 -> packed pair layout (pre-pass of the in-place API and of upload) ----
// Covers bodies [0, n_cover) of the destination starting at packed body index
// `off` (even); bodies >= n are written as zero-mass padding.
__global__ void pack_kernel(const float4* __restrict__ X, int n, float4* __restrict__ packed, int off, int n_cover)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;   // pair index
    if (2 * p >= n_cover) return;
    const int b0 = 2 * p, b1 = 2 * p + 1;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    const float4 v0 = b0 < n ? X[b0] : z;
    const float4 v1 = b1 < n ? X[b1] : z;
    packed[(size_t)off + b0] = make_float4(v0.x, v1.x, v0.y, v1.y);
    packed[(size_t)off + b0 + 1] = make_float4(v0.z, v1.z, v0.w, v1.w);
}

// ---- energy of the current state, double precision --------------------------------------
// part[2*b] = kinetic, part[2*b+1] = sum over ordered pairs i!=j of -m_i m_j / sqrt(r^2+eps2)
// (the host halves it).  Not on the step's hot path; O(N^2) in FP64.
__global__ void __launch_bounds__(256) energy_kernel(const float4* __restrict__ X, const float4* __restrict__ V, int n, double eps2, double* __restrict__ part)
{
    __shared__ float4 tile[256];
    __shared__ double red[2][256];
    const int i = blockIdx.x * 256 + threadIdx.x;
    const bool valid = i < n;
    const float4 xi = valid ? X[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    double pe = 0.0;
    for (int j0 = 0; j0 < n; j0 += 256) {
        const int j = j0 + threadIdx.x;
        tile[threadIdx.x] = j < n ? X[j] : make_float4(0.f, 0.f, 0.f, 0.f);
        __syncthreads();
        const int lim = min(256, n - j0);
        for (int jj = 0; jj < lim; ++jj) {
            const float4 b = tile[jj];
            const double rx = (double)b.x - (double)xi.x, ry = (double)b.y - (double)xi.y, rz = (double)b.z - (double)xi.z;
            const double d = rx * rx + ry * ry + rz * rz + eps2;
            if (j0 + jj != i) pe -= (double)b.w * rsqrt(d);
        }
        __syncthreads();
    }
    double ke = 0.0;
    if (valid) {
        const float4 v = V[i];
        ke = 0.5 * (double)xi.w * ((double)v.x * v.x + (double)v.y * v.y + (double)v.z * v.z);
        pe *= (double)xi.w;
    } else {
        pe = 0.0;
    }
    red[0][threadIdx.x] = ke; red[1][threadIdx.x] = pe;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (threadIdx.x < s) { red[0][threadIdx.x] += red[0][threadIdx.x + s]; red[1][threadIdx.x] += red[1][threadIdx.x + s]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { part[2 * blockIdx.x] = red[0][0]; part[2 * blockIdx.x + 1] = red[1][0]; }
}

}  // namespace nbody
