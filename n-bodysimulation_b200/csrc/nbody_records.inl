// nbody_records.inl -- fourth part of nbody_engine.cu: run records.
//
// The reference never emits its results: main.cpp:356-366 frees the device
// arrays without copying them back, and the "benchmark suite" its README
// names (README.md:15) is not in the repository.  These entry points give a
// run something to leave behind: a raw state snapshot that a later run can
// resume from (checkpoint / restart), and the step counter that goes with it.
//
// Snapshot file (little endian):
//   bytes 0..7    magic "NBODYB2\0"
//   u32 version (2) | u32 n_bodies | u64 steps_done
//   f32 eps2 | f32 dt | i32 integrator | u32 variant (kernel that wrote it; 0 = sharded / multi-device run)
//   f64 eps2_f64 | f64 dt_f64 | u64 FNV-1a of the payload     -> 64-byte header
// Version-1 files (no variant, no checksum) still load.  Load verifies the length, the checksum and that the
// physics in the header (eps2, dt, integrator) is the context's: a resumed run that silently continued with
// other parameters would carry steps_done over to a different system.  When the saving context and the loading
// one are both single-device and the loader resolved its kernel with NBODY_VARIANT_AUTO, the loader re-pins the
// saver's kernel variant, so a resumed run is bit-identical to the uninterrupted one even across sizes where the
// automatic choice could differ between library builds.
//   float4 X[n] (x, y, z, mass) | float4 V[n] (vx, vy, vz, pad)
// i.e. the reference's host arrays `bodies` and `velocity` (main.cpp:250-251)
// as they stand after `steps_done` steps.

namespace {

struct SnapshotHeader {
    char magic[8];
    unsigned version, n;
    unsigned long long steps_done;
    float eps2, dt;
    int integrator;
    unsigned variant;
    double eps2_f64, dt_f64;
    unsigned long long checksum;
};
static_assert(sizeof(SnapshotHeader) == 64, "snapshot header is 64 bytes");
const char kSnapshotMagic[8] = {'N', 'B', 'O', 'D', 'Y', 'B', '2', '\0'};

unsigned long long fnv1a(const void* data, size_t bytes, unsigned long long h = 0xcbf29ce484222325ull)
{
    // 8 bytes per round (the payload is float4 arrays): a corruption check, not a cryptographic digest
    const unsigned long long* w = static_cast<const unsigned long long*>(data);
    for (size_t i = 0; i < bytes / 8; ++i) { h ^= w[i]; h *= 0x100000001b3ull; }
    return h;
}

}  // namespace

extern "C" {

int nbody_steps_done(nbody_ctx* c, long long* steps)
{
    if (!c || !steps) return fail(NBODY_E_INVALID, "null argument");
    *steps = c->steps_done;
    return NBODY_OK;
}

int nbody_snapshot_save(nbody_ctx* c, const char* path)
{
    TRACE_CALL();
    if (!c || !path) return fail(NBODY_E_INVALID, "null argument");
    if (c->p.world != 1) return fail(NBODY_E_STATE, "nbody_snapshot_save needs the whole system: an unsharded or a multi-device context (gather the shards with nbody_download first)");
    const size_t nb = (size_t)c->n * sizeof(float4);
    std::vector<float4> x((size_t)c->n), v((size_t)c->n);
    int rc = nbody_download(c, reinterpret_cast<float*>(x.data()), reinterpret_cast<float*>(v.data()), nullptr);
    if (rc) return rc;
    SnapshotHeader h;
    memset(&h, 0, sizeof h);
    memcpy(h.magic, kSnapshotMagic, 8);
    h.version = 2; h.n = (unsigned)c->n; h.steps_done = (unsigned long long)c->steps_done;
    h.eps2 = c->p.eps2; h.dt = c->p.dt; h.integrator = c->p.integrator;
    // low 16 bits: kernel variant; bit 16: single-level summation
    h.variant = c->kids.empty() ? ((unsigned)c->var->id | (c->plan_full.single ? 0x10000u : 0u)) : 0u;
    h.eps2_f64 = c->p.eps2_f64; h.dt_f64 = c->p.dt_f64;
    h.checksum = fnv1a(v.data(), nb, fnv1a(x.data(), nb));
    FILE* f = fopen(path, "wb");
    if (!f) return fail(NBODY_E_INVALID, "cannot open %s for writing", path);
    const bool ok = fwrite(&h, sizeof h, 1, f) == 1 && fwrite(x.data(), nb, 1, f) == 1 && fwrite(v.data(), nb, 1, f) == 1;
    if (fclose(f) != 0 || !ok) return fail(NBODY_E_INVALID, "short write to %s", path);
    return NBODY_OK;
}

int nbody_snapshot_load(nbody_ctx* c, const char* path)
{
    TRACE_CALL();
    if (!c || !path) return fail(NBODY_E_INVALID, "null argument");
    FILE* f = fopen(path, "rb");
    if (!f) return fail(NBODY_E_INVALID, "cannot open %s", path);
    SnapshotHeader h;
    if (fread(&h, sizeof h, 1, f) != 1 || memcmp(h.magic, kSnapshotMagic, 8) != 0 || (h.version != 1 && h.version != 2)) {
        fclose(f);
        return fail(NBODY_E_INVALID, "%s is not a version-1/2 N-body snapshot", path);
    }
    if ((int)h.n != c->n) {
        fclose(f);
        return fail(NBODY_E_INVALID, "%s holds %u bodies, the context %d", path, h.n, c->n);
    }
    // same physics, or the step counter would carry over to a different system
    const bool f64 = c->p.integrator == NBODY_INTEG_F64;
    if (h.integrator != c->p.integrator || (f64 ? (h.eps2_f64 != c->p.eps2_f64 || h.dt_f64 != c->p.dt_f64) : (h.eps2 != c->p.eps2 || h.dt != c->p.dt))) {
        fclose(f);
        return fail(NBODY_E_INVALID, "%s was written with eps2=%g dt=%g integrator=%d (f64: %g, %g); the context has eps2=%g dt=%g integrator=%d (f64: %g, %g)",
                    path, (double)h.eps2, (double)h.dt, h.integrator, h.eps2_f64, h.dt_f64, (double)c->p.eps2, (double)c->p.dt, c->p.integrator, c->p.eps2_f64, c->p.dt_f64);
    }
    const size_t nb = (size_t)c->n * sizeof(float4);
    std::vector<float4> x((size_t)c->n), v((size_t)c->n);
    const bool ok = fread(x.data(), nb, 1, f) == 1 && fread(v.data(), nb, 1, f) == 1;
    char extra;
    const bool longer = ok && fread(&extra, 1, 1, f) == 1;
    fclose(f);
    if (!ok) return fail(NBODY_E_INVALID, "%s is truncated", path);
    if (longer) return fail(NBODY_E_INVALID, "%s is longer than a snapshot of %d bodies", path, c->n);
    if (h.version >= 2 && h.checksum != fnv1a(v.data(), nb, fnv1a(x.data(), nb)))
        return fail(NBODY_E_INVALID, "%s: payload checksum mismatch (corrupted file)", path);
    // bit-identical resume: run the kernel the saver ran (same tile shape and unit size = same summation order)
    // (the summation flag is compared with what THIS context would run on the saver's kernel: the small-N and row-split
    // families have one summation only and always write 0)
    if (h.version >= 2 && h.variant != 0 && c->kids.empty() && c->p.world == 1) {
        const Variant* sv = find_variant((int)(h.variant & 0xffffu));
        const bool would_be_single = c->p.summation == NBODY_SUM_SINGLE && (sv ? sv->kernel_single != nullptr : c->plan_full.single);
        if (((h.variant & 0x10000u) != 0) != would_be_single)
            return fail(NBODY_E_STATE, "%s was written with %s summation, the context uses the other one (nbody_params.summation)", path, (h.variant & 0x10000u) ? "single-level" : "two-level");
    }
    if (h.version >= 2 && h.variant != 0 && c->kids.empty() && c->p.world == 1 && (int)(h.variant & 0xffffu) != c->var->id) {
        const Variant* sv = find_variant((int)(h.variant & 0xffffu));
        const bool same_sums = sv && sv->small == c->var->small && sv->H == c->var->H && sv->T == c->var->T && sv->NT == c->var->NT && sv->UJ == c->var->UJ && sv->cs == c->var->cs && sv->cmode == c->var->cmode;
        if (!same_sums) {
            if (c->p.variant != NBODY_VARIANT_AUTO)
                return fail(NBODY_E_STATE, "%s was written by kernel variant %u; the context was created with variant %d (other summation order): create it with NBODY_VARIANT_AUTO or that variant", path, h.variant & 0xffffu, c->p.variant);
            if (!sv) return fail(NBODY_E_STATE, "%s was written by kernel variant %u, which this build of the library does not have", path, h.variant & 0xffffu);
            int rc = repin_variant(c, sv);
            if (rc) return rc;
        }
    }
    // a sharded context reads the whole file and keeps its own slice, like nbody_upload
    int rc = nbody_upload(c, reinterpret_cast<const float*>(x.data()), reinterpret_cast<const float*>(v.data()));
    if (rc) return rc;
    c->steps_done = (long long)h.steps_done;
    return NBODY_OK;
}

int nbody_snapshot_info(const char* path, int* n_bodies, long long* steps_done, float* eps2, float* dt)
{
    if (!path) return fail(NBODY_E_INVALID, "null path");
    FILE* f = fopen(path, "rb");
    if (!f) return fail(NBODY_E_INVALID, "cannot open %s", path);
    SnapshotHeader h;
    const bool ok = fread(&h, sizeof h, 1, f) == 1 && memcmp(h.magic, kSnapshotMagic, 8) == 0 && (h.version == 1 || h.version == 2);
    fclose(f);
    if (!ok) return fail(NBODY_E_INVALID, "%s is not a version-1/2 N-body snapshot", path);
    if (n_bodies) *n_bodies = (int)h.n;
    if (steps_done) *steps_done = (long long)h.steps_done;
    if (eps2) *eps2 = h.eps2;
    if (dt) *dt = h.dt;
    return NBODY_OK;
}

}  // extern "C"
