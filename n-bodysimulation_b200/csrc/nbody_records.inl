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
//   u32 version (1) | u32 n_bodies | u64 steps_done
//   f32 eps2 | f32 dt | i32 integrator | u32 reserved
//   f64 eps2_f64 | f64 dt_f64 | 8 bytes reserved             -> 64-byte header
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
    unsigned reserved0;
    double eps2_f64, dt_f64;
    unsigned char reserved1[8];
};
static_assert(sizeof(SnapshotHeader) == 64, "snapshot header is 64 bytes");
const char kSnapshotMagic[8] = {'N', 'B', 'O', 'D', 'Y', 'B', '2', '\0'};

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
    if (c->p.world != 1) return fail(NBODY_E_STATE, "nbody_snapshot_save needs an unsharded context (gather the shards with nbody_download first)");
    int rc = set_device(c);
    if (rc) return rc;
    const size_t nb = (size_t)c->n * sizeof(float4);
    std::vector<float4> x((size_t)c->n), v((size_t)c->n);
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaMemcpy(x.data(), c->X, nb, cudaMemcpyDeviceToHost));
    CU(cudaMemcpy(v.data(), c->V, nb, cudaMemcpyDeviceToHost));
    SnapshotHeader h;
    memset(&h, 0, sizeof h);
    memcpy(h.magic, kSnapshotMagic, 8);
    h.version = 1; h.n = (unsigned)c->n; h.steps_done = (unsigned long long)c->steps_done;
    h.eps2 = c->p.eps2; h.dt = c->p.dt; h.integrator = c->p.integrator;
    h.eps2_f64 = c->p.eps2_f64; h.dt_f64 = c->p.dt_f64;
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
    if (fread(&h, sizeof h, 1, f) != 1 || memcmp(h.magic, kSnapshotMagic, 8) != 0 || h.version != 1) {
        fclose(f);
        return fail(NBODY_E_INVALID, "%s is not a version-1 N-body snapshot", path);
    }
    if ((int)h.n != c->n) {
        fclose(f);
        return fail(NBODY_E_INVALID, "%s holds %u bodies, the context %d", path, h.n, c->n);
    }
    const size_t nb = (size_t)c->n * sizeof(float4);
    std::vector<float4> x((size_t)c->n), v((size_t)c->n);
    const bool ok = fread(x.data(), nb, 1, f) == 1 && fread(v.data(), nb, 1, f) == 1;
    fclose(f);
    if (!ok) return fail(NBODY_E_INVALID, "%s is truncated", path);
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
    const bool ok = fread(&h, sizeof h, 1, f) == 1 && memcmp(h.magic, kSnapshotMagic, 8) == 0 && h.version == 1;
    fclose(f);
    if (!ok) return fail(NBODY_E_INVALID, "%s is not a version-1 N-body snapshot", path);
    if (n_bodies) *n_bodies = (int)h.n;
    if (steps_done) *steps_done = (long long)h.steps_done;
    if (eps2) *eps2 = h.eps2;
    if (dt) *dt = h.dt;
    return NBODY_OK;
}

}  // extern "C"
