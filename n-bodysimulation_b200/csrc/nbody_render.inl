// nbody_render.inl -- third part of nbody_engine.cu: headless point-splat frame.
//
// Replaces what the reference draws after every step with OpenGL
// (TestProject/simulation_visualization.cpp:15-37 vertex shader, :40-56 fragment
// shader, :132-135 depth test + 1-pixel points, :165-169 glDrawArrays(GL_POINTS),
// uniforms :203-207, 800x600 window simulation_visualization.h:8-9; cleared
// every frame, main.cpp:48).  No graphics API is involved: the shader arithmetic
// runs as two CUDA kernels on the positions already resident in HBM and the
// frame is a device RGBA8 buffer (optionally written out as a PPM file).
//
//   splat  : one thread per body -> NDC, clip, pixel, depth; one 64-bit atomic per
//            visible body on a per-pixel key (depth bits << 32 | body index), so
//            GL_LESS with "first drawn wins ties" is an atomicMin.  With the
//            depth test off the key is the body index and the LAST body wins.
//   resolve: one thread per pixel -> colour of the winning body, RGBA8 store.
// Every float operation is an explicitly rounded intrinsic in the order the
// shader source states, so the frame is bit-identical to oracle_render().

namespace {

struct ViewArgs {
    int W, H;
    float mx, my, mz, min_w, max_w;
    int depth_test;
};

constexpr unsigned long long kKeyEmptyLess = ((unsigned long long)0x3f800000u << 32) | 0xffffffffull;   // depth 1.0, no body

__device__ __forceinline__ float ndc_of(float x, float m)
{
    // 2.0 * (x - (-m)) / (m - (-m)) - 1   (simulation_visualization.cpp:31-33)
    return __fsub_rn(__fdiv_rn(__fmul_rn(2.0f, __fsub_rn(x, -m)), __fsub_rn(m, -m)), 1.0f);
}

__global__ void render_clear_kernel(unsigned long long* __restrict__ keys, int npix, int depth_test)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p < npix) keys[p] = depth_test ? kKeyEmptyLess : 0ull;
}

__global__ void render_splat_kernel(const float4* __restrict__ X, int n, ViewArgs v, unsigned long long* __restrict__ keys)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 b = X[i];
    const float nx = ndc_of(b.x, v.mx), ny = ndc_of(b.y, v.my), nz = ndc_of(b.z, v.mz);
    if (!(nx >= -1.0f && nx <= 1.0f && ny >= -1.0f && ny <= 1.0f && nz >= -1.0f && nz <= 1.0f)) return;   // clipped
    const float xw = __fmul_rn(__fadd_rn(nx, 1.0f), __fmul_rn((float)v.W, 0.5f));
    const float yw = __fmul_rn(__fadd_rn(ny, 1.0f), __fmul_rn((float)v.H, 0.5f));
    const int px = (int)floorf(xw), py = (int)floorf(yw);
    if (px < 0 || px >= v.W || py < 0 || py >= v.H) return;
    const int p = (v.H - 1 - py) * v.W + px;   // image rows top-down, GL's y up
    if (v.depth_test) {
        const float depth = __fmul_rn(__fadd_rn(nz, 1.0f), 0.5f);
        if (!(depth < 1.0f)) return;            // GL_LESS against the cleared depth
        atomicMin(&keys[p], ((unsigned long long)__float_as_uint(depth) << 32) | (unsigned)i);
    } else {
        atomicMax(&keys[p], (unsigned long long)i + 1ull);
    }
}

__global__ void render_resolve_kernel(const float4* __restrict__ X, ViewArgs v, const unsigned long long* __restrict__ keys,
                                      uchar4* __restrict__ rgba)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= v.W * v.H) return;
    const unsigned long long k = keys[p];
    uchar4 out = make_uchar4(0, 0, 0, 0);      // glClear: colour (0,0,0,0)
    const bool hit = v.depth_test ? (k != kKeyEmptyLess) : (k != 0ull);
    if (hit) {
        const unsigned i = v.depth_test ? (unsigned)(k & 0xffffffffull) : (unsigned)(k - 1ull);
        const float w = X[i].w;
        const float wgt = __fdiv_rn(__fsub_rn(w, v.min_w), __fsub_rn(v.max_w, v.min_w));   // :27
        const float omw = __fsub_rn(1.0f, wgt);
        // mix(vec3(0,1,0), vec3(1,0,0), weight)  (:47-51)
        float r = __fadd_rn(__fmul_rn(0.0f, omw), __fmul_rn(1.0f, wgt));
        float g = __fadd_rn(__fmul_rn(1.0f, omw), __fmul_rn(0.0f, wgt));
        r = r < 0.0f ? 0.0f : (r > 1.0f ? 1.0f : r);
        g = g < 0.0f ? 0.0f : (g > 1.0f ? 1.0f : g);
        out.x = (unsigned char)(int)floorf(__fadd_rn(__fmul_rn(r, 255.0f), 0.5f));
        out.y = (unsigned char)(int)floorf(__fadd_rn(__fmul_rn(g, 255.0f), 0.5f));
        out.z = 0;
        out.w = 255;
    }
    rgba[p] = out;
}

std::mutex g_render_mu;
std::map<int, std::pair<unsigned long long*, size_t>> g_render_keys;   // device -> (keys, pixels)

void release_render_scratch()
{
    std::lock_guard<std::mutex> lk(g_render_mu);
    for (auto& kv : g_render_keys) {
        cudaSetDevice(kv.first);
        cudaFree(kv.second.first);
    }
    g_render_keys.clear();
}

int check_view(const nbody_view* v)
{
    if (!v) return fail(NBODY_E_INVALID, "null view");
    if (v->width <= 0 || v->height <= 0 || (long long)v->width * v->height > (1ll << 28)) return fail(NBODY_E_INVALID, "bad frame size %dx%d", v->width, v->height);
    if (!(v->max_x > 0.f && v->max_y > 0.f && v->max_z > 0.f) || !(v->max_w > v->min_w)) return fail(NBODY_E_INVALID, "bad view volume / weight range");
    return NBODY_OK;
}

// enqueue clear + splat + resolve on `st`; `keys` holds width*height 64-bit words and must not be
// shared with another frame in flight
int render_on_device(const float4* dX, int n, const nbody_view* v, unsigned long long* keys, unsigned char* d_rgba, cudaStream_t st)
{
    const int npix = v->width * v->height;
    ViewArgs a{v->width, v->height, v->max_x, v->max_y, v->max_z, v->min_w, v->max_w, v->depth_test ? 1 : 0};
    render_clear_kernel<<<(npix + 255) / 256, 256, 0, st>>>(keys, npix, a.depth_test);
    render_splat_kernel<<<(n + 255) / 256, 256, 0, st>>>(dX, n, a, keys);
    render_resolve_kernel<<<(npix + 255) / 256, 256, 0, st>>>(dX, a, keys, reinterpret_cast<uchar4*>(d_rgba));
    g_launches.fetch_add(3, std::memory_order_relaxed);
    CU(cudaGetLastError());
    return NBODY_OK;
}

int group_render(nbody_ctx* c, const nbody_view* v, unsigned char* h_rgba);   // nbody_group.inl

// frame of n bodies in device array X (on c's device) into a host buffer, on c's stream with c's frame scratch
int render_arrays(nbody_ctx* c, const float4* X, int n, const nbody_view* v, unsigned char* h_rgba)
{
    int rc = set_device(c);
    if (rc) return rc;
    const size_t npix = (size_t)v->width * v->height, bytes = npix * 4;
    if (c->frame_bytes < bytes) {   // frame + its keys live with the context (its stream orders their use)
        cudaFree(c->frame); cudaFree(c->frame_keys);
        c->frame = nullptr; c->frame_keys = nullptr; c->frame_bytes = 0;
        CU(cudaMalloc(&c->frame, bytes));
        CU(cudaMalloc(&c->frame_keys, npix * sizeof(unsigned long long)));
        c->frame_bytes = bytes;
    }
    rc = render_on_device(X, n, v, c->frame_keys, c->frame, c->stream);
    if (rc) return rc;
    CU(cudaMemcpyAsync(h_rgba, c->frame, bytes, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return NBODY_OK;
}

}  // namespace

extern "C" {

int nbody_view_default(nbody_view* v)
{
    if (!v) return fail(NBODY_E_INVALID, "null view");
    v->width = 800; v->height = 600;                             // simulation_visualization.h:8-9
    v->max_x = 200000.0f; v->max_y = 200000.0f; v->max_z = 200000.0f;   // MAX_VIEWX/Y/Z, constants.h:21-23
    v->min_w = 100000.0f; v->max_w = 1000000000.0f;              // MIN_W / MAX_W, constants.h:18-19
    v->depth_test = 1;                                           // setupOpenGL(), simulation_visualization.cpp:132-135
    return NBODY_OK;
}

int nbody_render_device(const float* d_bodies4, int n, const nbody_view* v, unsigned char* d_rgba)
{
    TRACE_CALL();
    if (!d_bodies4 || !d_rgba) return fail(NBODY_E_INVALID, "null argument");
    if (n <= 0) return fail(NBODY_E_INVALID, "N must be > 0 (got %d)", n);
    int rc = check_view(v);
    if (rc) return rc;
    const size_t npix = (size_t)v->width * v->height;
    int dev = 0;
    CU(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_render_mu);   // held until the frame is finished: the per-device keys are shared
    auto& kb = g_render_keys[dev];
    if (kb.second < npix) {
        cudaFree(kb.first);
        kb = std::make_pair((unsigned long long*)nullptr, (size_t)0);
        CU(cudaMalloc(&kb.first, npix * sizeof(unsigned long long)));
        kb.second = npix;
    }
    rc = render_on_device(reinterpret_cast<const float4*>(d_bodies4), n, v, kb.first, d_rgba, 0);
    if (rc) return rc;
    CU(cudaDeviceSynchronize());
    return NBODY_OK;
}

int nbody_render(nbody_ctx* c, const nbody_view* v, unsigned char* h_rgba)
{
    TRACE_CALL();
    if (!c || !h_rgba) return fail(NBODY_E_INVALID, "null argument");
    int rc = check_view(v);
    if (rc) return rc;
    if (!c->kids.empty()) return group_render(c, v, h_rgba);
    if (c->p.world != 1) return fail(NBODY_E_STATE, "nbody_render needs an unsharded context (a shard holds only its own bodies)");
    return render_arrays(c, c->X, c->n, v, h_rgba);
}

int nbody_write_ppm(const char* path, const unsigned char* rgba, int width, int height)
{
    if (!path || !rgba || width <= 0 || height <= 0) return fail(NBODY_E_INVALID, "bad argument");
    FILE* f = fopen(path, "wb");
    if (!f) return fail(NBODY_E_INVALID, "cannot open %s for writing", path);
    fprintf(f, "P6\n%d %d\n255\n", width, height);
    std::vector<unsigned char> row((size_t)width * 3);
    for (int y = 0; y < height; ++y) {
        for (int x = 0; x < width; ++x) {
            const unsigned char* p = rgba + 4 * ((size_t)y * width + x);
            row[3 * (size_t)x] = p[0]; row[3 * (size_t)x + 1] = p[1]; row[3 * (size_t)x + 2] = p[2];
        }
        if (fwrite(row.data(), 1, row.size(), f) != row.size()) { fclose(f); return fail(NBODY_E_INVALID, "short write to %s", path); }
    }
    fclose(f);
    return NBODY_OK;
}

}  // extern "C"
