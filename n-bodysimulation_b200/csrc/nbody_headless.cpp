// nbody_headless.cpp -- the reference's headless run, restated over the engine.
//
// Follows TestProject/main.cpp: pinned host state (:250-252), initial
// conditions (:254,271-272 -> utils.cpp:30-37), device copies (:275-283,352-354),
// simulationLoopNoVisual (:142-160: `steps` calls of simulate(), exceptions ->
// EXIT_FAILURE) and the frees (:358-366).  The three stdin prompts
// (main.cpp:163-228) become flags; the visual branches need OpenGL and are
// unavailable on this box.  Unlike the reference, which drops its results, the
// driver can print a state checksum and a device-timed throughput line.
//
//   nbody_headless [--n N] [--steps S] [--ic reference|uniform|plummer] [--seed K]
//                  [--eps2 E] [--dt D] [--resident] [--host-f3] [--json]
//                  [--flavour simulate|emb|emb-float3]   which launcher of kernel.cuh:2-7 (headless forms)
//                  [--ppm FILE]                          final frame (the reference's GL draw as CUDA)
//                  [--save FILE] [--load FILE]           snapshot / resume (with --resident)
//                  [--energy]                            with --resident: a second JSON line with KE, PE before / after and
//                                                        the relative drift of E_eff = KE + PE / 2 (BASELINE's drift bound)
//                  [--gpus G]                            G GPUs of this box from ONE process: with --resident ONE multi-device
//                                                        context (each device owns N/G targets; --exchange p2p|nccl), else
//                                                        simulate_set_devices(G): the reference's loop over simulate() with
//                                                        the caller's arrays on device 0 and the step spread over G devices
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <iostream>
#include <string>
#include <vector>

#include "../../include/kernel.cuh"
#include "../../include/nbody_b200.h"

namespace {

// TestProject/utils.cpp:6,30-37 with the unseeded rand() the reference relies on
float random_between(float lo, float hi) { return ((float)rand() / RAND_MAX) * (hi - lo) + lo; }
void fill_reference(float4* b, int n)
{
    for (int i = 0; i < n; ++i) {
        b[i].x = random_between(-MAX_X, MAX_X);
        b[i].y = random_between(-MAX_Y, MAX_Y);
        b[i].z = random_between(-MAX_Z, MAX_Z);
        b[i].w = random_between(MIN_W, MAX_W);
    }
}

// main.cpp:142-160
// `after_first` (optional) is recorded after the first call: that call also loads the kernel
// image and allocates the library's scratch, which the reference's launcher never had to do.
int simulationLoopNoVisual(float4* d_bodies, float4* d_accel, float4* d_vel, int n, int steps, cudaEvent_t after_first = nullptr)
{
    printf("Starting the simulation...\n");
    for (int counter = 0; counter < steps; ++counter) {
        try {
            simulate(d_bodies, d_accel, d_vel, n);
        } catch (const std::exception& e) {
            std::cerr << e.what() << std::endl;
            return EXIT_FAILURE;
        }
        if (counter == 0 && after_first) cudaEventRecord(after_first);
    }
    printf("Simulation complete\n");
    return 0;
}

double checksum(const float4* v, int n)
{
    double s = 0;
    for (int i = 0; i < n; ++i) s += (double)v[i].x + 2.0 * v[i].y + 3.0 * v[i].z;
    return s;
}

}  // namespace

int main(int argc, char** argv)
{
    int n = N_BODIES, steps = 10;
    unsigned long long seed = 12345;
    float eps2 = EPS2, dt = DT;
    std::string ic = "reference", flavour = "simulate", exchange = "p2p", ppm, save, load;
    bool resident = false, host_f3 = false, json = false, energy = false;
    int gpus = 1;
    for (int i = 1; i < argc; ++i) {
        auto need = [&](const char* f) { if (i + 1 >= argc) { fprintf(stderr, "%s needs a value\n", f); exit(2); } return argv[++i]; };
        if (!strcmp(argv[i], "--n")) n = atoi(need("--n"));
        else if (!strcmp(argv[i], "--steps")) steps = atoi(need("--steps"));
        else if (!strcmp(argv[i], "--ic")) ic = need("--ic");
        else if (!strcmp(argv[i], "--seed")) seed = strtoull(need("--seed"), nullptr, 10);
        else if (!strcmp(argv[i], "--eps2")) eps2 = (float)atof(need("--eps2"));
        else if (!strcmp(argv[i], "--dt")) dt = (float)atof(need("--dt"));
        else if (!strcmp(argv[i], "--resident")) resident = true;
        else if (!strcmp(argv[i], "--host-f3")) host_f3 = true;
        else if (!strcmp(argv[i], "--json")) json = true;
        else if (!strcmp(argv[i], "--energy")) energy = true;
        else if (!strcmp(argv[i], "--gpus")) gpus = atoi(need("--gpus"));
        else if (!strcmp(argv[i], "--flavour")) flavour = need("--flavour");
        else if (!strcmp(argv[i], "--exchange")) exchange = need("--exchange");
        else if (!strcmp(argv[i], "--ppm")) ppm = need("--ppm");
        else if (!strcmp(argv[i], "--save")) save = need("--save");
        else if (!strcmp(argv[i], "--load")) load = need("--load");
        else { fprintf(stderr, "unknown flag %s\n", argv[i]); return 2; }
    }
    if (n <= 0 || steps < 0) { fprintf(stderr, "bad --n / --steps\n"); return 2; }
    if (energy && !resident) { fprintf(stderr, "--energy needs --resident\n"); return 2; }
    if ((!save.empty() || !load.empty()) && !resident) { fprintf(stderr, "--save / --load need --resident\n"); return 2; }
    if (flavour != "simulate" && flavour != "emb" && flavour != "emb-float3") { fprintf(stderr, "unknown --flavour %s\n", flavour.c_str()); return 2; }
    if (exchange != "p2p" && exchange != "nccl") { fprintf(stderr, "unknown --exchange %s\n", exchange.c_str()); return 2; }
    if (gpus < 1 || gpus > NBODY_MAX_DEVICES) { fprintf(stderr, "bad --gpus\n"); return 2; }
    if (gpus > 1 && host_f3) { fprintf(stderr, "--host-f3 is the host-pointer flavour: one GPU\n"); return 2; }
    const size_t size4 = sizeof(float4) * (size_t)n;

    float4 *bodies, *velocity, *accelerations;   // pinned, main.cpp:250-252
    if (cudaMallocHost(&bodies, size4) != cudaSuccess || cudaMallocHost(&velocity, size4) != cudaSuccess ||
        cudaMallocHost(&accelerations, size4) != cudaSuccess) {
        fprintf(stderr, "cudaMallocHost failed: %s\n", cudaGetErrorString(cudaGetLastError()));
        return EXIT_FAILURE;
    }
    memset(velocity, 0, size4);
    memset(accelerations, 0, size4);
    if (ic == "reference") fill_reference(bodies, n);
    else if (ic == "uniform") nbody_ic_uniform((float*)bodies, n, seed, MAX_X, MIN_W, MAX_W);
    else if (ic == "plummer") nbody_ic_plummer((float*)bodies, (float*)velocity, n, seed);
    else { fprintf(stderr, "unknown --ic %s\n", ic.c_str()); return 2; }

    int rc = 0;
    float ms = -1.f, ms_first = -1.f, ms_steady = -1.f;
    double ke0 = 0, pe0 = 0, ke1 = 0, pe1 = 0;   // --energy: KE and pair potential before / after (conserved: KE + PE / 2)
    if (gpus > 1 && resident) {
        // ONE context spanning G devices (nbody_params.n_devices): the library owns the per-device shards, the peer
        // mapping, the exchange (--exchange p2p: fused into the step kernel, in-kernel flag barrier; nccl: ncclAllGather
        // on a second stream) -- the host code below is the single-GPU --resident branch with a device list.
        nbody_params p;
        nbody_params_default(&p);
        p.n_bodies = n; p.eps2 = eps2; p.dt = dt; p.n_devices = gpus;
        p.exchange = exchange == "nccl" ? NBODY_EXCHANGE_NCCL : NBODY_EXCHANGE_P2P;
        for (int g = 0; g < gpus && g < NBODY_MAX_DEVICES; ++g) p.devices[g] = g;
        nbody_ctx* ctx = nullptr;
        if (nbody_create(&ctx, &p) || (load.empty() ? nbody_upload(ctx, (float*)bodies, (float*)velocity) : nbody_snapshot_load(ctx, load.c_str()))) {
            std::cerr << nbody_last_error() << std::endl;
            return EXIT_FAILURE;
        }
        if (energy && nbody_energy(ctx, &ke0, &pe0)) { std::cerr << nbody_last_error() << std::endl; return EXIT_FAILURE; }
        printf("Starting the simulation...\n");
        if (nbody_step(ctx, steps)) { std::cerr << nbody_last_error() << std::endl; rc = EXIT_FAILURE; }
        else printf("Simulation complete\n");
        nbody_elapsed_ms(ctx, &ms);
        if (energy && rc == 0 && nbody_energy(ctx, &ke1, &pe1)) { std::cerr << nbody_last_error() << std::endl; rc = EXIT_FAILURE; }
        int timed_out = 0;
        nbody_shard_barrier_timeouts(ctx, &timed_out);
        if (timed_out) { std::cerr << "cross-GPU barrier timed out " << timed_out << " time(s): state invalid" << std::endl; rc = EXIT_FAILURE; }
        nbody_download(ctx, (float*)bodies, (float*)velocity, (float*)accelerations);
        if (!save.empty() && nbody_snapshot_save(ctx, save.c_str())) { std::cerr << nbody_last_error() << std::endl; rc = EXIT_FAILURE; }
        nbody_destroy(ctx);
    } else if (host_f3) {
        // headless-directory flavour: host pointers, float3 velocity (its constants: EPS2 0.002, DT 0.01)
        float3* v3 = new float3[n];
        float3* a3 = new float3[n];
        for (int i = 0; i < n; ++i) { v3[i] = make_float3(velocity[i].x, velocity[i].y, velocity[i].z); a3[i] = make_float3(0, 0, 0); }
        printf("Starting the simulation...\n");
        try {
            for (int s = 0; s < steps; ++s) simulate(bodies, a3, v3, n);
            printf("Simulation complete\n");
        } catch (const std::exception& e) { std::cerr << e.what() << std::endl; rc = EXIT_FAILURE; }
        for (int i = 0; i < n; ++i) { velocity[i].x = v3[i].x; velocity[i].y = v3[i].y; velocity[i].z = v3[i].z; }
        delete[] v3; delete[] a3;
    } else if (resident) {
        // state stays in HBM between steps (no per-step repack, device-timed)
        nbody_params p;
        nbody_params_default(&p);
        p.n_bodies = n; p.eps2 = eps2; p.dt = dt;
        nbody_ctx* ctx = nullptr;
        if (nbody_create(&ctx, &p) || (load.empty() ? nbody_upload(ctx, (float*)bodies, (float*)velocity) : nbody_snapshot_load(ctx, load.c_str()))) {
            std::cerr << nbody_last_error() << std::endl;
            return EXIT_FAILURE;
        }
        if (energy && nbody_energy(ctx, &ke0, &pe0)) { std::cerr << nbody_last_error() << std::endl; return EXIT_FAILURE; }
        printf("Starting the simulation...\n");
        if (nbody_step(ctx, steps)) { std::cerr << nbody_last_error() << std::endl; rc = EXIT_FAILURE; }
        else printf("Simulation complete\n");
        nbody_elapsed_ms(ctx, &ms);
        if (energy && rc == 0 && nbody_energy(ctx, &ke1, &pe1)) { std::cerr << nbody_last_error() << std::endl; rc = EXIT_FAILURE; }
        nbody_download(ctx, (float*)bodies, (float*)velocity, (float*)accelerations);
        if (!save.empty() && nbody_snapshot_save(ctx, save.c_str())) { std::cerr << nbody_last_error() << std::endl; rc = EXIT_FAILURE; }
        nbody_destroy(ctx);
    } else {
        // the reference's own sequence: main.cpp:275-283,352-356
        float4 *d_bodies, *d_velocity, *d_accelerations;
        cudaMalloc((void**)&d_velocity, size4);
        cudaMalloc((void**)&d_accelerations, size4);
        cudaMemcpy(d_velocity, velocity, size4, cudaMemcpyHostToDevice);
        cudaMemcpy(d_accelerations, accelerations, size4, cudaMemcpyHostToDevice);
        cudaMalloc((void**)&d_bodies, size4);
        cudaMemcpy(d_bodies, bodies, size4, cudaMemcpyHostToDevice);
        simulate_set_parameters(eps2, dt);
        try {
            simulate_set_devices(gpus);   // the one added line that lets main.cpp's loop use every GPU of the box
        } catch (const std::exception& e) { std::cerr << e.what() << std::endl; return EXIT_FAILURE; }
        cudaEvent_t e0, e1, ef;
        cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventCreate(&ef);
        cudaEventRecord(e0);
        if (flavour == "simulate") {
            rc = simulationLoopNoVisual(d_bodies, d_accelerations, d_velocity, n, steps, ef);
        } else {
            // main.cpp:20-106 (simulationLoopVisualEmb / _float3) without the window: the float3 flavour
            // views the same allocations as packed float3 arrays (they are large enough)
            printf("Starting the simulation...\n");
            try {
                for (int s = 0; s < steps; ++s) {
                    if (flavour == "emb") simulateHeadless_embParallel(d_bodies, d_accelerations, d_velocity, nullptr, n);
                    else simulateHeadless_embParallel_float3(d_bodies, (float3*)d_accelerations, (float3*)d_velocity, nullptr, n);
                    if (s == 0) cudaEventRecord(ef);
                }
                printf("Simulation complete\n");
            } catch (const std::exception& e) { std::cerr << e.what() << std::endl; rc = EXIT_FAILURE; }
        }
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1);
        if (steps > 1 && rc == 0) { cudaEventElapsedTime(&ms_first, e0, ef); ms_steady = (ms - ms_first) / (steps - 1); }
        cudaMemcpy(bodies, d_bodies, size4, cudaMemcpyDeviceToHost);
        cudaMemcpy(velocity, d_velocity, size4, cudaMemcpyDeviceToHost);
        cudaMemcpy(accelerations, d_accelerations, size4, cudaMemcpyDeviceToHost);
        cudaFree(d_bodies); cudaFree(d_velocity); cudaFree(d_accelerations);
    }
    if (!ppm.empty() && rc == 0) {
        // what renderBodies + the shaders would have put in the window after the last step
        nbody_view view;
        nbody_view_default(&view);
        const size_t fb = (size_t)view.width * view.height * 4;
        float4* d_x = nullptr;
        unsigned char* d_f = nullptr;
        unsigned char* h_f = (unsigned char*)malloc(fb);
        bool ok = cudaMalloc((void**)&d_x, size4) == cudaSuccess && cudaMalloc((void**)&d_f, fb) == cudaSuccess &&
                  cudaMemcpy(d_x, bodies, size4, cudaMemcpyHostToDevice) == cudaSuccess &&
                  nbody_render_device((const float*)d_x, n, &view, d_f) == NBODY_OK &&
                  cudaMemcpy(h_f, d_f, fb, cudaMemcpyDeviceToHost) == cudaSuccess &&
                  nbody_write_ppm(ppm.c_str(), h_f, view.width, view.height) == NBODY_OK;
        if (!ok) { std::cerr << "frame: " << nbody_last_error() << std::endl; rc = EXIT_FAILURE; }
        cudaFree(d_x); cudaFree(d_f); free(h_f);
    }
    if (json && rc == 0) {
        const double inter = (double)n * (double)n * steps;
        printf("{\"n\": %d, \"steps\": %d, \"ic\": \"%s\", \"flavour\": \"%s\", \"device_ms\": %.4f, \"g_interactions_per_s\": %.3f, "
               "\"first_call_ms\": %.4f, \"steady_ms_per_step\": %.4f, \"steady_g_interactions_per_s\": %.3f, "
               "\"checksum_x\": %.9e, \"checksum_v\": %.9e, \"launches\": %lld, \"gpus\": %d, \"exchange\": \"%s\"}\n",
               n, steps, ic.c_str(), host_f3 ? "host-f3" : resident ? "resident" : flavour.c_str(), ms, ms > 0 ? inter / (ms * 1e-3) / 1e9 : 0.0,
               ms_first, ms_steady, ms_steady > 0 ? (double)n * (double)n / (ms_steady * 1e-3) / 1e9 : 0.0,
               checksum(bodies, n), checksum(velocity, n), nbody_launch_count(), gpus, gpus > 1 && resident ? exchange.c_str() : gpus > 1 ? "set_devices" : "none");
        if (energy) {
            const double e0 = ke0 + 0.5 * pe0, e1 = ke1 + 0.5 * pe1;
            printf("{\"energy\": \"E_eff = KE + PE / 2\", \"ke0\": %.9e, \"pe0\": %.9e, \"ke1\": %.9e, \"pe1\": %.9e, \"e_eff0\": %.9e, \"e_eff1\": %.9e, \"rel_drift\": %.3e}\n",
                   ke0, pe0, ke1, pe1, e0, e1, e0 != 0 ? (e1 - e0) / (e0 < 0 ? -e0 : e0) : 0.0);
        }
    }
    nbody_release_cached();
    cudaFreeHost(bodies); cudaFreeHost(velocity); cudaFreeHost(accelerations);
    return rc;
}
