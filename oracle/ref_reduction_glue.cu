// ref_reduction_glue.cu -- TEST INFRASTRUCTURE.  Appended by oracle/Makefile to a /tmp copy of the reference's
// TestProject/kernel.cu (one translation unit), so that the reference's OWN reduction-flavour kernels can run on
// plain device memory: its launchers (simulateVisual_embParallel, kernel.cu:536-565, and _float3, :567-597) take an
// OpenGL-mapped buffer, which does not exist here.  Only the launch sequence is restated -- a 2-D grid of
// THREADS_PER_BLOCK^2 blocks for kernel_reduction[_float3], a device synchronisation, then
// inter_block_reduction[_float3] with one thread per body -- the kernels are the reference's, unchanged.
// Host arrays in and out; N must be a multiple of THREADS_PER_BLOCK (the kernels return before their barriers
// otherwise).  Returns the last CUDA error (0 = success).
template <class VA, class KR, class KI>
static int ref_reduction_step(float* h_x4, float* h_v, float* h_a, int N, KR reduce, KI integrate)
{
    const int tpb = THREADS_PER_BLOCK, bpg = (N + tpb - 1) / tpb;
    if (N <= 0 || N % tpb) return -1;
    float4* dx = nullptr;
    VA *dv = nullptr, *da = nullptr, *dr = nullptr;
    cudaMalloc(&dx, sizeof(float4) * N);
    cudaMalloc(&dv, sizeof(VA) * N);
    cudaMalloc(&da, sizeof(VA) * N);
    cudaMalloc(&dr, sizeof(VA) * (size_t)N * bpg);
    cudaMemcpy(dx, h_x4, sizeof(float4) * N, cudaMemcpyHostToDevice);
    cudaMemcpy(dv, h_v, sizeof(VA) * N, cudaMemcpyHostToDevice);
    cudaMemcpy(da, h_a, sizeof(VA) * N, cudaMemcpyHostToDevice);
    reduce<<<dim3(bpg, bpg), dim3(tpb, tpb), sizeof(float4) * tpb * tpb>>>(dx, dr, N);
    cudaDeviceSynchronize();
    integrate<<<bpg, tpb>>>(dx, da, dv, dr, N, bpg);
    cudaDeviceSynchronize();
    cudaMemcpy(h_x4, dx, sizeof(float4) * N, cudaMemcpyDeviceToHost);
    cudaMemcpy(h_v, dv, sizeof(VA) * N, cudaMemcpyDeviceToHost);
    cudaMemcpy(h_a, da, sizeof(VA) * N, cudaMemcpyDeviceToHost);
    cudaFree(dx); cudaFree(dv); cudaFree(da); cudaFree(dr);
    return (int)cudaGetLastError();
}

extern "C" int ref_embparallel_step(float* h_x4, float* h_v4, float* h_a4, int N)
{
    return ref_reduction_step<float4>(h_x4, h_v4, h_a4, N, kernel_reduction, inter_block_reduction);
}

extern "C" int ref_embparallel_float3_step(float* h_x4, float* h_v3, float* h_a3, int N)
{
    return ref_reduction_step<float3>(h_x4, h_v3, h_a3, N, kernel_reduction_float3, inter_block_reduction_float3);
}
