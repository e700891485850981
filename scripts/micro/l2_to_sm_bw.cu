// l2_to_sm_bw.cu — how many bytes per second arrive at the SMs from an L2-resident buffer (cp.async.bulk global -> shared,
// the path the fused kernels' TMA loads take).  The denominator of DESIGN.md 4.1c's "operand delivery" bound.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o nmf_gpu_b200/build/l2_to_sm_bw scripts/micro/l2_to_sm_bw.cu
//   ./l2_to_sm_bw [footprint MB = 48] [chunk KB = 16] [stages = 4] [CTAs per SM = 1]
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(128) k_stream(const char* __restrict__ buf, size_t footprint, int chunk, int stages, int iters,
                                                unsigned long long* sink) {
    extern __shared__ __align__(128) char smem[];
    uint64_t* bar = reinterpret_cast<uint64_t*>(smem);          // `stages` barriers, data from byte 1024
    char* data = smem + 1024;
    const size_t nchunks = footprint / (size_t)chunk;
    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(bar + s)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        size_t c = ((size_t)blockIdx.x * 7919u) % nchunks;          // CTAs start spread over the buffer and walk it with a stride
        for (int it = 0; it < iters + stages; ++it) {
            const int s = it % stages;
            if (it >= stages) {                                      // wait for the load issued `stages` trips ago
                const uint32_t par = ((it / stages) - 1) & 1;
                uint32_t done = 0;
                while (!done)
                    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                                 : "=r"(done) : "r"(smem_u32(bar + s)), "r"(par) : "memory");
            }
            if (it < iters) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar + s)), "r"(chunk) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(smem_u32(data + (size_t)s * chunk)), "l"(buf + c * (size_t)chunk), "r"(chunk), "r"(smem_u32(bar + s)) : "memory");
                c += gridDim.x; if (c >= nchunks) c -= nchunks;
            }
        }
        sink[blockIdx.x] = *reinterpret_cast<unsigned long long*>(data);
    }
}

int main(int argc, char** argv) {
    const size_t mb = argc > 1 ? atoi(argv[1]) : 48;
    const int chunk = (argc > 2 ? atoi(argv[2]) : 16) * 1024, stages = argc > 3 ? atoi(argv[3]) : 4, per_sm = argc > 4 ? atoi(argv[4]) : 1;
    cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0);
    const int grid = pr.multiProcessorCount * per_sm, iters = 20000;
    const size_t footprint = mb << 20;
    char* buf; unsigned long long* sink;
    cudaMalloc(&buf, footprint); cudaMemset(buf, 1, footprint); cudaMalloc(&sink, grid * 8);
    const int smem = 1024 + stages * chunk;
    cudaFuncSetAttribute(k_stream, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        k_stream<<<grid, 128, smem>>>(buf, footprint, chunk, stages, iters, sink);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep && ms < best) best = ms;
    }
    cudaError_t err = cudaGetLastError();
    const double bytes = (double)grid * iters * chunk;
    int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("{\"footprint_mb\": %zu, \"chunk_kb\": %d, \"stages\": %d, \"ctas\": %d, \"ms\": %.4f, \"tb_per_s\": %.3f, \"bytes_per_clk_per_sm_at_max_clock\": %.1f, \"err\": \"%s\"}\n",
           mb, chunk / 1024, stages, grid, best, bytes / (best * 1e-3) / 1e12, bytes / (best * 1e-3) / pr.multiProcessorCount / (clk * 1e3), cudaGetErrorString(err));
    return 0;
}
