// Microbenchmark: latency of one cp.async.bulk (global -> shared) of a given size, idle SM, L2-resident source.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__global__ void k(const uint8_t *src, int bytes, int reps, long long *out, int concurrent) {
    extern __shared__ __align__(1024) uint8_t sm[];
    __shared__ __align__(8) unsigned long long bar[4];
    if (threadIdx.x == 0) {
        for (int i = 0; i < 4; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[i])));
        asm volatile("fence.mbarrier_init.release.cluster;");
        long long best = 1LL << 60, sum = 0;
        for (int r = 0; r < reps; ++r) {
            long long t0 = clock64();
            for (int c = 0; c < concurrent; ++c) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar[c])), "r"(bytes));
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                                 smem_u32(sm + c * bytes)),
                             "l"(src + (size_t)((r * concurrent + c) % 64) * bytes ), "r"(bytes), "r"(smem_u32(&bar[c]))
                             : "memory");
            }
            for (int c = 0; c < concurrent; ++c) {
                uint32_t ok = 0;
                while (!ok)
                    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0,1,0,p;\n\t}"
                                 : "=r"(ok) : "r"(smem_u32(&bar[c])), "r"(r & 1) : "memory");
            }
            long long dt = clock64() - t0;
            if (r > 2) { sum += dt; if (dt < best) best = dt; }
        }
        out[blockIdx.x * 2] = best;
        out[blockIdx.x * 2 + 1] = sum / (reps - 3);
    }
}
int main() {
    const int bytes_list[3] = {4096, 18432, 36864};
    uint8_t *src; long long *out;
    cudaMalloc(&src, (size_t)148 * 64 * 36864);
    cudaMemset(src, 1, (size_t)148 * 64 * 36864);
    cudaMalloc(&out, 148 * 2 * sizeof(long long));
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 150000);
    for (int grid : {1, 148})
        for (int conc : {1, 2, 4})
            for (int b = 0; b < 3; ++b) {
                if (conc * bytes_list[b] > 150000) continue;
                k<<<grid, 32, conc * bytes_list[b]>>>(src, bytes_list[b], 40, out, conc);
                long long h[296];
                cudaMemcpy(h, out, grid * 2 * sizeof(long long), cudaMemcpyDeviceToHost);
                long long best = 1LL << 60, avg = 0;
                for (int i = 0; i < grid; ++i) { if (h[2 * i] < best) best = h[2 * i]; avg += h[2 * i + 1]; }
                printf("grid %3d  %d concurrent copies of %5d B: best %6lld cycles, mean %6lld cycles  (%.1f B/clk/SM)\n", grid, conc,
                       bytes_list[b], best, avg / grid, (double)conc * bytes_list[b] / (avg / grid));
            }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
