// Microbenchmark: tcgen05.ld (LDTM) throughput per SM.  W warps per SM (1, 2 or 4 per sub-partition) each load
// 32 lanes x 64 columns (8 KB) ITER times from their TMEM lane quarter; nothing else runs.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/micro/ldtm_bw tools/micro/ldtm_bw.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void ld64(uint32_t taddr, uint32_t (&v)[64]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "
        "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, "
        "%48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
          "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
          "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
          "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31]), "=r"(v[32]),
          "=r"(v[33]), "=r"(v[34]), "=r"(v[35]), "=r"(v[36]), "=r"(v[37]), "=r"(v[38]), "=r"(v[39]), "=r"(v[40]),
          "=r"(v[41]), "=r"(v[42]), "=r"(v[43]), "=r"(v[44]), "=r"(v[45]), "=r"(v[46]), "=r"(v[47]), "=r"(v[48]),
          "=r"(v[49]), "=r"(v[50]), "=r"(v[51]), "=r"(v[52]), "=r"(v[53]), "=r"(v[54]), "=r"(v[55]), "=r"(v[56]),
          "=r"(v[57]), "=r"(v[58]), "=r"(v[59]), "=r"(v[60]), "=r"(v[61]), "=r"(v[62]), "=r"(v[63])
        : "r"(taddr)
        : "memory");
}

__device__ __forceinline__ uint64_t desc_sw128(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) |
           ((uint64_t)2 << 61);
}

// mma != 0: warp 15 keeps the tensor pipe busy (M=128, N=128, K=16 fp16 MMAs into TMEM columns 256..383, operands
// = 32 KB of zeroed shared memory) until the loading warps are done.
template <int ITER>
__global__ void __launch_bounds__(512, 1) k(long long *out, uint32_t *sink, int warps, int mma, int mma_n) {
    __shared__ uint32_t slot;
    __shared__ volatile int done;
    extern __shared__ __align__(1024) unsigned char opnd[];
    for (int i = threadIdx.x; i < 32768 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(opnd)[i] = 0u;
    if (threadIdx.x == 0) done = 0;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = slot;
    long long t0 = 0, t1 = 0;
    uint32_t acc = 0;
    if (warp < warps) {
        const uint32_t tq = tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)((warp >> 2) * 64);
        uint32_t a[64], b[64];
        t0 = clock64();
        for (int i = 0; i < ITER; i += 2) {
            ld64(tq, a);
            ld64(tq + 256, b);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int j = 0; j < 64; ++j) acc += a[j] ^ b[j];
        }
        t1 = clock64();
    }
    if (warp < warps && lane == 0) atomicAdd((int *)&done, 1);
    if (mma && (warp == 15 || (mma == 3 && warp == 14))) {
        const uint32_t base = ((uint32_t)__cvta_generic_to_shared(opnd) + 1023u) & ~1023u;
        const uint64_t ad = desc_sw128(base), bd = desc_sw128(base + 16384);
        const uint32_t idesc = (1u << 4) | (((uint32_t)mma_n >> 3) << 17) | ((128u >> 4) << 24);
        long long n = 0;
        if (lane == 0) {
            while (done < warps) {
                for (int i = 0; i < 8; ++i) {
                    uint32_t dcol = 256u;
                    if (mma == 2) dcol = (i & 1) ? 384u : 256u;   // one lane, two accumulators in turn
                    if (mma == 3) dcol = warp == 15 ? 384u : 256u;  // two warps, one accumulator each
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem + dcol),
                                 "l"(ad + (uint64_t)(i & 3) * 2), "l"(bd + (uint64_t)(i & 3) * 2), "r"(idesc), "r"(1u)
                                 : "memory");
                }
                n += 8;
            }
            out[blockIdx.x * 16 + warp] = n;
        }
        __syncwarp();
    }
    if (lane == 0 && warp < warps) out[blockIdx.x * 16 + warp] = t1 - t0;
    if (acc == 0x12345678u) sink[0] = acc;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
}

int main() {
    long long *out;
    uint32_t *sink;
    cudaMalloc(&out, 148 * 16 * 8);
    cudaMalloc(&sink, 4);
    constexpr int ITER = 2048;
    cudaFuncSetAttribute(k<ITER>, cudaFuncAttributeMaxDynamicSharedMemorySize, 66560);
    for (int mma_n : {128})
    for (int mma : {1, 2, 3})
    for (int warps : {4}) {
        for (int rep = 0; rep < 2; ++rep) {
            cudaMemset(out, 0, 148 * 16 * 8);
            k<ITER><<<148, 512, 66560>>>(out, sink, warps, mma, mma_n);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
        }
        long long h[16];
        cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
        long long mx = 0;
        for (int w = 0; w < warps; ++w) mx = h[w] > mx ? h[w] : mx;
        const double bytes = (double)warps * ITER * 8192.0;
        if (mma) printf("mode %d (1: one lane, one accumulator; 2: one lane, two accumulators in turn; 3: two warps, one accumulator each): %lld MMAs M=128 N=%d K=16, %.1f cycles each: ", mma, h[15] + (mma == 3 ? h[14] : 0), mma_n, (double)mx / (h[15] + (mma == 3 ? h[14] : 0)));
        printf("%2d warps x %d loads of 32 lanes x 64 columns: %lld cycles -> %.1f B/clk/SM (%.1f cycles per 8 KB load per warp)\n", warps,
               ITER, mx, bytes / mx, (double)mx / ITER);
    }
    return 0;
}
