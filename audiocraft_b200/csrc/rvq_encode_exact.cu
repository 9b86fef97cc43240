// rvq_encode_exact.cu -- ResidualVectorQuantization.encode (core_vq.py:387-397), all stages fused,
// every code scored in fp32 on the CUDA cores in the canonical order (rvq_common.cuh).
//
// This is the reference-exact scan: it is the result the tensor-core path (rvq_encode_tc.cu) must
// reproduce bit for bit, the path taken for shapes / inputs the tensor-core search is not
// certified for (D % 64 != 0, non-finite codebooks, extreme scales), and RVQ_ALGO_EXACT.
//
// Work split: a warp owns F frames for all K stages (F chosen by the host from D and the frame count); lane l keeps elements 128*ch + 4*l + {0..3} of
// each frame's residual in registers (never written back).  A block of 8 warps streams the stage's
// codebook through shared memory in tiles of 32 rows, so each row is fetched from L2 once per
// 8*F frames.  Per tile every lane forms its 4*DCH-element partial of 32 inner products per frame;
// a transposed butterfly (31 shuffles per frame instead of 160) leaves code c0+l's inner product
// in lane l, with exactly the bits of the canonical xor butterfly.
#include <limits.h>
#include <math_constants.h>

#include "rvq_common.cuh"

#define RVQ_EX_WARPS 8
#define RVQ_EX_ROWS 32

template <int DCH, int F>
__global__ void __launch_bounds__(RVQ_EX_WARPS * 32)
rvq_encode_exact_kernel(const float *__restrict__ x, int64_t xs_b, int64_t xs_d, int64_t xs_t, int B, int D, int T,
                        const float *__restrict__ embed, const float *__restrict__ cnorm, int C, int K,
                        int64_t *__restrict__ codes, float *__restrict__ quantized, const RvqWire wire) {
    extern __shared__ float4 s_tile[];  // 2 x [RVQ_EX_ROWS][DCH * 32] float4
    const int lane = threadIdx.x & 31;
    const int wib = threadIdx.x >> 5;
    const int64_t N = (int64_t)B * T;
    const int64_t n0 = ((int64_t)blockIdx.x * RVQ_EX_WARPS + wib) * F;

    float r[F][DCH][4], q[F][DCH][4];
#pragma unroll
    for (int f = 0; f < F; ++f) {
        const int64_t n = n0 + f;
        const bool valid = n < N;
        const int64_t b = valid ? n / T : 0, t = valid ? n % T : 0;
#pragma unroll
        for (int ch = 0; ch < DCH; ++ch)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int d = ch * 128 + lane * 4 + i;
                r[f][ch][i] = (valid && d < D) ? x[b * xs_b + (int64_t)d * xs_d + t * xs_t] : 0.0f;
                q[f][ch][i] = 0.0f;  // quantized_out = 0.0 (core_vq.py:364)
            }
    }

    for (int k = 0; k < K; ++k) {
        const float *Ek = embed + (int64_t)k * C * D;
        const float *cnk = cnorm + (int64_t)k * C;
        float xx[F], best[F];
        int bidx[F];
#pragma unroll
        for (int f = 0; f < F; ++f) {
            float p = 0.0f;
#pragma unroll
            for (int ch = 0; ch < DCH; ++ch)
#pragma unroll
                for (int i = 0; i < 4; ++i) p = __fmaf_rn(r[f][ch][i], r[f][ch][i], p);
            xx[f] = rvq_butterfly(p);
            best[f] = -CUDART_INF_F;
            bidx[f] = INT_MAX;
        }

        // codebook tiles of 32 rows, double-buffered: the next tile travels (cp.async, zero-filled past C and D)
        // while this one is scored
        constexpr int TILE4 = RVQ_EX_ROWS * DCH * 32;  // float4 per tile
        auto issue_tile = [&](int c0, int buf) {
            for (int i = threadIdx.x; i < TILE4; i += RVQ_EX_WARPS * 32) {
                const int row = i / (DCH * 32), c4 = i % (DCH * 32);
                const int c = c0 + row, d = c4 * 4;
                const bool ok = c < C && d < D;
                const float *src = ok ? Ek + (int64_t)c * D + d : Ek;
                const uint32_t dst = (uint32_t)__cvta_generic_to_shared(s_tile + buf * TILE4 + i);
                asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(ok ? 16 : 0) : "memory");
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        };
        __syncthreads();  // the previous stage's last tile has been consumed by every warp
        issue_tile(0, 0);
        for (int c0 = 0, tix = 0; c0 < C; c0 += RVQ_EX_ROWS, ++tix) {
            const int buf = tix & 1;
            if (c0 + RVQ_EX_ROWS < C) {
                issue_tile(c0 + RVQ_EX_ROWS, buf ^ 1);
                asm volatile("cp.async.wait_group 1;" ::: "memory");
            } else {
                asm volatile("cp.async.wait_group 0;" ::: "memory");
            }
            __syncthreads();
            const float4 *tile = s_tile + buf * TILE4;

            float p[F][32];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
#pragma unroll
                for (int f = 0; f < F; ++f) p[f][j] = 0.0f;
#pragma unroll
                for (int ch = 0; ch < DCH; ++ch) {
                    const float4 e = tile[j * (DCH * 32) + ch * 32 + lane];
#pragma unroll
                    for (int f = 0; f < F; ++f) {
                        p[f][j] = __fmaf_rn(r[f][ch][0], e.x, p[f][j]);
                        p[f][j] = __fmaf_rn(r[f][ch][1], e.y, p[f][j]);
                        p[f][j] = __fmaf_rn(r[f][ch][2], e.z, p[f][j]);
                        p[f][j] = __fmaf_rn(r[f][ch][3], e.w, p[f][j]);
                    }
                }
            }
            const int col = c0 + lane;
            const float cc = col < C ? cnk[col] : 0.0f;
#pragma unroll
            for (int f = 0; f < F; ++f) {
                rvq_transposed_butterfly(p[f], lane);
                if (col < C) {
                    const float s = rvq_score(xx[f], p[f][0], cc);
                    if (rvq_better(s, col, best[f], bidx[f])) { best[f] = s; bidx[f] = col; }
                }
            }
            __syncthreads();  // this buffer is refilled by the next iteration's prefetch
        }

        // arg-max across lanes, residual update, quantized accumulation
#pragma unroll
        for (int f = 0; f < F; ++f) {
            float bs = best[f];
            int bi = bidx[f];
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) {
                float os = __shfl_xor_sync(0xffffffffu, bs, off);
                int oi = __shfl_xor_sync(0xffffffffu, bi, off);
                if (rvq_better(os, oi, bs, bi)) { bs = os; bi = oi; }
            }
            const int64_t n = n0 + f;
            if (n < N) {
                if (lane == 0) {
                    codes[(int64_t)k * N + n] = (int64_t)bi;  // [K][B][T]
                    if (wire.p) wire.p[(int64_t)k * wire.sk + (n / T) * wire.sb + (n % T)] = (int16_t)bi;
                }
#pragma unroll
                for (int ch = 0; ch < DCH; ++ch) {
                    const int d = ch * 128 + lane * 4;
                    if (d < D) {
                        const float4 e = __ldg(reinterpret_cast<const float4 *>(Ek + (int64_t)bi * D + d));
                        // residual = residual - quantized (core_vq.py:394); quantized_out += quantized (:376)
                        r[f][ch][0] = __fsub_rn(r[f][ch][0], e.x); q[f][ch][0] = __fadd_rn(q[f][ch][0], e.x);
                        r[f][ch][1] = __fsub_rn(r[f][ch][1], e.y); q[f][ch][1] = __fadd_rn(q[f][ch][1], e.y);
                        r[f][ch][2] = __fsub_rn(r[f][ch][2], e.z); q[f][ch][2] = __fadd_rn(q[f][ch][2], e.z);
                        r[f][ch][3] = __fsub_rn(r[f][ch][3], e.w); q[f][ch][3] = __fadd_rn(q[f][ch][3], e.w);
                    }
                }
            }
        }
    }

    if (quantized) {
#pragma unroll
        for (int f = 0; f < F; ++f) {
            const int64_t n = n0 + f;
            if (n >= N) continue;
#pragma unroll
            for (int ch = 0; ch < DCH; ++ch) {
                const int d = ch * 128 + lane * 4;
                if (d < D)
                    __stcs(reinterpret_cast<float4 *>(quantized + n * D + d),
                           make_float4(q[f][ch][0], q[f][ch][1], q[f][ch][2], q[f][ch][3]));
            }
        }
    }
}

template <int DCH, int F>
static int launch_exact(const float *x, const int64_t xs[3], int B, int D, int T, const RvqPackView &pv, int K,
                        int64_t *codes, float *quantized, const RvqWire &wire, cudaStream_t stream) {
    const int64_t N = (int64_t)B * T;
    const int64_t per_block = (int64_t)RVQ_EX_WARPS * F;
    const int64_t blocks = (N + per_block - 1) / per_block;
    const size_t smem = (size_t)2 * RVQ_EX_ROWS * DCH * 32 * sizeof(float4);
    auto kern = rvq_encode_exact_kernel<DCH, F>;
    if (smem > 48 * 1024) RVQ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<(unsigned)blocks, RVQ_EX_WARPS * 32, smem, stream>>>(x, xs[0], xs[1], xs[2], B, D, T, pv.embed, pv.cnorm,
                                                                pv.C, K, codes, quantized, wire);
    rvq_count_launch(1);
    return rvq_check_launch("rvq_encode_exact_kernel");
}

int rvq_encode_exact(const float *x, const int64_t xs[3], int B, int D, int T, const RvqPackView &pv, int K,
                     int64_t *codes, float *quantized, const RvqWire &wire, cudaStream_t stream) {
    RVQ_REQUIRE(D <= 512, RVQ_EUNSUPPORTED, "rvq_encode_f32: dimension %d > 512 is not implemented", D);
    RVQ_REQUIRE((int64_t)B * T <= (int64_t)INT_MAX * 8, RVQ_EUNSUPPORTED, "rvq_encode_f32: too many frames");
    // Frames per warp: every shared-memory read of a codebook row feeds 4 F multiply-adds, so more frames per warp
    // take the kernel off the shared-memory pipe (D = 256: 4.0 -> 2.5 ms on 24 000 frames) -- when there are enough
    // frames to keep every SM's eight warps busy with that many each.
    const int64_t N = (int64_t)B * T;
    const bool many = N >= (int64_t)8 * 148 * 8 * 2, some = N >= (int64_t)4 * 148 * 8 * 2;
#define RVQ_EX_GO(DCH, F) return launch_exact<DCH, F>(x, xs, B, D, T, pv, K, codes, quantized, wire, stream)
    if (D <= 128) { if (many) RVQ_EX_GO(1, 8); RVQ_EX_GO(1, 2); }
    if (D <= 256) { if (some) RVQ_EX_GO(2, 4); RVQ_EX_GO(2, 1); }
    if (D <= 384) { if (some) RVQ_EX_GO(3, 2); RVQ_EX_GO(3, 1); }
    if (some) RVQ_EX_GO(4, 2);
    RVQ_EX_GO(4, 1);
#undef RVQ_EX_GO
}
