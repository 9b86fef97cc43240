// rvq_pack.cu -- error plumbing + the codebook pack (derived, private copy of the `embed` buffers).
//
// Replaces, once per codebook change instead of once per call:
//   embed.pow(2).sum(0, keepdim=True)            core_vq.py:169
// and prepares the operand copies of the tensor-core search (rvq_encode_tc.cu).
#include <stdarg.h>
#include <string.h>

#include "rvq_common.cuh"

// ---------------------------------------------------------------- host-side plumbing
static thread_local char g_err[512] = "";
static thread_local int64_t g_launches = 0;

void rvq_set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

void rvq_count_launch(int n) { g_launches += n; }

int rvq_check_launch(const char *what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        rvq_set_error("launch of %s failed: %s", what, cudaGetErrorString(e));
        return RVQ_ECUDA;
    }
    return RVQ_OK;
}

extern "C" int rvq_b200_abi_version(void) { return RVQ_B200_ABI_VERSION; }
extern "C" const char *rvq_b200_last_error(void) { return g_err; }
extern "C" int64_t rvq_b200_launch_count(int reset) {
    int64_t v = g_launches;
    if (reset) g_launches = 0;
    return v;
}

extern "C" int64_t rvq_pack_bytes(int Ktot, int C, int D) {
    if (Ktot <= 0 || Ktot > RVQ_MAX_STAGES || C <= 0 || D <= 0) return RVQ_EINVAL;
    return rvq_pack_layout(Ktot, C, D).total_bytes;
}

// ---------------------------------------------------------------- kernels
// One warp per code: ||E_c||^2 in canonical order, plus per-stage maxima / non-finite flag.
__global__ void __launch_bounds__(256) rvq_cnorm_kernel(const float *__restrict__ embed, int Ktot, int C,
                                                        int D, float *__restrict__ cnorm,
                                                        RvqStageMeta *__restrict__ meta) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t total = (int64_t)Ktot * C;
    if (warp >= total) return;
    const float *row = embed + warp * D;
    float p = 0.0f, amax = 0.0f;
    int bad = 0;
    for (int base = 0; base < D; base += 128) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int d = base + 4 * lane + i;
            if (d < D) {
                float v = row[d];
                p = __fmaf_rn(v, v, p);
                float a = fabsf(v);
                if (a <= 3.0e38f) amax = fmaxf(amax, a); else bad = 1;  // NaN or Inf -> bad
            }
        }
    }
    p = rvq_butterfly(p);
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, off));
        bad |= __shfl_xor_sync(0xffffffffu, bad, off);
    }
    if (lane == 0) {
        cnorm[warp] = p;
        RvqStageMeta *m = meta + warp / C;
        // non-negative floats order like their bit patterns
        atomicMax((int *)&m->cmax_abs, __float_as_int(amax));
        if (p <= 3.0e38f) atomicMax((int *)&m->cnorm2_max, __float_as_int(p)); else bad = 1;
        if (bad) atomicOr(&m->nonfinite, 1);
    }
}

// scale_exp: cmax_abs * 2^scale_exp in [2^14, 2^15)  (0 for an all-zero codebook)
__global__ void rvq_scale_kernel(RvqStageMeta *__restrict__ meta, int Ktot) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= Ktot) return;
    float a = meta[k].cmax_abs;
    int e = 0;
    if (a > 0.0f) {
        int ex;
        frexpf(a, &ex);  // a = m * 2^ex, m in [0.5, 1) -> a in [2^(ex-1), 2^ex)
        e = 15 - ex;
    }
    if (e > 60 || e < -60) {  // outside the range the fp16 search is certified for
        atomicOr(&meta[k].nonfinite, 2);
        e = max(-60, min(60, e));
    }
    meta[k].scale_exp = e;
}

__global__ void __launch_bounds__(256) rvq_f16_kernel(const float *__restrict__ embed, int64_t per_stage,
                                                      int64_t total, const RvqStageMeta *__restrict__ meta,
                                                      __half *__restrict__ out) {
    int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i >= total) return;
    int k = (int)(i / per_stage);
    float s = ldexpf(1.0f, meta[k].scale_exp);
    float4 v = *reinterpret_cast<const float4 *>(embed + i);
    __half2 lo = __floats2half2_rn(v.x * s, v.y * s);
    __half2 hi = __floats2half2_rn(v.z * s, v.w * s);
    uint2 o;
    o.x = *reinterpret_cast<uint32_t *>(&lo);
    o.y = *reinterpret_cast<uint32_t *>(&hi);
    *reinterpret_cast<uint2 *>(out + i) = o;
}

// ---------------------------------------------------------------- entry point
extern "C" int rvq_pack_codebooks(const float *const *embed_ptrs, int Ktot, int C, int D, void *pack,
                                  int64_t pack_bytes, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    RVQ_REQUIRE(embed_ptrs && pack, RVQ_EINVAL, "rvq_pack_codebooks: null pointer");
    RVQ_REQUIRE(Ktot > 0 && Ktot <= RVQ_MAX_STAGES && C > 0 && D > 0, RVQ_EINVAL,
                "rvq_pack_codebooks: bad sizes Ktot=%d C=%d D=%d", Ktot, C, D);
    RVQ_REQUIRE(D % 4 == 0, RVQ_EUNSUPPORTED, "rvq_pack_codebooks: dimension %d is not a multiple of 4", D);
    RVQ_REQUIRE(((uintptr_t)pack & 255) == 0, RVQ_EINVAL, "rvq_pack_codebooks: pack must be 256-byte aligned");
    RvqPackHeader h = rvq_pack_layout(Ktot, C, D);
    RVQ_REQUIRE(pack_bytes >= h.total_bytes, RVQ_EINVAL, "rvq_pack_codebooks: pack too small (%lld < %lld)",
                (long long)pack_bytes, (long long)h.total_bytes);
    char *p = (char *)pack;
    RVQ_CUDA(cudaMemsetAsync(p, 0, 256 + rvq_align_up((int64_t)sizeof(RvqStageMeta) * Ktot, 256), stream));
    RVQ_CUDA(cudaMemcpyAsync(p, &h, sizeof(h), cudaMemcpyHostToDevice, stream));
    float *embed = (float *)(p + h.off_embed);
    const int64_t per_stage = (int64_t)C * D;
    for (int k = 0; k < Ktot; ++k) {
        RVQ_REQUIRE(embed_ptrs[k], RVQ_EINVAL, "rvq_pack_codebooks: embed_ptrs[%d] is null", k);
        RVQ_CUDA(cudaMemcpyAsync(embed + k * per_stage, embed_ptrs[k], per_stage * 4, cudaMemcpyDeviceToDevice,
                                 stream));
    }
    RvqStageMeta *meta = (RvqStageMeta *)(p + h.off_meta);
    const int64_t warps = (int64_t)Ktot * C;
    rvq_cnorm_kernel<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, stream>>>(
        embed, Ktot, C, D, (float *)(p + h.off_cnorm), meta);
    rvq_scale_kernel<<<1, 64, 0, stream>>>(meta, Ktot);
    const int64_t total = per_stage * Ktot;
    rvq_f16_kernel<<<(unsigned)((total / 4 + 255) / 256), 256, 0, stream>>>(embed, per_stage, total, meta,
                                                                           (__half *)(p + h.off_f16));
    rvq_count_launch(3);
    return rvq_check_launch("rvq_pack kernels");
}
