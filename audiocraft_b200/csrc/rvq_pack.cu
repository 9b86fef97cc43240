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
        if (bad) atomicOr(&m->flags, 1);
    }
}

// scale_exp / aug_exp: put the largest operand of each kind at [2^13, 2^14) in fp16
__global__ void rvq_scale_kernel(RvqStageMeta *__restrict__ meta, int Ktot) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= Ktot) return;
    int flags = meta[k].flags;
    int e = 0, ea = 0, ex;
    float a = meta[k].cmax_abs;
    if (a > 0.0f) { frexpf(a, &ex); e = 14 - ex; }   // a in [2^(ex-1), 2^ex)
    float h = 0.5f * meta[k].cnorm2_max;
    if (h > 0.0f) { frexpf(h, &ex); ea = 14 - ex; }
    if (e > 40 || e < -40 || ea > 40 || ea < -40) {  // outside the range the fp16 search is certified for
        flags |= 2;
        e = max(-40, min(40, e));
        ea = max(-40, min(40, ea));
    }
    meta[k].scale_exp = e;
    meta[k].aug_exp = ea;
    meta[k].flags = flags;
}

// One thread per 16-byte chunk of the tensor-core image: 18 chunks per code (16 data + 2 augmented).
__global__ void __launch_bounds__(256) rvq_image_kernel(const float *__restrict__ embed, const float *__restrict__ cnorm,
                                                        int Ktot, int C, const RvqStageMeta *__restrict__ meta,
                                                        uint8_t *__restrict__ img) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t total = (int64_t)Ktot * C * 18;
    if (i >= total) return;
    const int chunk = (int)(i % 18);
    const int64_t kc = i / 18;
    const int k = (int)(kc / C), c = (int)(kc % C);
    const int tile = c / RVQ_TC_TILE_CODES, r = c % RVQ_TC_TILE_CODES;
    uint8_t *base = img + ((int64_t)k * (C / RVQ_TC_TILE_CODES) + tile) * RVQ_TC_IMG_BYTES;
    __half v[8];
    size_t off;
    if (chunk < 16) {
        const float s = ldexpf(1.0f, meta[k].scale_exp);
        const float *src = embed + kc * 128 + chunk * 8;
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = __float2half_rn(src[j] * s);
        const int kb = chunk >> 3, cidx = chunk & 7;
        off = (size_t)kb * 16384 + (size_t)r * 128 + (size_t)((cidx ^ (r & 7)) << 4);
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = __float2half_rn(0.0f);
        if (chunk == 16) {
            const float t = -(0.5f * cnorm[kc]) * ldexpf(1.0f, meta[k].aug_exp);
            const __half hi = __float2half_rn(t);
            const float r1 = t - __half2float(hi);
            const __half mid = __float2half_rn(r1);
            const __half lo = __float2half_rn(r1 - __half2float(mid));
            v[0] = hi; v[1] = mid; v[2] = lo;
        }
        off = 32768 + (size_t)(r >> 3) * 256 + (size_t)(chunk - 16) * 128 + (size_t)(r & 7) * 16;
    }
    *reinterpret_cast<uint4 *>(base + off) = *reinterpret_cast<const uint4 *>(v);
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
    int launches = 2;
    if (D == 128 && C % RVQ_TC_TILE_CODES == 0) {
        const int64_t total = (int64_t)Ktot * C * 18;
        rvq_image_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(
            embed, (const float *)(p + h.off_cnorm), Ktot, C, meta, (uint8_t *)(p + h.off_f16));
        ++launches;
    }
    rvq_count_launch(launches);
    return rvq_check_launch("rvq_pack kernels");
}
