// rvq_common.cuh -- shared definitions of librvq_b200.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/rvq_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "librvq_b200 is written for sm_100a (B200) only"
#endif

#define RVQ_PACK_MAGIC 0x52565142u /* "RVQB" */

// Tensor-core operand image of one 128-code tile of a codebook (D == 128), laid out exactly as the
// UMMA shared-memory descriptors of rvq_encode_tc.cu read it, so that one 1-D bulk copy brings a tile in:
//   [0, 16384)      dims   0..63 : 128 rows x 128 B, 16-byte chunks XOR-swizzled with (row & 7)  (SWIZZLE_128B, K-major)
//   [16384, 32768)  dims 64..127 : same
//   [32768, 36864)  16 "augmented" K columns: 8x16-byte core matrices, no swizzle
//                   (row r, chunk c) at (r / 8) * 256 + c * 128 + (r % 8) * 16
// Values are fp16(E[c][d] * 2^scale_exp); augmented columns 0..2 hold the three-way fp16 split of
// -(||E_c||^2 / 2) * 2^aug_exp, the rest are zero.
#define RVQ_TC_TILE_CODES 128
#define RVQ_TC_IMG_BYTES 36864

// Per-stage metadata written by rvq_pack_codebooks.
struct RvqStageMeta {
    float cmax_abs;     // max |E[c][d]| over the codebook (finite part)
    float cnorm2_max;   // max ||E_c||^2 (canonical order)
    int32_t scale_exp;  // cmax_abs * 2^scale_exp in [2^13, 2^14)
    int32_t aug_exp;    // (cnorm2_max / 2) * 2^aug_exp in [2^13, 2^14)
    int32_t flags;      // bit 0: NaN/Inf in the codebook; bit 1: scale outside the certified range
    int32_t pad[3];
};

// First 256 bytes of a pack, followed by the sections at the byte offsets below.
struct RvqPackHeader {
    uint32_t magic;
    int32_t Ktot, C, D;
    int64_t off_meta;   // RvqStageMeta[Ktot]
    int64_t off_embed;  // float [Ktot][C][D]
    int64_t off_cnorm;  // float [Ktot][C]
    int64_t off_f16;    // tensor-core images [Ktot][C / 128][RVQ_TC_IMG_BYTES]; empty unless D == 128 and C % 128 == 0
    int64_t total_bytes;
};

static inline __host__ __device__ int64_t rvq_align_up(int64_t v, int64_t a) { return (v + a - 1) / a * a; }

static inline __host__ RvqPackHeader rvq_pack_layout(int Ktot, int C, int D) {
    RvqPackHeader h;
    h.magic = RVQ_PACK_MAGIC;
    h.Ktot = Ktot; h.C = C; h.D = D;
    int64_t off = 256;
    h.off_meta = off;  off = rvq_align_up(off + (int64_t)sizeof(RvqStageMeta) * Ktot, 256);
    h.off_embed = off; off = rvq_align_up(off + (int64_t)4 * Ktot * C * D, 256);
    h.off_cnorm = off; off = rvq_align_up(off + (int64_t)4 * Ktot * C, 256);
    off = rvq_align_up(off, 1024);
    h.off_f16 = off;
    if (D == 128 && C % RVQ_TC_TILE_CODES == 0) off += (int64_t)Ktot * (C / RVQ_TC_TILE_CODES) * RVQ_TC_IMG_BYTES;
    off = rvq_align_up(off, 1024);
    h.total_bytes = off;
    return h;
}

// Views into a pack (device pointers; computed on the host from the layout).
struct RvqPackView {
    const RvqStageMeta *meta;
    const float *embed;
    const float *cnorm;
    const uint8_t *img;  // tensor-core images (see RVQ_TC_IMG_BYTES)
    int Ktot, C, D;
};

static inline __host__ RvqPackView rvq_pack_view(const void *pack, int Ktot, int C, int D) {
    RvqPackHeader h = rvq_pack_layout(Ktot, C, D);
    const char *p = (const char *)pack;
    RvqPackView v;
    v.meta = (const RvqStageMeta *)(p + h.off_meta);
    v.embed = (const float *)(p + h.off_embed);
    v.cnorm = (const float *)(p + h.off_cnorm);
    v.img = (const uint8_t *)(p + h.off_f16);
    v.Ktot = Ktot; v.C = C; v.D = D;
    return v;
}

// Optional second output of the encode kernels: the codes as int16 in the layout a frame-sharded job puts on the
// wire (element (k, b, t) at p[k * sk + b * sb + t]; p == nullptr: not wanted).
struct RvqWire {
    int16_t *p;
    int64_t sk, sb;
};

// ---- error plumbing (host) ----
void rvq_set_error(const char *fmt, ...);
void rvq_count_launch(int n);
int rvq_check_launch(const char *what);  // cudaGetLastError -> RVQ_ECUDA

#define RVQ_REQUIRE(cond, code, ...)            \
    do {                                        \
        if (!(cond)) {                          \
            rvq_set_error(__VA_ARGS__);         \
            return (code);                      \
        }                                       \
    } while (0)

#define RVQ_CUDA(call)                                                                   \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess) {                                                         \
            rvq_set_error("%s failed: %s", #call, cudaGetErrorString(e_));               \
            return RVQ_ECUDA;                                                            \
        }                                                                                \
    } while (0)

// ---- canonical arithmetic (device) ----
// Canonical inner product (defined in DESIGN.md, "canonical order"): lane l of a warp owns elements
// 128*chunk + 4*l + {0,1,2,3}; partials advance with one fused multiply-add per element in
// increasing d; lanes are combined with the xor butterfly 16, 8, 4, 2, 1.  Every lane ends up with
// the same bits (fp32 add is commutative), so any lane's copy is "the" result.
__device__ __forceinline__ float rvq_butterfly(float p) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) p = __fadd_rn(p, __shfl_xor_sync(0xffffffffu, p, off));
    return p;
}

// The same sums for 32 columns at once: in place, p[0..31] (this lane's partials of 32 columns) -> p[0] = the
// canonical sum of column `lane` (31 shuffles instead of 160; the additions and their order are those of rvq_butterfly).
__device__ __forceinline__ void rvq_transposed_butterfly(float (&p)[32], int lane) {
#pragma unroll
    for (int half = 16; half >= 1; half >>= 1) {
        const bool up = (lane & half) != 0;
#pragma unroll
        for (int i = 0; i < half; ++i) {
            float send = up ? p[i] : p[i + half];
            float keep = up ? p[i + half] : p[i];
            p[i] = __fadd_rn(keep, __shfl_xor_sync(0xffffffffu, send, half));
        }
    }
}

// torch.max semantics (core_vq.py:171): NaN beats everything, first index wins ties.
__device__ __forceinline__ bool rvq_better(float s, int i, float best, int bi) {
    bool sn = s != s, bn = best != best;
    if (sn || bn) return (sn && !bn) || (sn && bn && i < bi);
    return (s > best) || (s == best && i < bi);
}

// score of a code: -((xx - 2*dot) + cc), one rounding per operation (core_vq.py:166-170)
__device__ __forceinline__ float rvq_score(float xx, float dot, float cc) {
    float two_dot = __fmul_rn(2.0f, dot);
    return -__fadd_rn(__fsub_rn(xx, two_dot), cc);
}
