// rvq_encode_tc.cu -- ResidualVectorQuantization.encode (core_vq.py:387-397) with the nearest-code
// search of EuclideanCodebook.quantize (core_vq.py:164-172) on the 5th-generation tensor cores.
//
// One persistent CTA per SM, all K stages fused, D == 128.
//
//   search (approximate)   acc[frame][code] ~= a * ( r.E_c - ||E_c||^2 / 2 ),  a > 0 per frame:
//       tcgen05.mma kind::f16, M = 128 frames x N = 128 codes x K = 144 (128 dims + 16 augmented
//       columns that carry -||E_c||^2/2 as a three-way fp16 split), fp16 operands scaled by powers
//       of two, fp32 accumulators in TMEM.  Codebook tiles arrive by 1-D bulk copies (UBLKCP) of a
//       pre-swizzled image (rvq_common.cuh) into a 3-deep ring; the residual operand is written to
//       shared memory by the warps that own the residual.
//   scan                   each of 128 threads owns one TMEM lane (= one frame) and reduces the 128
//       columns of an accumulator in groups of 16 codes.  A group is remembered when its maximum is
//       within 2*delta of the running maximum, delta being a rigorous bound on |acc - exact|.
//   resolve (exact)        the codes of the remembered groups whose approximate score is within
//       2*delta of the final maximum are the only possible winners.  One such code (9 frames in 10):
//       done.  Two to four: re-scored in fp32 in the canonical order (rvq_common.cuh), eight codes per
//       batch.  More, or a frame the bound cannot cover (non-finite values, extreme scales, list
//       overflow): whole groups or the whole codebook re-scored in place.  The emitted code is the
//       one the exact scan (rvq_encode_exact.cu) emits, bit for bit.
//   then the residual update r -= E_k[code] through the L2-resident scratch, and the fp16 operand
//   row of the next stage.  With a quantized output requested (eval forward), the tile's frames are
//   summed from their K winner rows in stage order after the last stage (template flag QOUT).
//
// A CTA runs two 128-frame tiles ("slots") in lock step: both are at the same codebook tile at the
// same time, so one copy of the tile serves 256 frames.  Warp roles (12 warps):
//   warp 0      producer: streams the codebook tiles of the CTA, in order, through the ring
//   warp 1, 2   tcgen05.mma issuers of slot 0 and slot 1 (warp 1 owns the TMEM allocation: 512 columns
//               = 2 slots x 2 accumulator buffers x 128)
//   warp 3      idle (setmaxnreg works on whole warpgroups)
//   warps 4-7   slot 0: residual owners / scanners / resolvers (warp w owns TMEM lanes 32*(w%4)..)
//   warps 8-11  slot 1
// profiles/r2_notes.md has the measurements behind this shape and the variants that lost.
#include <limits.h>
#include <math_constants.h>

#include <mutex>
#include <type_traits>

#include "rvq_common.cuh"

namespace {

constexpr int TILE_M = 128;
constexpr int NSLOT = 2;
constexpr int RING = 3;
constexpr int IMG = RVQ_TC_IMG_BYTES;  // 36864: kb0 | kb1 | aug
constexpr int LISTCAP = 16;
constexpr int THREADS = 384;

// shared memory map (offsets from a 1024-byte aligned base)
constexpr int SM_RING = 0;
constexpr int SM_A = SM_RING + RING * IMG;                 // 110592
constexpr int SM_LIST = SM_A + NSLOT * IMG;                // 184320
constexpr int SM_BAR = SM_LIST + NSLOT * LISTCAP * TILE_M * 4;  // 200704
constexpr int SM_TOTAL = SM_BAR + 256;
constexpr int SM_ALLOC = SM_TOTAL + 1024;                  // slack for manual alignment

// barrier indices (8 bytes each)
constexpr int BAR_BFULL = 0;                    // [RING]
constexpr int BAR_BEMPTY = BAR_BFULL + RING;    // [RING]
constexpr int BAR_ACCFULL = BAR_BEMPTY + RING;  // [NSLOT][2]
constexpr int BAR_ACCEMPTY = BAR_ACCFULL + 4;   // [NSLOT][2]
constexpr int BAR_AREADY = BAR_ACCEMPTY + 4;    // [NSLOT]
constexpr int BAR_COUNT = BAR_AREADY + NSLOT;   // 16
constexpr int SM_META = SM_BAR + BAR_COUNT * 8; // ring_meta[RING] (u32), tmem base (u32)

// instruction descriptor: kind::f16, A = B = F16, D = F32, both K-major, M = 128, N = 128
constexpr uint32_t IDESC = (1u << 4) | ((128u >> 3) << 17) | ((128u >> 4) << 24);

// ------------------------------------------------------------------------------------ PTX
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_test(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    while (!mbar_test(bar, parity)) {}
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
// one lane of a converged warp (the compiler then predicates tcgen05 / bulk-copy instructions directly
// instead of wrapping them in a per-thread serialisation loop, as it does under `if (lane == 0)`)
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "elect.sync _|p, 0xffffffff;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// descriptors as (low, high) words: only the low word (start address field) changes from MMA to MMA, so the
// caller's arithmetic stays 32-bit
__device__ __forceinline__ void umma_f16_lh(uint32_t d_tmem, uint32_t alo, uint32_t ahi, uint32_t blo, uint32_t bhi,
                                            uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
        "setp.ne.b32 p, %6, 0;\n\t"
        "mov.b64 da, {%1, %2};\n\t"
        "mov.b64 db, {%3, %4};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}" ::"r"(d_tmem),
        "r"(alo), "r"(ahi), "r"(blo), "r"(bhi), "r"(idesc), "r"(acc)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld64(uint32_t taddr, uint32_t (&v)[64]) {
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
// Makes a loop-invariant value opaque to the compiler, so that it stays in a register instead of being
// recomputed (special-register reads, 64-bit pointer arithmetic) at every use.
__device__ __forceinline__ uint32_t pin_u32(uint32_t v) {
    asm volatile("mov.b32 %0, %0;" : "+r"(v));
    return v;
}
template <typename T>
__device__ __forceinline__ T *pin_ptr(T *p) {
    asm volatile("mov.b64 %0, %0;" : "+l"(p));
    return p;
}
__device__ __forceinline__ float max3(float a, float b, float c) {
    float r;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}
__device__ __forceinline__ float group_max8(const uint32_t *v) {
    float m = max3(__uint_as_float(v[0]), __uint_as_float(v[1]), __uint_as_float(v[2]));
    m = max3(m, __uint_as_float(v[3]), __uint_as_float(v[4]));
    m = max3(m, __uint_as_float(v[5]), __uint_as_float(v[6]));
    return fmaxf(m, __uint_as_float(v[7]));
}
// Predicated "remember this group": its 16 approximate scores to the L2 scratch (two 256-bit stores: every global
// store of this kernel costs two extra instructions that move its memory descriptor into uniform registers, so
// fewer and wider stores matter), its packed list word to shared memory.  Straight-line code: a branch per group
// costs far more than the predicated-off instructions.
__device__ __forceinline__ void remember_store(bool on, float *gdst, const uint32_t *v, uint32_t sdst, uint32_t word) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %0, 0;\n\t"
        "@p st.global.v8.b32 [%1], {%2, %3, %4, %5, %6, %7, %8, %9};\n\t"
        "@p st.global.v8.b32 [%1+32], {%10, %11, %12, %13, %14, %15, %16, %17};\n\t"
        "@p st.shared.b32 [%18], %19;\n\t}"
        ::"r"((int)on), "l"(gdst), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
        "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(sdst), "r"(word)
        : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major SWIZZLE_128B operand: rows of 128 bytes, 8-row groups 1024 bytes apart (SBO), version 1
__device__ __forceinline__ uint64_t desc_sw128(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) |
           ((uint64_t)2 << 61);
}
// K-major, no swizzle: 8x16-byte core matrices; the two K chunks 128 bytes apart (LBO), 8-row groups 256 bytes apart (SBO)
__device__ __forceinline__ uint64_t desc_interleaved(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(128 >> 4) << 16) | ((uint64_t)(256 >> 4) << 32) |
           ((uint64_t)1 << 46);
}

// ------------------------------------------------------------------------------------ exact helpers
// 8 columns: after the three halving rounds lane l holds column (l >> 2) & 7; two plain rounds finish it.
__device__ __forceinline__ float transposed_butterfly8(float (&p)[8], int lane) {
#pragma unroll
    for (int half = 4; half >= 1; half >>= 1) {
        const int off = half * 4;  // lane distance 16, 8, 4
        const bool up = (lane & off) != 0;
#pragma unroll
        for (int i = 0; i < half; ++i) {
            float send = up ? p[i] : p[i + half];
            float keep = up ? p[i + half] : p[i];
            p[i] = __fadd_rn(keep, __shfl_xor_sync(0xffffffffu, send, off));
        }
    }
    float v = p[0];
    v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, 2));
    v = __fadd_rn(v, __shfl_xor_sync(0xffffffffu, v, 1));
    return v;
}

// The canonical combination of 32 lane partials (rvq_common.cuh: xor butterfly 16, 8, 4, 2, 1) when one thread
// holds all of them: p[i] += p[i + half].  Same additions, same order, same bits.
__device__ __forceinline__ float tree32(float (&p)[32]) {
#pragma unroll
    for (int half = 16; half >= 1; half >>= 1) {
#pragma unroll
        for (int i = 0; i < half; ++i) p[i] = __fadd_rn(p[i], p[i + half]);
    }
    return p[0];
}

__device__ __forceinline__ void warp_argmax(float &bs, int &bi) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        float os = __shfl_xor_sync(0xffffffffu, bs, off);
        int oi = __shfl_xor_sync(0xffffffffu, bi, off);
        if (rvq_better(os, oi, bs, bi)) { bs = os; bi = oi; }
    }
}

// canonical inner products of one frame with the 8 codes of a group; lane l gets column (l >> 2) & 7
__device__ __forceinline__ float group_dots(const float4 &rf, const float4 (&rows)[8], int lane) {
    float p[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        float a = __fmaf_rn(rf.x, rows[j].x, 0.0f);
        a = __fmaf_rn(rf.y, rows[j].y, a);
        a = __fmaf_rn(rf.z, rows[j].z, a);
        p[j] = __fmaf_rn(rf.w, rows[j].w, a);
    }
    return transposed_butterfly8(p, lane);
}

// Full exact scan of one frame by one warp (lane owns dims 4*lane..4*lane+3).  Same arithmetic as
// rvq_encode_exact.cu.  Used for frames the certified search cannot cover.
__device__ __noinline__ int exact_scan_frame(float4 rf, float xx, const float *__restrict__ Ek,
                                             const float *__restrict__ cnk, int C, int lane) {
    float best = -CUDART_INF_F;
    int bidx = INT_MAX;
    for (int c0 = 0; c0 < C; c0 += 32) {
        float p[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const float4 e = __ldg(reinterpret_cast<const float4 *>(Ek + (size_t)(c0 + j) * 128) + lane);
            float a = __fmaf_rn(rf.x, e.x, 0.0f);
            a = __fmaf_rn(rf.y, e.y, a);
            a = __fmaf_rn(rf.z, e.z, a);
            p[j] = __fmaf_rn(rf.w, e.w, a);
        }
        rvq_transposed_butterfly(p, lane);
        const int col = c0 + lane;
        const float s = rvq_score(xx, p[0], cnk[col]);
        if (rvq_better(s, col, best, bidx)) { best = s; bidx = col; }
    }
    warp_argmax(best, bidx);
    return bidx;
}

// Exact resolution of one frame, general case (many candidate codes, list overflow, or a frame the
// bound does not cover): re-score every remembered group (16 codes) that can still hold the winner.
// list: this frame's entries, entry e at list[e * TILE_M]; entry = (float bits & ~0x7FF) | group.
__device__ __noinline__ int resolve_slow(float4 rf, float xx, int cnt, float M, float twodelta, int flags,
                                          const uint32_t *list, const float *__restrict__ Ek,
                                          const float *__restrict__ cnk, int C, int lane) {
    if (flags != 0 || cnt > LISTCAP) return exact_scan_frame(rf, xx, Ek, cnk, C, lane);
    // entries were stored with 11 mantissa bits cleared: allow for that when filtering
    const float thr = M - twodelta - 0.0009765625f * (fabsf(M) + twodelta);
    float best = -CUDART_INF_F;
    int bidx = INT_MAX;
    for (int e = 0; e < cnt; ++e) {
        const uint32_t u = list[e * TILE_M];
        const float val = __uint_as_float(u & ~0x7FFu);
        if (!(val >= thr)) continue;  // warp uniform
#pragma unroll 1
        for (int half = 0; half < 2; ++half) {
            const int c0 = (int)(u & 0x7FFu) * 16 + half * 8;
            float4 rows[8];
            const float4 *base = reinterpret_cast<const float4 *>(Ek + (size_t)c0 * 128) + lane;
#pragma unroll
            for (int j = 0; j < 8; ++j) rows[j] = __ldg(base + j * 32);
            const float dot = group_dots(rf, rows, lane);
            const int col = c0 + ((lane >> 2) & 7);
            const float s = rvq_score(xx, dot, cnk[col]);
            if (rvq_better(s, col, best, bidx)) { best = s; bidx = col; }
        }
    }
    warp_argmax(best, bidx);
    if (bidx == INT_MAX) bidx = exact_scan_frame(rf, xx, Ek, cnk, C, lane);  // cannot happen; stay total
    return bidx;
}

// canonical score of one code for one frame (warp cooperative; every lane gets the same bits)
__device__ __forceinline__ float exact_score(const float4 &rf, float xx, const float *__restrict__ Ek,
                                             const float *__restrict__ cnk, int col, int lane) {
    const float4 e = __ldg(reinterpret_cast<const float4 *>(Ek + (size_t)col * 128) + lane);
    float a = __fmaf_rn(rf.x, e.x, 0.0f);
    a = __fmaf_rn(rf.y, e.y, a);
    a = __fmaf_rn(rf.z, e.z, a);
    a = __fmaf_rn(rf.w, e.w, a);
    return rvq_score(xx, rvq_butterfly(a), __ldg(cnk + col));
}

__device__ __forceinline__ float exp2i(int e) { return __uint_as_float((uint32_t)(127 + e) << 23); }  // 2^e, |e| <= 126

// Per-frame power-of-two scaling of the fp16 residual operand.
struct FrameScale {
    int flags;  // 0: covered by the certified search; else the frame takes the exact scan
    int ea;     // operand = residual * 2^ea, so that 2^ea * ||r|| lies in (0.35, 1]
};
__device__ __forceinline__ FrameScale frame_scale(float xx, const RvqStageMeta &mt) {
    FrameScale fs;
    fs.flags = mt.flags ? 4 : 0;
    fs.ea = 0;
    if (!(xx <= 3.0e38f)) fs.flags |= 1;  // NaN or Inf
    const uint32_t b = __float_as_uint(xx);
    const int E = (int)(b >> 23);         // biased exponent (xx >= 0)
    if (E != 0) {
        // xx = m * 2^e2 with m in [0.5, 1)  ->  ||r|| <= 2^ceil(e2/2)
        const int e2 = E - 126;
        fs.ea = -((e2 + 1) >> 1);
    } else if (b != 0) {
        fs.flags |= 2;  // subnormal ||r||^2
    }
    const int aexp = fs.ea + mt.scale_exp - mt.aug_exp;  // exponent of the augmented operand column
    if (aexp < -24 || aexp > 15) fs.flags |= 2;
    return fs;
}

// ||r||^2 (canonical order, every lane gets it) of the frame whose 4 dims per lane are `rf`, and its fp16
// operand row `row` of the SWIZZLE_128B image A (rvq_common.cuh) for the stage described by `mt`.
__device__ __forceinline__ float emit_row(uint8_t *A, int row, int lane, const float4 &rf, const RvqStageMeta &mt) {
    float a = __fmaf_rn(rf.x, rf.x, 0.0f);
    a = __fmaf_rn(rf.y, rf.y, a);
    a = __fmaf_rn(rf.z, rf.z, a);
    a = __fmaf_rn(rf.w, rf.w, a);
    const float xx = rvq_butterfly(a);
    const FrameScale fs = frame_scale(xx, mt);
    const float al = fs.flags ? 0.0f : exp2i(fs.ea);
    const __half2 h0 = __floats2half2_rn(rf.x * al, rf.y * al);
    const __half2 h1 = __floats2half2_rn(rf.z * al, rf.w * al);
    const int off = (lane >> 4) * 16384 + row * 128 + ((((lane & 15) >> 1) ^ (row & 7)) << 4) + (lane & 1) * 8;
    uint2 w;
    w.x = *reinterpret_cast<const uint32_t *>(&h0);
    w.y = *reinterpret_cast<const uint32_t *>(&h1);
    *reinterpret_cast<uint2 *>(A + off) = w;
    return xx;
}

struct TcParams {
    const float *x;
    int64_t xs_b, xs_d, xs_t;
    int B, T;
    int64_t N;            // frames
    int ntiles;           // ceil(N / 128)
    const uint8_t *img;   // [Ktot][NT][IMG]
    const float *embed;   // [Ktot][C][128]
    const float *cnorm;   // [Ktot][C]
    const RvqStageMeta *meta;
    int C, K, NT;         // NT = C / 128
    int64_t *codes;       // [K][N]
    RvqWire wire;         // optional int16 copy of the codes (frame-sharded jobs)
    float *quantized;     // optional quantized_out, element (b, t, d) at quantized[b * qs_b + t * qs_t + d]
    int64_t qs_b, qs_t;
    float *scratch;       // residuals of the tiles in flight: [grid][2 slots][128 frames][128] fp32
    float *vals;          // approximate scores of the remembered groups: [grid][2 slots][LISTCAP][128 frames][16]
    // diagnostics (optional): accumulators and per-frame scan state of CTA 0 / slot 0 / first tile
    float *dbg_acc;       // [128][C] raw accumulators of stage dbg_stage
    float *dbg_info;      // [128][8]
    long long *dbg_time;  // [12 warps][8] cycle counters of CTA 0
    int dbg_stage;
};

__device__ __forceinline__ int slot_tiles(int ntiles, int grid, int cta, int slot) {
    // slot s of CTA c owns tiles c + grid * (s + 2 m)
    const int first = cta + grid * slot;
    if (first >= ntiles) return 0;
    return (ntiles - first + 2 * grid - 1) / (2 * grid);
}

// ------------------------------------------------------------------------------------ kernel
template <bool DBG, bool TIMING, bool QOUT>
__global__ void __launch_bounds__(THREADS, 1) rvq_encode_tc_kernel(const TcParams P) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t *sm = smem_raw + (base - smem_u32(smem_raw));
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t bar0 = base + SM_BAR;
    auto BAR = [&](int i) { return bar0 + 8u * (uint32_t)i; };
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(sm + SM_META + 16);

    if (threadIdx.x == 0) {
        for (int i = 0; i < RING; ++i) { mbar_init(BAR(BAR_BFULL + i), 1); mbar_init(BAR(BAR_BEMPTY + i), NSLOT); }
        for (int i = 0; i < 4; ++i) { mbar_init(BAR(BAR_ACCFULL + i), 1); mbar_init(BAR(BAR_ACCEMPTY + i), 4); }
        for (int i = 0; i < NSLOT; ++i) mbar_init(BAR(BAR_AREADY + i), 4);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32((void *)tmem_slot)),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    const int grid = gridDim.x, cta = blockIdx.x;
    const int NT = P.NT, K = P.K;
    const int per_tile = K * NT;  // items per tile
    // diagnostics: per-role cycle counters of CTA 0 (DBG build only), written to P.dbg_time[warp][8]
    long long tm[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long t_last = TIMING ? clock64() : 0;
#ifdef RVQ_TIMELINE
#define RVQ_LAP(i)
#else
#define RVQ_LAP(i)                                  \
    if (TIMING) {                                   \
        const long long t_now = clock64();          \
        tm[i] += t_now - t_last;                    \
        t_last = t_now;                             \
    }
#endif

#ifdef RVQ_TIMELINE
    // tools/tc_timeline.py: (tag, clock) pairs of CTA 0, lane 0 of every warp, into the buffer passed as acc_out
    int tl_n = 0;
#define RVQ_TL(tag)                                                                        \
    if (cta == 0 && lane == 0 && P.dbg_acc) {                                              \
        long long *e = reinterpret_cast<long long *>(P.dbg_acc) + ((size_t)warp * 4096 + tl_n) * 2; \
        e[0] = (long long)(tag);                                                           \
        e[1] = clock64();                                                                  \
        if (tl_n < 4095) ++tl_n;                                                           \
    }
#else
#define RVQ_TL(tag)
#endif

    if (warp < 4) {
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        // The two slots of a CTA run in lock step: they are at the same codebook tile (stage k, tile j) at the
        // same time, so that one copy of the tile serves both (the L2 -> shared memory stream of codebook
        // tiles is what saturates first otherwise, and it slows every other L2 access of the kernel down).
        const int rounds0 = slot_tiles(P.ntiles, grid, cta, 0), rounds1 = slot_tiles(P.ntiles, grid, cta, 1);
        const int total = rounds0 * per_tile;  // codebook tiles this CTA streams; rounds1 <= rounds0
        if (warp == 0) {
            // ============================ producer (whole warp, one elected lane issues) ============================
            int ridx = 0, ii = 0;
            uint32_t rpar = 0;
            for (int it = 0; it < total; ++it) {
                RVQ_LAP(0)
                RVQ_TL(0x100)
                mbar_wait(BAR(BAR_BEMPTY + ridx), rpar ^ 1u);
                RVQ_TL(0x110)
                RVQ_LAP(1)
                if (elect_one()) {
                    mbar_arrive_expect_tx(BAR(BAR_BFULL + ridx), IMG);
                    bulk_g2s(base + SM_RING + ridx * IMG, P.img + (size_t)ii * IMG, IMG, BAR(BAR_BFULL + ridx));
                }
                __syncwarp();
                if (++ii == per_tile) ii = 0;
                if (++ridx == RING) { ridx = 0; rpar ^= 1u; }
            }
        } else if (warp <= NSLOT) {
            // ============================ MMA issuer of slot warp-1 (whole warp waits, one elected lane issues) ======
            const int s = warp - 1;
            const int mine = (s ? rounds1 : rounds0) * per_tile;  // tiles this slot consumes; the rest it only releases
            const uint64_t adesc = desc_sw128(base + SM_A) + (uint64_t)(s * (IMG >> 4));
            const uint64_t aaug = desc_interleaved(base + SM_A + 32768) + (uint64_t)(s * (IMG >> 4));
            const uint64_t bdesc0 = desc_sw128(base + SM_RING), baug0 = desc_interleaved(base + SM_RING + 32768);
            int j = 0, ridx = 0;
            uint32_t stage_par = 0, rpar = 0;
            for (int n = 0; n < total; ++n) {
                RVQ_LAP(0)
                RVQ_TL(0x200)
                mbar_wait(BAR(BAR_BFULL + ridx), rpar);
                RVQ_TL(0x210)
                RVQ_LAP(1)
                if (n < mine) {
                    if (j == 0) mbar_wait(BAR(BAR_AREADY + s), stage_par);
                    RVQ_LAP(2)
                    const int buf = n & 1;
                    mbar_wait(BAR(BAR_ACCEMPTY + s * 2 + buf), (((uint32_t)(n >> 1)) & 1u) ^ 1u);
                    RVQ_LAP(3)
                    RVQ_TL(0x220 | s)
                    tc_fence_after();
                    const uint32_t d = tmem + (uint32_t)(s * 256 + buf * 128);
                    // descriptors: only the 14-bit start address field (low word, 16-byte units) changes
                    const uint32_t blo = (uint32_t)bdesc0 + (uint32_t)(ridx * (IMG >> 4));
                    const uint32_t bauglo = (uint32_t)baug0 + (uint32_t)(ridx * (IMG >> 4));
                    if (elect_one()) {
#pragma unroll
                        for (int kb = 0; kb < 2; ++kb)
#pragma unroll
                            for (int ks = 0; ks < 4; ++ks)
                                umma_f16_lh(d, (uint32_t)adesc + (uint32_t)(kb * 1024 + ks * 2), (uint32_t)(adesc >> 32),
                                            blo + (uint32_t)(kb * 1024 + ks * 2), (uint32_t)(bdesc0 >> 32), IDESC,
                                            (kb | ks) ? 1u : 0u);
                        umma_f16_lh(d, (uint32_t)aaug, (uint32_t)(aaug >> 32), bauglo, (uint32_t)(baug0 >> 32), IDESC, 1u);
                        umma_commit(BAR(BAR_BEMPTY + ridx));
                        umma_commit(BAR(BAR_ACCFULL + s * 2 + buf));
                    }
                    __syncwarp();
                    RVQ_TL(0x230 | s)
                    if (++j == NT) { j = 0; stage_par ^= 1u; }
                } else {
                    // this slot has no tile in the last round: only hand the ring entry back
                    if (elect_one()) mbar_arrive(BAR(BAR_BEMPTY + ridx));
                    __syncwarp();
                }
                if (++ridx == RING) { ridx = 0; rpar ^= 1u; }
            }
        }
    } else {
        asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
        // ============================ frame warps ============================
        const int slot = (warp - 4) >> 2, q = warp & 3;
        uint8_t *A = sm + SM_A + slot * IMG;
        uint32_t *list = reinterpret_cast<uint32_t *>(sm + SM_LIST) + slot * LISTCAP * TILE_M;
        const int row_l = q * 32 + lane;  // the row (frame of the tile) this thread scans
        const int C = P.C;
        // this warp's 32 residual rows [frame][128] fp32 in the L2-resident scratch
        float *rs = P.scratch + ((size_t)(cta * NSLOT + slot) * TILE_M + q * 32) * 128;
        float4 *rs4 = reinterpret_cast<float4 *>(rs) + lane;  // + f * 32: this lane's 4 dims of frame f
        float *vals = P.vals + (size_t)(cta * NSLOT + slot) * LISTCAP * TILE_M * 16;
        uint32_t items = 0;  // accumulator uses of this slot so far
        const bool direct = P.xs_d == 1 && (P.xs_b & 3) == 0 && (P.xs_t & 3) == 0 && ((uintptr_t)P.x & 15) == 0;

        for (int tile = cta + grid * slot, first = 1; tile < P.ntiles; tile += 2 * grid, first = 0) {
            const int64_t n0 = (int64_t)tile * TILE_M + q * 32;
            RVQ_LAP(0)
            // lane = frame: where this lane's frame starts in x (-1: past the end)
            const int64_t n_l = n0 + lane;
            const int64_t xoff = n_l < P.N ? (n_l / P.T) * P.xs_b + (n_l % P.T) * P.xs_t : -1;
            const int64_t woff = (P.wire.p && n_l < P.N) ? (n_l / P.T) * P.wire.sb + (n_l % P.T) : -1;
            RvqStageMeta mt = P.meta[0];
            float xx = 0.0f;
            if (!direct) {
                // lane = frame reads (coalesced when T is the contiguous axis), the whole frame in flight.  The frame's
                // own thread takes ||x||^2 (the canonical tree) and writes the stage 0 operand row; the fp32 rows go to
                // the scratch transposed 32 dims at a time through this warp's own 4 KB of the A tile.
                float *st = reinterpret_cast<float *>(A + q * 4096);
                float v[4][32];
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int dd = 0; dd < 32; ++dd) v[c][dd] = xoff >= 0 ? P.x[xoff + (int64_t)(c * 32 + dd) * P.xs_d] : 0.0f;
                float p[32];
#pragma unroll
                for (int c = 0; c < 4; ++c) {
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        float a = __fmaf_rn(v[c][4 * i], v[c][4 * i], 0.0f);
                        a = __fmaf_rn(v[c][4 * i + 1], v[c][4 * i + 1], a);
                        a = __fmaf_rn(v[c][4 * i + 2], v[c][4 * i + 2], a);
                        p[8 * c + i] = __fmaf_rn(v[c][4 * i + 3], v[c][4 * i + 3], a);
                    }
                }
                xx = tree32(p);
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    __syncwarp();
#pragma unroll
                    for (int dd = 0; dd < 32; ++dd) st[dd * 32 + (lane ^ dd)] = v[c][dd];
                    __syncwarp();
#pragma unroll 8
                    for (int f = 0; f < 32; ++f) rs[f * 128 + c * 32 + lane] = st[lane * 32 + (f ^ lane)];
                }
                __syncwarp();
                const FrameScale fs0 = frame_scale(xx, mt);
                const float al = fs0.flags ? 0.0f : exp2i(fs0.ea);
#pragma unroll
                for (int c16 = 0; c16 < 16; ++c16) {  // 16-byte chunk = dims 8 c16 .. 8 c16 + 7
                    __half2 h[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        h[i] = __floats2half2_rn(v[c16 >> 2][(c16 & 3) * 8 + 2 * i] * al, v[c16 >> 2][(c16 & 3) * 8 + 2 * i + 1] * al);
                    uint4 w;
                    w.x = *reinterpret_cast<const uint32_t *>(&h[0]);
                    w.y = *reinterpret_cast<const uint32_t *>(&h[1]);
                    w.z = *reinterpret_cast<const uint32_t *>(&h[2]);
                    w.w = *reinterpret_cast<const uint32_t *>(&h[3]);
                    *reinterpret_cast<uint4 *>(A + (c16 >> 3) * 16384 + row_l * 128 + (((c16 & 7) ^ (row_l & 7)) << 4)) = w;
                }
            } else {
                // channels-last input: a frame is one 512-byte row; eight frames per iteration, the next eight requested
                // before these are converted; lane f keeps ||r_f||^2
                constexpr int NB = 8;
                auto fetch = [&](int f) {
                    const int64_t o = __shfl_sync(0xffffffffu, xoff, f);
                    return o >= 0 ? *(reinterpret_cast<const float4 *>(P.x + o) + lane) : make_float4(0.f, 0.f, 0.f, 0.f);
                };
                float4 rf[NB];
#pragma unroll
                for (int t = 0; t < NB; ++t) rf[t] = fetch(t);
#pragma unroll 1
                for (int f0 = 0; f0 < 32; f0 += NB) {
                    float4 cur[NB];
#pragma unroll
                    for (int t = 0; t < NB; ++t) {
                        cur[t] = rf[t];
                        rs4[(f0 + t) * 32] = cur[t];
                    }
                    if (f0 + NB < 32) {
#pragma unroll
                        for (int t = 0; t < NB; ++t) rf[t] = fetch(f0 + NB + t);
                    }
#pragma unroll
                    for (int t = 0; t < NB; ++t) {
                        const float xv = emit_row(A, q * 32 + f0 + t, lane, cur[t], mt);
                        if (lane == f0 + t) xx = xv;
                    }
                }
            }
            RVQ_LAP(1)
            for (int k = 0; k < K; ++k) {
                const float *Ek = P.embed + (size_t)k * C * 128;
                const float *cnk = P.cnorm + (size_t)k * C;
                // ---- this thread's frame (= lane): error bound, augmented operand column, hand-over
                const FrameScale fs = frame_scale(xx, mt);
                const int flags = fs.flags;
                const float scale_acc = exp2i(fs.ea + mt.scale_exp);  // acc ~= scale_acc * (r.c - cc/2)
                const float R = sqrtf(xx), cn_max = sqrtf(mt.cnorm2_max);
                const float e16 = 0.0009765625f * scale_acc * R * cn_max;  // 2^-10: fp16 rounding of both operands
                const float e32 = 2.384185791015625e-07f * (xx + 2.0f * R * cn_max + mt.cnorm2_max) +
                                  3.814697265625e-06f * R * cn_max;         // canonical fp32 evaluation
                const float twodelta = 2.0002f * (1.2505f * e16 + 0.5f * scale_acc * e32 + 1.0f);
                const __half a_aug = __float2half_rn(flags ? 0.0f : exp2i(fs.ea + mt.scale_exp - mt.aug_exp));
                {
                    const uint32_t h = (uint32_t)__half_as_ushort(a_aug);
                    uint8_t *ar = A + 32768 + (row_l >> 3) * 256 + (row_l & 7) * 16;
                    *reinterpret_cast<uint4 *>(ar) = make_uint4(h | (h << 16), h, 0u, 0u);
                    *reinterpret_cast<uint4 *>(ar + 128) = make_uint4(0u, 0u, 0u, 0u);
                }
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(BAR(BAR_AREADY + slot));
                RVQ_TL(0x300 | k)

                // ---- scan: this thread's frame = TMEM lane row_l.  A group = the 16 codes of one TMEM load.
                // A group whose maximum is within 2*delta of the running maximum is remembered: its packed
                // (max | group id) in shared memory, its 16 approximate scores in the L2 scratch.
                float M = -CUDART_INF_F, thr = -CUDART_INF_F;
                int cnt = 0;
                const bool dbg = DBG && cta == 0 && slot == 0 && first && k == P.dbg_stage;
                RVQ_LAP(2)
                // One accumulator = 128 codes = two TMEM loads of 64 columns, and one load is always in flight
                // while the other half is reduced: the second half of this accumulator during the first, the first
                // half of the next accumulator during the second.  The accumulator is handed back to the MMA warp
                // as soon as both halves are in registers.  (A half is judged against the maximum seen up to and
                // including itself: a lower threshold than the final one only remembers more groups.)
                // group maxima of one 64-column half (independent dependency chains)
                auto half_max = [&](const uint32_t(&v)[64], float(&gmx)[4]) {
#pragma unroll
                    for (int g = 0; g < 4; ++g) gmx[g] = fmaxf(group_max8(&v[16 * g]), group_max8(&v[16 * g + 8]));
                };
                // remember the groups of one half that are within 2*delta of the maximum seen so far
                float *vals_row = pin_ptr(vals + (size_t)row_l * 16);       // + cnt * TILE_M * 16
                const uint32_t list_row = pin_u32(smem_u32(list + row_l));  // + cnt * TILE_M * 4
                auto remember = [&](const uint32_t(&v)[64], const float(&gmx)[4], int j, int half) {
                    const uint32_t idb = (uint32_t)(j * 8 + half * 4);
#pragma unroll
                    for (int g = 0; g < 4; ++g) {
                        const uint32_t *vg = &v[16 * g];
                        if (DBG && dbg) {
#pragma unroll
                            for (int i = 0; i < 16; ++i)
                                P.dbg_acc[(size_t)row_l * C + j * 128 + half * 64 + g * 16 + i] = __uint_as_float(vg[i]);
                        }
                        const bool pass = gmx[g] >= thr;
                        remember_store(pass && cnt < LISTCAP, vals_row + (size_t)cnt * (TILE_M * 16), vg,
                                       list_row + (uint32_t)cnt * (TILE_M * 4),
                                       (__float_as_uint(gmx[g]) & ~0x7FFu) | (idb + (uint32_t)g));
                        cnt += pass ? 1 : 0;
                    }
                };
                uint32_t va[64], vb[64];
                const uint32_t tq = pin_u32(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(slot * 256));
                const uint32_t bar_full = pin_u32(BAR(BAR_ACCFULL + slot * 2)), bar_empty = pin_u32(BAR(BAR_ACCEMPTY + slot * 2));
                mbar_wait(bar_full + 8u * (items & 1u), (items >> 1) & 1u);
                tc_fence_after();
                tmem_ld64(tq + (items & 1) * 128, va);
                for (int j = 0; j < NT; ++j, ++items) {
                    float ga[4], gb[4];
                    const int buf = items & 1;
                    RVQ_TL(0x310)
                    tmem_wait_ld();
                    tmem_ld64(tq + buf * 128 + 64, vb);
                    RVQ_TL(0x320)
                    RVQ_LAP(3)
                    half_max(va, ga);
                    M = fmaxf(M, fmaxf(fmaxf(ga[0], ga[1]), fmaxf(ga[2], ga[3])));
                    thr = M - twodelta;
                    remember(va, ga, j, 0);
                    tmem_wait_ld();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_empty + 8u * (uint32_t)buf);
                    RVQ_LAP(4)
                    if (j + 1 < NT) {
                        // the first half of the next accumulator travels while the second half of this one is reduced
                        mbar_wait(bar_full + 8u * (uint32_t)(buf ^ 1), ((items + 1) >> 1) & 1u);
                        RVQ_TL(0x340)
                        tc_fence_after();
                        tmem_ld64(tq + (buf ^ 1) * 128, va);
                    }
                    half_max(vb, gb);
                    M = fmaxf(M, fmaxf(fmaxf(gb[0], gb[1]), fmaxf(gb[2], gb[3])));
                    thr = M - twodelta;
                    remember(vb, gb, j, 1);
                    RVQ_TL(0x330)
                }
                if (DBG && dbg) {
                    float *o = P.dbg_info + row_l * 8;
                    o[0] = xx; o[1] = flags ? 0.0f : exp2i(fs.ea); o[2] = twodelta; o[3] = M; o[4] = (float)cnt;
                    o[5] = (float)flags; o[6] = scale_acc; o[7] = __half2float(a_aug);
                }

                // The next tile of this slot comes from HBM: ask for it now (last stage: no residual update follows),
                // so that the tile load finds it in L2.
                if (k == K - 1 && tile + 2 * grid < P.ntiles) {
                    const int64_t nn = n_l + (int64_t)2 * grid * TILE_M;
                    if (nn < P.N) {
                        const float *xn = P.x + (nn / P.T) * P.xs_b + (nn % P.T) * P.xs_t;
                        if (P.xs_d == 1) {
#pragma unroll
                            for (int dd = 0; dd < 128; dd += 32) asm volatile("prefetch.global.L2 [%0];" ::"l"(xn + dd));
                        } else {
#pragma unroll 16
                            for (int dd = 0; dd < 128; ++dd) asm volatile("prefetch.global.L2 [%0];" ::"l"(xn + (int64_t)dd * P.xs_d));
                        }
                    }
                }

                // ---- resolve, phase 1 (lane = frame): the codes whose approximate score is within 2*delta of
                // the maximum are the only ones that can win.  Exactly one such code (the common case): it is
                // the winner, no exact arithmetic needed.  2..4: re-scored exactly below.  Otherwise resolve_slow.
                int ncand = 99, myidx = 0, c1 = 0, c2 = 0, c3 = 0;
                if (flags == 0 && cnt <= LISTCAP) {
                    const float keep = M - twodelta;
                    const float keep_packed = keep - 0.0009765625f * (fabsf(M) + twodelta);  // 11 mantissa bits were cleared
                    // pass A (shared memory only, straight-line: all LISTCAP words are read, those past cnt ignored):
                    // which remembered groups survive?  Up to 4 are examined here.
                    uint32_t sm_mask = 0;
#pragma unroll
                    for (int e = 0; e < LISTCAP; ++e) {
                        const uint32_t u = list[e * TILE_M + row_l];
                        const bool ok = e < cnt && __uint_as_float(u & ~0x7FFu) >= keep_packed;
                        sm_mask |= ok ? (1u << e) : 0u;
                    }
                    const int ne = __popc(sm_mask);
                    if (ne >= 1 && ne <= 4) {
                        int es[4];
                        uint32_t mrest = sm_mask;
#pragma unroll
                        for (int s4 = 0; s4 < 4; ++s4) {
                            es[s4] = mrest ? __ffs(mrest) - 1 : es[0];
                            mrest &= mrest - 1u;  // (0 stays 0)
                        }
                        // pass B: one round trip to the L2 scratch for all surviving groups
                        float4 t[4][4];
                        uint32_t us[4];
#pragma unroll
                        for (int s4 = 0; s4 < 4; ++s4) {
                            const float4 *src = reinterpret_cast<const float4 *>(vals + ((size_t)es[s4] * TILE_M + row_l) * 16);
#pragma unroll
                            for (int qd = 0; qd < 4; ++qd)
                                t[s4][qd] = s4 < ne ? src[qd] : make_float4(-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F);
                            us[s4] = list[es[s4] * TILE_M + row_l];
                        }
                        // 16-bit masks of the codes within 2*delta of the maximum, one per surviving group
                        unsigned long long mm = 0ull;
#pragma unroll
                        for (int s4 = 0; s4 < 4; ++s4) {
                            uint32_t mk = 0;
#pragma unroll
                            for (int qd = 0; qd < 4; ++qd) {
                                mk |= t[s4][qd].x >= keep ? (1u << (qd * 4 + 0)) : 0u;
                                mk |= t[s4][qd].y >= keep ? (1u << (qd * 4 + 1)) : 0u;
                                mk |= t[s4][qd].z >= keep ? (1u << (qd * 4 + 2)) : 0u;
                                mk |= t[s4][qd].w >= keep ? (1u << (qd * 4 + 3)) : 0u;
                            }
                            mm |= (unsigned long long)(s4 < ne ? mk : 0u) << (16 * s4);
                        }
                        ncand = __popcll(mm);
                        auto take = [&]() {  // lowest set position -> code index; clears it
                            const int pos = mm ? __ffsll((long long)mm) - 1 : 0;
                            mm &= mm - 1ull;
                            const uint32_t u = pos < 16 ? us[0] : (pos < 32 ? us[1] : (pos < 48 ? us[2] : us[3]));
                            return (int)(u & 0x7FFu) * 16 + (pos & 15);
                        };
                        myidx = take(); c1 = take(); c2 = take(); c3 = take();
                        if (ncand == 0 || ncand > 4) ncand = 99;  // 0 cannot happen (the maximum itself qualifies)
                    }
                }
                __syncwarp();  // lists (written by lane = frame) are read by every lane in resolve_slow
                RVQ_LAP(5)
                RVQ_TL(0x350)
#ifdef RVQ_TIMELINE
                if (P.dbg_acc && n0 + lane < P.N) {  // histogram of candidate counts over all frames (after the event area)
                    unsigned long long *hist = reinterpret_cast<unsigned long long *>(P.dbg_acc) + 16 * 4096 * 2;
                    atomicAdd(hist + k * 8 + (ncand > 4 ? 5 : ncand), 1ull);
                    atomicAdd(hist + 32 + k * 8 + min(cnt, 17) / 3, 1ull);
                }
#endif

                // ---- resolve, phase 2a: the frames with more than one candidate, exact fp32 in the canonical order.
                // Two frames per iteration where possible, so that their loads share one round trip to L2.
                auto best_of = [&](const float4 &rf, float xf, int nc, int j0, int j1, int j2, int j3) {
                    float bs = exact_score(rf, xf, Ek, cnk, j0, lane);
                    int idx = j0;
                    const float s1 = exact_score(rf, xf, Ek, cnk, j1, lane);
                    if (rvq_better(s1, j1, bs, idx)) { bs = s1; idx = j1; }
                    if (nc > 2) {
                        const float s2 = exact_score(rf, xf, Ek, cnk, j2, lane);
                        if (rvq_better(s2, j2, bs, idx)) { bs = s2; idx = j2; }
                    }
                    if (nc > 3) {
                        const float s3 = exact_score(rf, xf, Ek, cnk, j3, lane);
                        if (rvq_better(s3, j3, bs, idx)) { bs = s3; idx = j3; }
                    }
                    return idx;
                };
                // Two candidates (nearly all ambiguous frames): four frames = eight codes per iteration; three or four
                // candidates: two frames per iteration.  All rows of an iteration are requested together (one round
                // trip to L2), the eight canonical sums come from one transposed butterfly.  With fewer frames left
                // than fit, the last one is repeated; a frame with three candidates scores its unused fourth slot too
                // (a valid code that cannot win: its approximate score is more than 2*delta below the maximum).
                auto rescore_batches = [&](unsigned todo, auto cpf_tag) {
                    constexpr int CPF = decltype(cpf_tag)::value, FPI = 8 / CPF;
                    while (todo) {
                        int fr[FPI], jj[8];
                        float xf[FPI];
                        float4 rfv[FPI], rows[8];
#pragma unroll
                        for (int t = 0; t < FPI; ++t) {
                            fr[t] = todo ? __ffs(todo) - 1 : fr[t ? t - 1 : 0];
                            todo &= todo - 1u;  // (0 stays 0)
                            xf[t] = __shfl_sync(0xffffffffu, xx, fr[t]);
                            jj[t * CPF] = __shfl_sync(0xffffffffu, myidx, fr[t]);
                            jj[t * CPF + 1] = __shfl_sync(0xffffffffu, c1, fr[t]);
                            if (CPF == 4) {
                                jj[t * CPF + 2] = __shfl_sync(0xffffffffu, c2, fr[t]);
                                jj[t * CPF + 3] = __shfl_sync(0xffffffffu, c3, fr[t]);
                            }
                            rfv[t] = rs4[fr[t] * 32];
#pragma unroll
                            for (int i = 0; i < CPF; ++i)
                                rows[t * CPF + i] = __ldg(reinterpret_cast<const float4 *>(Ek + (size_t)jj[t * CPF + i] * 128) + lane);
                        }
                        float pc[8];
#pragma unroll
                        for (int c = 0; c < 8; ++c) {
                            const float4 &r = rfv[c / CPF];
                            float t = __fmaf_rn(r.x, rows[c].x, 0.0f);
                            t = __fmaf_rn(r.y, rows[c].y, t);
                            t = __fmaf_rn(r.z, rows[c].z, t);
                            pc[c] = __fmaf_rn(r.w, rows[c].w, t);
                        }
                        const float dot = transposed_butterfly8(pc, lane);  // lane l: column (l >> 2) & 7
                        const int col = (lane >> 2) & 7;
                        int jc = jj[0];
                        float xc = xf[0];
#pragma unroll
                        for (int c = 1; c < 8; ++c) {
                            if (col == c) { jc = jj[c]; xc = xf[c / CPF]; }
                        }
                        float sc = rvq_score(xc, dot, __ldg(cnk + jc));
                        // best of the frame's CPF columns (4 lanes per column): a symmetric exchange, so every lane
                        // of the frame's group ends with the same winner
#pragma unroll
                        for (int off = 4; off < 4 * CPF; off <<= 1) {
                            const float so = __shfl_xor_sync(0xffffffffu, sc, off);
                            const int jo = __shfl_xor_sync(0xffffffffu, jc, off);
                            if (rvq_better(so, jo, sc, jc)) { sc = so; jc = jo; }
                        }
#pragma unroll
                        for (int t = 0; t < FPI; ++t) {
                            const int wt = __shfl_sync(0xffffffffu, jc, 4 * CPF * t);
                            if (lane == fr[t]) { myidx = wt; ncand = 1; }
                        }
                    }
                };
                rescore_batches(__ballot_sync(0xffffffffu, ncand == 2), std::integral_constant<int, 2>());
                rescore_batches(__ballot_sync(0xffffffffu, ncand == 3 || ncand == 4), std::integral_constant<int, 4>());
                for (unsigned amb = __ballot_sync(0xffffffffu, ncand != 1); amb;) {
                    const int f = __ffs(amb) - 1;
                    amb &= amb - 1;
                    const int nc = __shfl_sync(0xffffffffu, ncand, f);
                    const float xf = __shfl_sync(0xffffffffu, xx, f);
                    const float4 rf = rs4[f * 32];
                    int idx;
                    if (nc <= 4) {
                        const int j0 = __shfl_sync(0xffffffffu, myidx, f), j1 = __shfl_sync(0xffffffffu, c1, f);
                        const int j2 = __shfl_sync(0xffffffffu, c2, f), j3 = __shfl_sync(0xffffffffu, c3, f);
                        const int g = amb ? __ffs(amb) - 1 : f;                 // the next ambiguous frame, if any
                        const int ncg = __shfl_sync(0xffffffffu, ncand, g);
                        if (amb && ncg <= 4) {
                            amb &= amb - 1;
                            const float xg = __shfl_sync(0xffffffffu, xx, g);
                            const float4 rg = rs4[g * 32];
                            const int i0 = __shfl_sync(0xffffffffu, myidx, g), i1 = __shfl_sync(0xffffffffu, c1, g);
                            const int i2 = __shfl_sync(0xffffffffu, c2, g), i3 = __shfl_sync(0xffffffffu, c3, g);
                            idx = best_of(rf, xf, nc, j0, j1, j2, j3);
                            const int idg = best_of(rg, xg, ncg, i0, i1, i2, i3);
                            if (lane == g) myidx = idg;
                        } else {
                            idx = best_of(rf, xf, nc, j0, j1, j2, j3);
                        }
                    } else {
                        const int cf = __shfl_sync(0xffffffffu, cnt, f);
                        const float Mf = __shfl_sync(0xffffffffu, M, f);
                        const float tf = __shfl_sync(0xffffffffu, twodelta, f);
                        const int ff = __shfl_sync(0xffffffffu, flags, f);
                        idx = resolve_slow(rf, xf, cf, Mf, tf, ff, list + q * 32 + f, Ek, cnk, C, lane);
                    }
                    if (lane == f) myidx = idx;
                }
                const int mycode = myidx;
                RVQ_LAP(6)
                RVQ_TL(0x360)

                // ---- resolve, phase 2b: residual update and the next stage's operand row.  Batches of four frames;
                // the winners' rows and the residuals of three batches are in flight (an L2 round trip lasts about as
                // long as three batches take to convert), in four register sets addressed statically (fully unrolled).
                const bool more = k + 1 < K;
                float xx_next = 0.0f;
                if (more) {
                    constexpr int NB = 4, NBATCH = 32 / NB, DEPTH = 3;
                    mt = P.meta[k + 1];
                    float4 w[DEPTH + 1][NB], rf[DEPTH + 1][NB];
#pragma unroll
                    for (int b = 0; b < DEPTH; ++b) {
#pragma unroll
                        for (int t = 0; t < NB; ++t) {
                            const int id = __shfl_sync(0xffffffffu, myidx, b * NB + t);
                            w[b][t] = __ldg(reinterpret_cast<const float4 *>(Ek + (size_t)id * 128) + lane);
                            rf[b][t] = rs4[(b * NB + t) * 32];
                        }
                    }
#pragma unroll
                    for (int b = 0; b < NBATCH; ++b) {
                        const int set = b % (DEPTH + 1);
                        if (b + DEPTH < NBATCH) {
                            const int sn = (b + DEPTH) % (DEPTH + 1);
#pragma unroll
                            for (int t = 0; t < NB; ++t) {
                                const int id = __shfl_sync(0xffffffffu, myidx, (b + DEPTH) * NB + t);
                                w[sn][t] = __ldg(reinterpret_cast<const float4 *>(Ek + (size_t)id * 128) + lane);
                                rf[sn][t] = rs4[((b + DEPTH) * NB + t) * 32];
                            }
                        }
                        float4 rn[NB];
#pragma unroll
                        for (int t = 0; t < NB; ++t) {
                            // residual = residual - quantized (core_vq.py:394)
                            rn[t].x = __fsub_rn(rf[set][t].x, w[set][t].x);
                            rn[t].y = __fsub_rn(rf[set][t].y, w[set][t].y);
                            rn[t].z = __fsub_rn(rf[set][t].z, w[set][t].z);
                            rn[t].w = __fsub_rn(rf[set][t].w, w[set][t].w);
                            rs4[(b * NB + t) * 32] = rn[t];
                        }
#pragma unroll
                        for (int t = 0; t < NB; ++t) {
                            const float xv = emit_row(A, q * 32 + b * NB + t, lane, rn[t], mt);
                            if (lane == b * NB + t) xx_next = xv;
                        }
                    }
                }
                if (n0 + lane < P.N) P.codes[(int64_t)k * P.N + n0 + lane] = (int64_t)mycode;
                if (woff >= 0) P.wire.p[(int64_t)k * P.wire.sk + woff] = (int16_t)mycode;
                xx = xx_next;
                __syncwarp();
                RVQ_TL(0x370)
                RVQ_LAP(7)
            }
            if (QOUT) {
                // quantized_out = 0.0 + E_0[code_0] + E_1[code_1] + ... in stage order (core_vq.py:363-385) for this warp's
                // 32 frames: warp = frame, lane = 4 dims; eight frames and two stages (16 rows) in flight
                // (four stages at once spill: the compiler gives this code the registers of the scan loop, not more).  The codes are
                // the ones this thread has just written (thread = frame), read back and passed round by shuffles.
                const int64_t qoff = n_l < P.N ? (n_l / P.T) * P.qs_b + (n_l % P.T) * P.qs_t : -1;
#pragma unroll 1
                for (int f0 = 0; f0 < 32; f0 += 8) {
                    float4 acc[8];
#pragma unroll
                    for (int t = 0; t < 8; ++t) acc[t] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
                    for (int k0 = 0; k0 < K; k0 += 2) {
                        const bool two = k0 + 1 < K;
                        const int ca = n_l < P.N ? (int)P.codes[(int64_t)k0 * P.N + n_l] : 0;
                        const int cb = (two && n_l < P.N) ? (int)P.codes[(int64_t)(k0 + 1) * P.N + n_l] : 0;
                        const float *Ea = P.embed + (size_t)k0 * C * 128 + lane * 4;
                        const float *Eb = Ea + (size_t)(two ? C : 0) * 128;
                        float4 ra[8], rb[8];
#pragma unroll
                        for (int t = 0; t < 8; ++t) {
                            ra[t] = __ldg(reinterpret_cast<const float4 *>(Ea + (size_t)__shfl_sync(0xffffffffu, ca, f0 + t) * 128));
                            rb[t] = __ldg(reinterpret_cast<const float4 *>(Eb + (size_t)__shfl_sync(0xffffffffu, cb, f0 + t) * 128));
                        }
#pragma unroll
                        for (int t = 0; t < 8; ++t) {
                            acc[t].x = __fadd_rn(acc[t].x, ra[t].x); acc[t].y = __fadd_rn(acc[t].y, ra[t].y);
                            acc[t].z = __fadd_rn(acc[t].z, ra[t].z); acc[t].w = __fadd_rn(acc[t].w, ra[t].w);
                            if (two) {
                                acc[t].x = __fadd_rn(acc[t].x, rb[t].x); acc[t].y = __fadd_rn(acc[t].y, rb[t].y);
                                acc[t].z = __fadd_rn(acc[t].z, rb[t].z); acc[t].w = __fadd_rn(acc[t].w, rb[t].w);
                            }
                        }
                    }
#pragma unroll
                    for (int t = 0; t < 8; ++t) {
                        const int64_t o = __shfl_sync(0xffffffffu, qoff, f0 + t);
                        if (o >= 0) __stcs(reinterpret_cast<float4 *>(P.quantized + o) + lane, acc[t]);
                    }
                }
            }
        }
    }
    if (TIMING && cta == 0 && lane == 0 && P.dbg_time) {
        for (int i = 0; i < 8; ++i) P.dbg_time[warp * 8 + i] = tm[i];
    }
#undef RVQ_LAP
#undef RVQ_TL

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
    }
}

}  // namespace

// ------------------------------------------------------------------------------------ host
bool rvq_tc_shape_supported(int D, int C) { return D == 128 && C % 128 == 0 && C >= 256 && C <= 16384; }

// SMs the persistent encode kernel leaves free (process-wide setting, rvq_b200_set_reserved_sms): a frame-sharded job
// keeps a few for the kernels of its code exchange, which otherwise find room only between two encodes.
static int g_reserved_sms = 0;

extern "C" int rvq_b200_set_reserved_sms(int n) {
    RVQ_REQUIRE(n >= 0 && n <= 64, RVQ_EINVAL, "rvq_b200_set_reserved_sms: %d is not in [0, 64]", n);
    g_reserved_sms = n;
    return RVQ_OK;
}

int rvq_reserved_sms() { return g_reserved_sms; }

static int device_sms(int *sms) {
    int dev = 0;
    RVQ_CUDA(cudaGetDevice(&dev));
    RVQ_CUDA(cudaDeviceGetAttribute(sms, cudaDevAttrMultiProcessorCount, dev));
    if (*sms - g_reserved_sms >= 8) *sms -= g_reserved_sms;
    return RVQ_OK;
}

static int tc_grid(int64_t N, int sms) {
    // one CTA per SM; with fewer tiles than SMs every tile gets a CTA of its own (its second slot stays idle),
    // which halves the latency of small inputs
    const int64_t ntiles = (N + TILE_M - 1) / TILE_M;
    int64_t grid = ntiles < sms ? ntiles : sms;
    return (int)(grid < 1 ? 1 : grid);
}

// workspace = [256 B: reserved] [grid x 128 KB: residuals] [grid x 256 KB: remembered approximate scores]
constexpr int64_t WS_HEADER = 256;
constexpr int64_t WS_R_PER_CTA = (int64_t)NSLOT * TILE_M * 128 * 4;              // residuals
constexpr int64_t WS_V_PER_CTA = (int64_t)NSLOT * LISTCAP * TILE_M * 16 * 4;   // remembered approximate scores
constexpr int64_t WS_PER_CTA = WS_R_PER_CTA + WS_V_PER_CTA;

int64_t rvq_tc_workspace_bytes(int B, int D, int T, int K, int C) {
    (void)D; (void)K; (void)C;
    int sms = 0;
    if (device_sms(&sms) != RVQ_OK) return RVQ_ECUDA;
    return WS_HEADER + WS_PER_CTA * tc_grid((int64_t)B * T, sms);
}

static int launch_tc(const float *x, const int64_t xs[3], int B, int T, const RvqPackView &pv, int K, int64_t *codes,
                     float *quantized, const RvqWire &wire, void *workspace, int64_t workspace_bytes, float *dbg_acc, float *dbg_info, int dbg_stage,
                     cudaStream_t stream, bool fit_grid_to_workspace = false) {
    TcParams P;
    P.x = x; P.xs_b = xs[0]; P.xs_d = xs[1]; P.xs_t = xs[2];
    P.B = B; P.T = T; P.N = (int64_t)B * T;
    RVQ_REQUIRE(P.N <= (int64_t)INT_MAX / 2, RVQ_EUNSUPPORTED, "rvq_encode_f32: too many frames for one call");
    P.ntiles = (int)((P.N + TILE_M - 1) / TILE_M);
    P.img = pv.img; P.embed = pv.embed; P.cnorm = pv.cnorm; P.meta = pv.meta;
    P.C = pv.C; P.K = K; P.NT = pv.C / 128;
    P.codes = codes;
    P.wire = wire;
    P.quantized = quantized; P.qs_b = (int64_t)T * 128; P.qs_t = 128;  // [B][T][D] contiguous
    P.dbg_acc = dbg_acc; P.dbg_info = dbg_info; P.dbg_stage = dbg_stage;
    P.dbg_time = dbg_info ? reinterpret_cast<long long *>(dbg_info + 128 * 8) : nullptr;  // info_out: 128*8 floats + 96 int64
    int sms = 0;
    int rc = device_sms(&sms);
    if (rc != RVQ_OK) return rc;
    int grid = tc_grid(P.N, sms);
    if (fit_grid_to_workspace) {  // fewer CTAs than SMs: every CTA simply walks more tiles
        const int64_t cap = (workspace_bytes - WS_HEADER) / WS_PER_CTA;
        if (cap < grid) grid = (int)cap;
    }
    RVQ_REQUIRE(workspace && grid >= 1 && workspace_bytes >= WS_HEADER + WS_PER_CTA * grid && ((uintptr_t)workspace & 15) == 0,
                RVQ_EINVAL, "rvq_encode_f32: a 16-byte aligned workspace of rvq_encode_workspace_bytes() bytes is required");
    P.scratch = reinterpret_cast<float *>((char *)workspace + WS_HEADER);
    P.vals = reinterpret_cast<float *>((char *)workspace + WS_HEADER + WS_R_PER_CTA * grid);
#ifdef RVQ_TIMELINE
    if (false) {
#else
    if (dbg_acc) {
#endif
        RVQ_CUDA(cudaFuncSetAttribute(rvq_encode_tc_kernel<true, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_ALLOC));
        rvq_encode_tc_kernel<true, true, false><<<grid, THREADS, SM_ALLOC, stream>>>(P);
    } else if (dbg_info) {  // cycle counters only: the production code path plus clock reads
        RVQ_CUDA(cudaFuncSetAttribute(rvq_encode_tc_kernel<false, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_ALLOC));
        rvq_encode_tc_kernel<false, true, false><<<grid, THREADS, SM_ALLOC, stream>>>(P);
    } else {
        if (quantized) {
            RVQ_CUDA(cudaFuncSetAttribute(rvq_encode_tc_kernel<false, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_ALLOC));
            rvq_encode_tc_kernel<false, false, true><<<grid, THREADS, SM_ALLOC, stream>>>(P);
        } else {
            RVQ_CUDA(cudaFuncSetAttribute(rvq_encode_tc_kernel<false, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_ALLOC));
            rvq_encode_tc_kernel<false, false, false><<<grid, THREADS, SM_ALLOC, stream>>>(P);
        }
    }
    rvq_count_launch(1);
    return rvq_check_launch("rvq_encode_tc_kernel");
}

int rvq_encode_tc(const float *x, const int64_t xs[3], int B, int D, int T, const void *pack, const RvqPackView &pv,
                  int K, int64_t *codes, float *quantized, const RvqWire &wire, void *workspace, int64_t workspace_bytes,
                  cudaStream_t stream, bool fit_grid_to_workspace) {
    (void)pack; (void)D;
    // quantized_out = 0 + q_0 + ... (core_vq.py:363-385) is written by the same kernel, tile by tile
    return launch_tc(x, xs, B, T, pv, K, codes, quantized, wire, workspace, workspace_bytes, nullptr, nullptr, 0, stream,
                     fit_grid_to_workspace);
}

// Diagnostics: raw tensor-core accumulators and scan state of the first 128 frames, stage `stage`.
extern "C" int rvq_debug_tc_scores(const float *x, const int64_t xs[3], int B, int D, int T, const void *pack, int Ktot,
                                   int C, int K, int stage, int64_t *codes, float *acc_out, float *info_out,
                                   void *workspace, int64_t workspace_bytes, void *stream_) {
    RVQ_REQUIRE(x && xs && pack && codes && info_out, RVQ_EINVAL, "rvq_debug_tc_scores: null pointer");
    RVQ_REQUIRE(rvq_tc_shape_supported(D, C) && K > 0 && K <= Ktot && stage >= 0 && stage < K, RVQ_EUNSUPPORTED,
                "rvq_debug_tc_scores: shape not covered by the tensor-core path");
    RvqPackView pv = rvq_pack_view(pack, Ktot, C, D);
    return launch_tc(x, xs, B, T, pv, K, codes, nullptr, RvqWire{nullptr, 0, 0}, workspace, workspace_bytes, acc_out, info_out, stage,
                     (cudaStream_t)stream_);
}

// ------------------------------------------------------------------------------------ no-allocation entry point
// Scratch for rvq_encode_f32_static: the same layout rvq_encode_f32 expects in its caller-provided workspace,
// sized for RVQ_STATIC_MAX_FRAMES frames (16 CTAs of two 128-frame tiles).  It lives in the module image, so
// every GPU that loads the library carries these ~6 MB until the process exits.
static_assert(RVQ_STATIC_MAX_FRAMES % (NSLOT * TILE_M) == 0, "the frame limit must be a whole number of CTA rounds");
static constexpr int STATIC_CTAS = RVQ_STATIC_MAX_FRAMES / (NSLOT * TILE_M);
static __device__ __align__(256) unsigned char g_static_scratch[WS_HEADER + WS_PER_CTA * STATIC_CTAS];

// In host code `g_static_scratch` is only a handle for the runtime, not an address a kernel can use: the device
// address has to be asked for, once per device, with that device current.
static int static_scratch_address(void **out) {
    static std::mutex mu;
    static void *cached[64] = {nullptr};
    int dev = 0;
    RVQ_CUDA(cudaGetDevice(&dev));
    RVQ_REQUIRE(dev >= 0 && dev < 64, RVQ_EUNSUPPORTED, "rvq_encode_f32_static: device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!cached[dev]) {
        void *p = nullptr;
        RVQ_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
        RVQ_REQUIRE(p != nullptr, RVQ_ECUDA, "rvq_encode_f32_static: no device address for the built-in scratch");
        cached[dev] = p;
    }
    *out = cached[dev];
    return RVQ_OK;
}

extern "C" int rvq_encode_f32_static(const float *x, const int64_t xs[3], int B, int D, int T, const void *pack, int Ktot,
                                     int C, int K, int64_t *codes, float *quantized, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    RVQ_REQUIRE(B >= 0 && T >= 0 && D > 0, RVQ_EINVAL, "rvq_encode_f32_static: bad sizes B=%d D=%d T=%d", B, D, T);
    RVQ_REQUIRE(Ktot > 0 && Ktot <= RVQ_MAX_STAGES && C > 0 && K >= 0 && K <= Ktot, RVQ_EINVAL,
                "rvq_encode_f32_static: bad codebook sizes Ktot=%d C=%d n_q=%d", Ktot, C, K);
    const int64_t N = (int64_t)B * T;
    RVQ_REQUIRE(N <= RVQ_STATIC_MAX_FRAMES, RVQ_EINVAL,
                "rvq_encode_f32_static: %lld frames exceed the built-in scratch (at most %d frames per call); "
                "use rvq_encode_f32 with a caller-provided workspace",
                (long long)N, RVQ_STATIC_MAX_FRAMES);
    if (N == 0 || K == 0) return RVQ_OK;
    RVQ_REQUIRE(x && xs && pack && codes, RVQ_EINVAL, "rvq_encode_f32_static: null pointer");
    RVQ_REQUIRE(rvq_tc_shape_supported(D, C), RVQ_EUNSUPPORTED,
                "rvq_encode_f32_static: only the tensor-core shapes (D = 128, C a multiple of 128) use scratch; "
                "for D=%d C=%d call rvq_encode_f32 with RVQ_ALGO_EXACT, which needs no workspace", D, C);
    RVQ_REQUIRE(!quantized || ((uintptr_t)quantized & 15) == 0, RVQ_EINVAL, "rvq_encode_f32_static: quantized must be 16-byte aligned");
    void *scratch = nullptr;
    int rc = static_scratch_address(&scratch);
    if (rc != RVQ_OK) return rc;
    RvqPackView pv = rvq_pack_view(pack, Ktot, C, D);
    // The array still holds what the previous call left: the residual / score areas are fully rewritten before
    // they are read.
    return rvq_encode_tc(x, xs, B, D, T, pack, pv, K, codes, quantized, RvqWire{nullptr, 0, 0}, scratch,
                         (int64_t)sizeof(g_static_scratch), stream,
                         /*fit_grid_to_workspace=*/true);
}
