// rvq_decode.cu -- ResidualVectorQuantization.decode (core_vq.py:399-405) as one gather-sum kernel.
//
//   out[b, :, t] = 0.0f + E_0[codes[b,0,t]] + E_1[codes[b,1,t]] + ...      (stage order, fp32)
//
// HBM-bound: per frame 8*K bytes of codes in and 4*D bytes out; the codebooks (K*C*D*4 bytes,
// 4 MB for MusicGen) stay L2 resident.  One warp owns 32 consecutive frames: it reads their codes
// coalesced (lane = frame), then walks the frames with lane = 4 consecutive dims so that every
// codebook row is one coalesced 128-bit-per-lane read and every output row one coalesced
// 128-bit-per-lane streaming store.  The K row reads of a frame are issued before the first add.
#include "rvq_common.cuh"

template <typename CodeT>
__device__ __forceinline__ long long load_code(const void *codes, int64_t idx) {
    return (long long)reinterpret_cast<const CodeT *>(codes)[idx];
}

__device__ __noinline__ void rvq_flag_bad_code(int64_t *status, long long flat) {
    atomicExch((unsigned long long *)&status[0], 1ull);
    // keep the smallest flat index; 0 means "none yet"
    unsigned long long *s1 = (unsigned long long *)&status[1];
    unsigned long long v = 1ull + (unsigned long long)flat, old = *s1;
    while (old == 0ull || v < old) {
        unsigned long long prev = atomicCAS(s1, old, v);
        if (prev == old) break;
        old = prev;
    }
}

// KT  = K when 1 <= K <= 8 (stage loop fully unrolled, codes of a frame fetched with one or two
//       128-bit shared loads), 0 = any K (chunks of 8).
// DCH = D / 128 when D is 128/256/512 (unrolled), else 0 -> generic chunk loop.
// Dynamic shared memory: int32 codes [warps][32 frames][KP] of the 32 frames a warp is working
// on; an out-of-range code is reported through `status` and replaced by 0.
// STAGED: the K codebooks are copied into shared memory first and every row comes from there (small models only:
// K*C*D*4 bytes must fit beside the codes; the debug model's 4 x 400 x 32 x 4 = 205 KB does).
// koff: per-stage extra offset (in elements) into `codes`: a delayed pattern is read in place (stage k's code of
// frame t sits koff.v[k] / cs_t steps later in the sequence, codebooks_patterns.py:339-356).
struct RvqStageOffsets { int64_t v[RVQ_MAX_STAGES]; };

template <bool STAGED>
__device__ __forceinline__ float4 load_row(const char *p) {
    if (STAGED) return *reinterpret_cast<const float4 *>(p);
    return __ldg(reinterpret_cast<const float4 *>(p));
}

template <typename CodeT, int KT, int DCH, bool STAGED>
__global__ void __launch_bounds__(256)
rvq_decode_kernel(const void *__restrict__ codes, int64_t cs_b, int64_t cs_k, int64_t cs_t, int B, int K, int T,
                  const float *__restrict__ embed_g, int C, int D, float *__restrict__ out, int64_t os_b,
                  int64_t os_d, int64_t os_t, int64_t *__restrict__ status, const RvqStageOffsets koff) {
    extern __shared__ __align__(16) int s_codes[];
    const float *embed = embed_g;
    if (STAGED) {
        // codebooks first (16-byte aligned at the start of dynamic shared memory), the warps' code areas after them
        float4 *cb = reinterpret_cast<float4 *>(s_codes);
        const int64_t n4 = (int64_t)K * C * D / 4;
        for (int64_t i = threadIdx.x; i < n4; i += blockDim.x) cb[i] = __ldg(reinterpret_cast<const float4 *>(embed_g) + i);
        __syncthreads();
        embed = reinterpret_cast<const float *>(cb);
    }
    const int lane = threadIdx.x & 31;
    const int wib = threadIdx.x >> 5;
    const int warps_per_block = blockDim.x >> 5;
    const int KP = KT > 0 ? ((KT + 3) & ~3) : ((K + 3) & ~3);  // codes per frame in smem, padded to 4
    int *wcodes = s_codes + (STAGED ? (size_t)K * C * D : 0) + (size_t)wib * KP * 32;
    const int64_t tblocks = (T + 31) / 32;  // 32-frame groups per batch item
    const int64_t ngroups = (int64_t)B * tblocks;
    const int nchunks = DCH > 0 ? DCH : (D + 127) / 128;
    const int Kc = KT > 0 ? KT : K;
    const uint32_t row_bytes = (uint32_t)D * 4u;
    const uint32_t stage_stride = (uint32_t)C * (uint32_t)D;  // elements per codebook (host checks < 2^31)

    for (int64_t g = (int64_t)blockIdx.x * warps_per_block + wib; g < ngroups;
         g += (int64_t)gridDim.x * warps_per_block) {
        const int64_t b = g / tblocks;
        const int t0 = (int)(g % tblocks) * 32;
        const int nf = min(32, T - t0);
        __syncwarp();
        // lane = frame: coalesced read of the K codes of this lane's frame
        if (lane < nf) {
            const int64_t base = b * cs_b + (int64_t)(t0 + lane) * cs_t;
#pragma unroll
            for (int k = 0; k < (KT > 0 ? KT : 1); ++k) {
                for (int kk = k; kk < Kc; kk += (KT > 0 ? KT : 1)) {
                    long long c = load_code<CodeT>(codes, base + kk * cs_k + koff.v[kk]);
                    if (c < 0 || c >= C) {
                        rvq_flag_bad_code(status, (b * K + kk) * (long long)T + t0 + lane);
                        c = 0;
                    }
                    wcodes[lane * KP + kk] = (int)c;
                }
            }
        }
        __syncwarp();
        // lane = 4 consecutive dims: walk the frames
        float *obase = out + b * os_b + (int64_t)t0 * os_t;
        for (int f = 0; f < nf; ++f) {
            const int *fc = wcodes + f * KP;
            for (int ch = 0; ch < nchunks; ++ch) {
                const int d = ch * 128 + lane * 4;
                if (DCH == 0 && d >= D) break;
                const char *ebase = reinterpret_cast<const char *>(embed + d);
                float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);  // quantized_out = 0.0 (core_vq.py:400)
                if (KT > 0) {
                    int code[8];
                    if ((KT & 3) == 0) {
#pragma unroll
                        for (int j = 0; j < KT; j += 4) {
                            int4 v = *reinterpret_cast<const int4 *>(fc + j);
                            code[j] = v.x; code[j + 1] = v.y; code[j + 2] = v.z; code[j + 3] = v.w;
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < KT; ++j) code[j] = fc[j];
                    }
                    float4 row[8];
#pragma unroll
                    for (int j = 0; j < KT; ++j)
                        row[j] = load_row<STAGED>(
                            ebase + (size_t)((uint32_t)j * stage_stride + (uint32_t)code[j] * (uint32_t)D) * 4u);
#pragma unroll
                    for (int j = 0; j < KT; ++j) {  // stage order (core_vq.py:401-404)
                        acc.x = __fadd_rn(acc.x, row[j].x);
                        acc.y = __fadd_rn(acc.y, row[j].y);
                        acc.z = __fadd_rn(acc.z, row[j].z);
                        acc.w = __fadd_rn(acc.w, row[j].w);
                    }
                } else {
                    for (int k0 = 0; k0 < K; k0 += 8) {
                        float4 row[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            row[j] = make_float4(0.f, 0.f, 0.f, 0.f);
                            if (k0 + j < K)
                                row[j] = load_row<STAGED>(
                                    ebase + ((size_t)(k0 + j) * stage_stride + (size_t)fc[k0 + j] * D) * 4u);
                        }
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            if (k0 + j < K) {
                                acc.x = __fadd_rn(acc.x, row[j].x);
                                acc.y = __fadd_rn(acc.y, row[j].y);
                                acc.z = __fadd_rn(acc.z, row[j].z);
                                acc.w = __fadd_rn(acc.w, row[j].w);
                            }
                        }
                    }
                }
                float *o = obase + (int64_t)f * os_t + (int64_t)d * os_d;
                if (os_d == 1) {
                    __stcs(reinterpret_cast<float4 *>(o), acc);  // streaming: keep the codebooks in L2
                } else {
                    o[0] = acc.x; o[os_d] = acc.y; o[2 * os_d] = acc.z; o[3 * os_d] = acc.w;
                }
            }
        }
        (void)row_bytes;
    }
}

template <typename CodeT, int KT>
static int launch_decode_k(const void *codes, const int64_t cs[3], int B, int K, int T, const RvqPackView &pv,
                           float *out, const int64_t os[3], int64_t *status, const RvqStageOffsets &koff,
                           cudaStream_t stream) {
    const int64_t ngroups = (int64_t)B * ((T + 31) / 32);
    int dev = 0, sms = 148;
    RVQ_CUDA(cudaGetDevice(&dev));
    RVQ_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    const int KP = (K + 3) & ~3;
    const int wpb = KP <= 32 ? 8 : 4;
    const size_t smem_codes = (size_t)wpb * KP * 32 * sizeof(int);
    int64_t blocks = (ngroups + wpb - 1) / wpb;
    // Small models: all K codebooks in shared memory, one persistent CTA per SM, when there are enough frames to
    // pay for the copy (each CTA reads the codebooks once from L2).
    const size_t cb_bytes = (size_t)K * pv.C * pv.D * sizeof(float);
    const bool staged = cb_bytes + smem_codes <= 220 * 1024 && ngroups * 32 >= 64 * (int64_t)K * pv.C;
    const size_t smem = smem_codes + (staged ? cb_bytes : 0);
    const int64_t cap = staged ? (int64_t)sms : (int64_t)sms * 8 * 8;  // grid-stride beyond a few waves of resident CTAs
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
#define RVQ_DEC_LAUNCH(DCH)                                                                                       \
    do {                                                                                                          \
        if (staged) {                                                                                             \
            auto kern = rvq_decode_kernel<CodeT, KT, DCH, true>;                                                  \
            RVQ_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));          \
            kern<<<(unsigned)blocks, wpb * 32, smem, stream>>>(codes, cs[0], cs[1], cs[2], B, K, T, pv.embed, pv.C, \
                                                               pv.D, out, os[0], os[1], os[2], status, koff);      \
        } else {                                                                                                  \
            rvq_decode_kernel<CodeT, KT, DCH, false><<<(unsigned)blocks, wpb * 32, smem, stream>>>(                \
                codes, cs[0], cs[1], cs[2], B, K, T, pv.embed, pv.C, pv.D, out, os[0], os[1], os[2], status, koff); \
        }                                                                                                         \
    } while (0)
    if (pv.D == 128) RVQ_DEC_LAUNCH(1);
    else if (pv.D == 256) RVQ_DEC_LAUNCH(2);
    else if (pv.D == 512) RVQ_DEC_LAUNCH(4);
    else RVQ_DEC_LAUNCH(0);
#undef RVQ_DEC_LAUNCH
    rvq_count_launch(1);
    return rvq_check_launch("rvq_decode_kernel");
}

template <typename CodeT>
static int launch_decode(const void *codes, const int64_t cs[3], int B, int K, int T, const RvqPackView &pv,
                         float *out, const int64_t os[3], int64_t *status, const RvqStageOffsets &koff,
                         cudaStream_t stream) {
    switch (K) {
        case 1: return launch_decode_k<CodeT, 1>(codes, cs, B, K, T, pv, out, os, status, koff, stream);
        case 2: return launch_decode_k<CodeT, 2>(codes, cs, B, K, T, pv, out, os, status, koff, stream);
        case 3: return launch_decode_k<CodeT, 3>(codes, cs, B, K, T, pv, out, os, status, koff, stream);
        case 4: return launch_decode_k<CodeT, 4>(codes, cs, B, K, T, pv, out, os, status, koff, stream);
        case 6: return launch_decode_k<CodeT, 6>(codes, cs, B, K, T, pv, out, os, status, koff, stream);
        case 8: return launch_decode_k<CodeT, 8>(codes, cs, B, K, T, pv, out, os, status, koff, stream);
        default: return launch_decode_k<CodeT, 0>(codes, cs, B, K, T, pv, out, os, status, koff, stream);
    }
}

// K == 0: the reference returns the scalar 0.0; the wrapper never calls with K == 0 on a
// non-empty output, but keep the ABI total: fill with +0.0f.
__global__ void rvq_fill_zero_kernel(float *__restrict__ out, int64_t os_b, int64_t os_d, int64_t os_t, int B,
                                     int D, int T) {
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t total = (int64_t)B * D * T;
    if (i >= total) return;
    int64_t d = i % D, t = (i / D) % T, b = i / ((int64_t)D * T);
    out[b * os_b + d * os_d + t * os_t] = 0.0f;
}

static int decode_common(const char *who, const void *codes, int code_bytes, const int64_t cs[3], int B, int K, int T,
                         const void *pack, int Ktot, int C, int D, float *out, const int64_t os[3], int64_t *status,
                         const int *steps_late, cudaStream_t stream) {
    RVQ_REQUIRE(B >= 0 && K >= 0 && T >= 0, RVQ_EINVAL, "%s: negative size", who);
    if ((int64_t)B * T == 0) return RVQ_OK;
    RVQ_REQUIRE(pack && out && os && status, RVQ_EINVAL, "%s: null pointer", who);
    RVQ_REQUIRE(Ktot > 0 && Ktot <= RVQ_MAX_STAGES && C > 0 && D > 0, RVQ_EINVAL, "%s: bad pack sizes", who);
    RVQ_REQUIRE(K <= Ktot, RVQ_EINVAL, "%s: %d code streams but only %d codebooks", who, K, Ktot);
    RVQ_REQUIRE(D % 4 == 0, RVQ_EUNSUPPORTED, "%s: dimension %d is not a multiple of 4", who, D);
    RVQ_REQUIRE(os[1] != 1 || (((uintptr_t)out & 15) == 0 && os[0] % 4 == 0 && os[2] % 4 == 0), RVQ_EINVAL,
                "%s: out must be 16-byte aligned with strides that are multiples of 4", who);
    RVQ_REQUIRE((int64_t)Ktot * C * D < ((int64_t)1 << 30), RVQ_EUNSUPPORTED,
                "%s: codebooks of %d x %d x %d elements exceed the 32-bit offsets of the kernel", who, Ktot, C, D);
    RvqPackView pv = rvq_pack_view(pack, Ktot, C, D);
    if (K == 0) {
        int64_t total = (int64_t)B * D * T;
        rvq_fill_zero_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(out, os[0], os[1], os[2], B, D, T);
        rvq_count_launch(1);
        return rvq_check_launch("rvq_fill_zero_kernel");
    }
    RVQ_REQUIRE(codes && cs, RVQ_EINVAL, "%s: null codes", who);
    RVQ_REQUIRE(code_bytes == 8 || code_bytes == 4 || code_bytes == 2, RVQ_EINVAL, "%s: codes must be int64, int32 or int16", who);
    RvqStageOffsets koff;
    for (int k = 0; k < RVQ_MAX_STAGES; ++k) koff.v[k] = 0;
    if (steps_late)
        for (int k = 0; k < K; ++k) {
            RVQ_REQUIRE(steps_late[k] >= 0, RVQ_EINVAL, "%s: negative delay", who);
            koff.v[k] = (int64_t)steps_late[k] * cs[2];
        }
    if (code_bytes == 8) return launch_decode<long long>(codes, cs, B, K, T, pv, out, os, status, koff, stream);
    if (code_bytes == 4) return launch_decode<int>(codes, cs, B, K, T, pv, out, os, status, koff, stream);
    return launch_decode<short>(codes, cs, B, K, T, pv, out, os, status, koff, stream);
}

extern "C" int rvq_decode_f32(const void *codes, int code_bytes, const int64_t cs[3], int B, int K, int T,
                              const void *pack, int Ktot, int C, int D, float *out, const int64_t os[3],
                              int64_t *status, void *stream_) {
    return decode_common("rvq_decode_f32", codes, code_bytes, cs, B, K, T, pack, Ktot, C, D, out, os, status, nullptr,
                         (cudaStream_t)stream_);
}

extern "C" int rvq_decode_delayed_f32(const void *seq, int code_bytes, const int64_t ss[3], int B, int K, int T,
                                      const int *steps_late, const void *pack, int Ktot, int C, int D, float *out,
                                      const int64_t os[3], int64_t *status, void *stream_) {
    RVQ_REQUIRE(steps_late || K == 0, RVQ_EINVAL, "rvq_decode_delayed_f32: null delays");
    return decode_common("rvq_decode_delayed_f32", seq, code_bytes, ss, B, K, T, pack, Ktot, C, D, out, os, status,
                         steps_late, (cudaStream_t)stream_);
}
