// rvq_codestream.cu -- the integer plumbing either side of the tokenizer: what happens to the codes between
// ResidualVectorQuantizer.encode and its consumers.  Pure index shuffles on int64 / int32 /
// int16 codes, bandwidth bound; every kernel reads and writes along the contiguous time axis.
//
//   rvq_widen_codes_i16     the all-gathered int16 codes of a frame-sharded job -> [K][B][T] int64 (sharded.py)
//   rvq_codes_convert       strided copy between code widths (int64 / int32 / int16); the compact cache format of
//                           the reference is `codes.to(torch.int16)` (modules/jasco_conditioners.py:136)
//   rvq_codes_pack_bits     n codes -> ceil(n * bits / 8) bytes (11 bits per code for 2048 bins), and back
//   rvq_stereo_interleave   models/encodec.py:474-479  'c b k t -> b (k c) t' / 'c b k t -> b k (t c)', and back (:483-486)
//   rvq_delay_build/revert  modules/codebooks_patterns.py:160-183 / :224-246 for the delayed patterns of :339-356
#include "rvq_common.cuh"

namespace {

constexpr int TPB = 256;

__device__ __forceinline__ int64_t load_code(const void *p, int bytes, int64_t i) {
    if (bytes == 8) return reinterpret_cast<const int64_t *>(p)[i];
    if (bytes == 4) return reinterpret_cast<const int32_t *>(p)[i];
    return reinterpret_cast<const int16_t *>(p)[i];
}
__device__ __forceinline__ void store_code(void *p, int bytes, int64_t i, int64_t v) {
    if (bytes == 8) reinterpret_cast<int64_t *>(p)[i] = v;
    else if (bytes == 4) reinterpret_cast<int32_t *>(p)[i] = (int32_t)v;
    else reinterpret_cast<int16_t *>(p)[i] = (int16_t)v;
}

// shard r of n items over W ranks holds base + (r < extra) items (sharded.py: shard_sizes)
__device__ __forceinline__ void shard_of(int c, int base, int extra, int &r, int &off) {
    const int big = extra * (base + 1);
    if (c < big) { r = c / (base + 1); off = c - r * (base + 1); }
    else { const int d = c - big; r = extra + d / base; off = d - (r - extra) * base; }
}

__global__ void __launch_bounds__(TPB) widen_kernel(const int16_t *__restrict__ recv, int W, int K, int B, int T, int axis,
                                                    int64_t *__restrict__ out) {
    const int n = axis == 0 ? B : T, base = n / W, extra = n % W, smax = base + (extra ? 1 : 0);
    const int64_t total = (int64_t)K * B * T;
    for (int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x; i < total; i += (int64_t)gridDim.x * TPB) {
        const int t = (int)(i % T);
        const int b = (int)((i / T) % B);
        const int k = (int)(i / ((int64_t)T * B));
        int r, off;
        int64_t src;
        if (axis == 0) {
            shard_of(b, base, extra, r, off);
            src = (((int64_t)r * K + k) * smax + off) * T + t;
        } else {
            shard_of(t, base, extra, r, off);
            src = (((int64_t)r * K + k) * B + b) * smax + off;
        }
        out[i] = (int64_t)recv[src];
    }
}

__global__ void __launch_bounds__(TPB) convert_kernel(const void *__restrict__ src, int sbytes, int64_t s0, int64_t s1,
                                                      int64_t s2, int B, int K, int T, void *__restrict__ dst, int dbytes,
                                                      int64_t d0, int64_t d1, int64_t d2) {
    const int64_t total = (int64_t)B * K * T;
    for (int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x; i < total; i += (int64_t)gridDim.x * TPB) {
        const int64_t t = i % T, k = (i / T) % K, b = i / ((int64_t)T * K);
        store_code(dst, dbytes, b * d0 + k * d1 + t * d2, load_code(src, sbytes, b * s0 + k * s1 + t * s2));
    }
}

// 8 codes -> `bits` bytes, little-endian bit order (code j of a group occupies bits [j*bits, (j+1)*bits))
__global__ void __launch_bounds__(TPB) pack_bits_kernel(const int64_t *__restrict__ codes, int64_t n, int bits,
                                                        uint8_t *__restrict__ out) {
    const int64_t groups = (n + 7) / 8;
    for (int64_t g = (int64_t)blockIdx.x * TPB + threadIdx.x; g < groups; g += (int64_t)gridDim.x * TPB) {
        unsigned __int128 acc = 0;
        for (int j = 0; j < 8; ++j) {
            const int64_t i = g * 8 + j;
            const unsigned long long v = i < n ? (unsigned long long)codes[i] & ((1ull << bits) - 1ull) : 0ull;
            acc |= (unsigned __int128)v << (j * bits);
        }
        const int64_t nbytes_total = (n * bits + 7) / 8;
        for (int j = 0; j < bits; ++j) {
            const int64_t o = g * bits + j;
            if (o < nbytes_total) out[o] = (uint8_t)(acc >> (8 * j));
        }
    }
}
__global__ void __launch_bounds__(TPB) unpack_bits_kernel(const uint8_t *__restrict__ in, int64_t n, int bits,
                                                          int64_t *__restrict__ codes) {
    const int64_t groups = (n + 7) / 8;
    const int64_t nbytes_total = (n * bits + 7) / 8;
    for (int64_t g = (int64_t)blockIdx.x * TPB + threadIdx.x; g < groups; g += (int64_t)gridDim.x * TPB) {
        unsigned __int128 acc = 0;
        for (int j = 0; j < bits; ++j) {
            const int64_t o = g * bits + j;
            if (o < nbytes_total) acc |= (unsigned __int128)in[o] << (8 * j);
        }
        for (int j = 0; j < 8; ++j) {
            const int64_t i = g * 8 + j;
            if (i < n) codes[i] = (int64_t)((unsigned long long)(acc >> (j * bits)) & ((1ull << bits) - 1ull));
        }
    }
}

// out[b][2k + c][t] (per_timestep == 0) or out[b][k][2t + c] (per_timestep == 1) = channel c's code (b, k, t)
__global__ void __launch_bounds__(TPB) stereo_kernel(const int64_t *__restrict__ c0, const int64_t *__restrict__ c1,
                                                     int64_t a0, int64_t a1, int64_t a2, int B, int K, int T,
                                                     int per_timestep, int64_t *__restrict__ joint, int split,
                                                     int64_t *__restrict__ o0, int64_t *__restrict__ o1) {
    const int64_t total = (int64_t)B * K * T * 2;
    for (int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x; i < total; i += (int64_t)gridDim.x * TPB) {
        // i indexes the joint tensor, contiguous
        int64_t b, k, t;
        int c;
        if (per_timestep) {
            c = (int)(i & 1);
            t = (i >> 1) % T; k = ((i >> 1) / T) % K; b = (i >> 1) / ((int64_t)T * K);
        } else {
            t = i % T;
            const int64_t kc = (i / T) % (2 * K);
            c = (int)(kc & 1); k = kc >> 1; b = i / ((int64_t)T * 2 * K);
        }
        if (!split) {
            joint[i] = (c ? c1 : c0)[b * a0 + k * a1 + t * a2];
        } else {
            (c ? o1 : o0)[(b * K + k) * T + t] = joint[i];
        }
    }
}

// build:  out[b][q][s] = z[b][q][s - offset - delay_q] when that timestep lies in [0, T), else `special`
// revert: out[b][q][t] = seq[b][q][t + offset + delay_q] when that step lies in [0, S), else `special`
// The [K][len] index and mask tables of the reference are written by the b == 0 slice.
struct Delays { int d[RVQ_MAX_STAGES]; };

__global__ void __launch_bounds__(TPB) delay_kernel(const int64_t *__restrict__ in, int64_t i0, int64_t i1, int64_t i2, int B,
                                                    int K, int Lin, int limit, int Lout, int offset, Delays delays,
                                                    int64_t special, int revert, int64_t *__restrict__ out, int64_t *__restrict__ indexes,
                                                    uint8_t *__restrict__ mask) {
    const int64_t total = (int64_t)B * K * Lout;
    for (int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x; i < total; i += (int64_t)gridDim.x * TPB) {
        const int64_t p = i % Lout, q = (i / Lout) % K, b = i / ((int64_t)Lout * K);
        const int64_t shift = (int64_t)offset + delays.d[q];
        const int64_t src = revert ? p + shift : p - shift;
        const bool ok = src >= 0 && src < limit;
        out[i] = ok ? in[b * i0 + q * i1 + src * i2] : special;
        if (b == 0) {
            if (indexes) indexes[q * Lout + p] = ok ? src + q * (int64_t)Lin : (int64_t)K * Lin;
            if (mask) mask[q * Lout + p] = ok ? 1 : 0;
        }
    }
}

unsigned grid_for(int64_t total) {
    int64_t g = (total + TPB - 1) / TPB;
    const int64_t cap = 148 * 16;  // a few waves; grid-stride beyond that
    return (unsigned)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace

extern "C" int rvq_widen_codes_i16(const int16_t *recv, int W, int K, int B, int T, int axis, int64_t *codes,
                                   void *stream) {
    RVQ_REQUIRE(W >= 1 && K >= 0 && B >= 0 && T >= 0 && (axis == 0 || axis == 2), RVQ_EINVAL,
                "rvq_widen_codes_i16: bad sizes W=%d K=%d B=%d T=%d axis=%d", W, K, B, T, axis);
    const int64_t total = (int64_t)K * B * T;
    if (total == 0) return RVQ_OK;
    RVQ_REQUIRE(recv && codes, RVQ_EINVAL, "rvq_widen_codes_i16: null pointer");
    RVQ_REQUIRE((axis == 0 ? B : T) >= W, RVQ_EINVAL, "rvq_widen_codes_i16: fewer items than ranks on the sharded axis");
    widen_kernel<<<grid_for(total), TPB, 0, (cudaStream_t)stream>>>(recv, W, K, B, T, axis, codes);
    rvq_count_launch(1);
    return rvq_check_launch("rvq_widen_codes_i16");
}

extern "C" int rvq_codes_convert(const void *src, int src_bytes, const int64_t ss[3], int B, int K, int T, void *dst,
                                 int dst_bytes, const int64_t ds[3], void *stream) {
    RVQ_REQUIRE((src_bytes == 8 || src_bytes == 4 || src_bytes == 2) && (dst_bytes == 8 || dst_bytes == 4 || dst_bytes == 2),
                RVQ_EINVAL, "rvq_codes_convert: code widths must be 2, 4 or 8 bytes");
    RVQ_REQUIRE(B >= 0 && K >= 0 && T >= 0, RVQ_EINVAL, "rvq_codes_convert: bad sizes");
    const int64_t total = (int64_t)B * K * T;
    if (total == 0) return RVQ_OK;
    RVQ_REQUIRE(src && dst && ss && ds, RVQ_EINVAL, "rvq_codes_convert: null pointer");
    convert_kernel<<<grid_for(total), TPB, 0, (cudaStream_t)stream>>>(src, src_bytes, ss[0], ss[1], ss[2], B, K, T, dst,
                                                                     dst_bytes, ds[0], ds[1], ds[2]);
    rvq_count_launch(1);
    return rvq_check_launch("rvq_codes_convert");
}

extern "C" int64_t rvq_codes_packed_bytes(int64_t n, int bits) {
    if (n < 0 || bits < 1 || bits > 16) return RVQ_EINVAL;
    return (n * bits + 7) / 8;
}

extern "C" int rvq_codes_pack_bits(const int64_t *codes, int64_t n, int bits, uint8_t *packed, void *stream) {
    RVQ_REQUIRE(n >= 0 && bits >= 1 && bits <= 16, RVQ_EINVAL, "rvq_codes_pack_bits: bad sizes n=%lld bits=%d", (long long)n, bits);
    if (n == 0) return RVQ_OK;
    RVQ_REQUIRE(codes && packed, RVQ_EINVAL, "rvq_codes_pack_bits: null pointer");
    pack_bits_kernel<<<grid_for((n + 7) / 8), TPB, 0, (cudaStream_t)stream>>>(codes, n, bits, packed);
    rvq_count_launch(1);
    return rvq_check_launch("rvq_codes_pack_bits");
}

extern "C" int rvq_codes_unpack_bits(const uint8_t *packed, int64_t n, int bits, int64_t *codes, void *stream) {
    RVQ_REQUIRE(n >= 0 && bits >= 1 && bits <= 16, RVQ_EINVAL, "rvq_codes_unpack_bits: bad sizes n=%lld bits=%d", (long long)n, bits);
    if (n == 0) return RVQ_OK;
    RVQ_REQUIRE(codes && packed, RVQ_EINVAL, "rvq_codes_unpack_bits: null pointer");
    unpack_bits_kernel<<<grid_for((n + 7) / 8), TPB, 0, (cudaStream_t)stream>>>(packed, n, bits, codes);
    rvq_count_launch(1);
    return rvq_check_launch("rvq_codes_unpack_bits");
}

extern "C" int rvq_stereo_interleave(const int64_t *c0, const int64_t *c1, const int64_t cs[3], int B, int K, int T,
                                     int per_timestep, int64_t *joint, void *stream) {
    RVQ_REQUIRE(B >= 0 && K >= 0 && T >= 0, RVQ_EINVAL, "rvq_stereo_interleave: bad sizes");
    const int64_t total = (int64_t)B * K * T * 2;
    if (total == 0) return RVQ_OK;
    RVQ_REQUIRE(c0 && c1 && cs && joint, RVQ_EINVAL, "rvq_stereo_interleave: null pointer");
    stereo_kernel<<<grid_for(total), TPB, 0, (cudaStream_t)stream>>>(c0, c1, cs[0], cs[1], cs[2], B, K, T, per_timestep ? 1 : 0,
                                                                    joint, 0, nullptr, nullptr);
    rvq_count_launch(1);
    return rvq_check_launch("rvq_stereo_interleave");
}

extern "C" int rvq_stereo_split(const int64_t *joint, int B, int K, int T, int per_timestep, int64_t *c0, int64_t *c1,
                                void *stream) {
    RVQ_REQUIRE(B >= 0 && K >= 0 && T >= 0, RVQ_EINVAL, "rvq_stereo_split: bad sizes");
    const int64_t total = (int64_t)B * K * T * 2;
    if (total == 0) return RVQ_OK;
    RVQ_REQUIRE(c0 && c1 && joint, RVQ_EINVAL, "rvq_stereo_split: null pointer");
    stereo_kernel<<<grid_for(total), TPB, 0, (cudaStream_t)stream>>>(nullptr, nullptr, 0, 0, 0, B, K, T, per_timestep ? 1 : 0,
                                                                    const_cast<int64_t *>(joint), 1, c0, c1);
    rvq_count_launch(1);
    return rvq_check_launch("rvq_stereo_split");
}

static int delay_common(const char *who, const int64_t *in, const int64_t is[3], int B, int K, int Lin, int limit, int Lout,
                        const int *delays, int offset, int64_t special, int revert, int64_t *out, int64_t *indexes,
                        uint8_t *mask, void *stream) {
    RVQ_REQUIRE(B >= 0 && K >= 1 && K <= RVQ_MAX_STAGES && Lin >= 0 && Lout >= 0 && offset >= 0, RVQ_EINVAL,
                "%s: bad sizes B=%d K=%d in=%d out=%d offset=%d", who, B, K, Lin, Lout, offset);
    RVQ_REQUIRE(delays, RVQ_EINVAL, "%s: null delays", who);
    Delays d;
    for (int q = 0; q < K; ++q) {
        RVQ_REQUIRE(delays[q] >= 0 && (q == 0 || delays[q] >= delays[q - 1]), RVQ_EINVAL,
                    "%s: delays must be non-negative and sorted (codebooks_patterns.py:336-337)", who);
        d.d[q] = delays[q];
    }
    const int64_t total = (int64_t)(B > 0 ? B : 1) * K * Lout;  // B == 0 still fills the tables
    if (total == 0 || (B == 0 && !indexes && !mask)) return RVQ_OK;
    RVQ_REQUIRE((in && is && out) || B == 0, RVQ_EINVAL, "%s: null pointer", who);
    if (B == 0) return RVQ_OK;
    delay_kernel<<<grid_for(total), TPB, 0, (cudaStream_t)stream>>>(in, is[0], is[1], is[2], B, K, Lin, limit < Lin ? limit : Lin, Lout, offset, d,
                                                                   special,
                                                                   revert, out, indexes, mask);
    rvq_count_launch(1);
    return rvq_check_launch(who);
}

extern "C" int rvq_delay_build(const int64_t *z, const int64_t zs[3], int B, int K, int T, int S, const int *delays,
                               int offset, int64_t special, int64_t *values, int64_t *indexes, uint8_t *mask,
                               void *stream) {
    return delay_common("rvq_delay_build", z, zs, B, K, T, T, S, delays, offset, special, 0, values, indexes, mask, stream);
}

extern "C" int rvq_delay_revert(const int64_t *seq, const int64_t ss[3], int B, int K, int S, int valid_steps, int T,
                                const int *delays, int offset, int64_t special, int64_t *values, int64_t *indexes, uint8_t *mask,
                                void *stream) {
    return delay_common("rvq_delay_revert", seq, ss, B, K, S, valid_steps, T, delays, offset, special, 1, values, indexes, mask, stream);
}
