// rvq_train_stats.cu -- the training-time statistics of EuclideanCodebook.forward (core_vq.py:205-217) without the
// [N, C] one-hot matrix and without atomics:
//
//   counts[c]  = #{n : code[n] == c}                       (reference: embed_onehot.sum(0), :209)
//   sums[c][:] = sum over those n, in increasing n, of x[n][:]   (reference: x.t() @ embed_onehot, :210)
//   cluster_size = cluster_size * decay + counts * (1 - decay)    (ema_inplace, core_vq.py:23-24)
//   embed_avg    = embed_avg * decay + sums * (1 - decay)
//   embed        = embed_avg / (laplace_smoothing(cluster_size, C, eps) * cluster_size.sum())   (:212-217)
//
// One CTA per code: it scans the N codes in order, compacts the matching frame indices into shared memory with a
// stable block scan (so the list is sorted), then thread d adds x[idx][d] over the list in that order.  The result
// is deterministic and independent of the launch geometry: the fp32 sum of a code's inputs in frame order (the
// reference's GEMM sums in an unspecified order; the tests allow 1e-5 relative).  A second small kernel forms
// cluster_size.sum() (fixed order, every CTA recomputes it) and the normalised codebook.
#include "rvq_common.cuh"

namespace {

constexpr int TPB = 256;
constexpr int LIST = 2048;  // matching frames gathered per round; more frames -> more rounds

template <typename CodeT>
__global__ void __launch_bounds__(TPB) stats_ema_kernel(const float *__restrict__ x, int64_t xs_n, int64_t xs_d,
                                                        const CodeT *__restrict__ codes, int64_t N, int C, int D, float decay,
                                                        float *__restrict__ cluster_size, float *__restrict__ embed_avg,
                                                        float *__restrict__ counts_out, float *__restrict__ sums_out) {
    __shared__ int s_list[LIST];
    __shared__ int s_warp[TPB / 32];
    __shared__ int s_fill;
    const int c = blockIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // threads own dims d = tid, tid + TPB, ... (D <= 4 * TPB)
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    long long count = 0;
    if (threadIdx.x == 0) s_fill = 0;
    __syncthreads();
    auto flush = [&](int n_items) {
        for (int i = 0; i < n_items; ++i) {
            const int64_t n = s_list[i];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int d = threadIdx.x + j * TPB;
                if (d < D) acc[j] = __fadd_rn(acc[j], x[n * xs_n + (int64_t)d * xs_d]);
            }
        }
    };
    for (int64_t n0 = 0; n0 < N; n0 += TPB) {
        const int64_t n = n0 + threadIdx.x;
        const bool hit = n < N && (long long)codes[n] == (long long)c;
        const unsigned ballot = __ballot_sync(0xffffffffu, hit);
        if (lane == 0) s_warp[warp] = __popc(ballot);
        __syncthreads();
        int before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < TPB / 32; ++w) {
            const int v = s_warp[w];
            before += w < warp ? v : 0;
            total += v;
        }
        const int fill = s_fill;
        if (fill + total > LIST) {  // uniform: make room first
            __syncthreads();
            flush(fill);
            __syncthreads();
            if (threadIdx.x == 0) s_fill = 0;
            __syncthreads();
        }
        const int base = s_fill;
        if (hit) s_list[base + before + __popc(ballot & ((1u << lane) - 1u))] = (int)n;
        __syncthreads();
        if (threadIdx.x == 0) s_fill = base + total;
        count += total;
        __syncthreads();
    }
    flush(s_fill);
    // EMA (ema_inplace: moving_avg.mul_(decay).add_(new, alpha=1 - decay))
    const float om = 1.0f - decay;
    if (threadIdx.x == 0) {
        const float cnt = (float)count;
        if (counts_out) counts_out[c] = cnt;
        cluster_size[c] = __fmaf_rn(cnt, om, __fmul_rn(cluster_size[c], decay));
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int d = threadIdx.x + j * TPB;
        if (d < D) {
            if (sums_out) sums_out[(int64_t)c * D + d] = acc[j];
            float *ea = embed_avg + (int64_t)c * D + d;
            *ea = __fmaf_rn(acc[j], om, __fmul_rn(*ea, decay));
        }
    }
}

__global__ void __launch_bounds__(TPB) normalise_kernel(const float *__restrict__ cluster_size, const float *__restrict__ embed_avg,
                                                        int C, int D, float eps, float *__restrict__ embed) {
    __shared__ float s_part[TPB];
    // cluster_size.sum(): thread t adds elements t, t + TPB, ... in order, then a fixed tree
    float p = 0.f;
    for (int i = threadIdx.x; i < C; i += TPB) p = __fadd_rn(p, cluster_size[i]);
    s_part[threadIdx.x] = p;
    __syncthreads();
    for (int off = TPB / 2; off >= 1; off >>= 1) {
        if ((int)threadIdx.x < off) s_part[threadIdx.x] = __fadd_rn(s_part[threadIdx.x], s_part[threadIdx.x + off]);
        __syncthreads();
    }
    const float total = s_part[0];
    const int c = blockIdx.x;
    // laplace_smoothing(x, n, eps) = (x + eps) / (x.sum() + n * eps); times x.sum()  (core_vq.py:32-33, 212-215)
    const float smoothed = __fmul_rn(__fdiv_rn(__fadd_rn(cluster_size[c], eps), __fadd_rn(total, __fmul_rn((float)C, eps))), total);
    for (int d = threadIdx.x; d < D; d += TPB) embed[(int64_t)c * D + d] = __fdiv_rn(embed_avg[(int64_t)c * D + d], smoothed);
}

}  // namespace

extern "C" int rvq_ema_update_f32(const float *x, const int64_t xs[2], const void *codes, int code_bytes, int64_t N, int C, int D,
                                  float decay, float eps, float *cluster_size, float *embed_avg, float *embed,
                                  float *counts_out, float *sums_out, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    RVQ_REQUIRE(N >= 0 && C > 0 && D > 0 && D <= 4 * TPB, RVQ_EINVAL, "rvq_ema_update_f32: bad sizes N=%lld C=%d D=%d",
                (long long)N, C, D);
    RVQ_REQUIRE(N < ((int64_t)1 << 31), RVQ_EUNSUPPORTED, "rvq_ema_update_f32: too many frames");
    RVQ_REQUIRE(cluster_size && embed_avg && embed && xs && (N == 0 || (x && codes)), RVQ_EINVAL, "rvq_ema_update_f32: null pointer");
    RVQ_REQUIRE(code_bytes == 8 || code_bytes == 4, RVQ_EINVAL, "rvq_ema_update_f32: codes must be int64 or int32");
    if (code_bytes == 8)
        stats_ema_kernel<long long><<<C, TPB, 0, stream>>>(x, xs[0], xs[1], (const long long *)codes, N, C, D, decay, cluster_size,
                                                           embed_avg, counts_out, sums_out);
    else
        stats_ema_kernel<int><<<C, TPB, 0, stream>>>(x, xs[0], xs[1], (const int *)codes, N, C, D, decay, cluster_size, embed_avg,
                                                     counts_out, sums_out);
    rvq_count_launch(1);
    int rc = rvq_check_launch("rvq_ema_update_f32 (statistics)");
    if (rc != RVQ_OK) return rc;
    normalise_kernel<<<C, TPB, 0, stream>>>(cluster_size, embed_avg, C, D, eps, embed);
    rvq_count_launch(1);
    return rvq_check_launch("rvq_ema_update_f32 (normalise)");
}
