// rvq_encode.cu -- C-ABI entry points of the encode path and the algorithm dispatch.
#include "rvq_common.cuh"

int rvq_encode_exact(const float *x, const int64_t xs[3], int B, int D, int T, const RvqPackView &pv, int K,
                     int64_t *codes, float *quantized, cudaStream_t stream);
// rvq_encode_tc.cu
bool rvq_tc_shape_supported(int D, int C);
int64_t rvq_tc_workspace_bytes(int B, int D, int T, int K, int C);
int rvq_encode_tc(const float *x, const int64_t xs[3], int B, int D, int T, const void *pack, const RvqPackView &pv,
                  int K, int64_t *codes, float *quantized, void *workspace, int64_t workspace_bytes,
                  cudaStream_t stream, bool fit_grid_to_workspace = false);

extern "C" int64_t rvq_encode_workspace_bytes(int B, int D, int T, int K, int C, int algo) {
    if (B < 0 || D <= 0 || T < 0 || K < 0 || C <= 0) return RVQ_EINVAL;
    if (algo == RVQ_ALGO_EXACT) return 0;
    if (!rvq_tc_shape_supported(D, C)) return algo == RVQ_ALGO_TC ? RVQ_EUNSUPPORTED : 0;
    return rvq_tc_workspace_bytes(B, D, T, K, C);
}

extern "C" int rvq_encode_f32(const float *x, const int64_t xs[3], int B, int D, int T, const void *pack, int Ktot,
                              int C, int K, int64_t *codes, float *quantized, void *workspace,
                              int64_t workspace_bytes, int algo, void *stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    RVQ_REQUIRE(B >= 0 && T >= 0 && D > 0, RVQ_EINVAL, "rvq_encode_f32: bad sizes B=%d D=%d T=%d", B, D, T);
    RVQ_REQUIRE(Ktot > 0 && Ktot <= RVQ_MAX_STAGES && C > 0, RVQ_EINVAL, "rvq_encode_f32: bad pack sizes");
    RVQ_REQUIRE(K >= 0 && K <= Ktot, RVQ_EINVAL, "rvq_encode_f32: n_q=%d but the pack holds %d codebooks", K, Ktot);
    RVQ_REQUIRE(algo == RVQ_ALGO_AUTO || algo == RVQ_ALGO_EXACT || algo == RVQ_ALGO_TC, RVQ_EINVAL,
                "rvq_encode_f32: unknown algo %d", algo);
    if ((int64_t)B * T == 0 || K == 0) return RVQ_OK;
    RVQ_REQUIRE(x && xs && pack && codes, RVQ_EINVAL, "rvq_encode_f32: null pointer");
    RVQ_REQUIRE(D % 4 == 0, RVQ_EUNSUPPORTED, "rvq_encode_f32: dimension %d is not a multiple of 4", D);
    RVQ_REQUIRE(!quantized || ((uintptr_t)quantized & 15) == 0, RVQ_EINVAL, "rvq_encode_f32: quantized must be 16-byte aligned");
    RvqPackView pv = rvq_pack_view(pack, Ktot, C, D);
    const bool tc_ok = rvq_tc_shape_supported(D, C);
    if (algo == RVQ_ALGO_TC)
        RVQ_REQUIRE(tc_ok, RVQ_EUNSUPPORTED, "rvq_encode_f32: the tensor-core path does not cover D=%d C=%d", D, C);
    if (algo == RVQ_ALGO_EXACT || !tc_ok) return rvq_encode_exact(x, xs, B, D, T, pv, K, codes, quantized, stream);
    return rvq_encode_tc(x, xs, B, D, T, pack, pv, K, codes, quantized, workspace, workspace_bytes, stream);
}
