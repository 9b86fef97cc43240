// rvq_encode.cu -- C-ABI entry points of the encode path and the algorithm dispatch.
#include "rvq_common.cuh"

int rvq_encode_exact(const float *x, const int64_t xs[3], int B, int D, int T, const RvqPackView &pv, int K,
                     int64_t *codes, float *quantized, const RvqWire &wire, cudaStream_t stream);
// rvq_encode_tc.cu
bool rvq_tc_shape_supported(int D, int C);
int64_t rvq_tc_workspace_bytes(int B, int D, int T, int K, int C);
int rvq_encode_tc(const float *x, const int64_t xs[3], int B, int D, int T, const void *pack, const RvqPackView &pv,
                  int K, int64_t *codes, float *quantized, const RvqWire &wire, void *workspace,
                  int64_t workspace_bytes, cudaStream_t stream, bool fit_grid_to_workspace = false);

extern "C" int64_t rvq_encode_workspace_bytes(int B, int D, int T, int K, int C, int algo) {
    if (B < 0 || D <= 0 || T < 0 || K < 0 || C <= 0) return RVQ_EINVAL;
    if (algo == RVQ_ALGO_EXACT) return 0;
    if (!rvq_tc_shape_supported(D, C)) return algo == RVQ_ALGO_TC ? RVQ_EUNSUPPORTED : 0;
    return rvq_tc_workspace_bytes(B, D, T, K, C);
}

static int encode_dispatch(const char *who, const float *x, const int64_t xs[3], int B, int D, int T, const void *pack,
                           int Ktot, int C, int K, int64_t *codes, float *quantized, const RvqWire &wire,
                           void *workspace, int64_t workspace_bytes, int algo, cudaStream_t stream) {
    RVQ_REQUIRE(B >= 0 && T >= 0 && D > 0, RVQ_EINVAL, "%s: bad sizes B=%d D=%d T=%d", who, B, D, T);
    RVQ_REQUIRE(Ktot > 0 && Ktot <= RVQ_MAX_STAGES && C > 0, RVQ_EINVAL, "%s: bad pack sizes", who);
    RVQ_REQUIRE(K >= 0 && K <= Ktot, RVQ_EINVAL, "%s: n_q=%d but the pack holds %d codebooks", who, K, Ktot);
    RVQ_REQUIRE(algo == RVQ_ALGO_AUTO || algo == RVQ_ALGO_EXACT || algo == RVQ_ALGO_TC, RVQ_EINVAL,
                "%s: unknown algo %d", who, algo);
    if ((int64_t)B * T == 0 || K == 0) return RVQ_OK;
    RVQ_REQUIRE(x && xs && pack && codes, RVQ_EINVAL, "%s: null pointer", who);
    RVQ_REQUIRE(D % 4 == 0, RVQ_EUNSUPPORTED, "%s: dimension %d is not a multiple of 4", who, D);
    RVQ_REQUIRE(!quantized || ((uintptr_t)quantized & 15) == 0, RVQ_EINVAL, "%s: quantized must be 16-byte aligned", who);
    RVQ_REQUIRE(!wire.p || C <= 32768, RVQ_EINVAL, "%s: %d bins do not fit the int16 wire format", who, C);
    RvqPackView pv = rvq_pack_view(pack, Ktot, C, D);
    const bool tc_ok = rvq_tc_shape_supported(D, C);
    if (algo == RVQ_ALGO_TC)
        RVQ_REQUIRE(tc_ok, RVQ_EUNSUPPORTED, "%s: the tensor-core path does not cover D=%d C=%d", who, D, C);
    if (algo == RVQ_ALGO_EXACT || !tc_ok) return rvq_encode_exact(x, xs, B, D, T, pv, K, codes, quantized, wire, stream);
    return rvq_encode_tc(x, xs, B, D, T, pack, pv, K, codes, quantized, wire, workspace, workspace_bytes, stream);
}

extern "C" int rvq_encode_f32(const float *x, const int64_t xs[3], int B, int D, int T, const void *pack, int Ktot,
                              int C, int K, int64_t *codes, float *quantized, void *workspace,
                              int64_t workspace_bytes, int algo, void *stream_) {
    return encode_dispatch("rvq_encode_f32", x, xs, B, D, T, pack, Ktot, C, K, codes, quantized, RvqWire{nullptr, 0, 0},
                           workspace, workspace_bytes, algo, (cudaStream_t)stream_);
}

extern "C" int rvq_encode_f32_wire(const float *x, const int64_t xs[3], int B, int D, int T, const void *pack, int Ktot,
                                   int C, int K, int64_t *codes, int16_t *codes16, const int64_t c16s[2],
                                   float *quantized, void *workspace, int64_t workspace_bytes, int algo,
                                   void *stream_) {
    RVQ_REQUIRE(codes16 && c16s, RVQ_EINVAL, "rvq_encode_f32_wire: null wire buffer");
    RVQ_REQUIRE(c16s[1] >= T && c16s[0] >= (int64_t)B * c16s[1], RVQ_EINVAL,
                "rvq_encode_f32_wire: wire strides (%lld, %lld) overlap for B=%d T=%d", (long long)c16s[0],
                (long long)c16s[1], B, T);
    return encode_dispatch("rvq_encode_f32_wire", x, xs, B, D, T, pack, Ktot, C, K, codes, quantized,
                           RvqWire{codes16, c16s[0], c16s[1]}, workspace, workspace_bytes, algo, (cudaStream_t)stream_);
}
