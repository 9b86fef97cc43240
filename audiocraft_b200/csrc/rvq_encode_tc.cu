// rvq_encode_tc.cu -- tensor-core (tcgen05) search path.  Placeholder until the kernel lands:
// no shape is reported as supported, so RVQ_ALGO_AUTO resolves to the exact scan.
#include "rvq_common.cuh"

bool rvq_tc_shape_supported(int D, int C) { (void)D; (void)C; return false; }
int64_t rvq_tc_workspace_bytes(int B, int D, int T, int K, int C) { (void)B; (void)D; (void)T; (void)K; (void)C; return 0; }
int rvq_encode_tc(const float *, const int64_t *, int, int, int, const void *, const RvqPackView &, int, int64_t *,
                  float *, void *, int64_t, cudaStream_t) {
    rvq_set_error("rvq_encode_tc: not built");
    return RVQ_EUNSUPPORTED;
}
