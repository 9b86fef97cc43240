/*
 * rvq_b200.h -- C ABI of librvq_b200.so: the B200 (sm_100a) implementation of the EnCodec
 * residual-vector-quantization tokenizer path of audiocraft/quantization.
 *
 * The reference path is Python (there is no FFI in the reference); every entry point below
 * names the reference function it replaces.  All pointers marked "device" are CUDA device
 * pointers on the CURRENT device; `stream` is a cudaStream_t passed as void* (NULL = default
 * stream).  No entry point allocates device memory: the caller owns every buffer.
 * Every function returns RVQ_OK (0) or a negative RVQ_E* code; rvq_b200_last_error() then
 * returns a thread-local, human readable message.  Launch errors are reported by the call that
 * made the launch; nothing synchronises the device.
 *
 * Shapes (names as in the reference): B batch, D `dimension`, T frames, C `bins`,
 * K active stages (`n_q`), Ktot layers in the module.
 */
#ifndef RVQ_B200_H
#define RVQ_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RVQ_B200_ABI_VERSION 2

#define RVQ_OK 0
#define RVQ_EINVAL (-1)      /* bad argument (shape, stride, null pointer, size) */
#define RVQ_EUNSUPPORTED (-2) /* valid for the reference but not implemented here (e.g. D % 4 != 0) */
#define RVQ_ECUDA (-3)       /* a CUDA runtime / driver call failed */
#define RVQ_EPACK (-4)       /* `pack` was not produced by rvq_pack_codebooks for these sizes */

#define RVQ_MAX_STAGES 64

/* encode algorithms */
#define RVQ_ALGO_AUTO 0   /* tensor-core path when the shape allows it, else exact */
#define RVQ_ALGO_EXACT 1  /* fp32 CUDA-core scan of every code (also the in-kernel fallback) */
#define RVQ_ALGO_TC 2     /* tcgen05 search + exact fp32 re-score; error if the shape is not covered */

int rvq_b200_abi_version(void);
const char *rvq_b200_last_error(void);

/*
 * Codebook pack: private, derived copy of the module's `embed` buffers
 * (core_vq.py:124-127; state_dict keys vq.layers.{i}._codebook.embed).  Holds the fp32 codebooks
 * as one [Ktot][C][D] block, ||E_c||^2 per code (replaces the per-call `embed.pow(2).sum(0)` of
 * core_vq.py:169), fp16 operand copies for the tensor-core search and per-stage metadata.
 * Must be rebuilt whenever an `embed` buffer changes.
 */
int64_t rvq_pack_bytes(int Ktot, int C, int D);

/* embed_ptrs: HOST array of Ktot device pointers, each a contiguous fp32 [C][D] codebook.
 * pack: device buffer of at least rvq_pack_bytes() bytes, 256-byte aligned. */
int rvq_pack_codebooks(const float *const *embed_ptrs, int Ktot, int C, int D, void *pack,
                       int64_t pack_bytes, void *stream);

/*
 * ResidualVectorQuantization.encode (core_vq.py:387-397), all K stages fused
 * (EuclideanCodebook.quantize core_vq.py:164-172 + dequantize :177-179 + residual subtract :394).
 *   pack       built by rvq_pack_codebooks for (Ktot, C, D); the first K <= Ktot stages are used
 *   x          device fp32, element (b,d,t) at x[b*xs[0] + d*xs[1] + t*xs[2]] (strides in elements)
 *   codes      device int64 [K][B][T] contiguous -- the torch.stack layout of core_vq.py:396;
 *              the caller returns codes.transpose(0,1) like vq.py:94
 *   quantized  NULL, or device fp32 [B][T][D] contiguous receiving 0.0f + q_0 + ... + q_{K-1}
 *              (eval ResidualVectorQuantization.forward, core_vq.py:363-385); left untouched when K == 0
 *              or B*T == 0 (the call returns at once: the caller zero-fills if it wants the empty sum)
 *   workspace  device scratch of rvq_encode_workspace_bytes() bytes, 16-byte aligned (may be NULL when
 *              that is 0).  The tensor-core path keeps the residuals of the frame tiles in flight there
 *              (128 KB per resident CTA, L2 resident); contents need not be preserved between calls,
 *              but two encodes that overlap on the GPU must not share it
 */
int64_t rvq_encode_workspace_bytes(int B, int D, int T, int K, int C, int algo);
int rvq_encode_f32(const float *x, const int64_t xs[3], int B, int D, int T, const void *pack,
                   int Ktot, int C, int K, int64_t *codes, float *quantized, void *workspace,
                   int64_t workspace_bytes, int algo, void *stream);

/*
 * rvq_encode_f32 with a second copy of the codes in the wire format of a frame-sharded job (sharded.py; the
 * reference's path is single-device, SURVEY.md 2a, so this output has no counterpart there): codes16 receives
 * the same codes as int16, element (k, b, t) at codes16[k * c16s[0] + b * c16s[1] + t].  The strides let every
 * rank write straight into the send buffer of the all-gather (padded to the largest shard), which removes the
 * narrowing pass between encode and collective.  Requires C <= 32768.  Everything else as rvq_encode_f32.
 */
int rvq_encode_f32_wire(const float *x, const int64_t xs[3], int B, int D, int T, const void *pack,
                        int Ktot, int C, int K, int64_t *codes, int16_t *codes16, const int64_t c16s[2],
                        float *quantized, void *workspace, int64_t workspace_bytes, int algo, void *stream);

/*
 * The other end of that exchange: the all-gathered buffer recv -- W ranks, each [K][smax][T] (axis == 0: the batch
 * axis is sharded, rank r holds items [lo_r, lo_r + n_r), n_r = B / W + (r < B % W), smax = max n_r) or
 * [K][B][smax] (axis == 2: the time axis is sharded the same way) -- becomes the int64 [K][B][T] block that
 * ResidualVectorQuantization.encode returns (core_vq.py:396).  One pass, no intermediate tensors.
 */
int rvq_widen_codes_i16(const int16_t *recv, int W, int K, int B, int T, int axis, int64_t *codes, void *stream);

/*
 * The same exchange without a collective library in the step (no counterpart in the reference): every rank owns a
 * "symmetric" buffer that its peers can address (sharded.py allocates it with torch's symmetric memory and passes
 * the W peer-mapped base addresses as a HOST array).  Layout: rvq_exchange_header_bytes() of flags, then `nslots`
 * slots of `slot_bytes`, each holding W shards of `shard_bytes` (rank r's int16 codes [K][smax][T] or [K][B][smax],
 * padded to a multiple of 16 bytes).  Step seq = 1, 2, ... uses slot seq % nslots; the encode kernel writes this
 * rank's shard of the slot directly (rvq_encode_f32_wire).
 *   rvq_exchange_push   stores this rank's shard into every peer's slot over NVLink and raises this rank's data flag
 *                       there (after waiting, for seq > nslots, until every peer has consumed step seq - nslots)
 *   rvq_exchange_widen  waits for the data flags of all ranks, writes codes int64 [K][B][T], acknowledges the slot
 * Waiting: spin_wait == 0 (normal): the waits are stream memory operations (cuStreamBatchMemOp, WAIT_VALUE_32), no
 * CTA occupies an SM while another GPU is late; spin_wait == 1: the kernels wait themselves and give up after about
 * 50 ms, writing status[0] = 1 (push) / 2 (widen) and status[1] = the rank waited for, and carry on.
 * scratch: 256 zeroed device bytes private to the caller; status: device int64[2], zeroed.
 */
int64_t rvq_exchange_header_bytes(void);
int rvq_exchange_push(void *const *peer_base, int W, int rank, int slot, int nslots, uint32_t seq,
                      int64_t slot_bytes, int64_t shard_bytes, void *scratch, int64_t *status, int spin_wait,
                      void *stream);
int rvq_exchange_widen(void *const *peer_base, int W, int rank, int slot, int nslots, uint32_t seq,
                       int64_t slot_bytes, int64_t shard_bytes, int K, int B, int T, int axis, int64_t *codes,
                       void *scratch, int64_t *status, int spin_wait, void *stream);

/*
 * rvq_encode_f32 for callers that may not allocate.  No counterpart in the reference.
 *
 * rvq_encode_f32 needs a workspace (the residuals and remembered scores of the frame tiles in flight), which
 * its caller normally takes from an allocator.  Some callers have none on the request path: a serving
 * process that planned all of its device memory at start-up and must not grow, an audio host that calls in
 * from a real-time thread where cudaMalloc's device-wide lock and implicit synchronisation are not
 * acceptable, or glue code that only holds raw device pointers owned by another runtime.  For them this
 * entry point uses a scratch array built into the library image instead (about 6 MB per GPU that loads the
 * library, held until the process exits).  It calls no cudaMalloc / cudaFree or any other allocating
 * routine; inputs and outputs are still the caller's.
 *
 *   - at most RVQ_STATIC_MAX_FRAMES frames (B*T) per call; more is refused with RVQ_EINVAL and a message
 *     that names the limit.  Only the tensor-core shapes (D == 128, C % 128 == 0, 256 <= C <= 16384) are
 *     accepted: the exact scan (rvq_encode_f32 with RVQ_ALGO_EXACT) needs no workspace in the first place.
 *   - the first call on each device looks up the device address of the built-in array
 *     (cudaGetSymbolAddress) and remembers it; it may also be the call that makes the CUDA runtime load the
 *     library's device code, array included.  Make that first call a warm-up, outside of the request path and
 *     outside of any stream capture; later calls touch no allocator.  The remembered address does not survive
 *     cudaDeviceReset.
 *   - all calls on a device share the one array, which keeps the previous call's contents (everything
 *     that has to start from zero is cleared by the call itself): two calls that overlap on the GPU --
 *     different streams, or different host threads -- race on it.  Serialise them on one stream.
 * Arguments and results are those of rvq_encode_f32 (codes and, optionally, the quantized sum), bit for bit.
 */
#define RVQ_STATIC_MAX_FRAMES 4096
int rvq_encode_f32_static(const float *x, const int64_t xs[3], int B, int D, int T, const void *pack,
                          int Ktot, int C, int K, int64_t *codes, float *quantized, void *stream);

/*
 * ResidualVectorQuantization.decode (core_vq.py:399-405; F.embedding gather core_vq.py:177-179):
 * out = 0.0f + E_0[c_0] + E_1[c_1] + ... in stage order.
 *   codes      device int64 (code_bytes == 8), int32 (4) or int16 (2, the compact cache format); element (b,k,t) at
 *              codes[b*cs[0] + k*cs[1] + t*cs[2]] (strides in elements)
 *   out        device fp32; element (b,d,t) at out[b*os[0] + d*os[1] + t*os[2]].  The reference's
 *              result is physically [B][T][D], i.e. os = {T*D, 1, D}
 *   status     device int64[2], zeroed by the caller before the first use.  An out-of-range code
 *              (the reference raises IndexError) sets status[0] = 1 and status[1] = 1 + the smallest
 *              flat index (b*K + k)*T + t seen; code 0 is used in its place (the output is then not meaningful:
 *              the host side raises IndexError from the status word).
 */
int rvq_decode_f32(const void *codes, int code_bytes, const int64_t cs[3], int B, int K, int T,
                   const void *pack, int Ktot, int C, int D, float *out, const int64_t os[3],
                   int64_t *status, void *stream);

/*
 * rvq_decode_f32 reading a delayed-pattern sequence in place: Pattern.revert_pattern_sequence
 * (modules/codebooks_patterns.py:224-246) fused into the gather.  seq is the language model's [B][K][S] token block
 * (lm.py:531,577); stage k's code of frame t is seq[b][k][t + steps_late[k]], steps_late[k] = offset + delays[k]
 * (HOST array of K ints; the caller guarantees t + steps_late[k] < S for t < T).  Special tokens (>= C) that a
 * truncated sequence still holds are reported through `status` like any out-of-range code.
 */
int rvq_decode_delayed_f32(const void *seq, int code_bytes, const int64_t ss[3], int B, int K, int T,
                           const int *steps_late, const void *pack, int Ktot, int C, int D, float *out,
                           const int64_t os[3], int64_t *status, void *stream);

/*
 * Code-stream plumbing after the tokenizer.  Integer index shuffles, one pass each.
 *
 * rvq_codes_convert: strided copy between code widths; element (b,k,t) at base[b*s[0] + k*s[1] + t*s[2]]
 *   (strides in elements), widths 8, 4 or 2 bytes.  The reference's compact cache format is
 *   `codes.to(torch.int16)` (modules/jasco_conditioners.py:136).
 * rvq_codes_pack_bits / rvq_codes_unpack_bits: n contiguous int64 codes <-> rvq_codes_packed_bytes(n, bits)
 *   bytes, `bits` bits per code (11 for 2048 bins), little-endian bit order within groups of 8 codes.
 *   No counterpart in the reference (it stores int64 or int16).
 */
int rvq_codes_convert(const void *src, int src_bytes, const int64_t ss[3], int B, int K, int T, void *dst,
                      int dst_bytes, const int64_t ds[3], void *stream);
int64_t rvq_codes_packed_bytes(int64_t n, int bits);
int rvq_codes_pack_bits(const int64_t *codes, int64_t n, int bits, uint8_t *packed, void *stream);
int rvq_codes_unpack_bits(const uint8_t *packed, int64_t n, int bits, int64_t *codes, void *stream);

/*
 * InterleaveStereoCompressionModel.encode / get_left_right_codes (models/encodec.py:474-486):
 *   per_timestep == 0: joint[b][2k + c][t] = c_c[b][k][t]   ('c b k t -> b (k c) t')
 *   per_timestep == 1: joint[b][k][2t + c] = c_c[b][k][t]   ('c b k t -> b k (t c)')
 * c0 / c1: int64, element (b,k,t) at c[b*cs[0] + k*cs[1] + t*cs[2]] (the [B,K,T] views encode returns);
 * joint and the outputs of rvq_stereo_split are contiguous.
 */
int rvq_stereo_interleave(const int64_t *c0, const int64_t *c1, const int64_t cs[3], int B, int K, int T,
                          int per_timestep, int64_t *joint, void *stream);
int rvq_stereo_split(const int64_t *joint, int B, int K, int T, int per_timestep, int64_t *c0, int64_t *c1,
                     void *stream);

/*
 * Pattern.build_pattern_sequence / revert_pattern_sequence (modules/codebooks_patterns.py:160-183, 224-246) for the
 * layouts DelayedPatternProvider.get_pattern produces with flatten_first == 0 (:339-356; ParallelPatternProvider is
 * the all-zero delay case): codebook q appears `offset + delays[q]` steps late, offset = 1 + empty_initial
 * (0 when the provider omits the special token).
 *   build:  values[b][q][s] = z[b][q][s - offset - delays[q]] if that timestep is in [0, T) else special
 *   revert: values[b][q][t] = seq[b][q][t + offset + delays[q]] if that step is in [0, min(S, valid_steps)) else special
 * z / seq: int64, strided like above; values contiguous [B][K][S] / [B][K][T].  indexes (int64 [K][S] / [K][T]) and
 * mask (uint8, same shape) are the reference's scatter tables (flattened source index, K*T or K*S where the special
 * token goes); either may be NULL.  delays: HOST array of K sorted non-negative ints.
 */
int rvq_delay_build(const int64_t *z, const int64_t zs[3], int B, int K, int T, int S, const int *delays,
                    int offset, int64_t special, int64_t *values, int64_t *indexes, uint8_t *mask, void *stream);
int rvq_delay_revert(const int64_t *seq, const int64_t ss[3], int B, int K, int S, int valid_steps, int T,
                     const int *delays, int offset, int64_t special, int64_t *values, int64_t *indexes,
                     uint8_t *mask, void *stream);

/*
 * Training-time codebook statistics, EuclideanCodebook.forward (core_vq.py:205-217) after the search:
 *   counts[c] = number of frames with code c (embed_onehot.sum(0), :209); sums[c] = sum of those frames' inputs in
 *   increasing frame order (x.t() @ embed_onehot, :210); cluster_size and embed_avg take their EMA step
 *   (ema_inplace, :23-24), embed becomes embed_avg / (laplace_smoothing(cluster_size, C, eps) * cluster_size.sum())
 *   (:212-217).  No [N, C] one-hot matrix, no atomics: results are deterministic.
 * x: device fp32, row n at x[n*xs[0] + d*xs[1]] (the flattened channels-last input of the layer); codes: device
 * int64 / int32 [N]; cluster_size [C], embed_avg [C][D], embed [C][D]: the module's buffers, updated in place;
 * counts_out [C] / sums_out [C][D]: optional copies of the raw statistics (NULL to skip).  D <= 1024.
 */
int rvq_ema_update_f32(const float *x, const int64_t xs[2], const void *codes, int code_bytes, int64_t N, int C, int D,
                       float decay, float eps, float *cluster_size, float *embed_avg, float *embed,
                       float *counts_out, float *sums_out, void *stream);

/*
 * Diagnostics (no counterpart in the reference): runs the tensor-core encode and also returns, for
 * the first 128 frames and stage `stage`, the raw approximate scores of every code
 * (acc_out, device fp32 [128][C]) and the per-frame scan state (info_out, device buffer of
 * 128*8 fp32 followed by 96 int64 per-role cycle counters of CTA 0; the fp32 part is [128][8]:
 * ||r||^2, alpha, 2*delta, running maximum, remembered groups, flags, score scale, augmented-column
 * value).  Used by the tests to check the error bound that certifies the exact re-scoring.
 * acc_out may be NULL: then only the cycle counters are collected, on the production code path.
 * workspace: as for rvq_encode_f32 with RVQ_ALGO_TC.
 */
int rvq_debug_tc_scores(const float *x, const int64_t xs[3], int B, int D, int T, const void *pack,
                        int Ktot, int C, int K, int stage, int64_t *codes, float *acc_out,
                        float *info_out, void *workspace, int64_t workspace_bytes, void *stream);

/* The tensor-core encode kernel is persistent, one CTA per SM, and each CTA takes its SM whole.  A frame-sharded job
 * sets aside `n` SMs (0..64, process-wide, default 0) for the small kernels of its code exchange, which could otherwise
 * only run in the gap between two encodes.  Takes effect at the next rvq_encode_workspace_bytes / rvq_encode_f32 pair
 * (call it before sizing the workspace).  No counterpart in the reference. */
int rvq_b200_set_reserved_sms(int n);

/* Measurement aid (no counterpart in the reference): one thread spins on the stream until *host_mapped_flag (an int
 * in pinned, device-visible host memory) becomes non-zero or timeout_ns (at most 2 s) pass.  bench.py starts the
 * timed region of every rank behind such a gate, opened at an agreed wall-clock time: the GPUs are awake and start
 * together (after an idle spell they begin to execute a millisecond or two apart otherwise). */
int rvq_debug_gate(const int *host_mapped_flag, int64_t timeout_ns, void *stream);

/* Introspection used by tests and the bench: number of kernels launched by this library on the
 * calling thread since the last reset (the bench's `gpu_launches`). */
int64_t rvq_b200_launch_count(int reset);

#ifdef __cplusplus
}
#endif
#endif /* RVQ_B200_H */
