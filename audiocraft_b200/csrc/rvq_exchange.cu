// rvq_exchange.cu -- the code exchange of a frame-sharded job without a collective library in the step: every rank
// stores its int16 codes straight into every peer's receive buffer over NVLink (peer-mapped symmetric memory) and
// raises a flag there; the widening kernel of each rank waits for the flags of all ranks and turns the gathered
// buffer into the [K][B][T] int64 block.  Two small launches per step on a side stream, no host round trip, no
// library kernel that has to find room beside the persistent encode kernel.  (The reference has no multi-GPU path:
// SURVEY.md 2a; design in SURVEY.md 5 "stretch" and 7.4.)
//
// Symmetric buffer of a rank (same layout on every rank, allocated and exchanged by sharded.py):
//   [0, 16384)                      flags, uint32, 64 bytes apart: data[slot][src rank] at byte 64 * (slot * W + src),
//                                   ack[slot][consumer rank] at byte 8192 + 64 * (slot * W + consumer)
//   16384 + slot * slot_bytes + r * shard_bytes   rank r's codes of the step that uses `slot`
// A step is numbered seq = 1, 2, ... and uses slot seq % nslots.  Flags only grow.
//   push(seq):  wait until every peer has acknowledged step seq - nslots (it has widened the slot's previous
//               contents), copy my shard into every rank's slot (mine included), then data[slot][me] = seq everywhere
//   widen(seq): wait until data[slot][r] >= seq for every r, widen, then ack[slot][me] = seq everywhere
// The waits are stream memory operations (cuStreamBatchMemOp, WAIT_VALUE_32 with GEQ): the stream itself holds the
// kernels back until the flags are there, so no CTA sits on an SM spinning while it waits for another GPU -- the
// persistent encode kernel of the next step needs every SM whole.  (spin_wait = 1 selects an in-kernel wait instead,
// which gives up after about 50 ms, records that in the status word and carries on.)
#include <cuda.h>

#include "rvq_common.cuh"

int rvq_reserved_sms();  // rvq_encode_tc.cu

namespace {

constexpr int MAXW = 16;
// The exchange kernels run on the SMs a sharded job keeps free of the persistent encode kernel
// (rvq_b200_set_reserved_sms): few enough blocks that all of them are resident at once, since every block waits for
// the flags before it starts to work.
constexpr int TPB = 256;
constexpr int BLOCKS_PER_SM = 8;
constexpr int64_t FLAG_BYTES = 16384;  // data flags in the first half, acknowledgements in the second
constexpr int64_t ACK_OFFSET = FLAG_BYTES / 2;
constexpr long long TIMEOUT_NS = 50LL * 1000 * 1000;

struct Peers {
    unsigned char *base[MAXW];  // peer-mapped base address of every rank's symmetric buffer (mine at [rank])
};

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ long long now_ns() {
    long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ uint32_t *data_flag(unsigned char *base, int W, int slot, int src) {
    return reinterpret_cast<uint32_t *>(base) + 16 * (slot * W + src);
}
__device__ __forceinline__ uint32_t *ack_flag(unsigned char *base, int W, int slot, int consumer) {
    return reinterpret_cast<uint32_t *>(base + ACK_OFFSET) + 16 * (slot * W + consumer);
}
// thread p < W of the block waits for flag p; everybody leaves together.  Returns false after a timeout.
__device__ bool wait_flags(bool acks, unsigned char *mine, int W, int slot, uint32_t want, int64_t *status, int what) {
    __shared__ int ok;
    if (threadIdx.x == 0) ok = 1;
    __syncthreads();
    if ((int)threadIdx.x < W && want > 0) {
        const uint32_t *f = acks ? ack_flag(mine, W, slot, (int)threadIdx.x) : data_flag(mine, W, slot, (int)threadIdx.x);
        const long long t0 = now_ns();
        while ((int32_t)(ld_acquire_sys(f) - want) < 0) {
            if (now_ns() - t0 > TIMEOUT_NS) {
                ok = 0;
                status[0] = what;                      // 1: push timed out waiting for an ack, 2: widen waiting for data
                status[1] = (int64_t)threadIdx.x;      // the rank that did not show up
                break;
            }
            __nanosleep(200);
        }
    }
    __syncthreads();
    return ok != 0;
}

// the last block to get here (device-scope ticket) runs `f`; every block has fenced its remote stores before
__device__ __forceinline__ bool last_block(unsigned int *ticket) {
    __shared__ int last;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int t = atomicAdd(ticket, 1u);
        last = (t == gridDim.x - 1);
        if (last) *ticket = 0;
    }
    __syncthreads();
    return last != 0;
}

__global__ void __launch_bounds__(TPB) push_kernel(Peers peers, int W, int rank, int slot, int nslots, uint32_t seq,
                                                   int64_t slot_bytes, int64_t shard_bytes, unsigned int *ticket,
                                                   int64_t *status, int spin) {
    unsigned char *mine = peers.base[rank];
    if (spin) wait_flags(true, mine, W, slot, seq > (uint32_t)nslots ? seq - (uint32_t)nslots : 0u, status, 1);
    const int64_t off = FLAG_BYTES + (int64_t)slot * slot_bytes + (int64_t)rank * shard_bytes;
    const uint4 *src = reinterpret_cast<const uint4 *>(mine + off);
    const int64_t n16 = shard_bytes / 16;
    for (int p = 0; p < W; ++p) {
        if (p == rank) continue;  // my own shard is already in place: the encode kernel wrote it there
        uint4 *dst = reinterpret_cast<uint4 *>(peers.base[p] + off);
        const int64_t stride = (int64_t)gridDim.x * TPB;
        for (int64_t i = (int64_t)blockIdx.x * TPB + threadIdx.x; i < n16; i += 4 * stride) {
            uint4 v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (i + j * stride < n16) v[j] = src[i + j * stride];
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (i + j * stride < n16) dst[i + j * stride] = v[j];
        }
    }
    if (last_block(ticket) && (int)threadIdx.x < W)
        st_release_sys(data_flag(peers.base[threadIdx.x], W, slot, rank), seq);
}

__device__ __forceinline__ void shard_of(int c, int base, int extra, int &r, int &off) {
    const int big = extra * (base + 1);
    if (c < big) { r = c / (base + 1); off = c - r * (base + 1); }
    else { const int d = c - big; r = extra + d / base; off = d - (r - extra) * base; }
}

// len int16 -> int64, one warp; four elements per lane and step when both sides are aligned for it
__device__ __forceinline__ void widen_row(int64_t *dst, const int16_t *src, int len, int lane) {
    if ((((uintptr_t)src & 7) | ((uintptr_t)dst & 15)) == 0) {
        const int n4 = len / 4;
        for (int i = lane; i < n4; i += 32) {
            const short4 v = reinterpret_cast<const short4 *>(src)[i];
            longlong2 a, c;
            a.x = v.x; a.y = v.y; c.x = v.z; c.y = v.w;
            reinterpret_cast<longlong2 *>(dst)[2 * i] = a;
            reinterpret_cast<longlong2 *>(dst)[2 * i + 1] = c;
        }
        for (int t = n4 * 4 + lane; t < len; t += 32) dst[t] = (int64_t)src[t];
    } else {
        for (int t = lane; t < len; t += 32) dst[t] = (int64_t)src[t];
    }
}

__global__ void __launch_bounds__(TPB) widen_wait_kernel(Peers peers, int W, int rank, int slot, uint32_t seq, int64_t slot_bytes,
                                                         int64_t shard_bytes, int K, int B, int T, int axis,
                                                         int64_t *__restrict__ out, unsigned int *ticket, int64_t *status, int spin) {
    unsigned char *mine = peers.base[rank];
    if (spin) wait_flags(false, mine, W, slot, seq, status, 2);
    const unsigned char *slot_base = mine + FLAG_BYTES + (int64_t)slot * slot_bytes;
    const int n = axis == 0 ? B : T, base = n / W, extra = n % W, smax = base + (extra ? 1 : 0);
    // one warp per row (k, b): the time axis is contiguous on both sides, no division per element
    const int64_t rows = (int64_t)K * B;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = TPB / 32;
    for (int64_t row = (int64_t)blockIdx.x * wpb + wib; row < rows; row += (int64_t)gridDim.x * wpb) {
        const int b = (int)(row % B), k = (int)(row / B);
        int64_t *dst = out + row * T;
        if (axis == 0) {
            int r, off;
            shard_of(b, base, extra, r, off);
            const int16_t *src = reinterpret_cast<const int16_t *>(slot_base + (int64_t)r * shard_bytes) + ((int64_t)k * smax + off) * T;
            widen_row(dst, src, T, lane);
        } else {
            int lo = 0;
            for (int r = 0; r < W; ++r) {
                const int len = base + (r < extra ? 1 : 0);
                const int16_t *src = reinterpret_cast<const int16_t *>(slot_base + (int64_t)r * shard_bytes) + ((int64_t)k * B + b) * smax;
                widen_row(dst + lo, src, len, lane);
                lo += len;
            }
        }
    }
    if (last_block(ticket) && (int)threadIdx.x < W)
        st_release_sys(ack_flag(peers.base[threadIdx.x], W, slot, rank), seq);
}

int64_t exchange_blocks() {
    const int sms = rvq_reserved_sms();
    return (int64_t)(sms > 0 ? sms : 4) * BLOCKS_PER_SM;
}

int fill_peers(const char *who, void *const *peer_base, int W, int rank, Peers &p) {
    RVQ_REQUIRE(peer_base && W >= 1 && W <= MAXW && rank >= 0 && rank < W, RVQ_EINVAL, "%s: bad ranks (W=%d rank=%d, at most %d)",
                who, W, rank, MAXW);
    for (int i = 0; i < MAXW; ++i) p.base[i] = nullptr;
    for (int i = 0; i < W; ++i) {
        RVQ_REQUIRE(peer_base[i] && ((uintptr_t)peer_base[i] & 15) == 0, RVQ_EINVAL, "%s: peer buffer %d is null or unaligned", who, i);
        p.base[i] = (unsigned char *)peer_base[i];
    }
    return RVQ_OK;
}

// stream-ordered wait until flag[p] >= want for every p != rank (cyclic comparison); flags live in MY buffer
int stream_wait_flags(const char *who, unsigned char *mine, bool acks, int W, int rank, int slot, uint32_t want, cudaStream_t stream) {
    if (want == 0 || W == 1) return RVQ_OK;
    CUstreamBatchMemOpParams ops[MAXW];
    int n = 0;
    for (int p = 0; p < W; ++p) {
        if (p == rank) continue;
        const uintptr_t addr = (uintptr_t)mine + (acks ? ACK_OFFSET : 0) + 64u * (uintptr_t)(slot * W + p);
        ops[n] = CUstreamBatchMemOpParams();
        ops[n].operation = CU_STREAM_MEM_OP_WAIT_VALUE_32;
        ops[n].waitValue.address = (CUdeviceptr)addr;
        ops[n].waitValue.value = want;
        ops[n].waitValue.flags = CU_STREAM_WAIT_VALUE_GEQ;
        ++n;
    }
    // the driver entry point is looked up through the runtime, so that the library itself does not depend on libcuda
    // (it must load, for its host-only entry points, on machines without a driver)
    typedef CUresult (*BatchMemOpFn)(CUstream, unsigned int, CUstreamBatchMemOpParams *, unsigned int);
    static BatchMemOpFn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuStreamBatchMemOp", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (BatchMemOpFn)p;
    }();
    RVQ_REQUIRE(fn != nullptr, RVQ_ECUDA, "%s: the driver does not offer cuStreamBatchMemOp", who);
    CUresult r = fn((CUstream)stream, (unsigned)n, ops, 0);
    RVQ_REQUIRE(r == CUDA_SUCCESS, RVQ_ECUDA, "%s: cuStreamBatchMemOp failed with CUresult %d", who, (int)r);
    return RVQ_OK;
}

}  // namespace

__global__ void gate_kernel(const volatile int *flag, long long timeout_ns) {
    const long long t0 = now_ns();
    while (*flag == 0 && now_ns() - t0 < timeout_ns) __nanosleep(500);
}

extern "C" int rvq_debug_gate(const int *host_mapped_flag, int64_t timeout_ns, void *stream) {
    RVQ_REQUIRE(host_mapped_flag && timeout_ns > 0 && timeout_ns <= 2000000000LL, RVQ_EINVAL, "rvq_debug_gate: bad arguments");
    gate_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(host_mapped_flag, timeout_ns);
    return rvq_check_launch("rvq_debug_gate");
}

extern "C" int64_t rvq_exchange_header_bytes(void) { return FLAG_BYTES; }

extern "C" int rvq_exchange_push(void *const *peer_base, int W, int rank, int slot, int nslots, uint32_t seq,
                                 int64_t slot_bytes, int64_t shard_bytes, void *scratch, int64_t *status, int spin_wait,
                                 void *stream) {
    Peers p;
    int rc = fill_peers("rvq_exchange_push", peer_base, W, rank, p);
    if (rc != RVQ_OK) return rc;
    RVQ_REQUIRE(nslots >= 1 && slot >= 0 && slot < nslots && (int64_t)nslots * W * 64 <= ACK_OFFSET && seq >= 1 && shard_bytes > 0 &&
                    shard_bytes % 16 == 0 && slot_bytes >= (int64_t)W * shard_bytes && slot_bytes % 16 == 0 && scratch && status,
                RVQ_EINVAL, "rvq_exchange_push: bad slot geometry");
    const int64_t n16 = shard_bytes / 16;
    int64_t blocks = (n16 + TPB * 4 - 1) / (TPB * 4);
    const int64_t cap = exchange_blocks();
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    const int spin = spin_wait ? 1 : 0;
    if (!spin) {
        rc = stream_wait_flags("rvq_exchange_push", p.base[rank], true, W, rank, slot,
                               seq > (uint32_t)nslots ? seq - (uint32_t)nslots : 0u, (cudaStream_t)stream);
        if (rc != RVQ_OK) return rc;
    }
    push_kernel<<<(unsigned)blocks, TPB, 0, (cudaStream_t)stream>>>(p, W, rank, slot, nslots, seq, slot_bytes, shard_bytes,
                                                                   (unsigned int *)scratch, status, spin);
    rvq_count_launch(1);
    return rvq_check_launch("rvq_exchange_push");
}

extern "C" int rvq_exchange_widen(void *const *peer_base, int W, int rank, int slot, int nslots, uint32_t seq,
                                  int64_t slot_bytes, int64_t shard_bytes, int K, int B, int T, int axis, int64_t *codes,
                                  void *scratch, int64_t *status, int spin_wait, void *stream) {
    Peers p;
    int rc = fill_peers("rvq_exchange_widen", peer_base, W, rank, p);
    if (rc != RVQ_OK) return rc;
    RVQ_REQUIRE(nslots >= 1 && slot >= 0 && slot < nslots && seq >= 1 && K >= 1 && B >= 1 && T >= 1 && (axis == 0 || axis == 2) &&
                    codes && scratch && status && (axis == 0 ? B : T) >= W,
                RVQ_EINVAL, "rvq_exchange_widen: bad arguments");
    const int64_t total = (int64_t)K * B * T;
    int64_t blocks = ((int64_t)K * B + TPB / 32 - 1) / (TPB / 32);
    const int64_t cap = exchange_blocks();
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    const int spin = spin_wait ? 1 : 0;
    if (!spin) {
        rc = stream_wait_flags("rvq_exchange_widen", p.base[rank], false, W, rank, slot, seq, (cudaStream_t)stream);
        if (rc != RVQ_OK) return rc;
    }
    widen_wait_kernel<<<(unsigned)blocks, TPB, 0, (cudaStream_t)stream>>>(p, W, rank, slot, seq, slot_bytes, shard_bytes, K, B, T,
                                                                         axis, codes, (unsigned int *)scratch + 16, status, spin);
    rvq_count_launch(1);
    return rvq_check_launch("rvq_exchange_widen");
}
