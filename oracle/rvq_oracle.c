/*
 * rvq_oracle.c -- CPU ORACLE for the EnCodec residual-vector-quantization path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * leg may load this library.  The product path (audiocraft_b200/) never does.
 *
 * What it restates (reference = /root/reference/audiocraft/quantization):
 *   - EuclideanCodebook.quantize       core_vq.py:164-172
 *   - EuclideanCodebook.dequantize     core_vq.py:177-179
 *   - ResidualVectorQuantization.encode core_vq.py:387-397
 *   - ResidualVectorQuantization.decode core_vq.py:399-405
 *   - (eval) ResidualVectorQuantization.forward core_vq.py:363-385
 *
 * Arithmetic.  The reference evaluates
 *        dist = -( sum(x^2) - (2*x) @ E^T + sum(E^2) )        (fp32, core_vq.py:166-170)
 * with the GEMM and the two norms in whatever summation order ATen/MKL picks.
 * That order is not part of the algorithm, so the oracle fixes ONE order (the
 * "canonical order", below) and is then PINNED against outputs of the unmodified
 * reference (tests/golden/, produced by tests/golden/make_golden.py): all golden
 * codes must be reproduced exactly.
 *
 * Canonical order of a length-D inner product <a,b>:
 *   32 partial sums p[0..31], all +0.0f.  Element d belongs to partial
 *   l = (d % 128) / 4.  Partials are advanced with single-rounded fmaf in
 *   increasing d.  Then five pairwise rounds p[l] = p[l] + p[l ^ off],
 *   off = 16, 8, 4, 2, 1 (all l at once); the result is p[0].
 * (It is the order a 32-lane GPU warp produces when each lane owns 4 consecutive
 *  elements of every 128-element chunk and the lanes finish with a butterfly.)
 *
 * Score of code c for residual r:   s = -(( n2(r) - 2*<r,E_c> ) + n2(E_c)),
 * each operation rounded to fp32 once.  2*<r,E_c> equals <2r,E_c> of the
 * reference exactly (scaling by two commutes with rounding).
 * arg-max: first index among equal scores; a NaN score beats every non-NaN
 * score and the first NaN wins (torch.max semantics, core_vq.py:171).
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off -mavx2 -mfma -pthread).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <pthread.h>
#include <unistd.h>

/* ---- minimal parallel-for over [0, n) with pthreads (libgomp is not in the image) ---- */
typedef void (*range_fn)(int64_t lo, int64_t hi, void *ctx);
typedef struct { range_fn fn; void *ctx; int64_t n, grain; int64_t *next; pthread_mutex_t *mu; } pf_job;

static int g_threads = 0; /* 0 = all online cores */
void rvq_oracle_set_threads(int n) { g_threads = n; }
int rvq_oracle_get_threads(void) {
    if (g_threads > 0) return g_threads;
    long n = sysconf(_SC_NPROCESSORS_ONLN);
    return n > 0 ? (int)n : 1;
}

static void *pf_worker(void *arg) {
    pf_job *j = (pf_job *)arg;
    for (;;) {
        pthread_mutex_lock(j->mu);
        int64_t lo = *j->next;
        *j->next = lo + j->grain;
        pthread_mutex_unlock(j->mu);
        if (lo >= j->n) break;
        int64_t hi = lo + j->grain < j->n ? lo + j->grain : j->n;
        j->fn(lo, hi, j->ctx);
    }
    return NULL;
}

static void parallel_for(int64_t n, int64_t grain, range_fn fn, void *ctx) {
    int nt = rvq_oracle_get_threads();
    if (nt > 64) nt = 64;
    if (n <= grain || nt <= 1) { fn(0, n, ctx); return; }
    int64_t next = 0;
    pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
    pf_job job = { fn, ctx, n, grain, &next, &mu };
    pthread_t th[64];
    int started = 0;
    for (int i = 0; i < nt - 1; ++i)
        if (pthread_create(&th[started], NULL, pf_worker, &job) == 0) ++started;
    pf_worker(&job);
    for (int i = 0; i < started; ++i) pthread_join(th[i], NULL);
}

#define LANES 32
#define CHUNK 128

/* number of padded elements per vector in the permuted layout */
static int64_t padded_dim(int D) { return ((int64_t)(D + CHUNK - 1) / CHUNK) * CHUNK; }

/* permuted layout: out[(chunk*4 + i)*32 + l] = in[chunk*128 + 4*l + i] (zero padded) */
static void permute_vec(const float *in, int64_t in_stride, int D, float *out) {
    int64_t Dp = padded_dim(D);
    memset(out, 0, sizeof(float) * (size_t)Dp);
    for (int d = 0; d < D; ++d) {
        int chunk = d / CHUNK, w = d % CHUNK, l = w / 4, i = w % 4;
        out[((int64_t)chunk * 4 + i) * LANES + l] = in[(int64_t)d * in_stride];
    }
}

/* the butterfly p[l] += p[l ^ off] evaluated for lane 0 only: by commutativity of the
 * fp32 add every lane of the butterfly holds the same bits, so halving is the same tree */
static inline float tree32(float *p) {
    for (int l = 0; l < 16; ++l) p[l] = p[l] + p[l + 16];
    for (int l = 0; l < 8; ++l) p[l] = p[l] + p[l + 8];
    for (int l = 0; l < 4; ++l) p[l] = p[l] + p[l + 4];
    for (int l = 0; l < 2; ++l) p[l] = p[l] + p[l + 2];
    return p[0] + p[1];
}

/* canonical inner product on permuted, zero padded vectors */
static inline float dot_perm(const float *a, const float *b, int64_t Dp) {
    float p[LANES];
    for (int l = 0; l < LANES; ++l) p[l] = 0.0f;
    for (int64_t j = 0; j < Dp; j += LANES)
        for (int l = 0; l < LANES; ++l) p[l] = fmaf(a[j + l], b[j + l], p[l]);
    return tree32(p);
}

static inline int better(float a, float b) { /* a replaces b ? */
    return (a > b) || (isnan(a) && !isnan(b));
}

/* ||E_c||^2 in canonical order.  embed: [K][C][D] contiguous. out: [K][C] */
typedef struct { const float *embed; int D; int64_t Dp; float *out; float *ep; } cn_ctx;

static void cn_range(int64_t lo, int64_t hi, void *vctx) {
    cn_ctx *c = (cn_ctx *)vctx;
    float *tmp = (float *)malloc(sizeof(float) * (size_t)c->Dp);
    for (int64_t kc = lo; kc < hi; ++kc) {
        float *dst = c->ep ? c->ep + kc * c->Dp : tmp;
        permute_vec(c->embed + kc * c->D, 1, c->D, dst);
        c->out[kc] = dot_perm(dst, dst, c->Dp);
    }
    free(tmp);
}

void rvq_oracle_cnorm(const float *embed, int K, int C, int D, float *out) {
    cn_ctx c = { embed, D, padded_dim(D), out, NULL };
    parallel_for((int64_t)K * C, 256, cn_range, &c);
}

/*
 * Encode (core_vq.py:387-397) and, optionally, the eval-forward quantized sum
 * (core_vq.py:363-385 / 399-405).
 *   x         : fp32, element (b,d,t) at x[b*sb + d*sd + t*st]
 *   embed     : [Ktot][C][D] contiguous, the first K stages are used
 *   codes     : int64 [K][B][T] contiguous (the reference's torch.stack layout)
 *   quantized : NULL or fp32 [B][T][D] contiguous: 0.0f + q_0 + q_1 + ... in stage order
 *   scores    : NULL or fp32 [K][B][T] winning score per stage (diagnostics)
 * returns 0.
 */
typedef struct {
    const float *x; int64_t sb, sd, st; int B, D, T; const float *embed; int K, C;
    int64_t *codes; float *quantized; float *scores; const float *ep; const float *cn; int64_t Dp;
} enc_ctx;

static void enc_range(int64_t lo, int64_t hi, void *vctx) {
    enc_ctx *e = (enc_ctx *)vctx;
    const int D = e->D, K = e->K, C = e->C, B = e->B, T = e->T;
    const int64_t Dp = e->Dp;
    float *r = (float *)malloc(sizeof(float) * (size_t)D);
    float *rp = (float *)malloc(sizeof(float) * (size_t)Dp);
    float *q = (float *)malloc(sizeof(float) * (size_t)D);
    for (int64_t n = lo; n < hi; ++n) {
        int64_t b = n / T, t = n % T;
        const float *xp = e->x + b * e->sb + t * e->st;
        for (int d = 0; d < D; ++d) { r[d] = xp[(int64_t)d * e->sd]; q[d] = 0.0f; }
        for (int k = 0; k < K; ++k) {
            permute_vec(r, 1, D, rp);
            float xx = dot_perm(rp, rp, Dp);
            const float *Ek = e->ep + (int64_t)k * C * Dp;
            float best = 0.0f; int64_t bi = -1;
            for (int c = 0; c < C; ++c) {
                float dot = 2.0f * dot_perm(rp, Ek + (int64_t)c * Dp, Dp);
                float s = -((xx - dot) + e->cn[(int64_t)k * C + c]);
                if (bi < 0 || better(s, best)) { best = s; bi = c; }
            }
            e->codes[((int64_t)k * B + b) * T + t] = bi;
            if (e->scores) e->scores[((int64_t)k * B + b) * T + t] = best;
            const float *row = e->embed + ((int64_t)k * C + bi) * D;
            for (int d = 0; d < D; ++d) { r[d] = r[d] - row[d]; q[d] = q[d] + row[d]; }
        }
        if (e->quantized) memcpy(e->quantized + n * D, q, sizeof(float) * (size_t)D);
    }
    free(r); free(rp); free(q);
}

int rvq_oracle_encode(const float *x, int64_t sb, int64_t sd, int64_t st, int B, int D, int T,
                      const float *embed, int K, int C, int64_t *codes, float *quantized,
                      float *scores) {
    const int64_t Dp = padded_dim(D);
    const int64_t N = (int64_t)B * T;
    if (N == 0 || K == 0) return 0;
    /* permuted codebooks + norms, shared by all threads */
    float *ep = (float *)malloc(sizeof(float) * (size_t)K * C * Dp);
    float *cn = (float *)malloc(sizeof(float) * (size_t)K * C);
    cn_ctx c = { embed, D, Dp, cn, ep };
    parallel_for((int64_t)K * C, 256, cn_range, &c);
    enc_ctx e = { x, sb, sd, st, B, D, T, embed, K, C, codes, quantized, scores, ep, cn, Dp };
    parallel_for(N, 16, enc_range, &e);
    free(ep); free(cn);
    return 0;
}

/*
 * Decode (core_vq.py:399-405): out = 0.0f; for k: out = out + E_k[code_k].
 *   codes : int64, element (b,k,t) at codes[b*cb + k*ck + t*ct]
 *   out   : fp32 [B][T][D] contiguous (the reference's physical layout)
 * returns 0, or 1 + flat index (b*K+k)*T+t of the first out-of-range code
 * (the reference raises IndexError there).
 */
typedef struct {
    const int64_t *codes; int64_t cb, ck, ct; int K, T; const float *embed; int C, D; float *out;
} dec_ctx;

static void dec_range(int64_t lo, int64_t hi, void *vctx) {
    dec_ctx *c = (dec_ctx *)vctx;
    const int D = c->D;
    for (int64_t n = lo; n < hi; ++n) {
        int64_t b = n / c->T, t = n % c->T;
        float *o = c->out + n * D;
        for (int d = 0; d < D; ++d) o[d] = 0.0f;
        for (int k = 0; k < c->K; ++k) {
            int64_t code = c->codes[b * c->cb + k * c->ck + t * c->ct];
            const float *row = c->embed + ((int64_t)k * c->C + code) * D;
            for (int d = 0; d < D; ++d) o[d] = o[d] + row[d];
        }
    }
}

int64_t rvq_oracle_decode(const int64_t *codes, int64_t cb, int64_t ck, int64_t ct, int B, int K,
                          int T, const float *embed, int C, int D, float *out) {
    for (int64_t b = 0; b < B; ++b)
        for (int64_t k = 0; k < K; ++k)
            for (int64_t t = 0; t < T; ++t) {
                int64_t c = codes[b * cb + k * ck + t * ct];
                if (c < 0 || c >= C) return 1 + (b * K + k) * T + t;
            }
    dec_ctx c = { codes, cb, ck, ct, K, T, embed, C, D, out };
    parallel_for((int64_t)B * T, 1024, dec_range, &c);
    return 0;
}

/* canonical inner product of two plain vectors (exposed for unit tests) */
float rvq_oracle_dot(const float *a, const float *b, int D) {
    int64_t Dp = padded_dim(D);
    float *ap = (float *)malloc(sizeof(float) * (size_t)Dp);
    float *bp = (float *)malloc(sizeof(float) * (size_t)Dp);
    permute_vec(a, 1, D, ap);
    permute_vec(b, 1, D, bp);
    float r = dot_perm(ap, bp, Dp);
    free(ap); free(bp);
    return r;
}
