/*
 * shelfi_oracle.c — CPU ORACLE (test infrastructure, never on the product path).
 * See shelfi_oracle.h for scope and parity status.  Citations "ckks.cpp:NN" are into
 * /root/reference/palisade_pybind/SHELFI_FHE/src/ckks.cpp; "[PAL]" marks PALISADE v1.11 semantics
 * restated from its published algorithm (the library itself is not in /root/reference).
 *
 * Build: see oracle/Makefile (-O3 -fopenmp -ffp-contract=off: PALISADE's FFT does not fuse
 * multiply-adds, and bit-equality of llround() results with the device depends on that).
 */
#include "shelfi_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef unsigned __int128 u128;
typedef __int128 i128;

/* ------------------------------------------------------------------------------------------- */
/* modular arithmetic                                                                           */
/* ------------------------------------------------------------------------------------------- */
static inline uint64_t mulmod(uint64_t a, uint64_t b, uint64_t q) { return (uint64_t)((u128)a * b % q); }
static inline uint64_t addmod(uint64_t a, uint64_t b, uint64_t q) {
    uint64_t s = a + b;
    return s >= q ? s - q : s;
}
static inline uint64_t submod(uint64_t a, uint64_t b, uint64_t q) { return a >= b ? a - b : a + q - b; }
static uint64_t powmod(uint64_t a, uint64_t e, uint64_t q) {
    uint64_t r = 1;
    a %= q;
    while (e) {
        if (e & 1) r = mulmod(r, a, q);
        a = mulmod(a, a, q);
        e >>= 1;
    }
    return r;
}
static inline uint64_t invmod(uint64_t a, uint64_t q) { return powmod(a, q - 2, q); }
/* Shoup companion floor(w * 2^64 / q) and the preconditioned product, fully reduced —
 * [PAL] NativeInteger::PrepModMulConst / ModMulFastConst */
static inline uint64_t shoup_pre(uint64_t w, uint64_t q) { return (uint64_t)(((u128)w << 64) / q); }
static inline uint64_t shoup_mul(uint64_t x, uint64_t w, uint64_t wp, uint64_t q) {
    uint64_t h = (uint64_t)(((u128)x * wp) >> 64);
    uint64_t r = x * w - h * q;
    return r >= q ? r - q : r;
}
static inline uint64_t signed_mod(int64_t x, uint64_t q) {
    if (x >= 0) return (uint64_t)x % q;
    uint64_t m = ((uint64_t)0 - (uint64_t)x) % q;
    return m ? q - m : 0;
}

static int is_prime64(uint64_t n) {
    if (n < 2) return 0;
    static const uint64_t small[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};
    for (size_t i = 0; i < sizeof small / sizeof *small; i++) {
        if (n == small[i]) return 1;
        if (n % small[i] == 0) return 0;
    }
    uint64_t d = n - 1;
    int s = 0;
    while (!(d & 1)) d >>= 1, s++;
    for (size_t i = 0; i < sizeof small / sizeof *small; i++) { /* deterministic for 64 bit */
        uint64_t x = powmod(small[i], d, n);
        if (x == 1 || x == n - 1) continue;
        int comp = 1;
        for (int r = 1; r < s; r++) {
            x = mulmod(x, x, n);
            if (x == n - 1) { comp = 0; break; }
        }
        if (comp) return 0;
    }
    return 1;
}

/* [PAL] FirstPrime<NativeInteger>(nBits, m): smallest prime q >= 2^nBits with q = 1 (mod m) */
uint64_t orc_first_prime(uint32_t bits, uint32_t m) {
    uint64_t r = (uint64_t)1 << bits;
    uint64_t q = r + (m - (r % m)) + 1; /* [PAL] qNew = 2^nBits + (m - 2^nBits mod m) + 1 */
    while (!is_prime64(q)) q += m;
    return q;
}
/* [PAL] PreviousPrime(q, m) */
uint64_t orc_previous_prime(uint64_t q, uint32_t m) {
    uint64_t p = q - m;
    while (!is_prime64(p)) p -= m;
    return p;
}
/* [PAL] RootOfUnity(m, q): of all primitive m-th roots mod q return the numerically smallest
 * (rule verified against both fixture contexts). m is a power of two here. */
uint64_t orc_min_root_of_unity(uint32_t m, uint64_t q) {
    uint64_t g = 0;
    for (uint64_t x = 2; x < q; x++) { /* any element of order exactly m */
        uint64_t c = powmod(x, (q - 1) / m, q);
        if (powmod(c, m / 2, q) == q - 1) { g = c; break; }
    }
    uint64_t g2 = mulmod(g, g, q), cur = g, best = g;
    for (uint32_t i = 1; i < m / 2; i++) { /* g^(odd) enumerates all primitive roots */
        cur = mulmod(cur, g2, q);
        if (cur < best) best = cur;
    }
    return best;
}
/* [PAL] LPAlgorithmParamsGenCKKS::ParamsGen ring-dimension choice (ckks.cpp:26 passes multDepth = 1),
 * HYBRID key switching, 60-bit first modulus and 60-bit auxiliary moduli, HEStd_128_classic, ternary
 * secrets (StdLatticeParm table).  numLargeDigits defaults to 3 above depth 3, 2 above depth 0, else 1. */
uint32_t orc_num_large_digits(uint32_t mult_depth) { return mult_depth > 3 ? 3 : mult_depth > 0 ? 2 : 1; }
uint32_t orc_select_ring_dim_ex(uint32_t scale_bits, uint32_t mult_depth) {
    static const uint32_t dims[] = {1024, 2048, 4096, 8192, 16384, 32768};
    static const uint32_t maxq[] = {27, 54, 109, 218, 438, 881};
    uint32_t dnum = orc_num_large_digits(mult_depth);
    uint32_t qb = 60 + mult_depth * scale_bits;
    uint32_t part = (qb + dnum - 1) / dnum; /* ceil(qBound / numPartQ) */
    qb += ((part + 59) / 60) * 60;
    uint32_t n = 0, prev = 0;
    for (int i = 0; i < 6; i++) {
        if (qb <= maxq[i] && qb > prev) { n = dims[i]; break; }
        prev = maxq[i];
    }
    if (!n) n = qb <= 27 ? 1024 : 65536;
    return n;
}
uint32_t orc_select_ring_dim(uint32_t scale_bits, uint32_t batch) {
    (void)batch; /* [PAL] batch > n/2 throws; callers check */
    return orc_select_ring_dim_ex(scale_bits, 1);
}
/* [PAL] ParamsGen modulus chain under EXACTRESCALE, firstModSize = 60: q_last = FirstPrime(bits, M); the inner
 * moduli alternate just below / just above the running scaling factor sf_{i} = sf_{i+1}^2 / q_{i+1} so the
 * scale stays close to 2^bits after every rescale; q_0 = PreviousPrime(FirstPrime(60, M), M).  For two limbs
 * (all that ckks.cpp:26 produces) this is the rule verified on the fixtures; the alternation for deeper chains
 * is restated from PALISADE v1.11 and has no fixture => parity unpinned. */
static uint64_t next_prime_step(uint64_t q, uint32_t m) { do q += m; while (!is_prime64(q)); return q; }
int orc_modulus_chain(uint32_t scale_bits, uint32_t mult_depth, uint32_t ring_dim, uint64_t *q) {
    uint32_t M = 2 * ring_dim;
    int L = (int)mult_depth + 1;
    if (L < 1 || L > ORC_MAX_LIMBS) return -1;
    q[L - 1] = orc_first_prime(scale_bits, M);
    if (L > 1) {
        double sf = (double)q[L - 1];
        uint32_t cnt = 0;
        for (int i = L - 2; i >= 1; i--) {
            sf = pow(sf, 2) / (double)q[i + 1];
            uint64_t sf_int = (uint64_t)llround(sf), rem = sf_int % M, cand;
            int clash = 1;
            if ((cnt % 2) == 0) {
                cand = sf_int - M - rem + 1;
                while (clash) { clash = 0; cand = orc_previous_prime(cand, M); for (int j = i + 1; j < L; j++) if (cand == q[j]) clash = 1; }
            } else {
                cand = sf_int + M - rem + 1;
                while (clash) { clash = 0; cand = next_prime_step(cand, M); for (int j = i + 1; j < L; j++) if (cand == q[j]) clash = 1; }
            }
            q[i] = cand;
            cnt++;
        }
        q[0] = orc_previous_prime(orc_first_prime(60, M), M);
    }
    return L;
}

/* ------------------------------------------------------------------------------------------- */
/* context                                                                                      */
/* ------------------------------------------------------------------------------------------- */
typedef struct {
    uint32_t M, N;
    int L;
    uint64_t q[ORC_MAX_LIMBS], psi[ORC_MAX_LIMBS];
    uint32_t bigQ[110];
    uint16_t bigQ_msb;
    uint64_t ptm;
    uint32_t batch;
    float sigma, alpha, hermite;
    int nints;
    uint32_t ints[9];
    uint32_t enabled;
} wire_params;

struct orc_ctx {
    uint32_t batch, scale_bits;
    char cryptodir[1024];
    char err[256];
    int have_cc, have_pk, have_sk;
    wire_params wp;
    char key_tag[33];
    uint64_t *pk_b, *pk_a, *sk_s;                /* L*N each, EVALUATION form */
    uint64_t *pk_bp, *pk_ap, *sk_sp;             /* Shoup companions */
    uint64_t *tw[ORC_MAX_LIMBS], *twp[ORC_MAX_LIMBS];   /* psi^bitrev(i) and companion */
    uint64_t *itw[ORC_MAX_LIMBS], *itwp[ORC_MAX_LIMBS]; /* psi^-bitrev(i) */
    uint64_t ninv[ORC_MAX_LIMBS], ninvp[ORC_MAX_LIMBS];
    uint32_t *rot_group; /* 5^j mod M, j < N/2 */
    double *ksi;         /* (cos,sin)(2 pi k / M), k <= M */
    uint64_t gauss_cdf[32];
    /* the three semantics the reference's fixtures do not pin, switchable here and in the engine alike */
    int opt_evalmult_round; /* 0: (int64)(c*Delta + 0.5)  [PAL v1.11 EvalMult(ct, double)];  1: llround(c*Delta) */
    int opt_decode_noise;   /* 0: off;  1: [PAL >= 1.10.6] Decode adds Gaussian noise before FFTSpecial */
    uint64_t noise_seed;    /* test key of the decode noise (the engine draws a fresh key per call unless seeded) */
};

static uint32_t bitrev(uint32_t x, int bits) {
    uint32_t r = 0;
    for (int i = 0; i < bits; i++) r |= ((x >> i) & 1u) << (bits - 1 - i);
    return r;
}
static int ilog2(uint32_t x) {
    int l = 0;
    while ((1u << l) < x) l++;
    return l;
}

static void free_tables(orc_ctx *c) {
    for (int i = 0; i < ORC_MAX_LIMBS; i++) {
        free(c->tw[i]); free(c->twp[i]); free(c->itw[i]); free(c->itwp[i]);
        c->tw[i] = c->twp[i] = c->itw[i] = c->itwp[i] = NULL;
    }
    free(c->rot_group); free(c->ksi);
    c->rot_group = NULL; c->ksi = NULL;
}

/* Discrete Gaussian CDF table shared with the device sampler:
 * cdf[k] = floor(2^63 * P(|X| <= k)), P(X = x) ~ exp(-x^2 / (2 sigma^2)), k = 0..31. */
static void build_gauss_cdf(double sigma, uint64_t cdf[32]) {
    double rho[64], total = 0.0;
    for (int k = 0; k < 64; k++) rho[k] = exp(-(double)k * (double)k / (2.0 * sigma * sigma));
    total = rho[0];
    for (int k = 1; k < 64; k++) total += 2.0 * rho[k];
    double acc = 0.0;
    for (int k = 0; k < 32; k++) {
        acc += (k == 0 ? rho[0] : 2.0 * rho[k]) / total;
        double v = ldexp(acc, 63);
        cdf[k] = (v >= 9223372036854775807.0) ? (UINT64_MAX >> 1) : (uint64_t)v;
    }
    cdf[31] = UINT64_MAX >> 1;
}

static void build_tables(orc_ctx *c) {
    free_tables(c);
    const wire_params *w = &c->wp;
    uint32_t N = w->N, M = w->M;
    int lg = ilog2(N);
    for (int l = 0; l < w->L; l++) {
        uint64_t q = w->q[l], psi = w->psi[l], ipsi = invmod(psi, q);
        c->tw[l] = malloc(N * 8); c->twp[l] = malloc(N * 8);
        c->itw[l] = malloc(N * 8); c->itwp[l] = malloc(N * 8);
        uint64_t p = 1, ip = 1;
        for (uint32_t i = 0; i < N; i++) { /* [PAL] rootOfUnityReverseTable: index bitrev(i) holds psi^i */
            uint32_t r = bitrev(i, lg);
            c->tw[l][r] = p; c->twp[l][r] = shoup_pre(p, q);
            c->itw[l][r] = ip; c->itwp[l][r] = shoup_pre(ip, q);
            p = mulmod(p, psi, q); ip = mulmod(ip, ipsi, q);
        }
        c->ninv[l] = invmod(N % q, q);
        c->ninvp[l] = shoup_pre(c->ninv[l], q);
    }
    /* [PAL] DiscreteFourierTransform::PreComputeTable(M): rotGroup[i] = 5^i mod M, ksiPows[j] = e^{2 pi i j / M} */
    c->rot_group = malloc((N / 2) * sizeof(uint32_t));
    uint32_t five = 1;
    for (uint32_t i = 0; i < N / 2; i++) { c->rot_group[i] = five; five = (uint32_t)(((uint64_t)five * 5) % M); }
    c->ksi = malloc((size_t)(M + 1) * 2 * sizeof(double));
    for (uint32_t j = 0; j < M; j++) {
        double angle = 2.0 * M_PI * j / M;
        c->ksi[2 * j] = cos(angle);
        c->ksi[2 * j + 1] = sin(angle);
    }
    c->ksi[2 * M] = c->ksi[0];
    c->ksi[2 * M + 1] = c->ksi[1];
    build_gauss_cdf((double)w->sigma, c->gauss_cdf);
}

orc_ctx *orc_create(uint32_t batch, uint32_t scale_bits, const char *cryptodir) {
    orc_ctx *c = calloc(1, sizeof *c);
    c->batch = batch;
    c->scale_bits = scale_bits;
    snprintf(c->cryptodir, sizeof c->cryptodir, "%s", cryptodir ? cryptodir : "");
    return c;
}
void orc_destroy(orc_ctx *c) {
    if (!c) return;
    free_tables(c);
    free(c->pk_b); free(c->pk_a); free(c->sk_s); free(c->pk_bp); free(c->pk_ap); free(c->sk_sp);
    free(c);
}
const char *orc_last_error(const orc_ctx *c) { return c->err; }
uint32_t orc_ring_dim(const orc_ctx *c) { return c->wp.N; }
uint32_t orc_num_limbs(const orc_ctx *c) { return (uint32_t)c->wp.L; }
uint32_t orc_slots(const orc_ctx *c) { return c->wp.batch; }
uint64_t orc_modulus(const orc_ctx *c, int l) { return c->wp.q[l]; }
uint64_t orc_psi(const orc_ctx *c, int l) { return c->wp.psi[l]; }
double orc_scaling_factor(const orc_ctx *c) { return (double)c->wp.q[c->wp.L - 1]; }
int orc_wire_variant(const orc_ctx *c) { return c->wp.nints; }
void orc_get_key_tag(const orc_ctx *c, char out[33]) { memcpy(out, c->key_tag, 33); }
int orc_get_public_key(const orc_ctx *c, uint64_t *b, uint64_t *a) {
    if (!c->have_pk) return 0;
    size_t n = (size_t)c->wp.L * c->wp.N * 8;
    memcpy(b, c->pk_b, n); memcpy(a, c->pk_a, n);
    return 1;
}
int orc_get_secret_key(const orc_ctx *c, uint64_t *s) {
    if (!c->have_sk) return 0;
    memcpy(s, c->sk_s, (size_t)c->wp.L * c->wp.N * 8);
    return 1;
}
void orc_free(void *p) { free(p); }

/* ------------------------------------------------------------------------------------------- */
/* NTT — [PAL] ChineseRemainderTransformFTT::ForwardTransformToBitReverse / InverseTransformFromBitReverse */
/* ------------------------------------------------------------------------------------------- */
/* Harvey lazy butterflies (values kept in [0,4q) forward, [0,2q) inverse, q < 2^60), branch-free, with one final
 * correction pass: the same arrangement PALISADE's native transform uses; results are the canonical residues. */
static inline uint64_t shoup_lazy(uint64_t x, uint64_t w, uint64_t wp, uint64_t q) { /* in [0, 2q) */
    return x * w - (uint64_t)(((u128)x * wp) >> 64) * q;
}
static inline uint64_t csub(uint64_t x, uint64_t m) { return x - (m & -(uint64_t)(x >= m)); }
static void ntt_fwd(uint64_t *a, uint32_t N, uint64_t q, const uint64_t *tw, const uint64_t *twp) {
    const uint64_t q2 = 2 * q;
    uint32_t t = N >> 1;
    for (uint32_t m = 1; m < N; m <<= 1, t >>= 1) {
        for (uint32_t i = 0; i < m; i++) {
            uint64_t w = tw[m + i], wp = twp[m + i];
            uint64_t *x = a + 2 * i * t, *y = x + t;
            for (uint32_t j = 0; j < t; j++) {
                uint64_t u = csub(x[j], q2), v = shoup_lazy(y[j], w, wp, q);
                x[j] = u + v;
                y[j] = u + q2 - v;
            }
        }
    }
    for (uint32_t j = 0; j < N; j++) a[j] = csub(csub(a[j], q2), q);
}
static void ntt_inv(uint64_t *a, uint32_t N, uint64_t q, const uint64_t *itw, const uint64_t *itwp, uint64_t ninv,
                    uint64_t ninvp) {
    const uint64_t q2 = 2 * q;
    uint32_t t = 1;
    for (uint32_t m = N >> 1; m >= 1; m >>= 1, t <<= 1) {
        for (uint32_t i = 0; i < m; i++) {
            uint64_t w = itw[m + i], wp = itwp[m + i];
            uint64_t *x = a + 2 * i * t, *y = x + t;
            for (uint32_t j = 0; j < t; j++) {
                uint64_t u = x[j], v = y[j];
                x[j] = csub(u + v, q2);
                y[j] = shoup_lazy(u + q2 - v, w, wp, q);
            }
        }
    }
    for (uint32_t j = 0; j < N; j++) a[j] = csub(shoup_lazy(a[j], ninv, ninvp, q), q);
}
void orc_ntt_forward(const orc_ctx *c, int l, uint64_t *a) { ntt_fwd(a, c->wp.N, c->wp.q[l], c->tw[l], c->twp[l]); }
void orc_ntt_inverse(const orc_ctx *c, int l, uint64_t *a) {
    ntt_inv(a, c->wp.N, c->wp.q[l], c->itw[l], c->itwp[l], c->ninv[l], c->ninvp[l]);
}

/* ------------------------------------------------------------------------------------------- */
/* canonical-embedding FFT — [PAL] DiscreteFourierTransform::FFTSpecial / FFTSpecialInv          */
/* ------------------------------------------------------------------------------------------- */
static void bit_reverse_complex(double *v, uint32_t size) {
    int lg = ilog2(size);
    for (uint32_t i = 0; i < size; i++) {
        uint32_t j = bitrev(i, lg);
        if (i < j) {
            double r = v[2 * i], im = v[2 * i + 1];
            v[2 * i] = v[2 * j]; v[2 * i + 1] = v[2 * j + 1];
            v[2 * j] = r; v[2 * j + 1] = im;
        }
    }
}
void orc_fft_special_inv(const orc_ctx *c, double *v, uint32_t size) {
    uint32_t M = c->wp.M;
    for (uint32_t len = size; len >= 2; len >>= 1) {
        uint32_t lenh = len >> 1, lenq = len << 2, gap = M / lenq;
        for (uint32_t i = 0; i < size; i += len) {
            for (uint32_t j = 0; j < lenh; j++) {
                uint32_t idx = (lenq - (c->rot_group[j] % lenq)) * gap;
                double wr = c->ksi[2 * idx], wi = c->ksi[2 * idx + 1];
                double *a = v + 2 * (i + j), *b = v + 2 * (i + j + lenh);
                double ur = a[0] + b[0], ui = a[1] + b[1];
                double dr = a[0] - b[0], di = a[1] - b[1];
                a[0] = ur; a[1] = ui;
                b[0] = dr * wr - di * wi;
                b[1] = dr * wi + di * wr;
            }
        }
    }
    bit_reverse_complex(v, size);
    for (uint32_t i = 0; i < size; i++) { v[2 * i] /= (double)size; v[2 * i + 1] /= (double)size; }
}
void orc_fft_special(const orc_ctx *c, double *v, uint32_t size) {
    uint32_t M = c->wp.M;
    bit_reverse_complex(v, size);
    for (uint32_t len = 2; len <= size; len <<= 1) {
        uint32_t lenh = len >> 1, lenq = len << 2, gap = M / lenq;
        for (uint32_t i = 0; i < size; i += len) {
            for (uint32_t j = 0; j < lenh; j++) {
                uint32_t idx = (c->rot_group[j] % lenq) * gap;
                double wr = c->ksi[2 * idx], wi = c->ksi[2 * idx + 1];
                double *a = v + 2 * (i + j), *b = v + 2 * (i + j + lenh);
                double vr = b[0] * wr - b[1] * wi, vi = b[0] * wi + b[1] * wr;
                double ur = a[0], ui = a[1];
                a[0] = ur + vr; a[1] = ui + vi;
                b[0] = ur - vr; b[1] = ui - vi;
            }
        }
    }
}

/* ------------------------------------------------------------------------------------------- */
/* sampler: ChaCha20 in counter mode (RFC 8439 block function), spec shared with the device      */
/* (fhe-fed_b200/csrc/ckks_core.cuh).  Written from the RFC, independently of the */
/* device code; tests/test_sampler_cpu.py checks both against the RFC's test vector.             */
/* ------------------------------------------------------------------------------------------- */
static inline uint32_t rotl(uint32_t x, int n) { return (x << n) | (x >> (32 - n)); }
static void quarter(uint32_t *s, int a, int b, int c, int d) {
    s[a] += s[b]; s[d] ^= s[a]; s[d] = rotl(s[d], 16);
    s[c] += s[d]; s[b] ^= s[c]; s[b] = rotl(s[b], 12);
    s[a] += s[b]; s[d] ^= s[a]; s[d] = rotl(s[d], 8);
    s[c] += s[d]; s[b] ^= s[c]; s[b] = rotl(s[b], 7);
}
void orc_chacha20_block(const uint32_t key[8], uint32_t counter, const uint32_t nonce[3], uint32_t out[16]) {
    uint32_t st[16] = {0x61707865u, 0x3320646eu, 0x79622d32u, 0x6b206574u};
    memcpy(st + 4, key, 32);
    st[12] = counter;
    memcpy(st + 13, nonce, 12);
    uint32_t w[16];
    memcpy(w, st, sizeof w);
    for (int r = 0; r < 10; r++) {
        quarter(w, 0, 4, 8, 12); quarter(w, 1, 5, 9, 13); quarter(w, 2, 6, 10, 14); quarter(w, 3, 7, 11, 15);
        quarter(w, 0, 5, 10, 15); quarter(w, 1, 6, 11, 12); quarter(w, 2, 7, 8, 13); quarter(w, 3, 4, 9, 14);
    }
    for (int i = 0; i < 16; i++) out[i] = w[i] + st[i];
}
/* test keys: the 64-bit seed of the tests laid into a 256-bit key per domain ("SHELFI-B200 seed") */
enum { KEY_ENCRYPT = 0, KEY_KEYGEN_SECRET = 1, KEY_KEYGEN_PUBLIC = 2, KEY_DECODE = 3 };
void orc_seed_key(uint64_t seed, uint32_t domain, uint32_t key[8]) {
    key[0] = (uint32_t)seed; key[1] = (uint32_t)(seed >> 32);
    key[2] = 0x4c454853u; key[3] = 0x422d4946u; key[4] = 0x20303032u; key[5] = 0x64656573u;
    key[6] = domain; key[7] = 0;
}
/* stream ids: nonce word 2 */
enum { STREAM_GAUSS = 0, STREAM_TERNARY = 1, STREAM_KEY_S = 2, STREAM_KEY_E = 3, STREAM_KEY_A = 4, STREAM_KEY_TAG = 5,
       STREAM_DECODE_NOISE = 6, STREAM_GAUSS_RETRY = 7, STREAM_TERNARY_RETRY = 8, STREAM_KEY_S_RETRY = 9,
       STREAM_KEY_E_RETRY = 10 };

/* Discrete Gaussian by CDF inversion on a 63-bit uniform U = (h >> 1) * 2^32 + l: magnitude = #{k < 31 : U >= cdf[k]},
 * sign = h & 1.  Written the plain way: l is always passed in.  The device decides from the top 31 bits and draws l only
 * when they tie with a table entry; the value is the same (tests/test_sampler_cpu.py checks the two against each other
 * on ties and non-ties), so here the caller simply fetches l from the retry stream whenever a tie is possible. */
static inline int64_t gauss_from_u63(const uint64_t cdf[32], uint32_t h, uint32_t l) {
    uint64_t u = ((uint64_t)(h >> 1) << 32) | l;
    int mag = 0;
    for (int k = 0; k < 31; k++)
        if (u >= cdf[k]) mag++;
    return (h & 1) ? -(int64_t)mag : (int64_t)mag;
}
static inline int gauss_needs_low_word(const uint64_t cdf[32], uint32_t h) {
    for (int k = 0; k < 31; k++)
        if ((uint32_t)(cdf[k] >> 32) == (h >> 1)) return 1;
    return 0;
}
int64_t orc_gauss_from_words(const orc_ctx *c, uint32_t h, uint32_t l) { return gauss_from_u63(c->gauss_cdf, h, l); }
/* one Gaussian draw: h from the main stream, l (if a tie makes it matter) word `which` of retry block j */
static int64_t gauss_draw(const uint64_t cdf[32], uint32_t h, const uint32_t key[8], const uint32_t nonce_retry[3], uint32_t j,
                          int which) {
    uint32_t l = 0;
    if (gauss_needs_low_word(cdf, h)) {
        uint32_t out[16];
        orc_chacha20_block(key, j, nonce_retry, out);
        l = out[which];
    }
    return gauss_from_u63(cdf, h, l);
}
/* uniform on {-1, 0, 1}: 2-bit attempts, low bits first, 3 = rejected; eight attempts in the 16-bit half word, then
 * the 256 attempts of retry block j */
static int64_t ternary_draw(uint32_t h16, const uint32_t key[8], const uint32_t nonce_retry[3], uint32_t j) {
    for (int a = 0; a < 8; a++) {
        uint32_t c = (h16 >> (2 * a)) & 3;
        if (c != 3) return (int64_t)c - 1;
    }
    uint32_t out[16];
    orc_chacha20_block(key, j, nonce_retry, out);
    for (int i = 0; i < 16; i++)
        for (int a = 0; a < 16; a++) {
            uint32_t c = (out[i] >> (2 * a)) & 3;
            if (c != 3) return (int64_t)c - 1;
        }
    return 0;
}
int64_t orc_ternary_from_half(uint32_t h16, const uint32_t key[8], const uint32_t nonce_retry[3], uint32_t j) {
    return ternary_draw(h16, key, nonce_retry, j);
}
static void ternary_stream(const uint32_t key[8], uint32_t n0, uint32_t n1, uint32_t stream, uint32_t retry_stream, uint32_t N,
                           int64_t *v) {
    uint32_t out[16], nonce[3] = {n0, n1, stream}, nr[3] = {n0, n1, retry_stream};
    for (uint32_t b = 0; b < N / 32; b++) {
        orc_chacha20_block(key, b, nonce, out);
        for (uint32_t i = 0; i < 32; i++)
            v[32 * b + i] = ternary_draw((out[i >> 1] >> (16 * (i & 1))) & 0xFFFF, key, nr, 32 * b + i);
    }
}

/* [PAL] LPAlgorithmCKKS::Encrypt draws v <- TernaryUniform (mode OPTIMIZED), e0,e1 <- DiscreteGaussian(sigma).
 * PALISADE's own PRNG is seeded from OS entropy, so only the distribution is pinned; this spec makes
 * oracle and device draw the same values for a given (key, ct_index):
 *   ternary block b, nonce (ct_lo, ct_hi, 1): v[32b + 2i], v[32b + 2i + 1] from the low / high half of word i
 *   gauss   block b, nonce (ct_lo, ct_hi, 0): e0[8b + i] from word 2i, e1[8b + i] from word 2i + 1
 *   retries: block j = coefficient, nonce (ct_lo, ct_hi, 8) for ternary, (ct_lo, ct_hi, 7) words 0 / 1 for e0 / e1 */
void orc_sample_encrypt_noise(const orc_ctx *c, uint64_t seed, uint64_t ct, int64_t *v, int64_t *e0, int64_t *e1) {
    uint32_t N = c->wp.N, out[16], key[8];
    orc_seed_key(seed, KEY_ENCRYPT, key);
    uint32_t ng[3] = {(uint32_t)ct, (uint32_t)(ct >> 32), STREAM_GAUSS};
    uint32_t ngr[3] = {(uint32_t)ct, (uint32_t)(ct >> 32), STREAM_GAUSS_RETRY};
    for (uint32_t b = 0; b < N / 8; b++) {
        orc_chacha20_block(key, b, ng, out);
        for (uint32_t i = 0; i < 8; i++) {
            e0[8 * b + i] = gauss_draw(c->gauss_cdf, out[2 * i], key, ngr, 8 * b + i, 0);
            e1[8 * b + i] = gauss_draw(c->gauss_cdf, out[2 * i + 1], key, ngr, 8 * b + i, 1);
        }
    }
    ternary_stream(key, (uint32_t)ct, (uint32_t)(ct >> 32), STREAM_TERNARY, STREAM_TERNARY_RETRY, N, v);
}

/* ------------------------------------------------------------------------------------------- */
/* encode / encrypt / decrypt / decode, one ciphertext                                          */
/* ------------------------------------------------------------------------------------------- */
/* [PAL] CKKSPackedEncoding::Encode under EXACTRESCALE, called from ckks.cpp:80,94 */
int orc_encode_coeffs(const orc_ctx *c, const double *x, uint32_t n, int64_t *coeff) {
    uint32_t N = c->wp.N, Nh = N / 2, slots = c->wp.batch;
    if (n > slots || slots > Nh || slots == 0) return -1;
    uint32_t gap = Nh / slots;
    double *z = calloc((size_t)slots * 2, sizeof(double));
    for (uint32_t i = 0; i < n; i++) z[2 * i] = x[i];
    orc_fft_special_inv(c, z, slots);
    double powP = orc_scaling_factor(c);
    memset(coeff, 0, (size_t)N * sizeof(int64_t));
    for (uint32_t i = 0; i < slots; i++) {
        coeff[gap * i] = llround(z[2 * i] * powP);
        coeff[gap * i + Nh] = llround(z[2 * i + 1] * powP);
    }
    free(z);
    return 0;
}

int orc_encrypt_ct(const orc_ctx *c, const double *x, uint32_t n, uint64_t seed, uint64_t ct_index, uint64_t *out) {
    if (!c->have_cc || !c->have_pk) return -1;
    uint32_t N = c->wp.N;
    int L = c->wp.L;
    int64_t *m = malloc(N * 8), *v = malloc(N * 8), *e0 = malloc(N * 8), *e1 = malloc(N * 8);
    uint64_t *tv = malloc(N * 8), *t0 = malloc(N * 8), *t1 = malloc(N * 8);
    int rc = orc_encode_coeffs(c, x, n, m);
    if (rc == 0) {
        orc_sample_encrypt_noise(c, seed, ct_index, v, e0, e1);
        for (int l = 0; l < L; l++) {
            uint64_t q = c->wp.q[l];
            /* one integer vector each, reduced into every limb; NTT is linear so m and e0 share one */
            for (uint32_t j = 0; j < N; j++) {
                tv[j] = signed_mod(v[j], q);
                t0[j] = addmod(signed_mod(m[j], q), signed_mod(e0[j], q), q);
                t1[j] = signed_mod(e1[j], q);
            }
            orc_ntt_forward(c, l, tv); orc_ntt_forward(c, l, t0); orc_ntt_forward(c, l, t1);
            const uint64_t *b = c->pk_b + (size_t)l * N, *bp = c->pk_bp + (size_t)l * N;
            const uint64_t *a = c->pk_a + (size_t)l * N, *ap = c->pk_ap + (size_t)l * N;
            uint64_t *c0 = out + (size_t)l * N, *c1 = out + (size_t)(L + l) * N;
            for (uint32_t j = 0; j < N; j++) { /* c0 = b*v + e0 + m ; c1 = a*v + e1 */
                c0[j] = addmod(shoup_mul(tv[j], b[j], bp[j], q), t0[j], q);
                c1[j] = addmod(shoup_mul(tv[j], a[j], ap[j], q), t1[j], q);
            }
        }
    }
    free(m); free(v); free(e0); free(e1); free(tv); free(t0); free(t1);
    return rc;
}

/* ---- multi-limb CRT helpers (chains deeper than two limbs): little-endian 64-bit words ---- */
#define ORC_CRT_WORDS 8 /* 8 limbs x 60 bits = 480 bits */
typedef struct { uint64_t w[ORC_CRT_WORDS]; } big_t;
static void big_mul_add(big_t *x, uint64_t m, uint64_t a) { /* x = x*m + a */
    u128 carry = a;
    for (int i = 0; i < ORC_CRT_WORDS; i++) {
        u128 cur = (u128)x->w[i] * m + carry;
        x->w[i] = (uint64_t)cur;
        carry = cur >> 64;
    }
}
static int big_cmp(const big_t *a, const big_t *b) {
    for (int i = ORC_CRT_WORDS - 1; i >= 0; i--)
        if (a->w[i] != b->w[i]) return a->w[i] > b->w[i] ? 1 : -1;
    return 0;
}
static void big_sub(big_t *r, const big_t *a, const big_t *b) { /* r = a - b (mod 2^512) */
    uint64_t borrow = 0;
    for (int i = 0; i < ORC_CRT_WORDS; i++) {
        uint64_t ai = a->w[i], bi = b->w[i], d = ai - bi - borrow;
        borrow = (ai < bi) || (ai == bi && borrow);
        r->w[i] = d;
    }
}
/* magnitude -> double with ONE round-to-nearest-even (what a correctly rounded BigInteger conversion gives;
 * PALISADE's own ConvertToDouble accumulates word by word and can differ in the last ulp: unpinned, ~1e-16) */
static double big_to_double(const big_t *m) {
    int top = ORC_CRT_WORDS - 1;
    while (top > 0 && !m->w[top]) top--;
    if (top == 0) return (double)m->w[0];
    int s = __builtin_clzll(m->w[top]);
    uint64_t hi = m->w[top], lo = m->w[top - 1];
    uint64_t win = s ? (hi << s) | (lo >> (64 - s)) : hi;
    int sticky = (s ? (lo << s) : lo) != 0; /* bits of the next word below the 64-bit window */
    for (int i = top - 2; i >= 0 && !sticky; i--) sticky |= m->w[i] != 0;
    if (sticky) win |= 1ull;
    return ldexp((double)win, 64 * top - s);
}
typedef struct {
    int L;
    uint64_t q[ORC_MAX_LIMBS];
    uint64_t inv[ORC_MAX_LIMBS][ORC_MAX_LIMBS]; /* inv[i][j] = q_j^-1 mod q_i, j < i */
    big_t Q, Qh;
} crt_plan;
static void crt_plan_init(crt_plan *p, const uint64_t *q, int L) {
    memset(p, 0, sizeof *p);
    p->L = L;
    p->Q.w[0] = 1;
    for (int i = 0; i < L; i++) {
        p->q[i] = q[i];
        for (int j = 0; j < i; j++) p->inv[i][j] = invmod(q[j] % q[i], q[i]);
        big_mul_add(&p->Q, q[i], 0);
    }
    for (int i = 0; i < ORC_CRT_WORDS; i++)
        p->Qh.w[i] = (p->Q.w[i] >> 1) | (i + 1 < ORC_CRT_WORDS ? p->Q.w[i + 1] << 63 : 0);
}
/* [PAL] DCRTPoly::CRTInterpolate for one coefficient: the unique x in [0, Q) with x = a_i (mod q_i), here by
 * Garner's mixed-radix digits + Horner; then the centred representative as sign + magnitude */
static void crt_centered(const crt_plan *p, const uint64_t *a, int *neg, big_t *mag) {
    uint64_t v[ORC_MAX_LIMBS] = {0};
    for (int i = 0; i < p->L; i++) {
        uint64_t qi = p->q[i], t = a[i];
        for (int j = 0; j < i; j++) t = mulmod(submod(t, v[j] % qi, qi), p->inv[i][j], qi);
        v[i] = t;
    }
    big_t x; memset(&x, 0, sizeof x);
    x.w[0] = v[p->L - 1];
    for (int i = p->L - 2; i >= 0; i--) big_mul_add(&x, p->q[i], v[i]);
    if (big_cmp(&x, &p->Qh) > 0) { *neg = 1; big_sub(mag, &p->Q, &x); }
    else { *neg = 0; *mag = x; }
}

/* [PAL] LPAlgorithmCKKS::Decrypt + DCRTPoly::CRTInterpolate + CKKSPackedEncoding::Decode (noise
 * flooding of PALISADE >= 1.10.6 left out: it is random, ~2^-40 relative, far below the 1e-6 tier).
 * words (optional): N x 8 little-endian words, two's complement of the centred integer. */
/* [PAL >= 1.10.6] CKKSPackedEncoding::Decode: before FFTSpecial, Gaussian noise is added to every slot (Li-Micciancio
 * key-recovery countermeasure).  The standard deviation is estimated from the part of the decrypted polynomial that an
 * exact real-valued message cannot have — m(X) - m(1/X), i.e. 2 Im(z) —, floored at sqrt(N)/8 and multiplied by
 * sqrt(M + 1) with M = 1.  Restated from the published source (ckkspackedencoding.cpp: Conjugate, StdDev, Decode);
 * PARITY UNPINNED: the reference holds no decrypt output.  v: slots complex values (re, im interleaved), natural order,
 * already at scale 2^p.  Returns the standard deviation of the noise that was added. */
static double decode_stddev(const double *v, uint32_t n, uint32_t ring_dim) {
    if (n == 1) return fabs(v[1]);
    /* cv[i] = vec[i] - Conjugate(vec)[i], Conjugate[0] = conj(vec[0]), Conjugate[i] = (-Im vec[n-i], -Re vec[n-i]) */
    double cv0_im = 2.0 * v[1];
    double mean = 0.0;
    for (uint32_t i = 1; i < n / 2; i++) mean += (v[2 * i] + v[2 * (n - i) + 1]) + (v[2 * i + 1] + v[2 * (n - i)]);
    double half_re = v[2 * (n / 2)] + v[2 * (n - n / 2) + 1];
    mean = 2.0 * mean + cv0_im + 2.0 * half_re;
    mean /= (double)(2 * n) - 1.0;
    double var = 0.0;
    for (uint32_t i = 1; i < n / 2; i++) {
        double re = v[2 * i] + v[2 * (n - i) + 1] - mean, im = v[2 * i + 1] + v[2 * (n - i)] - mean;
        var += re * re + im * im;
    }
    var = 2.0 * var + (cv0_im - mean) * (cv0_im - mean) + 2.0 * (half_re - mean) * (half_re - mean);
    var /= (double)(2 * n) - 2.0;
    double sd = 0.5 * sqrt(var);
    double floor_sd = 0.125 * sqrt((double)ring_dim);
    return sd < floor_sd ? floor_sd : sd;
}
static void normal_pair(const uint32_t w[4], double *z0, double *z1) { /* Box-Muller on two 53-bit uniforms */
    double u1 = (double)(((((uint64_t)w[1] << 32) | w[0]) >> 11) + 1) * 0x1p-53; /* (0, 1] */
    double u2 = (double)((((uint64_t)w[3] << 32) | w[2]) >> 11) * 0x1p-53;        /* [0, 1) */
    double r = sqrt(-2.0 * log(u1)), a = 6.283185307179586476925286766559 * u2;
    *z0 = r * cos(a); *z1 = r * sin(a);
}
double orc_decode_noise(const orc_ctx *c, double *v, uint32_t n, const uint32_t key[8], uint64_t ct_index) {
    double sd = 1.4142135623730951 * decode_stddev(v, n, c->wp.N); /* sqrt(M + 1), M = 1 */
    uint32_t nonce[3] = {(uint32_t)ct_index, (uint32_t)(ct_index >> 32), STREAM_DECODE_NOISE}, out[16];
    for (uint32_t i = 0; i < n; i++) { /* slot i: words 4 (i & 3) .. + 3 of block i >> 2 */
        if ((i & 3) == 0) orc_chacha20_block(key, i >> 2, nonce, out);
        double z0, z1;
        normal_pair(out + 4 * (i & 3), &z0, &z1);
        v[2 * i] += sd * z0; v[2 * i + 1] += sd * z1;
    }
    return sd;
}

static int decrypt_ct_impl_at(const orc_ctx *c, const uint64_t *ct, int L, double sf, double *out, int64_t *coeff_hi,
                              uint64_t *coeff_lo, uint64_t *words, uint64_t ct_index);
static int decrypt_ct_impl(const orc_ctx *c, const uint64_t *ct, int L, double sf, double *out, int64_t *coeff_hi,
                           uint64_t *coeff_lo, uint64_t *words) {
    return decrypt_ct_impl_at(c, ct, L, sf, out, coeff_hi, coeff_lo, words, 0);
}
static int decrypt_ct_impl_at(const orc_ctx *c, const uint64_t *ct, int L, double sf, double *out, int64_t *coeff_hi,
                              uint64_t *coeff_lo, uint64_t *words, uint64_t ct_index) {
    if (!c->have_cc || !c->have_sk || L < 1 || L > c->wp.L) return -1;
    uint32_t N = c->wp.N, Nh = N / 2, slots = c->wp.batch;
    /* L = towers of this ciphertext: the first L towers of the secret key and of the chain ([PAL] Decrypt drops the
     * key's extra towers to match a rescaled ciphertext) */
    uint32_t gap = Nh / slots;
    uint64_t *t = malloc((size_t)L * N * 8);
    for (int l = 0; l < L; l++) {
        uint64_t q = c->wp.q[l];
        const uint64_t *c0 = ct + (size_t)l * N, *c1 = ct + (size_t)(L + l) * N;
        const uint64_t *s = c->sk_s + (size_t)l * N, *sp = c->sk_sp + (size_t)l * N;
        uint64_t *tl = t + (size_t)l * N;
        for (uint32_t j = 0; j < N; j++) tl[j] = addmod(c0[j], shoup_mul(c1[j], s[j], sp[j], q), q);
        orc_ntt_inverse(c, l, tl);
    }
    double p = (double)c->wp.ptm;
    double pre = pow(sf, -1) * pow(2, p), powP = pow(2, -p);
    double *z = malloc((size_t)slots * 2 * sizeof(double));
    crt_plan plan;
    crt_plan_init(&plan, c->wp.q, L);
    for (uint32_t j = 0; j < N; j++) {
        uint64_t a[ORC_MAX_LIMBS];
        for (int l = 0; l < L; l++) a[l] = t[(size_t)l * N + j];
        int neg; big_t mag;
        crt_centered(&plan, a, &neg, &mag);
        if (coeff_hi || words) {
            big_t zero, tc; memset(&zero, 0, sizeof zero);
            if (neg) big_sub(&tc, &zero, &mag); else tc = mag;
            if (coeff_hi) { coeff_hi[j] = (int64_t)tc.w[1]; coeff_lo[j] = tc.w[0]; }
            if (words) memcpy(words + (size_t)j * ORC_CRT_WORDS, tc.w, sizeof tc.w);
        }
        uint32_t idx = j < Nh ? j : j - Nh;
        if (idx % gap == 0) {
            uint32_t slot = idx / gap;
            double d = big_to_double(&mag);
            z[2 * slot + (j < Nh ? 0 : 1)] = (neg ? -d : d) * pre;
        }
    }
    if (c->opt_decode_noise) {
        uint32_t key[8];
        orc_seed_key(c->noise_seed, KEY_DECODE, key);
        orc_decode_noise(c, z, slots, key, ct_index);
    }
    orc_fft_special(c, z, slots);
    /* [PAL] brings the values to scale 2^p (`pre`), adds its flooding noise there (opt_decode_noise), and
     * multiplies by 2^-p after the FFT; the two powers of two are exact. */
    for (uint32_t i = 0; i < slots; i++) out[i] = z[2 * i] * powP;
    free(t); free(z);
    return 0;
}
int orc_decrypt_ct(const orc_ctx *c, const uint64_t *ct, double sf, double *out, int64_t *coeff_hi, uint64_t *coeff_lo) {
    return decrypt_ct_impl(c, ct, c->wp.L, sf, out, coeff_hi, coeff_lo, NULL);
}
int orc_decrypt_ct_words(const orc_ctx *c, const uint64_t *ct, double sf, double *out, uint64_t *words) {
    return decrypt_ct_impl(c, ct, c->wp.L, sf, out, NULL, NULL, words);
}
int orc_decrypt_ct_level(const orc_ctx *c, const uint64_t *ct, uint32_t limbs, double sf, double *out, uint64_t *words) {
    return decrypt_ct_impl(c, ct, (int)limbs, sf, out, NULL, NULL, words);
}

/* [PAL] NativeVector::SwitchModulus for one value: the centred representative of n (mod q_old) reduced mod q_new.
 * Written with PALISADE's two branches (larger / smaller new modulus). */
static inline uint64_t switch_modulus(uint64_t n, uint64_t q_old, uint64_t q_new) {
    uint64_t half = q_old >> 1;
    if (q_new > q_old) return n > half ? n + (q_new - q_old) : n;
    uint64_t diff = (q_old - q_new) % q_new, a = n % q_new;
    return n > half ? submod(a, diff, q_new) : a;
}
/* big = big / d, returns the remainder */
static uint64_t big_divmod_u64(big_t *x, uint64_t d) {
    u128 rem = 0;
    for (int i = ORC_CRT_WORDS - 1; i >= 0; i--) {
        u128 cur = (rem << 64) | x->w[i];
        x->w[i] = (uint64_t)(cur / d);
        rem = cur % d;
    }
    return (uint64_t)rem;
}
/* [PAL] the two per-tower constants of DropLastElementAndScale for a polynomial of Lin towers (l = Lin - 1):
 *   QlQlInvModqlDivqlModq[i] = ((Q_l/q_l) * ((Q_l/q_l)^-1 mod q_l)) / q_l  mod q_i,   qlInvModq[i] = q_l^-1 mod q_i */
void orc_rescale_constants(const orc_ctx *c, uint32_t Lin, uint64_t *t, uint64_t *ql_inv) {
    const uint64_t *q = c->wp.q;
    int l = (int)Lin - 1;
    uint64_t p_mod_ql = 1;
    for (int j = 0; j < l; j++) p_mod_ql = mulmod(p_mod_ql, q[j] % q[l], q[l]);
    uint64_t p_inv = invmod(p_mod_ql, q[l]);
    big_t x; memset(&x, 0, sizeof x);
    x.w[0] = 1;
    for (int j = 0; j < l; j++) big_mul_add(&x, q[j], 0);
    big_mul_add(&x, p_inv, 0);
    big_divmod_u64(&x, q[l]); /* floor */
    for (int i = 0; i < l; i++) {
        big_t y = x;
        t[i] = big_divmod_u64(&y, q[i]);
        ql_inv[i] = invmod(q[l] % q[i], q[i]);
    }
}
/* [PAL] ModReduce under EXACTRESCALE = DCRTPoly::DropLastElementAndScale on both polynomials:
 * in [2][Lin][N] -> out [2][Lin-1][N], EVALUATION form throughout.  The last tower goes to coefficient form, is
 * switched into every remaining modulus, multiplied by its constant, transformed back, and added to the tower scaled
 * by q_l^-1 — i.e. (c - [c]_{q_l}) / q_l with the centred remainder. */
int orc_rescale_ct(const orc_ctx *c, const uint64_t *in, uint32_t Lin, uint64_t *out) {
    if (!c->have_cc || Lin < 2 || (int)Lin > c->wp.L) return -1;
    uint32_t N = c->wp.N;
    int l = (int)Lin - 1;
    uint64_t t[ORC_MAX_LIMBS], qi[ORC_MAX_LIMBS];
    orc_rescale_constants(c, Lin, t, qi);
    uint64_t *last = malloc((size_t)N * 8), *tmp = malloc((size_t)N * 8);
    for (int p = 0; p < 2; p++) {
        memcpy(last, in + ((size_t)p * Lin + l) * N, (size_t)N * 8);
        orc_ntt_inverse(c, l, last);
        for (int i = 0; i < l; i++) {
            uint64_t q = c->wp.q[i];
            for (uint32_t j = 0; j < N; j++) tmp[j] = mulmod(switch_modulus(last[j], c->wp.q[l], q), t[i], q);
            orc_ntt_forward(c, i, tmp);
            const uint64_t *src = in + ((size_t)p * Lin + i) * N;
            uint64_t *dst = out + ((size_t)p * l + i) * N;
            for (uint32_t j = 0; j < N; j++) dst[j] = addmod(mulmod(src[j], qi[i], q), tmp[j], q);
        }
    }
    free(last); free(tmp);
    return 0;
}

/* [PAL] EvalMult(ct, pt) with pt = MakeCKKSPackedPlaintext(w) (fresh: depth 1, level 0, scale Delta), then EvalAdd:
 * acc (+)= ct (*) NTT(Encode(w)) on both polynomials of every limb.  This is the per-slot plaintext weighting of the
 * TenSEAL/FLARE-style aggregation (code/benchmark_nvidia.py:439,461-465 multiply each ciphertext by a plaintext
 * vector); the result has depth 2 and scale Delta^2 like EvalMult(ct, double), and is not rescaled (EXACTRESCALE). */
int orc_ct_mul_plain_accumulate(const orc_ctx *c, const uint64_t *ct, const double *w, uint32_t n, int first, uint64_t *acc) {
    if (!c->have_cc) return -1;
    uint32_t N = c->wp.N;
    int L = c->wp.L;
    int64_t *m = malloc((size_t)N * 8);
    uint64_t *pt = malloc((size_t)N * 8);
    int rc = orc_encode_coeffs(c, w, n, m);
    for (int l = 0; rc == 0 && l < L; l++) {
        uint64_t q = c->wp.q[l];
        for (uint32_t j = 0; j < N; j++) pt[j] = signed_mod(m[j], q);
        orc_ntt_forward(c, l, pt);
        for (int p = 0; p < 2; p++) {
            const uint64_t *src = ct + ((size_t)p * L + l) * N;
            uint64_t *dst = acc + ((size_t)p * L + l) * N;
            for (uint32_t j = 0; j < N; j++) {
                uint64_t v = mulmod(src[j], pt[j], q);
                dst[j] = first ? v : addmod(dst[j], v, q);
            }
        }
    }
    free(m); free(pt);
    return rc;
}

int64_t orc_weight_constant(const orc_ctx *c, float w) {
    double sc = (double)w; /* ckks.cpp:287 narrows the Python float to float32, [PAL] widens it back */
    if (c->opt_evalmult_round == 1) return (int64_t)llround(sc * orc_scaling_factor(c)); /* the other reading */
    return (int64_t)(sc * orc_scaling_factor(c) + 0.5);
}
/* option ids shared with the engine's shelfi_set_option */
int orc_set_option(orc_ctx *c, int option, int64_t value) {
    switch (option) {
        case 1: if (value != 0 && value != 1) return -1; c->opt_evalmult_round = (int)value; return 0;
        case 2: if (value != 0 && value != 1) return -1; c->opt_decode_noise = (int)value; return 0;
        case 4: c->noise_seed = (uint64_t)value; return 0;
        default: return -1;
    }
}

/* ------------------------------------------------------------------------------------------- */
/* cereal PortableBinary wire format (grammar: SURVEY §8c, verified on every fixture)           */
/* ------------------------------------------------------------------------------------------- */
#define W_SAME 0x40000000u
#define W_NEW 0x80000000u
enum { /* C++ types whose class-version word is written on first occurrence only */
    V_CC, V_CKKSP, V_RLWEP, V_LPP, V_DCRTP, V_EPBIG, V_BIGINT, V_NATP, V_EPNAT, V_NATINT, V_ENCP, V_SCHEME,
    V_SCHEMEBASE, V_KEY0, V_KEY1, V_KEY2, V_CT, V_CRYPTOOBJ, V_DCRTPOLY, V_NATPOLY, V_NATVEC, V_COUNT
};
static const uint32_t ver_value[V_COUNT] = {1, 0, 0, 0, 1, 1, 1, 1, 1, 1, 1, 0, 0, 0, 0, 0, 1, 0, 1, 1, 1};
static const char NAME_PARAMS[] = "lbcrypto::LPCryptoParametersCKKS<lbcrypto::DCRTPoly>";
static const char NAME_SCHEME[] = "lbcrypto::LPPublicKeyEncryptionSchemeCKKS<lbcrypto::DCRTPoly>";

typedef struct { const uint8_t *b; size_t n, p; int err; uint32_t seen; } rd_t;
typedef struct { uint8_t *b; size_t n, cap; uint32_t seen; uint32_t next_id; } wr_t;

static void rd_raw(rd_t *r, void *dst, size_t k) {
    if (r->err || r->p + k > r->n) { r->err = 1; memset(dst, 0, k); return; }
    memcpy(dst, r->b + r->p, k);
    r->p += k;
}
static uint8_t rd_u8(rd_t *r) { uint8_t v; rd_raw(r, &v, 1); return v; }
static uint16_t rd_u16(rd_t *r) { uint16_t v; rd_raw(r, &v, 2); return v; }
static uint32_t rd_u32(rd_t *r) { uint32_t v; rd_raw(r, &v, 4); return v; }
static uint64_t rd_u64(rd_t *r) { uint64_t v; rd_raw(r, &v, 8); return v; }
static float rd_f32(rd_t *r) { float v; rd_raw(r, &v, 4); return v; }
static double rd_f64(rd_t *r) { double v; rd_raw(r, &v, 8); return v; }
static void rd_expect32(rd_t *r, uint32_t want) { if (rd_u32(r) != want) r->err = 2; }
static void rd_ver(rd_t *r, int t) {
    if (r->seen & (1u << t)) return;
    r->seen |= 1u << t;
    if (rd_u32(r) != ver_value[t]) r->err = 3;
}
static void rd_name(rd_t *r, const char *name, uint32_t poly_id) {
    uint32_t id = rd_u32(r);
    if (id == (W_NEW | poly_id)) {
        uint64_t len = rd_u64(r);
        char buf[96];
        if (len != strlen(name) || len >= sizeof buf) { r->err = 4; return; }
        rd_raw(r, buf, len);
        if (memcmp(buf, name, len)) r->err = 4;
    } else if (id != poly_id) r->err = 4;
}

static void wr_raw(wr_t *w, const void *src, size_t k) {
    if (w->n + k > w->cap) {
        size_t nc = w->cap ? w->cap * 2 : 4096;
        while (nc < w->n + k) nc *= 2;
        w->b = realloc(w->b, nc);
        w->cap = nc;
    }
    memcpy(w->b + w->n, src, k);
    w->n += k;
}
static void wr_u8(wr_t *w, uint8_t v) { wr_raw(w, &v, 1); }
static void wr_u16(wr_t *w, uint16_t v) { wr_raw(w, &v, 2); }
static void wr_u32(wr_t *w, uint32_t v) { wr_raw(w, &v, 4); }
static void wr_u64(wr_t *w, uint64_t v) { wr_raw(w, &v, 8); }
static void wr_f32(wr_t *w, float v) { wr_raw(w, &v, 4); }
static void wr_f64(wr_t *w, double v) { wr_raw(w, &v, 8); }
static void wr_ver(wr_t *w, int t) {
    if (w->seen & (1u << t)) return;
    w->seen |= 1u << t;
    wr_u32(w, ver_value[t]);
}
static uint32_t wr_new(wr_t *w) { uint32_t id = w->next_id++; wr_u32(w, W_NEW | id); return id; }
static void wr_name(wr_t *w, const char *name, uint32_t poly_id, int first) {
    if (first) { wr_u32(w, W_NEW | poly_id); wr_u64(w, strlen(name)); wr_raw(w, name, strlen(name)); }
    else wr_u32(w, poly_id);
}

/* ids assigned while reading/writing one archive */
typedef struct { uint32_t cc, dcrt, native[ORC_MAX_LIMBS]; } wire_ids;

static void rd_bigint(rd_t *r, uint32_t words[110], uint16_t *msb) {
    for (int i = 0; i < 110; i++) words[i] = rd_u32(r);
    *msb = rd_u16(r);
}
static void rd_bigint_zero(rd_t *r) {
    uint32_t w[110]; uint16_t m;
    rd_bigint(r, w, &m);
    for (int i = 0; i < 110; i++) if (w[i]) r->err = 5;
    if (m) r->err = 5;
}
static void wr_bigint(wr_t *w, const uint32_t words[110], uint16_t msb) {
    for (int i = 0; i < 110; i++) wr_u32(w, words[i]);
    wr_u16(w, msb);
}
static void wr_bigint_zero(wr_t *w) { static const uint32_t z[110]; wr_bigint(w, z, 0); }

/* CryptoContext := SAME NEW(id) ... (full) | SAME REF(id) */
static void rd_context(rd_t *r, wire_params *p, wire_ids *ids, uint32_t *next_id) {
    rd_expect32(r, W_SAME);
    uint32_t id = rd_u32(r);
    if (!(id & W_NEW)) { if (id != ids->cc) r->err = 6; return; }
    ids->cc = id & ~W_NEW; (*next_id)++;
    rd_ver(r, V_CC);
    rd_name(r, NAME_PARAMS, 1);
    rd_u32(r); (*next_id)++; /* NEW(params) */
    rd_ver(r, V_CKKSP); rd_ver(r, V_RLWEP); rd_ver(r, V_LPP);
    /* DCRTParams */
    rd_expect32(r, W_SAME);
    ids->dcrt = rd_u32(r) & ~W_NEW; (*next_id)++;
    rd_ver(r, V_DCRTP); rd_ver(r, V_EPBIG);
    p->M = rd_u32(r); p->N = rd_u32(r);
    if (rd_u8(r) != 1) r->err = 7;
    rd_ver(r, V_BIGINT);
    rd_bigint(r, p->bigQ, &p->bigQ_msb);
    rd_bigint_zero(r); rd_bigint_zero(r); rd_bigint_zero(r);
    uint64_t L = rd_u64(r);
    if (L < 1 || L > ORC_MAX_LIMBS) { r->err = 8; return; }
    p->L = (int)L;
    for (int l = 0; l < p->L; l++) {
        rd_expect32(r, W_SAME);
        ids->native[l] = rd_u32(r) & ~W_NEW; (*next_id)++;
        rd_ver(r, V_NATP); rd_ver(r, V_EPNAT);
        if (rd_u32(r) != p->M) r->err = 9;
        if (rd_u32(r) != p->N) r->err = 9;
        if (rd_u8(r) != 1) r->err = 9;
        rd_ver(r, V_NATINT);
        p->q[l] = rd_u64(r); p->psi[l] = rd_u64(r);
        if (rd_u64(r) || rd_u64(r)) r->err = 9; /* bigModulus, bigRootOfUnity */
    }
    rd_bigint_zero(r); /* originalModulus */
    /* EncodingParams */
    rd_expect32(r, W_SAME);
    rd_u32(r); (*next_id)++;
    rd_ver(r, V_ENCP);
    p->ptm = rd_u64(r);
    if (rd_u64(r) || rd_u64(r) || rd_u64(r) || rd_u32(r)) r->err = 10;
    p->batch = rd_u32(r);
    p->sigma = rd_f32(r); p->alpha = rd_f32(r); p->hermite = rd_f32(r);
    /* 8- or 9-int RLWE/CKKS block; the scheme name marker tells which */
    size_t rest = r->p;
    int n = -1;
    for (int k = 8; k <= 9 && !r->err; k++) {
        size_t at = rest + 4u * k;
        if (at + 4 <= r->n) {
            uint32_t v; memcpy(&v, r->b + at, 4);
            if (v == (W_NEW | 2u)) { n = k; break; }
        }
    }
    if (n < 0) { r->err = 11; return; }
    p->nints = n;
    for (int i = 0; i < n; i++) p->ints[i] = rd_u32(r);
    rd_name(r, NAME_SCHEME, 2);
    rd_u32(r); (*next_id)++; /* NEW(scheme) */
    rd_ver(r, V_SCHEME); rd_ver(r, V_SCHEMEBASE);
    p->enabled = rd_u32(r);
    if (rd_u64(r) != 4) r->err = 12;
    char s[4]; rd_raw(r, s, 4);
    if (memcmp(s, "CKKS", 4)) r->err = 12;
}
static void wr_context(wr_t *w, const wire_params *p, wire_ids *ids, int full) {
    wr_u32(w, W_SAME);
    if (!full) { wr_u32(w, ids->cc); return; }
    ids->cc = wr_new(w);
    wr_ver(w, V_CC);
    wr_name(w, NAME_PARAMS, 1, 1);
    wr_new(w);
    wr_ver(w, V_CKKSP); wr_ver(w, V_RLWEP); wr_ver(w, V_LPP);
    wr_u32(w, W_SAME);
    ids->dcrt = wr_new(w);
    wr_ver(w, V_DCRTP); wr_ver(w, V_EPBIG);
    wr_u32(w, p->M); wr_u32(w, p->N); wr_u8(w, 1);
    wr_ver(w, V_BIGINT);
    wr_bigint(w, p->bigQ, p->bigQ_msb);
    wr_bigint_zero(w); wr_bigint_zero(w); wr_bigint_zero(w);
    wr_u64(w, (uint64_t)p->L);
    for (int l = 0; l < p->L; l++) {
        wr_u32(w, W_SAME);
        ids->native[l] = wr_new(w);
        wr_ver(w, V_NATP); wr_ver(w, V_EPNAT);
        wr_u32(w, p->M); wr_u32(w, p->N); wr_u8(w, 1);
        wr_ver(w, V_NATINT);
        wr_u64(w, p->q[l]); wr_u64(w, p->psi[l]); wr_u64(w, 0); wr_u64(w, 0);
    }
    wr_bigint_zero(w);
    wr_u32(w, W_SAME);
    wr_new(w);
    wr_ver(w, V_ENCP);
    wr_u64(w, p->ptm); wr_u64(w, 0); wr_u64(w, 0); wr_u64(w, 0); wr_u32(w, 0); wr_u32(w, p->batch);
    wr_f32(w, p->sigma); wr_f32(w, p->alpha); wr_f32(w, p->hermite);
    for (int i = 0; i < p->nints; i++) wr_u32(w, p->ints[i]);
    wr_name(w, NAME_SCHEME, 2, 1);
    wr_new(w);
    wr_ver(w, V_SCHEME); wr_ver(w, V_SCHEMEBASE);
    wr_u32(w, p->enabled);
    wr_u64(w, 4); wr_raw(w, "CKKS", 4);
}

static void bigint_from_product(const uint64_t *q, int L, uint32_t words[110], uint16_t *msb);

/* DCRTPoly := [VER] u64 Lc Tower*Lc u32 format SAME (REF(dcrt) | reduced parameters) ; data goes to dst[l*N..]
 * Lc = the context's tower count for a level-0 ciphertext.  After a rescale, [PAL] DCRTPolyImpl::DropLastElement gives
 * the polynomial a parameter object of its own (a copy of the context's ILDCRTParams with the last tower popped and the
 * modulus divided by it); cereal then writes that object in line — NEW(id), the ElemParams fields, and the towers as
 * back references to the ILNativeParams the context already carried.  No fixture pins this form ("parity unpinned"). */
static void rd_dcrtpoly(rd_t *r, const wire_params *p, const wire_ids *ids, uint32_t *next_id, int *ct_limbs,
                        uint64_t *dst, uint32_t *format) {
    rd_ver(r, V_DCRTPOLY);
    uint64_t Lc = rd_u64(r);
    if (Lc < 1 || Lc > (uint64_t)p->L || (*ct_limbs && Lc != (uint64_t)*ct_limbs)) { r->err = 20; return; }
    *ct_limbs = (int)Lc;
    for (int l = 0; l < (int)Lc && !r->err; l++) {
        rd_ver(r, V_NATPOLY);
        rd_expect32(r, W_SAME);
        if (rd_u8(r) != 1) r->err = 21;
        rd_ver(r, V_NATVEC);
        if (rd_u64(r) != p->N) { r->err = 22; return; }
        if (dst) rd_raw(r, dst + (size_t)l * p->N, (size_t)p->N * 8);
        else if (r->p + (size_t)p->N * 8 <= r->n) r->p += (size_t)p->N * 8;
        else r->err = 1;
        if (rd_u64(r) != p->q[l]) r->err = 23;
        uint32_t f = rd_u32(r);
        if (format) *format = f;
        rd_expect32(r, W_SAME);
        if (rd_u32(r) != ids->native[l]) r->err = 24;
    }
    uint32_t f = rd_u32(r);
    if (format) *format = f;
    rd_expect32(r, W_SAME);
    uint32_t id = rd_u32(r);
    if ((int)Lc == p->L) { if (id != ids->dcrt) r->err = 25; return; }
    if (id != (W_NEW | *next_id)) { r->err = 26; return; }
    (*next_id)++;
    rd_ver(r, V_DCRTP); rd_ver(r, V_EPBIG);
    if (rd_u32(r) != p->M || rd_u32(r) != p->N || rd_u8(r) != 1) r->err = 27;
    rd_ver(r, V_BIGINT);
    uint32_t want[110], got[110]; uint16_t want_msb, got_msb;
    bigint_from_product(p->q, (int)Lc, want, &want_msb);
    rd_bigint(r, got, &got_msb);
    if (memcmp(want, got, sizeof want) || want_msb != got_msb) r->err = 28;
    rd_bigint_zero(r); rd_bigint_zero(r); rd_bigint_zero(r);
    if (rd_u64(r) != Lc) r->err = 29;
    for (int l = 0; l < (int)Lc && !r->err; l++) {
        rd_expect32(r, W_SAME);
        if (rd_u32(r) != ids->native[l]) r->err = 29;
    }
    rd_bigint_zero(r);
}
static void wr_dcrtpoly(wr_t *w, const wire_params *p, const wire_ids *ids, int Lc, const uint64_t *src, uint32_t format) {
    wr_ver(w, V_DCRTPOLY);
    wr_u64(w, (uint64_t)Lc);
    for (int l = 0; l < Lc; l++) {
        wr_ver(w, V_NATPOLY);
        wr_u32(w, W_SAME); wr_u8(w, 1);
        wr_ver(w, V_NATVEC);
        wr_u64(w, p->N);
        wr_raw(w, src + (size_t)l * p->N, (size_t)p->N * 8);
        wr_u64(w, p->q[l]); wr_u32(w, format);
        wr_u32(w, W_SAME); wr_u32(w, ids->native[l]);
    }
    wr_u32(w, format);
    wr_u32(w, W_SAME);
    if (Lc == p->L) { wr_u32(w, ids->dcrt); return; }
    wr_new(w);
    wr_ver(w, V_DCRTP); wr_ver(w, V_EPBIG);
    wr_u32(w, p->M); wr_u32(w, p->N); wr_u8(w, 1);
    wr_ver(w, V_BIGINT);
    uint32_t big[110]; uint16_t msb;
    bigint_from_product(p->q, Lc, big, &msb);
    wr_bigint(w, big, msb);
    wr_bigint_zero(w); wr_bigint_zero(w); wr_bigint_zero(w);
    wr_u64(w, (uint64_t)Lc);
    for (int l = 0; l < Lc; l++) { wr_u32(w, W_SAME); wr_u32(w, ids->native[l]); }
    wr_bigint_zero(w);
}

typedef struct { uint64_t depth, level; double sf; uint32_t enc; char tag[33]; int limbs; } ct_meta;

/* Ciphertext := SAME NEW(id) [VER VER] CryptoContext u64 32 tag u64 2 DCRTPoly DCRTPoly depth level sf enc NEW(id) u64 0 */
static void rd_ciphertext(rd_t *r, wire_params *p, wire_ids *ids, uint32_t *next_id, uint64_t *dst, ct_meta *m,
                          uint64_t *first_data_off) {
    rd_expect32(r, W_SAME);
    if (rd_u32(r) != (W_NEW | *next_id)) r->err = 30;
    (*next_id)++;
    rd_ver(r, V_CT); rd_ver(r, V_CRYPTOOBJ);
    rd_context(r, p, ids, next_id);
    if (rd_u64(r) != 32) { r->err = 31; return; }
    rd_raw(r, m->tag, 32); m->tag[32] = 0;
    if (rd_u64(r) != 2) { r->err = 32; return; }
    if (r->err) return;
    for (int k = 0; k < 2 && !r->err; k++) {
        if (k == 0 && first_data_off) {
            /* offset of the first coefficient run: after [VER] u64 L [VER] SAME u8 [VER] u64 N */
            rd_t probe = *r;
            rd_ver(&probe, V_DCRTPOLY); rd_u64(&probe); rd_ver(&probe, V_NATPOLY); rd_u32(&probe); rd_u8(&probe);
            rd_ver(&probe, V_NATVEC); rd_u64(&probe);
            *first_data_off = probe.p;
        }
        uint32_t fmt = 0;
        if (k == 0) m->limbs = 0;
        rd_dcrtpoly(r, p, ids, next_id, &m->limbs, dst ? dst + (size_t)k * m->limbs * p->N : NULL, &fmt);
        if (fmt != 0) r->err = 33; /* EVALUATION */
    }
    m->depth = rd_u64(r); m->level = rd_u64(r); m->sf = rd_f64(r); m->enc = rd_u32(r);
    if (rd_u32(r) != (W_NEW | *next_id)) r->err = 34;
    (*next_id)++;
    if (rd_u64(r) != 0) r->err = 35; /* metadata map must be empty */
}
static void wr_ciphertext(wr_t *w, const wire_params *p, wire_ids *ids, int full_ctx, const uint64_t *src,
                          const ct_meta *m) {
    wr_u32(w, W_SAME); wr_new(w);
    wr_ver(w, V_CT); wr_ver(w, V_CRYPTOOBJ);
    wr_context(w, p, ids, full_ctx);
    wr_u64(w, 32); wr_raw(w, m->tag, 32);
    wr_u64(w, 2);
    int Lc = m->limbs ? m->limbs : p->L;
    wr_dcrtpoly(w, p, ids, Lc, src, 0);
    wr_dcrtpoly(w, p, ids, Lc, src + (size_t)Lc * p->N, 0);
    wr_u64(w, m->depth); wr_u64(w, m->level); wr_f64(w, m->sf); wr_u32(w, m->enc);
    wr_new(w); wr_u64(w, 0);
}

/* parsed vector<Ciphertext> */
/* cL = towers per ciphertext polynomial: p.L at level 0, fewer after a rescale */
typedef struct { wire_params p; uint64_t n_ct; uint64_t *data; ct_meta *meta; uint64_t first_off, stride; int cL; } ct_vec;
static void ct_vec_free(ct_vec *v) { free(v->data); free(v->meta); v->data = NULL; v->meta = NULL; }

static int parse_ct_blob(const uint8_t *blob, uint64_t len, int single, ct_vec *out, int want_data) {
    memset(out, 0, sizeof *out);
    rd_t r = {blob, len, 0, 0, 0};
    if (rd_u8(&r) != 1) return -1;
    uint64_t n_ct = single ? 1 : rd_u64(&r);
    if (r.err || n_ct == 0 || n_ct > (len / 64)) return -2;
    wire_ids ids; memset(&ids, 0, sizeof ids);
    uint32_t next_id = 1;
    out->n_ct = n_ct;
    out->meta = calloc(n_ct, sizeof(ct_meta));
    size_t ct_words = 0;
    uint64_t start1 = 0;
    for (uint64_t j = 0; j < n_ct; j++) {
        size_t start = r.p;
        if (j == 0) {
            /* need L,N before allocating: parse ct 0 without data first */
            rd_t probe = r; wire_ids pid = ids; uint32_t pn = next_id; ct_meta pm;
            rd_ciphertext(&probe, &out->p, &pid, &pn, NULL, &pm, &out->first_off);
            if (probe.err) { ct_vec_free(out); return -3; }
            out->cL = pm.limbs;
            ct_words = (size_t)2 * out->cL * out->p.N;
            if (want_data) out->data = malloc(n_ct * ct_words * 8);
        }
        rd_ciphertext(&r, &out->p, &ids, &next_id, want_data ? out->data + j * ct_words : NULL, &out->meta[j], NULL);
        if (r.err || out->meta[j].limbs != out->cL) { ct_vec_free(out); return -4; }
        if (j == 1) start1 = start;
        if (j == 2) out->stride = start - start1;
    }
    if (n_ct == 2) out->stride = r.p - start1;
    if (r.p != len) { ct_vec_free(out); return -5; }
    return 0;
}
static void emit_ct_blob(wr_t *w, const wire_params *p, uint64_t n_ct, const uint64_t *data, const ct_meta *meta,
                         int single, int same_meta) {
    memset(w, 0, sizeof *w);
    w->next_id = 1;
    wr_u8(w, 1);
    if (!single) wr_u64(w, n_ct);
    wire_ids ids; memset(&ids, 0, sizeof ids);
    size_t ct_words = (size_t)2 * (meta[0].limbs ? meta[0].limbs : p->L) * p->N;
    for (uint64_t j = 0; j < n_ct; j++) wr_ciphertext(w, p, &ids, j == 0, data + j * ct_words, &meta[same_meta ? 0 : j]);
}

int orc_blob_parse(const uint8_t *blob, uint64_t len, int single, orc_blob_info *info, uint64_t *payload, uint64_t cap) {
    ct_vec v;
    int rc = parse_ct_blob(blob, len, single, &v, payload != NULL);
    if (rc) return rc;
    size_t words = (size_t)v.n_ct * 2 * v.cL * v.p.N;
    if (payload) {
        if (cap < words) { ct_vec_free(&v); return -6; }
        memcpy(payload, v.data, words * 8);
    }
    if (info) {
        info->n_ct = v.n_ct; info->n_polys = 2; info->n_limbs = (uint32_t)v.cL; info->ring_dim = v.p.N;
        info->depth = v.meta[0].depth; info->level = v.meta[0].level; info->scaling_factor = v.meta[0].sf;
        info->encoding_type = v.meta[0].enc;
        memcpy(info->key_tag, v.meta[0].tag, 33);
        info->first_data_offset = v.first_off; info->ct_stride = v.stride;
    }
    ct_vec_free(&v);
    return 0;
}
int orc_blob_reemit(const uint8_t *blob, uint64_t len, int single, uint8_t **out, uint64_t *out_len) {
    ct_vec v;
    int rc = parse_ct_blob(blob, len, single, &v, 1);
    if (rc) return rc;
    wr_t w;
    emit_ct_blob(&w, &v.p, v.n_ct, v.data, v.meta, single, 0);
    *out = w.b; *out_len = w.n;
    ct_vec_free(&v);
    return 0;
}

/* key files: Key-header := SAME NEW(1) VER VER VER CryptoContext u64 32 tag ; pk: u64 2 poly poly ; sk: poly */
static int parse_key_file(const uint8_t *in, uint64_t len, int kind, wire_params *p, char tag[33], uint64_t **polys) {
    rd_t r = {in, len, 0, 0, 0};
    wire_ids ids; memset(&ids, 0, sizeof ids);
    uint32_t next_id = 1;
    if (rd_u8(&r) != 1) return -1;
    if (kind == 0) {
        rd_context(&r, p, &ids, &next_id);
        return (r.err || r.p != len) ? -2 : 0;
    }
    rd_expect32(&r, W_SAME);
    if (rd_u32(&r) != (W_NEW | 1u)) return -3;
    next_id++;
    rd_ver(&r, V_KEY0); rd_ver(&r, V_KEY1); rd_ver(&r, V_KEY2);
    rd_context(&r, p, &ids, &next_id);
    if (rd_u64(&r) != 32) return -4;
    rd_raw(&r, tag, 32); tag[32] = 0;
    if (r.err) return -5;
    int np = kind == 1 ? 2 : 1;
    if (kind == 1 && rd_u64(&r) != 2) return -6;
    size_t words = (size_t)p->L * p->N;
    *polys = malloc(np * words * 8);
    for (int k = 0; k < np; k++) {
        uint32_t fmt = 0;
        int key_limbs = p->L; /* keys always carry the full chain */
        rd_dcrtpoly(&r, p, &ids, &next_id, &key_limbs, *polys + k * words, &fmt);
        if (fmt != 0) r.err = 40;
    }
    if (r.err || r.p != len) { free(*polys); *polys = NULL; return -7; }
    return 0;
}
static void emit_key_file(wr_t *w, int kind, const wire_params *p, const char tag[33], const uint64_t *polys) {
    memset(w, 0, sizeof *w);
    w->next_id = 1;
    wire_ids ids; memset(&ids, 0, sizeof ids);
    wr_u8(w, 1);
    if (kind == 0) { wr_context(w, p, &ids, 1); return; }
    wr_u32(w, W_SAME); wr_new(w);
    wr_ver(w, V_KEY0); wr_ver(w, V_KEY1); wr_ver(w, V_KEY2);
    wr_context(w, p, &ids, 1);
    wr_u64(w, 32); wr_raw(w, tag, 32);
    size_t words = (size_t)p->L * p->N;
    if (kind == 1) { wr_u64(w, 2); wr_dcrtpoly(w, p, &ids, p->L, polys, 0); wr_dcrtpoly(w, p, &ids, p->L, polys + words, 0); }
    else wr_dcrtpoly(w, p, &ids, p->L, polys, 0);
}
int orc_file_reemit(const uint8_t *in, uint64_t len, int kind, uint8_t **out, uint64_t *out_len) {
    wire_params p; char tag[33] = {0}; uint64_t *polys = NULL;
    int rc = parse_key_file(in, len, kind, &p, tag, &polys);
    if (rc) return rc;
    wr_t w;
    emit_key_file(&w, kind, &p, tag, polys);
    free(polys);
    *out = w.b; *out_len = w.n;
    return 0;
}

/* ------------------------------------------------------------------------------------------- */
/* loading (ckks.cpp:11-23)                                                                     */
/* ------------------------------------------------------------------------------------------- */
static uint8_t *read_file(const char *path, size_t *len) {
    FILE *f = fopen(path, "rb");
    if (!f) return NULL;
    fseek(f, 0, SEEK_END);
    long n = ftell(f);
    fseek(f, 0, SEEK_SET);
    uint8_t *b = malloc(n > 0 ? (size_t)n : 1);
    if (fread(b, 1, (size_t)n, f) != (size_t)n) { free(b); fclose(f); return NULL; }
    fclose(f);
    *len = (size_t)n;
    return b;
}
static uint64_t *shoup_array(const uint64_t *x, const wire_params *p) {
    uint64_t *o = malloc((size_t)p->L * p->N * 8);
    for (int l = 0; l < p->L; l++)
        for (uint32_t j = 0; j < p->N; j++) o[(size_t)l * p->N + j] = shoup_pre(x[(size_t)l * p->N + j], p->q[l]);
    return o;
}
int orc_load_from_memory(orc_ctx *c, const uint8_t *cc, size_t cc_len, const uint8_t *pk, size_t pk_len,
                         const uint8_t *sk, size_t sk_len) {
    int status = 0;
    wire_params p; char tag[33]; uint64_t *polys = NULL;
    c->have_cc = c->have_pk = c->have_sk = 0;
    if (cc && parse_key_file(cc, cc_len, 0, &p, tag, &polys) == 0) { c->wp = p; c->have_cc = 1; build_tables(c); }
    else status |= 1;
    if (c->have_cc && pk && parse_key_file(pk, pk_len, 1, &p, tag, &polys) == 0) {
        size_t words = (size_t)p.L * p.N;
        free(c->pk_b); free(c->pk_a); free(c->pk_bp); free(c->pk_ap);
        c->pk_b = malloc(words * 8); c->pk_a = malloc(words * 8);
        memcpy(c->pk_b, polys, words * 8); memcpy(c->pk_a, polys + words, words * 8);
        c->pk_bp = shoup_array(c->pk_b, &p); c->pk_ap = shoup_array(c->pk_a, &p);
        memcpy(c->key_tag, tag, 33);
        free(polys); polys = NULL;
        c->have_pk = 1;
    } else status |= 2;
    if (c->have_cc && sk && parse_key_file(sk, sk_len, 2, &p, tag, &polys) == 0) {
        free(c->sk_s); free(c->sk_sp);
        c->sk_s = polys; polys = NULL;
        c->sk_sp = shoup_array(c->sk_s, &p);
        if (!c->have_pk) memcpy(c->key_tag, tag, 33);
        c->have_sk = 1;
    } else status |= 4;
    return status;
}
int orc_load_crypto_params(orc_ctx *c) {
    char path[1200];
    size_t n[3] = {0, 0, 0};
    uint8_t *b[3];
    static const char *names[3] = {"cryptocontext.txt", "key-public.txt", "key-private.txt"};
    for (int i = 0; i < 3; i++) { /* ckks.cpp:12,16,20: cryptodir is concatenated raw */
        snprintf(path, sizeof path, "%s%s", c->cryptodir, names[i]);
        b[i] = read_file(path, &n[i]);
    }
    int rc = orc_load_from_memory(c, b[0], n[0], b[1], n[1], b[2], n[2]);
    for (int i = 0; i < 3; i++) free(b[i]);
    return rc;
}

/* ------------------------------------------------------------------------------------------- */
/* blob level (ckks.cpp:61-104, :170-213, :264-320)                                             */
/* ------------------------------------------------------------------------------------------- */
uint64_t orc_num_ciphertexts(const orc_ctx *c, uint64_t n) { return (n + c->batch - 1) / c->batch; }

static void default_meta(const orc_ctx *c, ct_meta *m, uint64_t depth, uint64_t level, double sf) {
    m->depth = depth; m->level = level; m->sf = sf; m->enc = 4; /* CKKSPacked */
    m->limbs = c->wp.L - (int)level; /* EXACTRESCALE: one tower dropped per level */
    memcpy(m->tag, c->key_tag, 33);
}
int orc_blob_build(const orc_ctx *c, uint64_t n_ct, const uint64_t *payload, uint64_t depth, uint64_t level, double sf,
                   uint8_t **out, uint64_t *out_len) {
    ct_meta m; default_meta(c, &m, depth, level, sf);
    wr_t w;
    emit_ct_blob(&w, &c->wp, n_ct, payload, &m, 0, 1);
    *out = w.b; *out_len = w.n;
    return 0;
}
uint64_t orc_blob_size(const orc_ctx *c, uint64_t n_ct) {
    /* header sizes do not depend on the payload: build a 1- and 2-ct blob of zeros once */
    size_t words = (size_t)2 * c->wp.L * c->wp.N;
    uint64_t *z = calloc(2 * words, 8);
    uint8_t *b; uint64_t l1, l2;
    orc_blob_build(c, 1, z, 1, 0, 1.0, &b, &l1); free(b);
    orc_blob_build(c, 2, z, 1, 0, 1.0, &b, &l2); free(b);
    free(z);
    return l1 + (n_ct - 1) * (l2 - l1);
}

int orc_encrypt(orc_ctx *c, const double *x, uint64_t n, uint64_t seed, uint8_t **blob, uint64_t *len) {
    if (!c->have_cc || !c->have_pk) { snprintf(c->err, sizeof c->err, "crypto params not loaded"); return -1; }
    if (n == 0) { snprintf(c->err, sizeof c->err, "empty input (undefined behaviour in ckks.cpp:66,95)"); return -2; }
    if (c->batch > c->wp.batch) { snprintf(c->err, sizeof c->err, "batchSize exceeds context slots"); return -3; }
    uint64_t n_ct = orc_num_ciphertexts(c, n); /* ckks.cpp:65 */
    size_t words = (size_t)2 * c->wp.L * c->wp.N;
    uint64_t *data = malloc(n_ct * words * 8);
    int bad = 0;
#pragma omp parallel for schedule(dynamic) /* ckks.cpp:70 */
    for (uint64_t i = 0; i < n_ct; i++) {
        uint64_t first = i * c->batch, last = first + c->batch < n ? first + c->batch : n; /* ckks.cpp:72 */
        if (orc_encrypt_ct(c, x + first, (uint32_t)(last - first), seed, i, data + i * words)) bad = 1;
    }
    if (bad) { free(data); return -4; }
    int rc = orc_blob_build(c, n_ct, data, 1, 0, orc_scaling_factor(c), blob, len); /* ckks.cpp:98-101 */
    free(data);
    return rc;
}

int orc_weighted_average(orc_ctx *c, const uint8_t *const *blobs, const uint64_t *lens, const float *w, uint32_t k,
                         uint8_t **blob, uint64_t *len) {
    if (!c->have_cc) { snprintf(c->err, sizeof c->err, "crypto params not loaded"); return -1; }
    if (k == 0) { snprintf(c->err, sizeof c->err, "no learners"); return -2; }
    ct_vec acc; memset(&acc, 0, sizeof acc);
    double sf0 = orc_scaling_factor(c);
    for (uint32_t i = 0; i < k; i++) { /* ckks.cpp:273: serial learner loop */
        ct_vec v;
        if (parse_ct_blob(blobs[i], lens[i], 0, &v, 1)) { /* ckks.cpp:278-281 */
            snprintf(c->err, sizeof c->err, "learner %u: malformed ciphertext blob", i);
            ct_vec_free(&acc); return -3;
        }
        if (v.p.N != c->wp.N || v.p.L != c->wp.L || memcmp(v.p.q, c->wp.q, sizeof(uint64_t) * v.p.L) ||
            v.cL != v.p.L || (i > 0 && v.n_ct != acc.n_ct)) {
            snprintf(c->err, sizeof c->err, "learner %u: ciphertext parameters differ", i);
            ct_vec_free(&v); ct_vec_free(&acc); return -4;
        }
        int64_t sc = orc_weight_constant(c, w[i]); /* ckks.cpp:287-288 + [PAL] EvalMult(ct,double) */
        int L = v.p.L; uint32_t N = v.p.N;
        size_t words = (size_t)2 * L * N;
        for (uint64_t j = 0; j < v.n_ct; j++) {
            if (v.meta[j].depth != 1 || v.meta[j].level != 0) {
                snprintf(c->err, sizeof c->err, "learner %u ct %llu: only fresh depth-1 level-0 inputs supported", i,
                         (unsigned long long)j);
                ct_vec_free(&v); ct_vec_free(&acc); return -5;
            }
            uint64_t *ct = v.data + j * words;
#pragma omp parallel for /* [PAL] DCRTPoly ops parallelise over towers */
            for (int pl = 0; pl < 2 * L; pl++) {
                int l = pl % L;
                uint64_t q = v.p.q[l], cm = signed_mod(sc, q), cp = shoup_pre(cm, q);
                uint64_t *xx = ct + (size_t)pl * N;
                for (uint32_t t = 0; t < N; t++) xx[t] = shoup_mul(xx[t], cm, cp, q);
                if (i > 0) { /* ckks.cpp:294-297 EvalAdd */
                    uint64_t *aa = acc.data + j * words + (size_t)pl * N;
                    for (uint32_t t = 0; t < N; t++) aa[t] = addmod(aa[t], xx[t], q);
                }
            }
            v.meta[j].depth += 1;      /* [PAL] depth + 1 */
            v.meta[j].sf *= sf0;       /* [PAL] scalingFactor * SF(level 0), in double */
        }
        if (i == 0) acc = v; /* ckks.cpp:291-292 */
        else ct_vec_free(&v);
    }
    wr_t wr;
    emit_ct_blob(&wr, &acc.p, acc.n_ct, acc.data, acc.meta, 0, 0); /* ckks.cpp:308-310 */
    *blob = wr.b; *len = wr.n;
    ct_vec_free(&acc);
    return 0;
}

int orc_decrypt(orc_ctx *c, const uint8_t *blob, uint64_t len, uint64_t n, double *out) {
    if (!c->have_cc || !c->have_sk) { snprintf(c->err, sizeof c->err, "crypto params not loaded"); return -1; }
    ct_vec v;
    if (parse_ct_blob(blob, len, 0, &v, 1)) { snprintf(c->err, sizeof c->err, "malformed ciphertext blob"); return -3; }
    if (v.p.N != c->wp.N || v.p.L != c->wp.L || memcmp(v.p.q, c->wp.q, sizeof(uint64_t) * v.p.L)) {
        ct_vec_free(&v); snprintf(c->err, sizeof c->err, "ciphertext parameters differ"); return -4;
    }
    uint64_t batch = c->batch, slots = c->wp.batch;
    if (n > v.n_ct * batch || n + batch <= v.n_ct * batch || batch > slots) {
        ct_vec_free(&v); snprintf(c->err, sizeof c->err, "data_dimensions inconsistent with blob"); return -5;
    }
    size_t words = (size_t)2 * v.cL * v.p.N;
    int bad = 0;
#pragma omp parallel for schedule(dynamic) /* ckks.cpp:186 */
    for (uint64_t i = 0; i < v.n_ct; i++) {
        double *tmp = malloc(slots * sizeof(double));
        if (decrypt_ct_impl_at(c, v.data + i * words, v.cL, v.meta[i].sf, tmp, NULL, NULL, NULL, i)) bad = 1;
        uint64_t length = (i == v.n_ct - 1) ? n - i * batch : batch; /* ckks.cpp:192-196 */
        memcpy(out + i * batch, tmp, length * sizeof(double));        /* ckks.cpp:200-204 (64-bit index) */
        free(tmp);
    }
    ct_vec_free(&v);
    return bad ? -6 : 0;
}

/* Rescale-before-transport (not something ckks.cpp does — it leaves the average at depth 2):
 * [PAL] cc->ModReduce on every ciphertext of a blob.  depth - 1, level + 1, scalingFactor / q_l (in double). */
int orc_rescale_blob(orc_ctx *c, const uint8_t *blob, uint64_t len, uint8_t **out, uint64_t *out_len) {
    if (!c->have_cc) { snprintf(c->err, sizeof c->err, "crypto params not loaded"); return -1; }
    ct_vec v;
    if (parse_ct_blob(blob, len, 0, &v, 1)) { snprintf(c->err, sizeof c->err, "malformed ciphertext blob"); return -3; }
    if (v.p.N != c->wp.N || v.p.L != c->wp.L || memcmp(v.p.q, c->wp.q, sizeof(uint64_t) * v.p.L) || v.cL < 2) {
        ct_vec_free(&v); snprintf(c->err, sizeof c->err, "ciphertext parameters differ or no tower left to drop"); return -4;
    }
    int Lin = v.cL;
    size_t win = (size_t)2 * Lin * v.p.N, wout = (size_t)2 * (Lin - 1) * v.p.N;
    uint64_t *data = malloc(v.n_ct * wout * 8);
    int bad = 0;
#pragma omp parallel for schedule(dynamic)
    for (uint64_t j = 0; j < v.n_ct; j++) {
        if (orc_rescale_ct(c, v.data + j * win, (uint32_t)Lin, data + j * wout)) bad = 1;
        v.meta[j].depth -= 1; v.meta[j].level += 1;
        v.meta[j].sf /= (double)c->wp.q[Lin - 1];
        v.meta[j].limbs = Lin - 1;
    }
    if (!bad) {
        wr_t wr;
        emit_ct_blob(&wr, &v.p, v.n_ct, data, v.meta, 0, 0);
        *out = wr.b; *out_len = wr.n;
    }
    free(data); ct_vec_free(&v);
    return bad ? -6 : 0;
}

/* ------------------------------------------------------------------------------------------- */
/* genCryptoContextAndKeyGen (ckks.cpp:25-59)                                                   */
/* ------------------------------------------------------------------------------------------- */
static void bigint_from_product(const uint64_t *q, int L, uint32_t words[110], uint16_t *msb) {
    uint32_t le[110] = {0}; /* little-endian accumulator */
    le[0] = 1;
    for (int l = 0; l < L; l++) {
        u128 carry = 0;
        for (int i = 0; i < 110; i++) {
            u128 cur = (u128)le[i] * q[l] + carry;
            le[i] = (uint32_t)cur;
            carry = cur >> 32;
        }
    }
    int top = 109;
    while (top > 0 && !le[top]) top--;
    int bits = 0;
    for (uint32_t v = le[top]; v; v >>= 1) bits++;
    *msb = (uint16_t)(top * 32 + bits);
    for (int i = 0; i < 110; i++) words[i] = le[109 - i]; /* most-significant word first */
}

static int write_file(const char *dir, const char *name, const uint8_t *b, size_t n) {
    char path[1200];
    snprintf(path, sizeof path, "%s%s", dir, name);
    FILE *f = fopen(path, "wb");
    if (!f) return 0;
    int ok = fwrite(b, 1, n, f) == n;
    fclose(f);
    return ok;
}

int orc_gen_context_and_keys(uint32_t batch, uint32_t scale_bits, const char *cryptodir, uint64_t seed, int nine_int) {
    return orc_gen_context_and_keys_ex(batch, scale_bits, 1, 0, cryptodir, seed, nine_int); /* ckks.cpp:26 */
}
/* the same with multDepth and ring dimension free (ring_dim 0: PALISADE's choice): contexts the reference's
 * hard-coded multDepth = 1 cannot produce (BASELINE configs 3-4) */
int orc_gen_context_and_keys_ex(uint32_t batch, uint32_t scale_bits, uint32_t mult_depth, uint32_t ring_dim,
                                const char *cryptodir, uint64_t seed, int nine_int) {
    wire_params p; memset(&p, 0, sizeof p);
    uint32_t N = ring_dim ? ring_dim : orc_select_ring_dim_ex(scale_bits, mult_depth);
    if (batch > N / 2 || batch == 0 || (batch & (batch - 1)) || (N & (N - 1))) return 0;
    p.N = N; p.M = 2 * N;
    p.L = orc_modulus_chain(scale_bits, mult_depth, N, p.q); /* multDepth + 1 limbs */
    if (p.L < 1) return 0;
    for (int l = 0; l < p.L; l++) p.psi[l] = orc_min_root_of_unity(p.M, p.q[l]);
    bigint_from_product(p.q, p.L, p.bigQ, &p.bigQ_msb);
    p.ptm = scale_bits; p.batch = batch;
    p.sigma = 3.19f; p.alpha = 9.0f; p.hermite = 1.006f;
    /* relinWindow, depth, maxDepth, [extra 0], mode(OPTIMIZED), stdLevel, rsTech(EXACT), ksTech(HYBRID), dnum */
    uint32_t dnum = orc_num_large_digits(mult_depth);
    if (nine_int) { uint32_t v[9] = {0, 1, 2, 0, 1, 0, 2, 1, dnum}; memcpy(p.ints, v, sizeof v); p.nints = 9; }
    else { uint32_t v[8] = {0, 1, 2, 1, 0, 2, 1, dnum}; memcpy(p.ints, v, sizeof v); p.nints = 8; }
    p.enabled = 5; /* ENCRYPTION | SHE (ckks.cpp:32-33) */

    orc_ctx *c = orc_create(batch, scale_bits, cryptodir);
    c->wp = p; c->have_cc = 1;
    build_tables(c);
    wr_t w;
    emit_key_file(&w, 0, &p, NULL, NULL);
    int ok = write_file(cryptodir, "cryptocontext.txt", w.b, w.n); /* ckks.cpp:41 */
    free(w.b);
    if (!ok) { orc_destroy(c); return 0; }

    /* [PAL] LPAlgorithmCKKS::KeyGen: a <- uniform (EVALUATION), s <- ternary, e <- Gaussian, b = e - a*s.
     * s and e are drawn under the secret test key, a and the key tag under the public one:
     *   s: the ternary stream of nonce (0, 0, STREAM_KEY_S);  e[16b + i]: word i of block b, nonce (0, 0, STREAM_KEY_E)
     *   a_l[j]: attempt t: words (2(j&7), 2(j&7)+1) of block j >> 3, nonce (l, t, STREAM_KEY_A), top bits(q_l) bits, accepted below q_l */
    size_t words = (size_t)p.L * N;
    uint64_t *pk = malloc(2 * words * 8), *sk = malloc(words * 8), *e = malloc(words * 8);
    uint32_t out[16], ksec[8], kpub[8];
    orc_seed_key(seed, KEY_KEYGEN_SECRET, ksec);
    orc_seed_key(seed, KEY_KEYGEN_PUBLIC, kpub);
    {
        int64_t *sv = malloc((size_t)N * 8);
        ternary_stream(ksec, 0, 0, STREAM_KEY_S, STREAM_KEY_S_RETRY, N, sv);
        for (uint32_t j = 0; j < N; j++)
            for (int l = 0; l < p.L; l++) sk[(size_t)l * N + j] = signed_mod(sv[j], p.q[l]);
        free(sv);
    }
    uint32_t n_e[3] = {0, 0, STREAM_KEY_E}, n_er[3] = {0, 0, STREAM_KEY_E_RETRY};
    for (uint32_t b = 0; b < N / 16; b++) {
        orc_chacha20_block(ksec, b, n_e, out);
        for (uint32_t i = 0; i < 16; i++) {
            int64_t ev = gauss_draw(c->gauss_cdf, out[i], ksec, n_er, 16 * b + i, 0);
            for (int l = 0; l < p.L; l++) e[(size_t)l * N + 16 * b + i] = signed_mod(ev, p.q[l]);
        }
    }
    for (int l = 0; l < p.L; l++) {
        orc_ntt_forward(c, l, sk + (size_t)l * N);
        orc_ntt_forward(c, l, e + (size_t)l * N);
        int bits = 64 - __builtin_clzll(p.q[l]);
        for (uint32_t j = 0; j < N; j++) {
            uint64_t a;
            uint32_t tries = 0;
            do { /* rejection sampling of a 64-bit word below q */
                uint32_t n_a[3] = {(uint32_t)l, tries++, STREAM_KEY_A};
                orc_chacha20_block(kpub, j >> 3, n_a, out);
                a = (((uint64_t)out[2 * (j & 7) + 1] << 32) | out[2 * (j & 7)]) >> (64 - bits);
            } while (a >= p.q[l]);
            pk[words + (size_t)l * N + j] = a;
            pk[(size_t)l * N + j] = submod(e[(size_t)l * N + j], mulmod(a, sk[(size_t)l * N + j], p.q[l]), p.q[l]);
        }
    }
    char tag[33]; /* [PAL] GenerateUniqueKeyID: 32 hex chars */
    uint32_t n_t[3] = {0, 0, STREAM_KEY_TAG};
    orc_chacha20_block(kpub, 0, n_t, out);
    snprintf(tag, sizeof tag, "%08x%08x%08x%08x", out[0], out[1], out[2], out[3]);
    emit_key_file(&w, 1, &p, tag, pk);
    ok = write_file(cryptodir, "key-public.txt", w.b, w.n); /* ckks.cpp:48 */
    free(w.b);
    if (ok) {
        emit_key_file(&w, 2, &p, tag, sk);
        ok = write_file(cryptodir, "key-private.txt", w.b, w.n); /* ckks.cpp:53 */
        free(w.b);
    }
    free(pk); free(sk); free(e);
    orc_destroy(c);
    return ok ? 1 : 0;
}
