/*
 * shelfi_oracle.h — CPU ORACLE for the SHELFI_FHE CKKS path.  TEST INFRASTRUCTURE ONLY.
 *
 * This is a plain-C restatement of what the reference's hot path computes
 * (palisade_pybind/SHELFI_FHE/src/ckks.cpp:11-23, :25-59, :61-104, :170-213, :264-320)
 * including the PALISADE semantics it calls into (PALISADE is a third-party dependency that
 * is NOT vendored in /root/reference: Dockerfile:27 clones palisade-release at HEAD, the
 * commented alternative at Dockerfile:22 is release-v1.11.2).
 *
 * PARITY STATUS: the wire format, modulus chain, psi, NTT ordering and key relation are
 * pinned by the reference's binary fixtures (tests/golden, see tests/test_oracle_fixtures.py).
 * The numeric behaviour of encode/decode/EvalMult(double) is restated from PALISADE v1.11's
 * published algorithm; the reference has no asserting tests for it => "parity unpinned" for
 * those three.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (fhe-fed_b200/) never links or imports it.
 */
#ifndef SHELFI_ORACLE_H
#define SHELFI_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_MAX_LIMBS 8

typedef struct orc_ctx orc_ctx;

/* ---- lifecycle (ckks.cpp:5-9, :11-23) ---- */
/* batch = the CKKS ctor's batchSize (chunk length); cryptodir must end in '/' like the reference */
orc_ctx *orc_create(uint32_t batch, uint32_t scale_bits, const char *cryptodir);
void orc_destroy(orc_ctx *c);
/* returns 0 ok; bit0: context missing, bit1: public key missing, bit2: secret key missing */
int orc_load_crypto_params(orc_ctx *c);
/* load from memory images of the three files (any may be NULL) */
int orc_load_from_memory(orc_ctx *c, const uint8_t *cc, size_t cc_len, const uint8_t *pk, size_t pk_len,
                         const uint8_t *sk, size_t sk_len);
/* ckks.cpp:25-59: returns 1 ok / 0 on failure, writes the three files, does NOT load them.
 * seed makes the key material reproducible (PALISADE uses OS entropy). nine_int selects the newer
 * 9-int RLWE block wire variant. */
int orc_gen_context_and_keys(uint32_t batch, uint32_t scale_bits, const char *cryptodir, uint64_t seed,
                             int nine_int);
/* extension: multDepth and ring dimension free (ring_dim 0 = PALISADE's automatic choice); mult_depth = 1,
 * ring_dim = 0 is orc_gen_context_and_keys */
int orc_gen_context_and_keys_ex(uint32_t batch, uint32_t scale_bits, uint32_t mult_depth, uint32_t ring_dim,
                                const char *cryptodir, uint64_t seed, int nine_int);
const char *orc_last_error(const orc_ctx *c);

/* ---- introspection ---- */
uint32_t orc_ring_dim(const orc_ctx *c);
uint32_t orc_num_limbs(const orc_ctx *c);
uint32_t orc_slots(const orc_ctx *c); /* encoding-params batch size of the loaded context */
uint64_t orc_modulus(const orc_ctx *c, int limb);
uint64_t orc_psi(const orc_ctx *c, int limb);
double orc_scaling_factor(const orc_ctx *c); /* q_last as double */
int orc_wire_variant(const orc_ctx *c);      /* 8 or 9 */
/* copies L*N words each (EVALUATION form, limb-major); returns 0 if the key is not loaded */
int orc_get_public_key(const orc_ctx *c, uint64_t *b, uint64_t *a);
int orc_get_secret_key(const orc_ctx *c, uint64_t *s);
void orc_get_key_tag(const orc_ctx *c, char out[33]);

/* ---- parameter generation rules ---- */
uint64_t orc_first_prime(uint32_t bits, uint32_t m);          /* smallest prime >= 2^bits, = 1 mod m */
uint64_t orc_previous_prime(uint64_t q, uint32_t m);          /* next smaller prime = 1 mod m        */
uint64_t orc_min_root_of_unity(uint32_t m, uint64_t q);       /* smallest primitive m-th root mod q  */
/* ring dimension PALISADE's genCryptoContextCKKS picks for (multDepth=1, bits, batch) */
uint32_t orc_select_ring_dim(uint32_t scale_bits, uint32_t batch);
uint32_t orc_select_ring_dim_ex(uint32_t scale_bits, uint32_t mult_depth);
uint32_t orc_num_large_digits(uint32_t mult_depth);
/* EXACTRESCALE modulus chain q[0..mult_depth]; returns the limb count or -1 */
int orc_modulus_chain(uint32_t scale_bits, uint32_t mult_depth, uint32_t ring_dim, uint64_t *q);

/* ---- primitives (PALISADE ChineseRemainderTransformFTT / DiscreteFourierTransform) ---- */
void orc_ntt_forward(const orc_ctx *c, int limb, uint64_t *a); /* natural -> bit-reversed, in place */
void orc_ntt_inverse(const orc_ctx *c, int limb, uint64_t *a); /* bit-reversed -> natural          */
void orc_fft_special(const orc_ctx *c, double *re_im, uint32_t size);     /* interleaved re,im */
void orc_fft_special_inv(const orc_ctx *c, double *re_im, uint32_t size);
/* RFC 8439 ChaCha20 block function and the test-key layout shared with the device sampler */
void orc_chacha20_block(const uint32_t key[8], uint32_t counter, const uint32_t nonce[3], uint32_t out[16]);
void orc_seed_key(uint64_t seed, uint32_t domain, uint32_t key[8]);
/* the unpinned semantics, switchable: option 1 = EvalMult rounding (0: (int64)(c*Delta + 0.5), 1: llround),
 * option 2 = decode noise (0 off, 1 on), option 4 = test seed of the decode noise */
int orc_set_option(orc_ctx *c, int option, int64_t value);
/* [PAL >= 1.10.6] Decode's noise flooding on slots complex values at scale 2^p; returns the noise's standard deviation */
double orc_decode_noise(const orc_ctx *c, double *v, uint32_t n, const uint32_t key[8], uint64_t ct_index);
/* the two primitive samplers on caller-chosen random words (unit tests of the rare paths) */
int64_t orc_gauss_from_words(const orc_ctx *c, uint32_t h, uint32_t l);
int64_t orc_ternary_from_half(uint32_t h16, const uint32_t key[8], const uint32_t nonce_retry[3], uint32_t j);
/* fills v (ternary), e0, e1 (discrete Gaussian) for ciphertext index ct_index: N values each */
void orc_sample_encrypt_noise(const orc_ctx *c, uint64_t seed, uint64_t ct_index, int64_t *v, int64_t *e0,
                              int64_t *e1);

/* MakeCKKSPackedPlaintext (ckks.cpp:80): n <= slots real values -> N signed coefficients */
int orc_encode_coeffs(const orc_ctx *c, const double *x, uint32_t n, int64_t *coeff);
/* per-ciphertext Encrypt (ckks.cpp:81): out = [c0 limb0..L-1][c1 limb0..L-1], EVALUATION form */
int orc_encrypt_ct(const orc_ctx *c, const double *x, uint32_t n, uint64_t seed, uint64_t ct_index,
                   uint64_t *out);
/* per-ciphertext Decrypt+Decode (ckks.cpp:189-199). ct = 2*L*N words; sf = stored scaling factor.
 * If coeff_hi/coeff_lo are non-NULL they receive the centred CRT integer of every coefficient
 * (two's-complement 128 bit) — the bit-exact integer stage. out gets `slots` doubles. */
int orc_decrypt_ct(const orc_ctx *c, const uint64_t *ct, double sf, double *out, int64_t *coeff_hi,
                   uint64_t *coeff_lo);
/* any limb count: words = N x 8 little-endian u64, two's complement of the centred CRT integer */
int orc_decrypt_ct_words(const orc_ctx *c, const uint64_t *ct, double sf, double *out, uint64_t *words);
/* acc (+)= ct (*) NTT(Encode(w)): EvalMult(ct, plaintext vector) + EvalAdd, one ciphertext [2][L][N] */
int orc_ct_mul_plain_accumulate(const orc_ctx *c, const uint64_t *ct, const double *w, uint32_t n, int first, uint64_t *acc);
/* ---- rescale-before-transport: PALISADE ModReduce under EXACTRESCALE ---- */
/* the per-tower constants of DCRTPoly::DropLastElementAndScale for a polynomial of Lin towers */
void orc_rescale_constants(const orc_ctx *c, uint32_t Lin, uint64_t *t, uint64_t *ql_inv);
/* one ciphertext [2][Lin][N] -> [2][Lin-1][N] */
int orc_rescale_ct(const orc_ctx *c, const uint64_t *in, uint32_t Lin, uint64_t *out);
/* Decrypt+Decode of a ciphertext that kept only its first `limbs` towers ([2][limbs][N]); words as above */
int orc_decrypt_ct_level(const orc_ctx *c, const uint64_t *ct, uint32_t limbs, double sf, double *out, uint64_t *words);
/* every ciphertext of a blob rescaled by one level; metadata: depth - 1, level + 1, scalingFactor / q_l */
int orc_rescale_blob(orc_ctx *c, const uint8_t *blob, uint64_t len, uint8_t **out, uint64_t *out_len);
/* EvalMult(ct,(float)w) constant: (int64)((double)w * sf0 + 0.5)   (ckks.cpp:287-288) */
int64_t orc_weight_constant(const orc_ctx *c, float w);

/* ---- blob level, following ckks.cpp control flow ---- */
uint64_t orc_num_ciphertexts(const orc_ctx *c, uint64_t n); /* ceil(n / batch) */
uint64_t orc_blob_size(const orc_ctx *c, uint64_t n_ct);
/* ckks.cpp:61-104. *blob is malloc'd; release with orc_free. */
int orc_encrypt(orc_ctx *c, const double *x, uint64_t n, uint64_t seed, uint8_t **blob, uint64_t *len);
/* ckks.cpp:264-320. Size mismatch is the caller's concern (lists have one length here). */
int orc_weighted_average(orc_ctx *c, const uint8_t *const *blobs, const uint64_t *lens, const float *w,
                         uint32_t k, uint8_t **blob, uint64_t *len);
/* ckks.cpp:170-213 */
int orc_decrypt(orc_ctx *c, const uint8_t *blob, uint64_t len, uint64_t n, double *out);
void orc_free(void *p);

/* ---- wire helpers for tests ---- */
typedef struct {
    uint64_t n_ct;
    uint32_t n_polys, n_limbs, ring_dim;
    uint64_t depth, level;
    double scaling_factor;
    uint32_t encoding_type;
    char key_tag[33];
    uint64_t first_data_offset; /* byte offset of ct0.c0.limb0 */
    uint64_t ct_stride;         /* byte distance between consecutive ciphertexts (>= 2nd) */
} orc_blob_info;
/* parse a vector<Ciphertext> blob (single != 0: a lone Ciphertext as in CT1.txt). payload (may be
 * NULL) receives n_ct*2*L*N words. */
int orc_blob_parse(const uint8_t *blob, uint64_t len, int single, orc_blob_info *info, uint64_t *payload,
                   uint64_t payload_cap_words);
/* re-emit what orc_blob_parse read, byte for byte (round-trip check K1) */
int orc_blob_reemit(const uint8_t *blob, uint64_t len, int single, uint8_t **out, uint64_t *out_len);
/* parse + re-emit a key or context file: kind 0 = context, 1 = public key, 2 = private key */
int orc_file_reemit(const uint8_t *in, uint64_t len, int kind, uint8_t **out, uint64_t *out_len);
/* build a blob from dense payload using the loaded context's header */
int orc_blob_build(const orc_ctx *c, uint64_t n_ct, const uint64_t *payload, uint64_t depth, uint64_t level,
                   double sf, uint8_t **out, uint64_t *out_len);

#ifdef __cplusplus
}
#endif
#endif
