// kernels.h — host-side launchers of the sm_100a kernels (kernels.cu).  Plain C++ signatures, no
// CUDA types beyond cudaStream_t, so engine.cpp stays ordinary host code.
#pragma once
#include <cuda_runtime.h>
#include "ckks_core.cuh"

namespace shelfi {

cudaError_t kernels_init();  // opt-in shared-memory sizes; call once per device

// transforms of `rows` limb polynomials laid out [row][N], limb = row % L.  Forward: in == out is allowed only for
// ring dimension 8192 (one tile per polynomial); inverse is in place.
void launch_ntt_forward(const RingConst &rc, const u64 *in, u64 *out, u32 rows, cudaStream_t s);
void launch_ntt_inverse(const RingConst &rc, u64 *data, u32 rows, cudaStream_t s);
// experiment: last stage through a lane exchange instead of shared memory (ring dimension 8192; profiles/README.md)
void launch_ntt_forward_shfl(const RingConst &rc, const u64 *in, u64 *out, u32 rows, cudaStream_t s);

// encode + sample (rings above one tile): x[n] -> m_e0[n_ct][N] (m + e0, signed) and ve[n_ct][N] (pack_ve bytes)
// `sel` (optional, here and below): selective encryption — chunk element i is x[sel[i]] on the way in and
// out[sel[i]] on the way out, so gather and scatter cost no pass of their own
void launch_encode_sample(const RingConst &rc, const EncodeConst &ec, const SamplerConst &sc, const double *x, u64 n,
                          u32 n_ct, i64 *m_e0, uint8_t *ve, cudaStream_t s, const u32 *sel = nullptr, bool f32 = false);
// encrypt: three forward NTTs per (ciphertext, limb) + public-key combine -> ct[n_ct][2][L][N]
void launch_encrypt(const RingConst &rc, const i64 *m_e0, const uint8_t *ve, const tw_t *pk_b, const tw_t *pk_a,
                    u64 *ct, u32 n_ct, cudaStream_t s);
// encode + sample + encrypt in one kernel, one CTA per ciphertext, ring dimension 8192 (no scratch: the draws stay in
// shared memory and m + e0 is parked in an output row of the ciphertext itself)
void launch_encrypt_fused(const RingConst &rc, const EncodeConst &ec, const SamplerConst &sc, const EncryptBatch &eb,
                          u64 n, const tw_t *pk_b, const tw_t *pk_a, cudaStream_t s);
// key generation: s, e (coefficient form, rows[2L][N]) and uniform a[L][N]; after launch_ntt_forward over the 2L rows,
// b = e^ - a s^
void launch_keygen_sample(const RingConst &rc, const SamplerConst &secret, const SamplerKey &pub, u64 *rows, u64 *a,
                          cudaStream_t s);
void launch_keygen_combine(const RingConst &rc, const u64 *rows_hat, const u64 *a, u64 *b, cudaStream_t s);
void launch_chacha20_block(const SamplerKey &key, u32 counter, u32 n0, u32 n1, u32 n2, u32 *out, cudaStream_t s);
// decrypt stage 1: c0 + c1*s, inverse NTT -> coeff[n_ct][L][N]
void launch_decrypt_intt(const RingConst &rc, const u64 *ct, const tw_t *sk, u64 *coeff, u32 n_ct, cudaStream_t s);
// rescale by one level: ct[n_ct][2][L][N] -> out[n_ct][2][L-1][N] (rc.limbs = L = towers of the input);
// t = scratch of n_ct * 2 * N words (the dropped tower in coefficient form)
void launch_rescale(const RingConst &rc, const RescaleConst &rs, const u64 *ct, u64 *t, u64 *out, u32 n_ct,
                    cudaStream_t s);
// decrypt stage 2: CRT, scale, FFTSpecial -> out[n] (chunk length `batch`, last chunk short)
void launch_decode(const RingConst &rc, const EncodeConst &ec, const DecodeConst &dc, const u64 *coeff, u32 n_ct,
                   double *out, u64 n, cudaStream_t s, const u32 *sel = nullptr);
// both stages in one launch (ring dimension 8192, at most two limbs): same results as launch_decrypt_intt + launch_decode
void launch_decrypt_fused(const RingConst &rc, const EncodeConst &ec, const DecodeConst &dc, const u64 *ct,
                          const tw_t *sk, u64 *coeff, u32 n_ct, double *out, u64 n, cudaStream_t s,
                          const u32 *sel = nullptr);
// decrypt (both stages) with the scratch in a file-scope device array of kStaticScratchWords u64 (8 MiB) instead of
// caller- or handle-owned memory: allocates nothing.  general = the launch_crt + launch_decode_fft route (any ring and
// limb count); otherwise launch_decode's fused route (ring dimension 8192, at most two limbs).
// static_scratch_words() is what a call needs; beyond the array launch_decrypt_static returns cudaErrorInvalidValue.
constexpr u64 kStaticScratchWords = (8ull << 20) / sizeof(u64);
u64 static_scratch_words(const RingConst &rc, const EncodeConst &ec, u64 n_ct, bool general);
cudaError_t launch_decrypt_static(const RingConst &rc, const EncodeConst &ec, const DecodeConst &dc, const CrtNConst &cn,
                                  bool general, const u64 *ct, const tw_t *sk, u32 n_ct, double *out, u64 n,
                                  cudaStream_t s);
// general decode: CRT of any limb count -> vals[n_ct][N] doubles (x dc.pre), optional taps of the centred integer
// (tap8: 8 two's-complement words per coefficient; tap_hi/tap_lo: its low 128 bits)
void launch_crt(const RingConst &rc, const DecodeConst &dc, const CrtNConst &cn, const u64 *coeff, u32 n_ct,
                double *vals, u64 *tap8, i64 *tap_hi, u64 *tap_lo, cudaStream_t s);
// [PAL >= 1.10.6] Decode's noise flooding on vals (between launch_crt and launch_decode_fft), one key per call
void launch_decode_noise(const RingConst &rc, const EncodeConst &ec, const SamplerKey &key, u64 ct_index0, double *vals,
                         u32 n_ct, cudaStream_t s);
// ... then FFTSpecial per 4096-slot tile (+ tail pass for more slots; zs = scratch of n_ct * slots complex)
void launch_decode_fft(const RingConst &rc, const EncodeConst &ec, const DecodeConst &dc, const double *vals, cplx *zs,
                       u32 n_ct, double *out, u64 n, cudaStream_t s, const u32 *sel = nullptr);
// debug/parity taps of the two halves above
void launch_encode_only(const RingConst &rc, const EncodeConst &ec, const double *x, u64 n, u32 n_ct, i64 *m,
                        cudaStream_t s, const u32 *sel = nullptr, bool f32 = false);
void launch_crt_only(const RingConst &rc, const DecodeConst &dc, const u64 *coeff, u32 n_ct, i64 *hi, u64 *lo,
                     cudaStream_t s);

// out = sum_k c_k * in_k mod q over dense payloads of n_ct ciphertexts
void launch_weighted_sum(const RingConst &rc, const WavgArgs &wa, u64 *out, u64 n_ct, cudaStream_t s);

// EvalMult(ct, plaintext vector): encoded weights m[n_ct][N] -> evaluation form pt_hat[n_ct][L][N], then
// out (+)= ct (*) pt_hat on both polynomials
void launch_plain_ntt(const RingConst &rc, const i64 *m, u64 *pt_hat, u32 n_ct, cudaStream_t s);
void launch_ct_pt_mac(const RingConst &rc, const u64 *ct, const u64 *pt_hat, u64 *out, u64 n_ct, bool accumulate,
                      cudaStream_t s);
// out[i] (+)= sum_k w_k * x_k[i] in plaintext (the unencrypted part of a selectively encrypted model)
void launch_plain_weighted_sum(const PlainWavgArgs &pa, u64 n, double *out, cudaStream_t s);

// serialized blob image <-> dense payload
void launch_unpack(const RingConst &rc, const BlobRuns &br, const u64 *image, u64 *dense, u64 n_ct, cudaStream_t s);
void launch_pack(const RingConst &rc, const BlobRuns &br, const u64 *dense, u64 *image, u64 n_ct, cudaStream_t s);
// acc (+)= c * unpack(image) : one learner streamed into the resident accumulator
void launch_unpack_scale_accumulate(const RingConst &rc, const BlobRuns &br, const u64 *image, const LimbScalars &cs,
                                    u32 first, u64 *acc, u64 n_ct, cudaStream_t s);

// serialized image: bookkeeping bytes (everything between the coefficient runs) copied in from a packed table
struct SegDesc { unsigned long long dst; unsigned src, len; };   // image byte offset, offset into `bytes`, length
void launch_write_segments(const SegDesc *desc, const unsigned char *bytes, unsigned char *image, u32 n_seg, cudaStream_t s);

// Barrier between the GPUs of one box, on the stream: rank `me` publishes `epoch` in slot `me` of every rank's flag
// array (peer-mapped) and waits until its own array shows `epoch` in every slot.  flags[r] -> rank r's array of >= 9
// u64 (slot 8 = status: set to `epoch` if the wait gave up after `timeout_ns`).
struct PeerFlags { unsigned long long *flags[8]; };
// poison (optional): on a timeout waiting for rank r, poison[r * poison_stride] is set to NaN (the first value of the
// slice that rank owed this GPU), so a consumer that never looks at the status word still cannot mistake the result
void launch_peer_barrier(const PeerFlags &pf, u32 n_ranks, u32 me, u64 epoch, u64 timeout_ns, double *poison,
                         u64 poison_stride, cudaStream_t s);

}  // namespace shelfi
