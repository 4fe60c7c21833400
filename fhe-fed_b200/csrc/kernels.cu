// kernels.cu — the sm_100a kernels of the SHELFI CKKS path.  Build: nvcc -gencode arch=compute_100a,code=sm_100a.
//
// Data layout in HBM: a ciphertext batch is dense limb-major SoA u64 [n_ct][2][L][N];
// every kernel maps one CTA to one 64 KiB limb polynomial (or a slice of one) so all global traffic is
// contiguous 8/16-byte-per-lane runs.  NTT working set lives in a swizzled 64 KiB shared-memory tile,
// three CTAs per SM; butterflies are Shoup/Harvey lazy 64-bit modmuls on the IMAD/ALU pipes.
#include "kernels.h"

namespace shelfi {

#define NTT_SMEM_BYTES (kN * 8)

// ------------------------------------------------------------------------------------------
// standalone forward / inverse NTT (parity taps, keygen, and the "NTT HBM GB/s" metric)
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kNttThreads, 3) ntt_forward_kernel(const __grid_constant__ RingConst rc, u64 *data) {
    extern __shared__ __align__(16) u64 sm[];
    const int tid = threadIdx.x;
    const LimbConst lc = rc.limb[blockIdx.x % rc.limbs];
    u64 *row = data + (size_t)blockIdx.x * kN;
    fwd_phase_a(tid, sm, lc, LoadGlobal{row});
    __syncthreads();
    fwd_phase_mid<7>(tid, sm, lc);
    __syncthreads();
    fwd_phase_mid<4>(tid, sm, lc);
    __syncthreads();
    fwd_phase_mid<1>(tid, sm, lc);
    __syncthreads();
    fwd_phase_e(tid, sm, lc, StoreFullReducePair{row, &lc});
}

__global__ void __launch_bounds__(kNttThreads, 3) ntt_inverse_kernel(const __grid_constant__ RingConst rc, u64 *data) {
    extern __shared__ __align__(16) u64 sm[];
    const int tid = threadIdx.x;
    const LimbConst lc = rc.limb[blockIdx.x % rc.limbs];
    u64 *row = data + (size_t)blockIdx.x * kN;
    inv_phase_e(tid, sm, lc, LoadPairGlobal{row});
    __syncthreads();
    inv_phase_mid<1>(tid, sm, lc);
    __syncthreads();
    inv_phase_mid<4>(tid, sm, lc);
    __syncthreads();
    inv_phase_mid<7>(tid, sm, lc);
    __syncthreads();
    inv_phase_last(tid, sm, lc, StoreGlobal{row});
}

// ------------------------------------------------------------------------------------------
// encode (+ sampler): one CTA per ciphertext, FFT tile in shared memory
// ------------------------------------------------------------------------------------------
template <bool SAMPLE>
__global__ void __launch_bounds__(kFftThreads) encode_kernel(const __grid_constant__ EncodeConst ec,
                                                             const __grid_constant__ SamplerConst sc, const double *x,
                                                             u64 n, i64 *m_e0, uint16_t *ve) {
    extern __shared__ __align__(16) unsigned char smraw[];
    cplx *z = reinterpret_cast<cplx *>(smraw);
    __shared__ u64 cdf[32];
    const int tid = threadIdx.x;
    const u64 ct = blockIdx.x;
    if (SAMPLE && tid < 32) cdf[tid] = sc.gauss_cdf[tid];
    u64 first = ct * ec.batch;
    u64 n_valid = n - first < ec.batch ? n - first : ec.batch;
    encode_load(tid, kFftThreads, z, ec, x + first, n_valid);
    __syncthreads();
    if (ec.slots == kFftFast) {
        fft_inv_pass8<9>(tid, z, ec.tw_inv); __syncthreads();
        fft_inv_pass8<6>(tid, z, ec.tw_inv); __syncthreads();
        fft_inv_pass8<3>(tid, z, ec.tw_inv); __syncthreads();
        fft_inv_pass8<0>(tid, z, ec.tw_inv); __syncthreads();
    } else {
        for (u32 len = ec.slots; len >= 2; len >>= 1) {
            fft_inv_stage(tid, kFftThreads, z, ec.slots, len, ec.tw_inv);
            __syncthreads();
        }
    }
    encode_finalize(tid, kFftThreads, z, ec);
    __syncthreads();
    i64 *mrow = m_e0 + ct * (u64)kN;
    if (SAMPLE) {
        encode_sample_emit(tid, kFftThreads, z, ec, sc, cdf, sc.ct_index0 + ct, mrow, ve + ct * (u64)kN);
    } else {
        for (u32 j = tid; j < (u32)kN; j += kFftThreads) mrow[j] = encode_coeff(z, ec, j);
    }
}

// ------------------------------------------------------------------------------------------
// encrypt: per (ciphertext, limb) CTA, three forward NTTs through one shared-memory tile.
//   NTT(v)      -> v^ parked in the c1 row (re-read by the same thread, L2-resident)
//   NTT(m + e0) -> c0 = b*v^ + .     NTT(e1) -> c1 = a*v^ + .  (overwrites the parked row)
// ------------------------------------------------------------------------------------------
template <class Load, class Store>
__device__ __forceinline__ void ntt_forward_tile(int tid, u64 *sm, const LimbConst &lc, Load ld, Store st) {
    fwd_phase_a(tid, sm, lc, ld);
    __syncthreads();
    fwd_phase_mid<7>(tid, sm, lc);
    __syncthreads();
    fwd_phase_mid<4>(tid, sm, lc);
    __syncthreads();
    fwd_phase_mid<1>(tid, sm, lc);
    __syncthreads();
    fwd_phase_e(tid, sm, lc, st);
}

__global__ void __launch_bounds__(kNttThreads, 3) encrypt_kernel(const __grid_constant__ RingConst rc, const i64 *m_e0,
                                                                 const uint16_t *ve, const tw_t *pk_b, const tw_t *pk_a,
                                                                 u64 *ct_out) {
    extern __shared__ __align__(16) u64 sm[];
    const int tid = threadIdx.x;
    const u32 L = rc.limbs;
    const u32 l = blockIdx.x % L;
    const u64 ct = blockIdx.x / L;
    const LimbConst lc = rc.limb[l];
    const uint16_t *ve_row = ve + ct * (u64)kN;
    const i64 *m_row = m_e0 + ct * (u64)kN;
    u64 *c0 = ct_out + ((ct * 2 + 0) * L + l) * (u64)kN;
    u64 *c1 = ct_out + ((ct * 2 + 1) * L + l) * (u64)kN;
    ntt_forward_tile(tid, sm, lc, LoadTernary{ve_row, lc.q},
                     StoreLazyPair{c1});
    __syncthreads();
    ntt_forward_tile(tid, sm, lc, LoadSigned{m_row, &lc}, StoreMulAdd{pk_b + (size_t)l * kN, c1, c0, &lc});
    __syncthreads();
    ntt_forward_tile(tid, sm, lc, LoadGaussHi{ve_row, lc.q}, StoreMulAdd{pk_a + (size_t)l * kN, c1, c1, &lc});
}

// Fused encrypt: one CTA per ciphertext does encode + sampling (FFT tile), parks m+e0 / (v,e1) in global scratch
// (L2-resident, 80 KiB per ciphertext), then reuses the same 64 KiB tile for the 3 NTTs of every limb.  Compared
// with encode_kernel + encrypt_kernel this samples once per ciphertext, removes one launch and its partial
// wave, and lets the latency-bound FFT/sampler phases of one CTA hide under the integer work of its two
// co-resident CTAs.
__global__ void __launch_bounds__(kNttThreads, 3) encrypt_fused_kernel(const __grid_constant__ RingConst rc,
                                                                       const __grid_constant__ EncodeConst ec,
                                                                       const __grid_constant__ SamplerConst sc,
                                                                       const __grid_constant__ EncryptBatch eb,
                                                                       u64 n, i64 *m_e0, uint16_t *ve,
                                                                       const tw_t *pk_b, const tw_t *pk_a) {
    extern __shared__ __align__(16) u64 sm[];
    cplx *z = reinterpret_cast<cplx *>(sm);
    __shared__ u64 cdf[32];
    const int tid = threadIdx.x;
    const u32 learner = blockIdx.x / eb.n_ct;
    const u64 ct = blockIdx.x % eb.n_ct;
    const double *x = eb.x[learner];
    u64 *ct_out = eb.ct[learner];
    SamplerConst sl;   // per-learner seed; the CDF lives in shared memory
    sl.seed = eb.seed[learner]; sl.ct_index0 = sc.ct_index0;
    if (tid < 32) cdf[tid] = sc.gauss_cdf[tid];
    const u64 first = ct * ec.batch;
    const u64 n_valid = n - first < ec.batch ? n - first : ec.batch;
    i64 *m_row = m_e0 + (u64)blockIdx.x * (u64)kN;
    uint16_t *ve_row = ve + (u64)blockIdx.x * (u64)kN;
    encode_load(tid, kNttThreads, z, ec, x + first, n_valid);
    __syncthreads();
    if (ec.slots == kFftFast) {
        fft_inv_pass8<9>(tid, z, ec.tw_inv); __syncthreads();
        fft_inv_pass8<6>(tid, z, ec.tw_inv); __syncthreads();
        fft_inv_pass8<3>(tid, z, ec.tw_inv); __syncthreads();
        fft_inv_pass8<0>(tid, z, ec.tw_inv); __syncthreads();
    } else {
        for (u32 len = ec.slots; len >= 2; len >>= 1) {
            fft_inv_stage(tid, kNttThreads, z, ec.slots, len, ec.tw_inv);
            __syncthreads();
        }
    }
    encode_finalize(tid, kNttThreads, z, ec);
    __syncthreads();
    encode_sample_emit(tid, kNttThreads, z, ec, sl, cdf, sl.ct_index0 + ct, m_row, ve_row);
    __syncthreads();  // scratch rows written by this block are visible to all of its threads; tile is free
    const u32 L = rc.limbs;
    for (u32 l = 0; l < L; l++) {
        const LimbConst lc = rc.limb[l];
        u64 *c0 = ct_out + ((ct * 2 + 0) * L + l) * (u64)kN;
        u64 *c1 = ct_out + ((ct * 2 + 1) * L + l) * (u64)kN;
        ntt_forward_tile(tid, sm, lc, LoadTernary{ve_row, lc.q},
                         StoreLazyPair{c1});
        __syncthreads();
        ntt_forward_tile(tid, sm, lc, LoadSigned{m_row, &lc}, StoreMulAdd{pk_b + (size_t)l * kN, c1, c0, &lc});
        __syncthreads();
        ntt_forward_tile(tid, sm, lc, LoadGaussHi{ve_row, lc.q}, StoreMulAdd{pk_a + (size_t)l * kN, c1, c1, &lc});
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------
// decrypt stage 1: t = c0 + c1*s per limb, inverse NTT to coefficients
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kNttThreads, 3) decrypt_intt_kernel(const __grid_constant__ RingConst rc,
                                                                      const u64 *ct_in, const tw_t *sk, u64 *coeff) {
    extern __shared__ __align__(16) u64 sm[];
    const int tid = threadIdx.x;
    const u32 L = rc.limbs;
    const u32 l = blockIdx.x % L;
    const u64 ct = blockIdx.x / L;
    const LimbConst lc = rc.limb[l];
    const u64 *c0 = ct_in + ((ct * 2 + 0) * L + l) * (u64)kN;
    const u64 *c1 = ct_in + ((ct * 2 + 1) * L + l) * (u64)kN;
    inv_phase_e(tid, sm, lc, LoadDecryptPair{c0, c1, sk + (size_t)l * kN, lc.q});
    __syncthreads();
    inv_phase_mid<1>(tid, sm, lc);
    __syncthreads();
    inv_phase_mid<4>(tid, sm, lc);
    __syncthreads();
    inv_phase_mid<7>(tid, sm, lc);
    __syncthreads();
    inv_phase_last(tid, sm, lc, StoreGlobal{coeff + (ct * L + l) * (u64)kN});
}

// ------------------------------------------------------------------------------------------
// decrypt stage 2: CRT -> double -> FFTSpecial -> real parts
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kFftThreads) decode_kernel(const __grid_constant__ EncodeConst ec,
                                                             const __grid_constant__ DecodeConst dc, u32 limbs,
                                                             const u64 *coeff, double *out, u64 n, i64 *tap_hi,
                                                             u64 *tap_lo) {
    extern __shared__ __align__(16) unsigned char smraw[];
    cplx *z = reinterpret_cast<cplx *>(smraw);
    const int tid = threadIdx.x;
    const u64 ct = blockIdx.x;
    decode_load(tid, kFftThreads, z, ec, dc, limbs, coeff + ct * limbs * (u64)kN,
                tap_hi ? tap_hi + ct * (u64)kN : nullptr, tap_lo ? tap_lo + ct * (u64)kN : nullptr);
    if (!out) return;
    __syncthreads();
    if (ec.slots == kFftFast) {
        fft_fwd_pass8<0>(tid, z, ec.tw_fwd); __syncthreads();
        fft_fwd_pass8<3>(tid, z, ec.tw_fwd); __syncthreads();
        fft_fwd_pass8<6>(tid, z, ec.tw_fwd); __syncthreads();
        fft_fwd_pass8<9>(tid, z, ec.tw_fwd); __syncthreads();
    } else {
        for (u32 len = 2; len <= ec.slots; len <<= 1) {
            fft_fwd_stage(tid, kFftThreads, z, ec.slots, len, ec.tw_fwd);
            __syncthreads();
        }
    }
    u64 first = ct * ec.batch;
    if (first >= n) return;
    u64 length = n - first < ec.batch ? n - first : ec.batch;
    decode_store(tid, kFftThreads, z, dc, out + first, length);
}

// ------------------------------------------------------------------------------------------
// weighted sum over learners: pure streaming, 128-bit lazy accumulation, one reduction per coefficient
// ------------------------------------------------------------------------------------------
constexpr int kWavgThreads = 256;
constexpr int kWavgPairsPerThread = 4;                                   // 4 x 16 B per learner in flight
constexpr int kWavgElemsPerCta = kWavgThreads * 2 * kWavgPairsPerThread;  // 2048: a quarter of a limb run

__device__ __forceinline__ ulonglong2 ld_stream(const ulonglong2 *p) {
    ulonglong2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p));
    return v;
}

__global__ void __launch_bounds__(kWavgThreads) weighted_sum_kernel(const __grid_constant__ RingConst rc,
                                                                    const __grid_constant__ WavgArgs wa, u64 *out) {
    const u64 base = (u64)blockIdx.x * kWavgElemsPerCta;
    const u32 l = (u32)((base / kN) % rc.limbs);
    const LimbConst lc = rc.limb[l];
    acc128 acc[kWavgPairsPerThread][2];
#pragma unroll
    for (int p = 0; p < kWavgPairsPerThread; p++) {
        acc[p][0] = acc128{0, 0};
        acc[p][1] = acc128{0, 0};
    }
    const u64 pair0 = (base >> 1) + threadIdx.x;
    if (wa.accumulate) {
#pragma unroll
        for (int p = 0; p < kWavgPairsPerThread; p++) {
            ulonglong2 v = reinterpret_cast<const ulonglong2 *>(out)[pair0 + (u64)p * kWavgThreads];
            acc[p][0].lo = v.x; acc[p][1].lo = v.y;
        }
    }
    for (u32 k = 0; k < wa.k; k++) {
        const ulonglong2 *src = reinterpret_cast<const ulonglong2 *>(wa.in[k]) + pair0;
        const u64 c = wa.c[k][l];
        ulonglong2 v[kWavgPairsPerThread];
#pragma unroll
        for (int p = 0; p < kWavgPairsPerThread; p++) v[p] = ld_stream(src + (u64)p * kWavgThreads);
#pragma unroll
        for (int p = 0; p < kWavgPairsPerThread; p++) {
            mac128(acc[p][0], v[p].x, c);
            mac128(acc[p][1], v[p].y, c);
        }
    }
#pragma unroll
    for (int p = 0; p < kWavgPairsPerThread; p++) {
        ulonglong2 r = make_ulonglong2(reduce128(acc[p][0], lc), reduce128(acc[p][1], lc));
        reinterpret_cast<ulonglong2 *>(out)[pair0 + (u64)p * kWavgThreads] = r;
    }
}

// ------------------------------------------------------------------------------------------
// serialized image <-> dense payload (host-blob API only; the device API never leaves dense form)
// ------------------------------------------------------------------------------------------
constexpr int kPackThreads = 256;

__global__ void __launch_bounds__(kPackThreads) unpack_kernel(const __grid_constant__ BlobRuns br, const u64 *image,
                                                              u64 *dense) {
    const u64 run = blockIdx.x;                       // ct * runs + r
    const u64 ct = run / br.runs;
    const u32 r = (u32)(run % br.runs);
    const u64 a0 = run_offset(br, ct, r);
    u64 *dst = dense + run * (u64)kN;
    for (u32 i = blockIdx.y * kPackThreads + threadIdx.x; i < (u32)kN; i += gridDim.y * kPackThreads)
        dst[i] = load_unaligned(image, a0 + 8ull * i);
}

__global__ void __launch_bounds__(kPackThreads) pack_kernel(const __grid_constant__ BlobRuns br, const u64 *dense,
                                                            u64 *image) {
    const u64 run = blockIdx.x;
    const u64 ct = run / br.runs;
    const u32 r = (u32)(run % br.runs);
    const u64 a0 = run_offset(br, ct, r);
    const u64 w0 = a0 >> 3, w1 = (a0 + 8ull * kN - 1) >> 3;
    const u64 *src = dense + run * (u64)kN;
    for (u64 w = w0 + blockIdx.y * kPackThreads + threadIdx.x; w <= w1; w += (u64)gridDim.y * kPackThreads)
        pack_word(image, a0, w, src);
}

// acc = (first ? 0 : acc) + c * unpack(image) mod q: one learner folded into the resident accumulator
__global__ void __launch_bounds__(kPackThreads) unpack_scale_accumulate_kernel(const __grid_constant__ RingConst rc,
                                                                               const __grid_constant__ BlobRuns br,
                                                                               const u64 *image,
                                                                               const __grid_constant__ LimbScalars cs,
                                                                               u32 first, u64 *acc) {
    const u64 run = blockIdx.x;
    const u64 ct = run / br.runs;
    const u32 r = (u32)(run % br.runs);
    const u32 l = r % rc.limbs;
    const LimbConst lc = rc.limb[l];
    const u64 a0 = run_offset(br, ct, r);
    const u64 c = cs.c[l];
    u64 *dst = acc + run * (u64)kN;
    for (u32 i = blockIdx.y * kPackThreads + threadIdx.x; i < (u32)kN; i += gridDim.y * kPackThreads) {
        acc128 a{first ? 0ull : dst[i], 0ull};
        mac128(a, load_unaligned(image, a0 + 8ull * i), c);
        dst[i] = reduce128(a, lc);
    }
}

// ------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------
cudaError_t kernels_init() {
    cudaError_t e;
#define OPT_IN(k, bytes)                                                                      \
    e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes));  \
    if (e != cudaSuccess) return e;
    OPT_IN(ntt_forward_kernel, NTT_SMEM_BYTES)
    OPT_IN(ntt_inverse_kernel, NTT_SMEM_BYTES)
    OPT_IN(encrypt_kernel, NTT_SMEM_BYTES)
    OPT_IN(encrypt_fused_kernel, NTT_SMEM_BYTES)
    OPT_IN(decrypt_intt_kernel, NTT_SMEM_BYTES)
    OPT_IN(encode_kernel<true>, (kN / 2) * sizeof(cplx))
    OPT_IN(encode_kernel<false>, (kN / 2) * sizeof(cplx))
    OPT_IN(decode_kernel, (kN / 2) * sizeof(cplx))
#undef OPT_IN
    return cudaSuccess;
}

void launch_ntt_forward(const RingConst &rc, u64 *data, u32 rows, cudaStream_t s) {
    if (rows) ntt_forward_kernel<<<rows, kNttThreads, NTT_SMEM_BYTES, s>>>(rc, data);
}
void launch_ntt_inverse(const RingConst &rc, u64 *data, u32 rows, cudaStream_t s) {
    if (rows) ntt_inverse_kernel<<<rows, kNttThreads, NTT_SMEM_BYTES, s>>>(rc, data);
}
void launch_encode_sample(const RingConst &, const EncodeConst &ec, const SamplerConst &sc, const double *x, u64 n,
                          u32 n_ct, i64 *m_e0, uint16_t *ve, cudaStream_t s) {
    if (n_ct) encode_kernel<true><<<n_ct, kFftThreads, ec.slots * sizeof(cplx), s>>>(ec, sc, x, n, m_e0, ve);
}
void launch_encode_only(const RingConst &, const EncodeConst &ec, const double *x, u64 n, u32 n_ct, i64 *m,
                        cudaStream_t s) {
    SamplerConst sc{};
    if (n_ct) encode_kernel<false><<<n_ct, kFftThreads, ec.slots * sizeof(cplx), s>>>(ec, sc, x, n, m, nullptr);
}
void launch_encrypt(const RingConst &rc, const i64 *m_e0, const uint16_t *ve, const tw_t *pk_b, const tw_t *pk_a,
                    u64 *ct, u32 n_ct, cudaStream_t s) {
    if (n_ct) encrypt_kernel<<<n_ct * rc.limbs, kNttThreads, NTT_SMEM_BYTES, s>>>(rc, m_e0, ve, pk_b, pk_a, ct);
}
void launch_encrypt_fused(const RingConst &rc, const EncodeConst &ec, const SamplerConst &sc, const EncryptBatch &eb,
                          u64 n, i64 *m_e0, uint16_t *ve, const tw_t *pk_b, const tw_t *pk_a, cudaStream_t s) {
    if (eb.n_ct && eb.k)
        encrypt_fused_kernel<<<eb.n_ct * eb.k, kNttThreads, NTT_SMEM_BYTES, s>>>(rc, ec, sc, eb, n, m_e0, ve, pk_b, pk_a);
}
void launch_decrypt_intt(const RingConst &rc, const u64 *ct, const tw_t *sk, u64 *coeff, u32 n_ct, cudaStream_t s) {
    if (n_ct) decrypt_intt_kernel<<<n_ct * rc.limbs, kNttThreads, NTT_SMEM_BYTES, s>>>(rc, ct, sk, coeff);
}
void launch_decode(const RingConst &rc, const EncodeConst &ec, const DecodeConst &dc, const u64 *coeff, u32 n_ct,
                   double *out, u64 n, cudaStream_t s) {
    if (n_ct)
        decode_kernel<<<n_ct, kFftThreads, ec.slots * sizeof(cplx), s>>>(ec, dc, rc.limbs, coeff, out, n, nullptr,
                                                                        nullptr);
}
void launch_crt_only(const RingConst &rc, const DecodeConst &dc, const u64 *coeff, u32 n_ct, i64 *hi, u64 *lo,
                     cudaStream_t s) {
    EncodeConst ec{};
    ec.slots = kN / 2; ec.gap = 1; ec.log_slots = kLogN - 1; ec.batch = kN / 2;
    if (n_ct)
        decode_kernel<<<n_ct, kFftThreads, ec.slots * sizeof(cplx), s>>>(ec, dc, rc.limbs, coeff, nullptr, 0, hi, lo);
}
void launch_weighted_sum(const RingConst &rc, const WavgArgs &wa, u64 *out, u64 n_ct, cudaStream_t s) {
    u64 elems = n_ct * 2 * rc.limbs * (u64)kN;
    if (elems) weighted_sum_kernel<<<(unsigned)(elems / kWavgElemsPerCta), kWavgThreads, 0, s>>>(rc, wa, out);
}
static dim3 run_grid(const BlobRuns &br, u64 n_ct) { return dim3((unsigned)(n_ct * br.runs), 4); }
void launch_unpack(const RingConst &, const BlobRuns &br, const u64 *image, u64 *dense, u64 n_ct, cudaStream_t s) {
    if (n_ct) unpack_kernel<<<run_grid(br, n_ct), kPackThreads, 0, s>>>(br, image, dense);
}
void launch_pack(const RingConst &, const BlobRuns &br, const u64 *dense, u64 *image, u64 n_ct, cudaStream_t s) {
    if (n_ct) pack_kernel<<<run_grid(br, n_ct), kPackThreads, 0, s>>>(br, dense, image);
}
void launch_unpack_scale_accumulate(const RingConst &rc, const BlobRuns &br, const u64 *image, const LimbScalars &cs,
                                    u32 first, u64 *acc, u64 n_ct, cudaStream_t s) {
    if (n_ct)
        unpack_scale_accumulate_kernel<<<run_grid(br, n_ct), kPackThreads, 0, s>>>(rc, br, image, cs, first,
                                                                                  acc);
}

}  // namespace shelfi
