// emu.cpp — CPU replay harness for the device code.  TEST INFRASTRUCTURE ONLY (tests/ is the only user).
//
// Compiles fhe-fed_b200/csrc/ckks_core.cuh with g++ (HD functions become plain inline functions) and
// replays each kernel of kernels.cu phase by phase: every phase runs for all thread ids before the
// next one starts, which is exactly what __syncthreads() guarantees on the GPU.  It lets the CPU-only
// test suite check index math, lazy-reduction bounds and wire-format code bit for bit against the
// oracle.  It is NOT a fallback: nothing in the product links or loads it.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../fhe-fed_b200/csrc/ckks_core.cuh"
#include "../../fhe-fed_b200/csrc/numtheory.h"
#include "../../fhe-fed_b200/csrc/tables.h"
#include "../../fhe-fed_b200/csrc/wire.h"

using namespace shelfi;

struct Emu {
    RingConst rc;
    EncodeConst ec;
    Crt2Const crt;
    CrtNConst crtn;
    uint64_t cdf[32];
    uint32_t cdf_hi[32];
    GaussTable gauss() const { return GaussTable{cdf_hi, cdf}; }
    std::vector<tw_t> tw[kMaxLimbs], itw[kMaxLimbs];
    std::vector<double> dtw[kMaxLimbs];
    bool fp64 = true;
    std::vector<cplx> fft_inv, fft_fwd;
    std::string err;
};

#define FOR_THREADS(n) for (int tid = 0; tid < (n); tid++)

template <class Load, class Store>
static void replay_forward(u64 *sm, const LimbConst &lc, u32 hb, Load ld, Store st) {
    FOR_THREADS(kNttThreads) fwd_phase_a(tid, sm, lc, hb, ld);
    FOR_THREADS(kNttThreads) fwd_phase_mid<7>(tid, sm, lc, hb);
    FOR_THREADS(kNttThreads) fwd_phase_mid<4>(tid, sm, lc, hb);
    FOR_THREADS(kNttThreads) fwd_phase_mid<1>(tid, sm, lc, hb);
    FOR_THREADS(kNttThreads) fwd_phase_e(tid, sm, lc, hb, st);
}
// the FP64-pipe variant (ntt_forward_tile_d)
template <class Load, class Store>
static void replay_forward_d(u64 *sm, const LimbConst &lc, u32 hb, Load ld, Store st) {
    FOR_THREADS(kNttThreads) dfwd_phase_a(tid, sm, lc, hb, ld);
    FOR_THREADS(kNttThreads) dfwd_phase_mid<7>(tid, sm, lc, hb);
    FOR_THREADS(kNttThreads) dfwd_phase_mid<4>(tid, sm, lc, hb);
    FOR_THREADS(kNttThreads) dfwd_phase_mid<1>(tid, sm, lc, hb);
    FOR_THREADS(kNttThreads) dfwd_phase_e(tid, sm, lc, hb, st);
}
// intt_tile<D>
template <int D, class Load>
static void replay_inverse(u64 *sm, const LimbConst &lc, u32 hb, Load ld, u64 *dst) {
    if (lc.ditw) {   // the FP64-pipe inverse transform
        FOR_THREADS(kNttThreads) dinv_phase_e(tid, sm, lc, hb, ld);
        FOR_THREADS(kNttThreads) dinv_phase_mid<1>(tid, sm, lc, hb);
        FOR_THREADS(kNttThreads) dinv_phase_mid<4>(tid, sm, lc, hb);
        FOR_THREADS(kNttThreads) dinv_phase_mid<7>(tid, sm, lc, hb);
        if (D == 0) { FOR_THREADS(kNttThreads) dinv_phase_last(tid, sm, lc, StoreGlobal{dst}); }
        else { FOR_THREADS(kNttThreads) dinv_phase_last_lazy(tid, sm, lc, hb, StoreGlobal{dst}); }
        return;
    }
    FOR_THREADS(kNttThreads) inv_phase_e(tid, sm, lc, hb, ld);
    FOR_THREADS(kNttThreads) inv_phase_mid<1>(tid, sm, lc, hb);
    FOR_THREADS(kNttThreads) inv_phase_mid<4>(tid, sm, lc, hb);
    FOR_THREADS(kNttThreads) inv_phase_mid<7>(tid, sm, lc, hb);
    if (D == 0) { FOR_THREADS(kNttThreads) inv_phase_last(tid, sm, lc, StoreGlobal{dst}); }
    else { FOR_THREADS(kNttThreads) inv_phase_last_lazy(tid, sm, lc, hb, StoreGlobal{dst}); }
}
// intt_tail_kernel<D> over one row
template <int D>
static void replay_tail(const LimbConst &lc, u64 *row) {
    for (u32 i = 0; i < (u32)kN; i++) {
        u64 a[1 << D];
        for (int k = 0; k < (1 << D); k++) a[k] = row[i + (size_t)k * kN];
        inv_tail<D>(a, lc);
        for (int k = 0; k < (1 << D); k++) row[i + (size_t)k * kN] = a[k];
    }
}
#define EMU_D(d, ...)                                         \
    switch (d) {                                              \
        case 0: { constexpr int D = 0; __VA_ARGS__; break; }  \
        case 1: { constexpr int D = 1; __VA_ARGS__; break; }  \
        default: { constexpr int D = 2; __VA_ARGS__; break; } \
    }

extern "C" {

Emu *emu_create_ex(const uint64_t *q, const uint64_t *psi, uint32_t limbs, uint32_t logn, uint32_t slots, uint32_t batch,
                   double sigma) {
    Emu *e = new Emu();
    std::memset(&e->rc, 0, sizeof e->rc);
    const uint32_t n = 1u << logn;
    e->rc.n = n; e->rc.logn = logn; e->rc.log_blocks = logn - kLogN; e->rc.limbs = limbs; e->rc.slots = slots;
    for (uint32_t l = 0; l < limbs; l++) {
        e->err = make_limb_const(q[l], psi[l], logn, &e->rc.limb[l], &e->tw[l], &e->itw[l], &e->dtw[l]);
        if (!e->err.empty()) return e;
        e->rc.limb[l].tw = e->tw[l].data();
        e->rc.limb[l].itw = e->itw[l].data();
        e->rc.limb[l].dtw = e->dtw[l].empty() ? nullptr : e->dtw[l].data();
        e->rc.limb[l].ditw = e->dtw[l].empty() ? nullptr : e->dtw[l].data() + n;
    }
    make_fft_tables(2 * n, slots, &e->fft_inv, &e->fft_fwd);
    std::memset(&e->ec, 0, sizeof e->ec);
    e->ec.slots = slots; e->ec.gap = (n / 2) / slots; e->ec.batch = batch;
    while ((1u << e->ec.log_slots) < slots) e->ec.log_slots++;
    e->ec.pow_p = (double)q[limbs - 1];
    e->ec.inv_size = 1.0 / (double)slots;
    e->ec.tw_inv = e->fft_inv.data();
    e->ec.tw_fwd = e->fft_fwd.data();
    e->crt = make_crt2(q[0], limbs >= 2 ? q[1] : 0);
    e->crtn = make_crtn(q, limbs);
    build_gauss_cdf(sigma, e->cdf);
    for (int k = 0; k < 32; k++) e->cdf_hi[k] = (uint32_t)(e->cdf[k] >> 32);
    return e;
}
Emu *emu_create(const uint64_t *q, const uint64_t *psi, uint32_t limbs, uint32_t slots, uint32_t batch, double sigma) {
    return emu_create_ex(q, psi, limbs, kLogN, slots, batch, sigma);
}
void emu_destroy(Emu *e) { delete e; }
const char *emu_error(const Emu *e) { return e->err.c_str(); }
uint32_t emu_fwd_mask(const Emu *e, uint32_t l) { return e->rc.limb[l].fwd_mask; }
uint32_t emu_rmode(const Emu *e, uint32_t l) { return e->rc.limb[l].rmode; }

// ntt_forward_kernel<D> / ntt_inverse_kernel<D> (+ intt_tail_kernel<D>), one row of rc.n values
void emu_ntt_forward(const Emu *e, uint32_t limb, uint64_t *row) {
    std::vector<u64> sm(kN), src(row, row + e->rc.n);
    const LimbConst lc = e->rc.limb[limb];
    const u32 d = e->rc.log_blocks;
    for (u32 blk = 0; blk < (1u << d); blk++) {
        u64 *dst = row + (size_t)blk * kN;
        if (lc.dtw) {
            EMU_D(d, replay_forward_d(sm.data(), lc, head_bits(d, blk),
                                      DHeadLoad<D, DLoadGlobal>{DLoadGlobal{src.data(), lc.q}, &lc, blk},
                                      StoreFullReducePair{dst, &lc}))
            continue;
        }
        EMU_D(d, replay_forward(sm.data(), lc, head_bits(d, blk), HeadLoad<D, LoadGlobal>{LoadGlobal{src.data()}, &lc, blk},
                                StoreFullReducePair{dst, &lc}))
    }
}
// 1 if limb l takes the FP64-pipe forward transform
int emu_limb_fp64(const Emu *e, uint32_t l) { return e->rc.limb[l].dtw != nullptr; }
// switch the FP64 path off / on for this context (the integer path must give the same bits)
void emu_set_fp64(Emu *e, int on) {
    for (u32 l = 0; l < e->rc.limbs; l++) {
        e->rc.limb[l].dtw = (on && !e->dtw[l].empty()) ? e->dtw[l].data() : nullptr;
        e->rc.limb[l].ditw = e->rc.limb[l].dtw ? e->rc.limb[l].dtw + e->rc.n : nullptr;
    }
}
void emu_ntt_inverse(const Emu *e, uint32_t limb, uint64_t *row) {
    std::vector<u64> sm(kN);
    const LimbConst lc = e->rc.limb[limb];
    const u32 d = e->rc.log_blocks;
    for (u32 blk = 0; blk < (1u << d); blk++) {
        u64 *t = row + (size_t)blk * kN;
        EMU_D(d, replay_inverse<D>(sm.data(), lc, head_bits(d, blk), LoadPairGlobal{t}, t))
    }
    if (d == 1) replay_tail<1>(lc, row);
    if (d == 2) replay_tail<2>(lc, row);
}

// rescale_intt_kernel<D> (+ intt_tail_kernel<D>) then rescale_ntt_kernel<D>: one ciphertext [2][L][N] -> [2][L-1][N]
void emu_rescale(const Emu *e, const uint64_t *ct, uint64_t *out) {
    const u32 L = e->rc.limbs, d = e->rc.log_blocks;
    const size_t n = e->rc.n;
    RescaleConst rs{};
    rs.q_last = e->rc.limb[L - 1].q;
    for (u32 i = 0; i + 1 < L; i++) {
        const u64 q = e->rc.limb[i].q;
        rs.inv[i] = nt::inv_mod(rs.q_last % q, q);
        rs.inv_s[i] = nt::shoup(rs.inv[i], q);
    }
    std::vector<u64> sm(kN), t(n);
    for (u32 p = 0; p < 2; p++) {
        const LimbConst last = e->rc.limb[L - 1];
        for (u32 blk = 0; blk < (1u << d); blk++) {
            const size_t off = (size_t)blk * kN;
            EMU_D(d, replay_inverse<D>(sm.data(), last, head_bits(d, blk), LoadPairGlobal{ct + (p * L + (L - 1)) * n + off},
                                       t.data() + off))
        }
        if (d == 1) replay_tail<1>(last, t.data());
        if (d == 2) replay_tail<2>(last, t.data());
        for (u32 i = 0; i + 1 < L; i++) {
            const LimbConst lc = e->rc.limb[i];
            for (u32 blk = 0; blk < (1u << d); blk++) {
                const size_t off = (size_t)blk * kN;
                const StoreRescale st{ct + (p * L + i) * n + off, out + (p * (L - 1) + i) * n + off, &lc, rs.inv[i], rs.inv_s[i]};
                if (lc.dtw) {
                    EMU_D(d, replay_forward_d(sm.data(), lc, head_bits(d, blk),
                                              DHeadLoad<D, DLoadSwitch>{DLoadSwitch{t.data(), rs.q_last, &lc}, &lc, blk}, st))
                } else {
                    EMU_D(d, replay_forward(sm.data(), lc, head_bits(d, blk),
                                            HeadLoad<D, LoadSwitch>{LoadSwitch{t.data(), rs.q_last, &lc}, &lc, blk}, st))
                }
            }
        }
    }
}

// selective encryption: when set, chunk element i is x[sel[i]] on encode and out[sel[i]] on decode
static const u32 *g_sel = nullptr;
void emu_set_selection(const uint32_t *sel) { g_sel = sel; }

// encode_kernel<SAMPLE> (ring dimension 8192) or encode_block_kernel<DF> (+ sample_add_kernel): one ciphertext
static void replay_encode(const Emu *e, const double *x, uint64_t n_valid, std::vector<cplx> &z) {
    z.assign(e->ec.slots, cplx{0, 0});
    FOR_THREADS(kFftThreads) encode_load(tid, kFftThreads, z.data(), e->ec, x, n_valid, g_sel);
    if (e->ec.slots == kFftFast) {
        FOR_THREADS(kFftThreads) fft_inv_pass8<9>(tid, z.data(), e->ec.tw_inv);
        FOR_THREADS(kFftThreads) fft_inv_pass8<6>(tid, z.data(), e->ec.tw_inv);
        FOR_THREADS(kFftThreads) fft_inv_pass8<3>(tid, z.data(), e->ec.tw_inv);
        FOR_THREADS(kFftThreads) fft_inv_pass8<0>(tid, z.data(), e->ec.tw_inv);
    } else {
        for (u32 len = e->ec.slots; len >= 2; len >>= 1)
            FOR_THREADS(kFftThreads) fft_inv_stage(tid, kFftThreads, z.data(), e->ec.slots, len, e->ec.tw_inv);
    }
    FOR_THREADS(kFftThreads) encode_finalize(tid, kFftThreads, z.data(), e->ec);
}
extern "C++" {
template <int DF>
static void replay_encode_block(const Emu *e, const double *x, uint64_t n_valid, u32 fb, int64_t *m) {
    const EncodeConst &ec = e->ec;
    std::vector<cplx> z(kFftFast, cplx{0, 0});
    FOR_THREADS(kFftThreads) encode_load_block<DF>(tid, kFftThreads, z.data(), ec, x, n_valid, fb, g_sel);
    EncodeConst et = ec;
    if (ec.slots >= kFftFast) {
        et.slots = kFftFast; et.log_slots = 12;
        FOR_THREADS(kFftThreads) fft_inv_pass8<9>(tid, z.data(), ec.tw_inv);
        FOR_THREADS(kFftThreads) fft_inv_pass8<6>(tid, z.data(), ec.tw_inv);
        FOR_THREADS(kFftThreads) fft_inv_pass8<3>(tid, z.data(), ec.tw_inv);
        FOR_THREADS(kFftThreads) fft_inv_pass8<0>(tid, z.data(), ec.tw_inv);
    } else {
        for (u32 len = ec.slots; len >= 2; len >>= 1)
            FOR_THREADS(kFftThreads) fft_inv_stage(tid, kFftThreads, z.data(), ec.slots, len, ec.tw_inv);
    }
    FOR_THREADS(kFftThreads) encode_finalize(tid, kFftThreads, z.data(), et);
    FOR_THREADS(kFftThreads) encode_emit_block<DF>(tid, kFftThreads, z.data(), ec, fb, e->rc.n >> 1, (i64 *)m);
}
}  // extern "C++"
void emu_encode(const Emu *e, const double *x, uint64_t n_valid, int64_t *m) {
    if (e->rc.log_blocks == 0) {
        std::vector<cplx> z;
        replay_encode(e, x, n_valid, z);
        FOR_THREADS(kFftThreads)
            for (u32 j = tid; j < (u32)kN; j += kFftThreads) m[j] = encode_coeff(z.data(), e->ec, j);
        return;
    }
    const u32 df = e->ec.log_slots > 12 ? e->ec.log_slots - 12 : 0;
    for (u32 fb = 0; fb < (1u << df); fb++) { EMU_D(df, replay_encode_block<D>(e, x, n_valid, fb, m)) }
}
// the sampler phases of encrypt_fused_kernel (ring dimension 8192) / sample_add_kernel (larger rings) under the test
// key of `seed`: m_e0[N] = m + e0, ve[N] = pack_ve(v, e1)
void emu_encode_sample(const Emu *e, const double *x, uint64_t n_valid, uint64_t seed, uint64_t ct_index, int64_t *m_e0,
                       uint8_t *ve) {
    const SamplerKey key = seed_key(seed, kKeyDomainEncrypt);
    FOR_THREADS(kNttThreads) sample_ternary_phase(tid, kNttThreads, key, ct_index, e->rc.n, ve);
    if (e->rc.log_blocks == 0) {
        std::vector<cplx> z;
        replay_encode(e, x, n_valid, z);
        FOR_THREADS(kNttThreads)
            sample_gauss_phase(tid, kNttThreads, z.data(), e->ec, key, e->gauss(), ct_index, e->rc.n, (i64 *)m_e0, ve);
        return;
    }
    emu_encode(e, x, n_valid, m_e0);
    FOR_THREADS(kNttThreads)
        sample_gauss_phase(tid, kNttThreads, nullptr, e->ec, key, e->gauss(), ct_index, e->rc.n, (i64 *)m_e0, ve);
}
// one ChaCha20 block through the device code's implementation (RFC 8439 known-answer test)
void emu_chacha20_block(const uint32_t *key, uint32_t counter, const uint32_t *nonce, uint32_t *out) {
    SamplerKey k;
    std::memcpy(k.k, key, sizeof k.k);
    u32 o[16];
    chacha20_block(k, counter, nonce[0], nonce[1], nonce[2], o);
    std::memcpy(out, o, sizeof o);
}
// the primitive samplers on caller-chosen random words: gauss_lazy with l supplied (used only on a tie of the top 31
// bits; *used_retry reports it), ternary16 with the retry value supplied
int emu_gauss_lazy(const Emu *e, uint32_t h, uint32_t l, int *used_retry) {
    int used = 0;
    int r = gauss_lazy(e->gauss(), h, [&] { used = 1; return l; });
    if (used_retry) *used_retry = used;
    return r;
}
int emu_ternary16(uint32_t h16, int retry_value, int *used_retry) {
    int used = 0;
    int r = ternary16(h16, [&] { used = 1; return retry_value; });
    if (used_retry) *used_retry = used;
    return r;
}
int emu_retry_ternary(const uint32_t *key, uint64_t ct, uint32_t stream, uint32_t j) {
    SamplerKey k;
    std::memcpy(k.k, key, sizeof k.k);
    return retry_ternary(k, ct, stream, j);
}
void emu_gauss_table(const Emu *e, uint64_t *cdf63) { std::memcpy(cdf63, e->cdf, sizeof e->cdf); }
// the device keygen draws under the test keys of `seed`: s[N], e[N] (signed), a[L][N]
void emu_keygen_draws(const Emu *e, uint64_t seed, int64_t *s, int64_t *err, uint64_t *a) {
    const SamplerKey ks = seed_key(seed, kKeyDomainKeySecret), kp = seed_key(seed, kKeyDomainKeyPublic);
    for (u32 b = 0; b < e->rc.n / 32; b++) {
        u32 w[8];
        sample_ternary_block(ks, 0, b, kStreamKeyS, kStreamKeySRetry, w);
        for (int i = 0; i < 32; i++) s[32 * b + i] = (int)((w[i >> 2] >> (8 * (i & 3))) & 3u) - 1;
    }
    for (u32 b = 0; b < e->rc.n / 16; b++) {
        int v[16];
        keygen_error_block(ks, e->gauss(), b, v);
        for (int i = 0; i < 16; i++) err[16 * b + i] = v[i];
    }
    for (u32 l = 0; l < e->rc.limbs; l++)
        for (u32 j = 0; j < e->rc.n; j++) a[(size_t)l * e->rc.n + j] = keygen_uniform(kp, l, j, e->rc.limb[l].q);
}

// encrypt_kernel<D>: one ciphertext, all limbs; pk_b / pk_a: [L][N] values; out: [2][L][N]
void emu_encrypt(const Emu *e, const int64_t *m_e0, const uint8_t *ve, const uint64_t *pk_b, const uint64_t *pk_a,
                 uint64_t *out) {
    const u32 L = e->rc.limbs, d = e->rc.log_blocks;
    const size_t n = e->rc.n;
    std::vector<u64> sm(kN);
    for (u32 l = 0; l < L; l++) {
        const LimbConst lc = e->rc.limb[l];
        std::vector<tw_t> b(n), a(n);
        for (size_t j = 0; j < n; j++) {
            b[j] = tw_t{pk_b[l * n + j], nt::shoup(pk_b[l * n + j], lc.q)};
            a[j] = tw_t{pk_a[l * n + j], nt::shoup(pk_a[l * n + j], lc.q)};
        }
        for (u32 blk = 0; blk < (1u << d); blk++) {
            const size_t off = (size_t)blk * kN;
            u64 *c0 = out + (0 * L + l) * n + off, *c1 = out + (1 * L + l) * n + off;
            const u32 hb = head_bits(d, blk);
            if (lc.dtw) {
                EMU_D(d, {
                    replay_forward_d(sm.data(), lc, hb, DHeadLoad<D, DLoadTernary>{DLoadTernary{ve}, &lc, blk}, StoreLazyPair{c1, 0});
                    replay_forward_d(sm.data(), lc, hb, DHeadLoad<D, DLoadSigned>{DLoadSigned{(const i64 *)m_e0, &lc}, &lc, blk},
                                     StoreMulAdd{b.data() + off, c1, c0, &lc, false});
                    replay_forward_d(sm.data(), lc, hb, DHeadLoad<D, DLoadGaussHi>{DLoadGaussHi{ve}, &lc, blk},
                                     StoreMulAdd{a.data() + off, c1, c1, &lc, false});
                })
                continue;
            }
            EMU_D(d, {
                replay_forward(sm.data(), lc, hb, HeadLoad<D, LoadTernary>{LoadTernary{ve, lc.q}, &lc, blk}, StoreLazyPair{c1, 0});
                replay_forward(sm.data(), lc, hb, HeadLoad<D, LoadSigned>{LoadSigned{(const i64 *)m_e0, &lc}, &lc, blk},
                               StoreMulAdd{b.data() + off, c1, c0, &lc, false});
                replay_forward(sm.data(), lc, hb, HeadLoad<D, LoadGaussHi>{LoadGaussHi{ve, lc.q}, &lc, blk},
                               StoreMulAdd{a.data() + off, c1, c1, &lc, false});
            })
        }
    }
}

extern "C++" {
template <int L>
static void replay_crt(const Emu *e, const DecodeConst &dc, const u64 *coeff, double *vals, u64 *tap8) {
    for (u32 j = 0; j < e->rc.n; j++) {
        u64 t[8];
        double v = decode_value<L>(coeff, e->rc.n, j, dc.crt, e->crtn, tap8 ? t : nullptr);
        vals[j] = shelfi::fmul(v, dc.pre);
        if (tap8) std::memcpy(tap8 + (size_t)j * 8, t, sizeof t);
    }
}
template <int DF>
static void replay_decode_fft(const Emu *e, const DecodeConst &dc, const double *vals, cplx *zs, double *out, u64 length) {
    const EncodeConst &ec = e->ec;
    for (u32 fb = 0; fb < (1u << DF); fb++) {
        std::vector<cplx> z(kFftFast);
        FOR_THREADS(kFftThreads) decode_load_block<DF>(tid, kFftThreads, z.data(), ec, fb, e->rc.n, vals);
        if (ec.slots >= kFftFast) {
            FOR_THREADS(kFftThreads) fft_fwd_pass8<0>(tid, z.data(), ec.tw_fwd);
            FOR_THREADS(kFftThreads) fft_fwd_pass8<3>(tid, z.data(), ec.tw_fwd);
            FOR_THREADS(kFftThreads) fft_fwd_pass8<6>(tid, z.data(), ec.tw_fwd);
            FOR_THREADS(kFftThreads) fft_fwd_pass8<9>(tid, z.data(), ec.tw_fwd);
        } else {
            for (u32 len = 2; len <= ec.slots; len <<= 1)
                FOR_THREADS(kFftThreads) fft_fwd_stage(tid, kFftThreads, z.data(), ec.slots, len, ec.tw_fwd);
        }
        if (DF == 0) { FOR_THREADS(kFftThreads) decode_store(tid, kFftThreads, z.data(), dc, out, length, g_sel); }
        else for (u32 i = 0; i < kFftFast; i++) zs[(size_t)fb * kFftFast + i] = z[zsw(i)];
    }
    if (DF > 0) for (u32 i = 0; i < kFftFast; i++) decode_tail<DF>(i, zs, ec, dc, out, length, g_sel);
}

}  // extern "C++"
// decrypt_intt_kernel<D> (+ tail) then decode_kernel (8192, <= 2 limbs) or crt_kernel<L> + decode_fft_kernel<DF> (+ tail):
// one ciphertext; sk: [L][N]; out: slots doubles (first `length` written); tap8: N x 8 words or NULL
void emu_decrypt(const Emu *e, const uint64_t *ct, const uint64_t *sk, double pre, double pow_p, double *out,
                 uint64_t length, int64_t *tap_hi, uint64_t *tap_lo, uint64_t *tap8) {
    const u32 L = e->rc.limbs, d = e->rc.log_blocks;
    const size_t n = e->rc.n;
    std::vector<u64> sm(kN), coeff(L * n);
    for (u32 l = 0; l < L; l++) {
        const LimbConst lc = e->rc.limb[l];
        std::vector<tw_t> s(n);
        for (size_t j = 0; j < n; j++) s[j] = tw_t{sk[l * n + j], nt::shoup(sk[l * n + j], lc.q)};
        for (u32 blk = 0; blk < (1u << d); blk++) {
            const size_t off = (size_t)blk * kN;
            EMU_D(d, replay_inverse<D>(sm.data(), lc, head_bits(d, blk),
                                       LoadDecryptPair{ct + (0 * L + l) * n + off, ct + (1 * L + l) * n + off, s.data() + off, lc.q},
                                       coeff.data() + l * n + off))
        }
        if (d == 1) replay_tail<1>(lc, coeff.data() + l * n);
        if (d == 2) replay_tail<2>(lc, coeff.data() + l * n);
    }
    DecodeConst dc{};
    dc.crt = e->crt; dc.pre = pre; dc.pow_p = pow_p;
    if (d == 0 && L <= 2 && !tap8) {
        std::vector<cplx> z(e->ec.slots);
        FOR_THREADS(kFftThreads)
            decode_load(tid, kFftThreads, z.data(), e->ec, dc, L, coeff.data(), (i64 *)tap_hi, (u64 *)tap_lo);
        if (e->ec.slots == kFftFast) {
            FOR_THREADS(kFftThreads) fft_fwd_pass8<0>(tid, z.data(), e->ec.tw_fwd);
            FOR_THREADS(kFftThreads) fft_fwd_pass8<3>(tid, z.data(), e->ec.tw_fwd);
            FOR_THREADS(kFftThreads) fft_fwd_pass8<6>(tid, z.data(), e->ec.tw_fwd);
            FOR_THREADS(kFftThreads) fft_fwd_pass8<9>(tid, z.data(), e->ec.tw_fwd);
        } else {
            for (u32 len = 2; len <= e->ec.slots; len <<= 1)
                FOR_THREADS(kFftThreads) fft_fwd_stage(tid, kFftThreads, z.data(), e->ec.slots, len, e->ec.tw_fwd);
        }
        FOR_THREADS(kFftThreads) decode_store(tid, kFftThreads, z.data(), dc, out, length, g_sel);
        return;
    }
    std::vector<double> vals(n);
    std::vector<u64> t8(tap8 || tap_hi ? n * 8 : 0);
    u64 *tp = t8.empty() ? nullptr : t8.data();
    switch (L) {
        case 1: replay_crt<1>(e, dc, coeff.data(), vals.data(), tp); break;
        case 2: replay_crt<2>(e, dc, coeff.data(), vals.data(), tp); break;
        case 3: replay_crt<3>(e, dc, coeff.data(), vals.data(), tp); break;
        case 4: replay_crt<4>(e, dc, coeff.data(), vals.data(), tp); break;
        case 5: replay_crt<5>(e, dc, coeff.data(), vals.data(), tp); break;
        case 6: replay_crt<6>(e, dc, coeff.data(), vals.data(), tp); break;
        case 7: replay_crt<7>(e, dc, coeff.data(), vals.data(), tp); break;
        default: replay_crt<8>(e, dc, coeff.data(), vals.data(), tp); break;
    }
    if (tap8) std::memcpy(tap8, t8.data(), n * 8 * sizeof(u64));
    if (tap_hi) for (size_t j = 0; j < n; j++) { tap_hi[j] = (int64_t)t8[j * 8 + 1]; tap_lo[j] = t8[j * 8]; }
    const u32 df = e->ec.log_slots > 12 ? e->ec.log_slots - 12 : 0;
    std::vector<cplx> zs(e->ec.slots);
    EMU_D(df, replay_decode_fft<D>(e, dc, vals.data(), zs.data(), out, length))
}

// weighted_sum_kernel arithmetic on one limb row: out[j] = sum_k c[k] * in[k][j] mod q
void emu_weighted_sum(const Emu *e, uint32_t limb, const uint64_t *const *in, const uint64_t *c, uint32_t k,
                      uint64_t *out, uint32_t count) {
    const LimbConst lc = e->rc.limb[limb];
    for (u32 j = 0; j < count; j++) {
        acc128 a{0, 0};
        for (u32 i = 0; i < k; i++) mac128(a, in[i][j], c[i]);
        out[j] = reduce128(a, lc);
    }
}

// pack_kernel / unpack_kernel over a whole blob image
static BlobRuns runs_of(const uint64_t *run0, const uint64_t *run1, uint64_t stride, uint32_t runs, uint32_t n) {
    BlobRuns br;
    std::memset(&br, 0, sizeof br);
    br.runs = runs; br.stride = stride; br.n = n;
    for (u32 r = 0; r < runs; r++) { br.off0[r] = run0[r]; br.off1[r] = run1[r]; }
    return br;
}
void emu_unpack_n(const uint64_t *run0, const uint64_t *run1, uint64_t stride, uint32_t runs, uint32_t n, uint64_t n_ct,
                  const uint64_t *image, uint64_t *dense) {
    BlobRuns br = runs_of(run0, run1, stride, runs, n);
    for (u64 run = 0; run < n_ct * runs; run++) {
        u64 a0 = run_offset(br, run / runs, (u32)(run % runs));
        for (u32 i = 0; i < n; i++) dense[run * n + i] = load_unaligned(image, a0 + 8ull * i);
    }
}
void emu_pack_n(const uint64_t *run0, const uint64_t *run1, uint64_t stride, uint32_t runs, uint32_t n, uint64_t n_ct,
                const uint64_t *dense, uint64_t *image) {
    BlobRuns br = runs_of(run0, run1, stride, runs, n);
    for (u64 run = 0; run < n_ct * runs; run++) {
        u64 a0 = run_offset(br, run / runs, (u32)(run % runs));
        for (u64 w = a0 >> 3; w <= (a0 + 8ull * n - 1) >> 3; w++) pack_word(image, a0, w, dense + run * n, n);
    }
}
void emu_unpack(const uint64_t *run0, const uint64_t *run1, uint64_t stride, uint32_t runs, uint64_t n_ct,
                const uint64_t *image, uint64_t *dense) {
    emu_unpack_n(run0, run1, stride, runs, kN, n_ct, image, dense);
}
void emu_pack(const uint64_t *run0, const uint64_t *run1, uint64_t stride, uint32_t runs, uint64_t n_ct,
              const uint64_t *dense, uint64_t *image) {
    emu_pack_n(run0, run1, stride, runs, kN, n_ct, dense, image);
}

// ---- wire format (wire.cpp is plain host code; exercised here without the CUDA library) ----
// kind 0 context, 1 public key, 2 private key: parse then emit; returns malloc'd bytes or NULL (+ error text)
uint8_t *emu_wire_reemit(int kind, const uint8_t *in, uint64_t len, uint64_t *out_len, char *err, size_t err_cap) {
    std::string e;
    std::vector<uint8_t> out;
    if (kind == 0) {
        wire::ContextParams p;
        e = wire::parse_context(in, len, &p);
        if (e.empty()) out = wire::emit_context(p);
    } else {
        wire::KeyFile k;
        e = kind == 1 ? wire::parse_public_key(in, len, &k) : wire::parse_private_key(in, len, &k);
        if (e.empty()) out = kind == 1 ? wire::emit_public_key(k) : wire::emit_private_key(k);
    }
    if (!e.empty()) { std::snprintf(err, err_cap, "%s", e.c_str()); return nullptr; }
    uint8_t *p = (uint8_t *)std::malloc(out.size() ? out.size() : 1);
    std::memcpy(p, out.data(), out.size());
    *out_len = out.size();
    return p;
}
// map a blob; fills runs (2L each), stride, n_ct, first ciphertext's metadata
int emu_wire_map(const uint8_t *in, uint64_t len, int single, uint64_t *n_ct, uint32_t *limbs, uint64_t *run0,
                 uint64_t *run1, uint64_t *stride, uint64_t *depth, uint64_t *level, double *sf, char *tag, char *err,
                 size_t err_cap) {
    wire::BlobMap m;
    std::string e = wire::map_ciphertext_blob(in, len, single != 0, &m);
    if (!e.empty()) { std::snprintf(err, err_cap, "%s", e.c_str()); return -1; }
    *n_ct = m.n_ct; *limbs = m.ct_limbs; *stride = m.stride;
    std::memcpy(run0, m.run0, sizeof m.run0);
    std::memcpy(run1, m.run1, sizeof m.run1);
    *depth = m.meta[0].depth; *level = m.meta[0].level; *sf = m.meta[0].scaling_factor;
    std::memcpy(tag, m.meta[0].tag, 33);
    return 0;
}
// re-create a blob from its own map: plan headers, pack payload with the replayed pack kernel
uint8_t *emu_wire_rebuild(const uint8_t *in, uint64_t len, uint64_t *out_len, char *err, size_t err_cap) {
    wire::BlobMap m;
    std::string e = wire::map_ciphertext_blob(in, len, false, &m);
    if (!e.empty()) { std::snprintf(err, err_cap, "%s", e.c_str()); return nullptr; }
    const uint32_t runs = 2 * m.ct_limbs;   // fewer than the context's towers after a rescale
    const uint32_t n = m.ctx.ring_dim;
    std::vector<uint64_t> image((len + 7) / 8 + 1, 0), dense(m.n_ct * runs * (size_t)n);
    std::memcpy(image.data(), in, len);
    emu_unpack_n(m.run0, m.run1, m.stride, runs, n, m.n_ct, image.data(), dense.data());
    wire::BlobPlan plan = wire::plan_ciphertext_blob(m.ctx, m.n_ct, m.meta);
    std::vector<uint64_t> out_image((plan.size + 7) / 8 + 1, 0xA5A5A5A5A5A5A5A5ull);
    emu_pack_n(plan.run0, plan.run1, plan.stride, runs, n, m.n_ct, dense.data(), out_image.data());
    uint8_t *out = (uint8_t *)std::malloc(plan.size);
    std::memcpy(out, out_image.data(), plan.size);
    for (const auto &seg : plan.segments) std::memcpy(out + seg.offset, seg.bytes.data(), seg.bytes.size());
    *out_len = plan.size;
    return out;
}
uint64_t emu_wire_blob_size(const uint8_t *ctx_file, uint64_t len, uint64_t n_ct) {
    wire::ContextParams p;
    if (!wire::parse_context(ctx_file, len, &p).empty()) return 0;
    return wire::ciphertext_blob_size(p, n_ct);
}
uint64_t emu_wire_blob_size_level(const uint8_t *ctx_file, uint64_t len, uint64_t n_ct, uint32_t limbs) {
    wire::ContextParams p;
    if (!wire::parse_context(ctx_file, len, &p).empty()) return 0;
    return wire::ciphertext_blob_size(p, n_ct, limbs);
}
void emu_free(void *p) { std::free(p); }

// parameter generation rules
uint64_t emu_first_prime(uint32_t bits, uint32_t m) { return nt::first_prime(bits, m); }
uint64_t emu_previous_prime(uint64_t q, uint32_t m) { return nt::previous_prime(q, m); }
uint64_t emu_min_primitive_root(uint32_t m, uint64_t q) { return nt::min_primitive_root(m, q); }
uint32_t emu_select_ring_dim(uint32_t bits) { return nt::select_ring_dim(bits); }
uint32_t emu_select_ring_dim_ex(uint32_t bits, uint32_t mult_depth) { return nt::select_ring_dim(bits, mult_depth); }
int emu_modulus_chain(uint32_t bits, uint32_t mult_depth, uint32_t ring_dim, uint64_t *q) {
    return nt::modulus_chain(bits, mult_depth, ring_dim, q);
}

}  // extern "C"
