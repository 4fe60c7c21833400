// tables.cpp — see tables.h.
#include "tables.h"

#include <cmath>
#include <cstring>

#include "numtheory.h"

namespace shelfi {

namespace {
typedef unsigned __int128 u128;
uint32_t bit_reverse(uint32_t x, int bits) {
    uint32_t r = 0;
    for (int i = 0; i < bits; i++) r |= ((x >> i) & 1u) << (bits - 1 - i);
    return r;
}
int bit_length(uint64_t x) { return x ? 64 - __builtin_clzll(x) : 0; }
}  // namespace

void build_gauss_cdf(double sigma, uint64_t cdf[32]) {
    double rho[64], total;
    for (int k = 0; k < 64; k++) rho[k] = std::exp(-(double)k * (double)k / (2.0 * sigma * sigma));
    total = rho[0];
    for (int k = 1; k < 64; k++) total += 2.0 * rho[k];
    double acc = 0.0;
    for (int k = 0; k < 32; k++) {
        acc += (k == 0 ? rho[0] : 2.0 * rho[k]) / total;
        double v = std::ldexp(acc, 63);
        cdf[k] = (v >= 9223372036854775807.0) ? (UINT64_MAX >> 1) : (uint64_t)v;
    }
    cdf[31] = UINT64_MAX >> 1;
}

std::string make_limb_const(uint64_t q, uint64_t psi, uint32_t logn, LimbConst *lc, std::vector<tw_t> *tw,
                            std::vector<tw_t> *itw, std::vector<double> *dtw) {
    if (logn < (uint32_t)kLogN || logn > (uint32_t)(kLogN + kMaxLogBlocks)) return "ring dimension must be 8192, 16384 or 32768";
    const uint32_t n = 1u << logn, d = logn - kLogN;
    if (q >= ((uint64_t)1 << 60) || q < 64 || !(q & 1)) return "limb modulus must be an odd value in [64, 2^60)";
    if ((q - 1) % (2 * (uint64_t)n)) return "limb modulus is not 1 mod 2N";
    if (nt::pow_mod(psi, n, q) != q - 1) return "root of unity is not a primitive 2N-th root";
    std::memset(lc, 0, sizeof *lc);
    lc->q = q;
    lc->q2 = 2 * q;
    lc->nq = (uint64_t)0 - q;
    lc->mu = (uint64_t)((((u128)1) << 64) / q);
    uint64_t ipsi = nt::inv_mod(psi, q);
    tw->assign(n, tw_t{0, 0});
    itw->assign(n, tw_t{0, 0});
    uint64_t p = 1, ip = 1;
    for (uint32_t i = 0; i < n; i++) {  // PALISADE's reverse tables: index bitrev(i) holds psi^i
        uint32_t r = bit_reverse(i, (int)logn);
        (*tw)[r] = tw_t{p, nt::shoup(p, q)};
        (*itw)[r] = tw_t{ip, nt::shoup(ip, q)};
        p = nt::mul_mod(p, psi, q);
        ip = nt::mul_mod(ip, ipsi, q);
    }
    lc->ninv = nt::inv_mod(n % q, q);
    lc->ninv_s = nt::shoup(lc->ninv, q);
    lc->wninv = nt::mul_mod((*itw)[1].w, lc->ninv, q);
    lc->wninv_s = nt::shoup(lc->wninv, q);
    lc->r64 = (uint64_t)((((u128)1) << 64) % q);
    lc->r64_s = nt::shoup(lc->r64, q);
    // quick_reduce flavour: must send every 64-bit value into [0, 2q)
    int b = bit_length(q);
    u128 two_b = (u128)1 << b, kmax0 = ((u128)1 << (64 - b)) - 1;
    u128 delta1 = (u128)q - ((u128)1 << (b - 1)), kmax1 = ((u128)1 << (65 - b)) - 1;
    if (b >= 33 && two_b + kmax0 * (two_b - q) <= (u128)2 * q) {
        lc->rmode = 0; lc->rshift = (uint32_t)b; lc->radd = 0;   // q just below 2^b:  x - (x >> b) q
    } else if (b >= 34 && kmax1 * delta1 <= (u128)q) {
        lc->rmode = 1; lc->rshift = (uint32_t)(b - 1); lc->radd = q;   // q just above 2^(b-1): x - (x >> (b-1)) q + q
    } else {
        lc->rmode = 2; lc->rshift = 0;                    // anything else: Barrett quotient by mulhi
    }
    // forward NTT: every stage adds 3q to the bound (truncated Shoup quotient: products < 3q); reduce at the end of
    // pass p-1 when pass p would overflow.
    // The tile's loader has already applied d stages (HeadLoad): inputs are below (1 + 2d) q.
    const uint64_t H = UINT64_MAX / q;
    const int stages[5] = {3, 3, 3, 3, 1};
    uint64_t B = 1 + 2 * (uint64_t)d;
    lc->fwd_mask = 0;
    for (int pass = 0; pass < 5; pass++) {
        if (B + kBflyGrowth * stages[pass] > H) {
            if (pass == 0 || 2 + kBflyGrowth * (uint64_t)stages[pass] > H) return "modulus leaves no headroom for lazy butterflies";
            lc->fwd_mask |= 1u << (pass - 1);
            B = 2;
        }
        B += kBflyGrowth * stages[pass];
    }
    if (H < 16) return "modulus leaves no headroom for the inverse transform";
    // FP64-pipe forward transform: exact for q <= 1.125 * 2^52 (ckks_core.cuh: the lazy first stage of a radix-8 pass
    // needs 1.5 q + 2^50 < 2^53)
    if (dtw) {
        dtw->clear();
        if (q <= (((uint64_t)9) << 49) && q >= 1024) {
            auto centred = [q](uint64_t w) { return w > (q >> 1) ? -(double)(q - w) : (double)w; };
            dtw->assign(2 * (size_t)n, 0.0);          // forward twiddles, then the inverse ones (LimbConst::ditw = dtw + n)
            for (uint32_t i = 0; i < n; i++) {
                (*dtw)[i] = centred((*tw)[i].w);
                (*dtw)[n + i] = centred((*itw)[i].w);
            }
            lc->d_ninv = centred(lc->ninv);
            lc->d_wninv = centred(lc->wninv);
            lc->qd = (double)q;
            lc->qinv = 1.0 / lc->qd;
            double thr = 0.49 * lc->qd;
            uint64_t bits;
            std::memcpy(&bits, &thr, 8);
            lc->dthr = (uint32_t)(bits >> 32);
        }
    }
    return "";
}

void make_fft_tables(uint32_t M, uint32_t slots, std::vector<cplx> *inv, std::vector<cplx> *fwd) {
    std::vector<uint32_t> rot(M / 4);
    uint32_t five = 1;
    for (uint32_t i = 0; i < M / 4; i++) { rot[i] = five; five = (uint32_t)(((uint64_t)five * 5) % M); }
    std::vector<cplx> ksi(M + 1);
    for (uint32_t j = 0; j < M; j++) {
        double angle = 2.0 * M_PI * j / M;
        ksi[j].re = std::cos(angle);
        ksi[j].im = std::sin(angle);
    }
    ksi[M] = ksi[0];
    inv->assign(slots, cplx{0, 0});
    fwd->assign(slots, cplx{0, 0});
    for (uint32_t len = 2; len <= slots; len <<= 1) {
        uint32_t lenh = len >> 1, lenq = len << 2, gap = M / lenq;
        for (uint32_t j = 0; j < lenh; j++) {
            (*inv)[lenh - 1 + j] = ksi[(lenq - (rot[j] % lenq)) * gap];
            (*fwd)[lenh - 1 + j] = ksi[(rot[j] % lenq) * gap];
        }
    }
}

CrtNConst make_crtn(const uint64_t *q, uint32_t limbs) {
    CrtNConst c;
    std::memset(&c, 0, sizeof c);
    c.limbs = limbs;
    uint64_t Q[kMaxLimbs + 1] = {1};
    for (uint32_t i = 0; i < limbs; i++) {
        c.q[i] = q[i];
        c.mu[i] = (uint64_t)((((u128)1) << 64) / q[i]);
        for (uint32_t j = 0; j < i; j++) {
            c.inv[i][j] = nt::inv_mod(q[j] % q[i], q[i]);
            c.inv_s[i][j] = nt::shoup(c.inv[i][j], q[i]);
        }
        u128 carry = 0;
        for (int k = 0; k <= kMaxLimbs; k++) {
            u128 cur = (u128)Q[k] * q[i] + carry;
            Q[k] = (uint64_t)cur;
            carry = cur >> 64;
        }
    }
    for (int k = 0; k < kMaxLimbs; k++) {
        c.Q[k] = Q[k];
        c.Qh[k] = (Q[k] >> 1) | (Q[k + 1] << 63);
    }
    return c;
}

Crt2Const make_crt2(uint64_t q0, uint64_t q1) {
    Crt2Const c;
    std::memset(&c, 0, sizeof c);
    c.q0 = q0;
    if (!q1) return c;
    c.q1 = q1;
    c.q0inv = nt::inv_mod(q0 % q1, q1);
    c.q0inv_s = nt::shoup(c.q0inv, q1);
    c.mu1 = (uint64_t)((((u128)1) << 64) / q1);
    u128 Q = (u128)q0 * q1, Qh = Q >> 1;
    c.Q_hi = (uint64_t)(Q >> 64); c.Q_lo = (uint64_t)Q;
    c.Qh_hi = (uint64_t)(Qh >> 64); c.Qh_lo = (uint64_t)Qh;
    return c;
}

}  // namespace shelfi
