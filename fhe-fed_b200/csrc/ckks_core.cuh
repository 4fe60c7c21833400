// ckks_core.cuh — per-thread building blocks of the SHELFI CKKS hot path (sm_100a).
//
// Everything here is a `__host__ __device__` function of (thread id, shared-memory tile, limb
// constants): the CUDA kernels in kernels.cu call these between __syncthreads(), and the test
// harness tests/emu/ (g++, no GPU) replays the same phases thread by thread so index math and
// modular arithmetic can be checked bit for bit on a CPU-only box.  The harness is test
// infrastructure; the product path is the kernels.
//
// Reference semantics being reproduced (PALISADE, called from
// palisade_pybind/SHELFI_FHE/src/ckks.cpp:80-81,189,288,296):
//   * ChineseRemainderTransformFTT::ForwardTransformToBitReverse  -> fwd_* phases
//   * ChineseRemainderTransformFTT::InverseTransformFromBitReverse -> inv_* phases
//   * DiscreteFourierTransform::FFTSpecial / FFTSpecialInv        -> fft_* phases
//   * DCRTPoly::CRTInterpolate + CKKSPackedEncoding::Decode        -> crt2_centered / i128_to_double
#pragma once
#include <stdint.h>
#include <math.h>
#include <string.h>

#ifdef __CUDACC__
#define HD __host__ __device__ __forceinline__
#define HD_COLD static __host__ __device__ __noinline__    // rare paths: one out-of-line copy, not one per call site
#else
#define HD inline
#define HD_COLD inline
#endif

namespace shelfi {

typedef uint64_t u64;
typedef int64_t i64;
typedef unsigned int u32;
typedef int i32;

constexpr int kMaxLimbs = 8;
// The NTT working set is one 8192-point tile (64 KiB of shared memory, 3 CTAs per SM).  Ring dimension 8192 —
// all that SHELFI's multDepth = 1 contexts produce — is exactly one tile.  Ring dimensions 2^(13+d), d = 1, 2
// (BASELINE configs 3-4) split every limb polynomial into 2^d tiles: the d outermost Cooley-Tukey stages are
// applied by the tile's loader (HeadLoad), the d outermost Gentleman-Sande stages by a streaming tail pass,
// and the 13 inner stages run on the tile with twiddle indices offset by the tile number (`hb` below).
constexpr int kLogN = 13;
constexpr int kN = 1 << kLogN;       // tile points (NOT the ring dimension: that is RingConst::n)
constexpr int kMaxLogBlocks = 2;     // ring dimension <= 32768
constexpr int kNttThreads = 256;     // one CTA per tile, 32 coefficients per thread
constexpr int kFftThreads = 256;

struct __attribute__((aligned(16))) tw_t { u64 w, ws; };  // twiddle and its Shoup companion floor(w*2^64/q)
struct __attribute__((aligned(16))) u64x2 { u64 x, y; };
#ifndef __CUDACC__
struct __attribute__((aligned(16))) double2 { double x, y; };
#endif

// L2 residency hints (fused encrypt).  A CTA parks values in global memory for its own later passes (m + e0 in a
// not-yet-written output row, v^ in the c1 row) while it streams finished rows that nobody reads again; without hints
// the streamed rows push the parked ones out of L2 before they are re-read or overwritten (round 1: 1.46x the
// algorithmic DRAM bytes).  keep = a createpolicy evict_last descriptor (0: plain store), stream = st.global.cs.
HD u64 make_keep_policy() {
#ifdef __CUDA_ARCH__
    u64 pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
    return pol;
#else
    return 0;
#endif
}
HD void st_pair(u64x2 *p, u64x2 v, u64 keep, bool stream) {
#ifdef __CUDA_ARCH__
    if (keep) {
        asm volatile("st.global.L2::cache_hint.v2.u64 [%0], {%1, %2}, %3;" ::"l"(p), "l"(v.x), "l"(v.y), "l"(keep) : "memory");
        return;
    }
    if (stream) {
        asm volatile("st.global.cs.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(v.x), "l"(v.y) : "memory");
        return;
    }
#else
    (void)keep; (void)stream;
#endif
    *p = v;
}
// Per-limb constants, passed to kernels by value inside RingConst.
struct LimbConst {
    u64 q;          // modulus (<= 60 bits)
    u64 q2;         // 2q
    u64 nq;         // 2^64 - q: x*w + h*nq == x*w - h*q (mod 2^64) with the subtraction folded into the multiply-add
    u64 mu;         // floor(2^64 / q): generic Barrett quotient estimate
    u64 ninv, ninv_s;       // N^-1 mod q and Shoup companion
    u64 wninv, wninv_s;     // itw[1] * N^-1 mod q (last inverse stage with the scaling folded in)
    u64 r64, r64_s;         // 2^64 mod q and Shoup companion (128-bit accumulator reduction)
    u64 radd;       // quick-reduce addend: 0 (mode 0) or q (mode 1)
    u32 rshift;     // quick-reduce shift (>= 32 in modes 0 and 1)
    u32 rmode;      // 0: q just below 2^rshift; 1: q just above 2^rshift; 2: generic (mulhi by mu)
    u32 fwd_mask;   // bit p: quick-reduce to [0,2q) at the end of forward pass p (0..3)
    u32 dthr;       // FP64 path: high word of the double 0.49 q (renormalisation threshold)
    const tw_t *tw;   // forward table, index 1..N-1: psi^bitrev(i)      (PALISADE rootOfUnityReverseTable)
    const tw_t *itw;  // inverse table:                psi^-bitrev(i)
    // FP64-pipe forward transform (limbs with q < 2^53, ring dimension 8192): centred twiddles as doubles, or null
    const double *dtw;
    double qd, qinv;  // q and 1/q as doubles
    // ... and the inverse transform: centred psi^-bitrev(i), N^-1 and itw[1] * N^-1 as doubles (null / 0 with dtw)
    const double *ditw;
    double d_ninv, d_wninv;
};

struct RingConst {
    u32 n, logn, limbs, slots;
    u32 log_blocks;   // d = logn - 13: tiles per limb polynomial = 1 << d
    u32 pad;
    LimbConst limb[kMaxLimbs];
};
// "head bits" of tile blk of a polynomial split into 2^d tiles: the twiddle index of stage s inside the tile is
// (hb << s) + (local index bits), hb = 2^d | blk; one tile per polynomial gives hb = 1 (the plain tables).
HD u32 head_bits(u32 d, u32 blk) { return (1u << d) | blk; }

// ------------------------------------------------------------------------------------------
// 64-bit modular arithmetic on the integer pipes
// ------------------------------------------------------------------------------------------
HD u64 mulhi(u64 a, u64 b) {
#ifdef __CUDA_ARCH__
    return __umul64hi(a, b);
#else
    return (u64)(((unsigned __int128)a * b) >> 64);
#endif
}
HD u32 mulhi32(u32 a, u32 b) {
#ifdef __CUDA_ARCH__
    return __umulhi(a, b);
#else
    return (u32)(((uint64_t)a * b) >> 32);
#endif
}
// Shoup product, lazy: x*w - floor(x*ws/2^64)*q, in [0, 2q) for ANY 64-bit x (w < q).
HD u64 shoup_lazy(u64 x, u64 w, u64 ws, u64 q) { return x * w - mulhi(x, ws) * q; }
// same value, written as two multiply-adds (nq = 2^64 - q): no negate / 64-bit subtract on the ALU pipe
HD u64 shoup_lazy_nq(u64 x, u64 w, u64 ws, u64 nq) {
#ifdef __CUDA_ARCH__
    // 2 IMAD.WIDE + 4 IMAD on the fma pipe, nothing on the ALU pipe: both low products accumulate into one
    // 64-bit register, the four cross terms ride on IMAD's 32-bit addend.
    u64 h = __umul64hi(x, ws);
    u32 xl = (u32)x, xh = (u32)(x >> 32), wl = (u32)w, wh = (u32)(w >> 32);
    u32 hl = (u32)h, hh = (u32)(h >> 32), nl = (u32)nq, nh = (u32)(nq >> 32);
    u64 r;
    asm("{\n\t"
        ".reg .u64 t;\n\t"
        ".reg .u32 lo, hi;\n\t"
        "mul.wide.u32 t, %1, %3;\n\t"
        "mad.wide.u32 t, %5, %7, t;\n\t"
        "mov.b64 {lo, hi}, t;\n\t"
        "mad.lo.u32 hi, %1, %4, hi;\n\t"
        "mad.lo.u32 hi, %2, %3, hi;\n\t"
        "mad.lo.u32 hi, %5, %8, hi;\n\t"
        "mad.lo.u32 hi, %6, %7, hi;\n\t"
        "mov.b64 %0, {lo, hi};\n\t"
        "}"
        : "=l"(r)
        : "r"(xl), "r"(xh), "r"(wl), "r"(wh), "r"(hl), "r"(hh), "r"(nl), "r"(nh));
    return r;
#else
    return x * w + mulhi(x, ws) * nq;
#endif
}
// floor(x*ws / 2^64) or one less: the 2^0 partial product xl*sl is dropped (it contributes < 1), the two middle
// products are summed exactly.  3 IMAD.WIDE instead of the 4 (+ carry chain) of __umul64hi.
HD u64 mulhi_trunc(u64 x, u64 ws) {
    u32 xl = (u32)x, xh = (u32)(x >> 32), sl = (u32)ws, sh = (u32)(ws >> 32);
#ifdef __CUDA_ARCH__
    u64 r;
    asm("{\n\t"
        ".reg .u64 t1, t2, a;\n\t"
        ".reg .u32 t1l, t1h, t2l, t2h, c, ch;\n\t"
        "mul.wide.u32 t1, %2, %3;\n\t"
        "mov.b64 {t1l, t1h}, t1;\n\t"
        "cvt.u64.u32 a, t1l;\n\t"
        "mad.wide.u32 t2, %1, %4, a;\n\t"
        "mov.b64 {t2l, t2h}, t2;\n\t"
        "add.cc.u32 c, t1h, t2h;\n\t"
        "addc.u32 ch, 0, 0;\n\t"
        "mov.b64 a, {c, ch};\n\t"
        "mad.wide.u32 %0, %2, %4, a;\n\t"
        "}"
        : "=l"(r)
        : "r"(xl), "r"(xh), "r"(sl), "r"(sh));
    return r;
#else
    u64 t1 = (u64)xh * sl;
    u64 t2 = (u64)xl * sh + (u32)t1;
    return (u64)xh * sh + (t1 >> 32) + (t2 >> 32);
#endif
}
// Shoup product with the truncated quotient: x*w - Q'*q with Q' in {Q - 1, Q}, so the result lies in [0, 3q) for ANY
// 64-bit x.  The butterflies use this one; their bounds grow by 3q per stage instead of 2q (tables.cpp schedules the
// quick-reduces accordingly), which costs the 60-bit limb two more quick-reduces per transform and saves one
// IMAD.WIDE plus the carry chain in each of the 53,248 butterflies.
HD u64 shoup_lazy3_nq(u64 x, u64 w, u64 ws, u64 nq) {
    u64 h = mulhi_trunc(x, ws);
#ifdef __CUDA_ARCH__
    u32 xl = (u32)x, xh = (u32)(x >> 32), wl = (u32)w, wh = (u32)(w >> 32);
    u32 hl = (u32)h, hh = (u32)(h >> 32), nl = (u32)nq, nh = (u32)(nq >> 32);
    u64 r;
    asm("{\n\t"
        ".reg .u64 t;\n\t"
        ".reg .u32 lo, hi;\n\t"
        "mul.wide.u32 t, %1, %3;\n\t"
        "mad.wide.u32 t, %5, %7, t;\n\t"
        "mov.b64 {lo, hi}, t;\n\t"
        "mad.lo.u32 hi, %1, %4, hi;\n\t"
        "mad.lo.u32 hi, %2, %3, hi;\n\t"
        "mad.lo.u32 hi, %5, %8, hi;\n\t"
        "mad.lo.u32 hi, %6, %7, hi;\n\t"
        "mov.b64 %0, {lo, hi};\n\t"
        "}"
        : "=l"(r)
        : "r"(xl), "r"(xh), "r"(wl), "r"(wh), "r"(hl), "r"(hh), "r"(nl), "r"(nh));
    return r;
#else
    return x * w + h * nq;
#endif
}
HD u64 csub(u64 x, u64 m) { return x >= m ? x - m : x; }
// x (any 64-bit value) -> [0, 2q).  PALISADE's primes sit next to a power of two, so the quotient estimate is a shift:
// modes 0 / 1 (q just below 2^rshift / just above it, rshift >= 32): x - (x >> rshift) q (+ q), written as
// x + radd + hi * (2^64 - q) = one funnel shift, one IMAD.WIDE, one IMAD, no branch per value; mode 2 (anything
// else): Barrett quotient by mulhi.  tables.cpp proves the range for every 64-bit x before choosing a mode.
HD u64 quick_reduce_shift(u64 x, const LimbConst &lc) {
    u32 hi = (u32)(x >> lc.rshift);
#ifdef __CUDA_ARCH__
    u32 nl = (u32)lc.nq, nh = (u32)(lc.nq >> 32);
    u64 r, base = x + lc.radd;
    asm("{\n\t"
        ".reg .u64 t;\n\t"
        ".reg .u32 lo, hi;\n\t"
        "mad.wide.u32 t, %1, %2, %4;\n\t"
        "mov.b64 {lo, hi}, t;\n\t"
        "mad.lo.u32 hi, %1, %3, hi;\n\t"
        "mov.b64 %0, {lo, hi};\n\t"
        "}"
        : "=l"(r)
        : "r"(hi), "r"(nl), "r"(nh), "l"(base));
    return r;
#else
    return x + lc.radd + (u64)hi * lc.nq;
#endif
}
HD u64 quick_reduce(u64 x, const LimbConst &lc) {
    if (lc.rmode == 2) return x - mulhi(x, lc.mu) * lc.q;
    return quick_reduce_shift(x, lc);
}
// the same over the 8 values of a radix-8 pass with the mode test hoisted out
HD void quick_reduce8(u64 (&a)[8], const LimbConst &lc) {
    if (lc.rmode == 2) {
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = a[k] - mulhi(a[k], lc.mu) * lc.q;
    } else {
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = quick_reduce_shift(a[k], lc);
    }
}
HD u64 full_reduce(u64 x, const LimbConst &lc) { return csub(quick_reduce(x, lc), lc.q); }
// any 64-bit x -> [0, q)
HD u64 barrett_reduce(u64 x, const LimbConst &lc) { return csub(x - mulhi(x, lc.mu) * lc.q, lc.q); }
HD u64 signed_to_mod(i64 x, const LimbConst &lc) {
    u64 mag = x < 0 ? (u64)0 - (u64)x : (u64)x;
    u64 r = csub(quick_reduce(mag, lc), lc.q);   // quick_reduce is valid for every 64-bit value (tables.cpp)
    return (x < 0 && r) ? lc.q - r : r;
}

// shared-memory swizzle: XOR address bits 3..1 with bits 6..4.  Keeps (2i, 2i+1) adjacent (128-bit
// accesses in the radix-2 pass) and makes every radix-8 pass's 64-bit accesses conflict-free.
HD u32 swz(u32 i) { return i ^ ((i >> 3) & 0xEu); }

// Shared-memory tile addressing.  All offsets are BYTE offsets into the 64 KiB tile; with the swizzle above every
// access of a pass is `thread constant (+/^) compile-time constant`, so the hardware addressing modes absorb it:
//   pass A/B (LO >= 7): bits 6..4 of the element index come from the thread only -> one XOR per thread, then
//                       immediate offsets k << (LO+3) and round * 2048*8 (A) / 16384 (B)
//   pass C   (LO == 4): the register index k sits in bits 6..4 -> offset ^ (k << 4), one LOP3 per access
//   pass D   (LO == 1): k sits in bits 3..1 and is XORed with a thread constant -> (base | h3 << 4) ^ (k << 4)
//   pass E            : pairs (2vt, 2vt+1): 16*vt ^ (((vt >> 3) & 7) << 4), rounds add 4096
HD u64 &tile(char *sm, u32 byte_off) { return *reinterpret_cast<u64 *>(sm + byte_off); }
HD const u64 &tile(const char *sm, u32 byte_off) { return *reinterpret_cast<const u64 *>(sm + byte_off); }
template <int LO> HD u32 tile_base(u32 vt) {  // byte offset of the k = 0 element of virtual thread vt
    if (LO >= 7) {
        u32 low = vt & ((1u << LO) - 1u), high = vt >> LO;
        u32 e = (high << (LO + 3)) | low;
        return (e ^ ((e >> 3) & 0xEu)) << 3;
    } else if (LO == 4) {
        return (((vt >> 4) << 7) | (vt & 15u)) << 3;
    } else {  // LO == 1
        u32 high = vt >> 1;
        return (((high << 4) | (vt & 1u)) << 3) | ((high & 7u) << 4);
    }
}
template <int LO> HD u32 tile_off(u32 base, int k) {
    if (LO >= 7) return base + ((u32)k << (LO + 3));
    if (LO == 4) return (base ^ ((u32)k << 4)) + ((u32)k << 7);
    return base ^ ((u32)k << 4);
}
HD u32 tile_pair(u32 vt) { return (vt << 4) ^ (((vt >> 3) & 7u) << 4); }

// Cooley-Tukey butterfly, Harvey lazy form: no per-butterfly correction, bounds grow by 3q per stage (truncated
// Shoup quotient) and are brought back by quick_reduce at the pass boundaries named in LimbConst::fwd_mask.
// q3 = 3q.
// Measured on B200 (profiles/README.md): the truncated quotient is NOT faster — the 60-bit limb needs three
// quick-reduces per transform instead of one and ptxas already schedules __umul64hi's carry chain well
// (8140-row forward NTT 0.541 ms truncated vs 0.533 ms exact).  Kept behind a switch, off by default.
#ifndef SHELFI_TRUNC_SHOUP
#define SHELFI_TRUNC_SHOUP 0
#endif
constexpr int kBflyGrowth = SHELFI_TRUNC_SHOUP ? 3 : 2;   // bound growth per forward stage, in units of q
HD u64 bfly_mul(u64 x, u64 w, u64 ws, u64 nq) {
    return SHELFI_TRUNC_SHOUP ? shoup_lazy3_nq(x, w, ws, nq) : shoup_lazy_nq(x, w, ws, nq);
}
HD void ct_bfly(u64 &x, u64 &y, const tw_t &w, u64 nq, u64 q3) {
    u64 t = bfly_mul(y, w.w, w.ws, nq);
    y = x + q3 - t;
    x = x + t;
}
// radix-8 register pass: a[k], k = (b2 b1 b0) with b2 the address bit of the first stage.
HD void ct_radix8(u64 (&a)[8], const tw_t *__restrict__ tw, u32 t0, u64 nq, u64 q3) {
    tw_t w = tw[t0];
#pragma unroll
    for (int k = 0; k < 4; k++) ct_bfly(a[k], a[k + 4], w, nq, q3);
    tw_t w0 = tw[2 * t0], w1 = tw[2 * t0 + 1];
    ct_bfly(a[0], a[2], w0, nq, q3); ct_bfly(a[1], a[3], w0, nq, q3);
    ct_bfly(a[4], a[6], w1, nq, q3); ct_bfly(a[5], a[7], w1, nq, q3);
#pragma unroll
    for (int k = 0; k < 4; k++) ct_bfly(a[2 * k], a[2 * k + 1], tw[4 * t0 + k], nq, q3);
}

// ---- forward NTT phases (natural order in, bit-reversed order out), 8192 points, 256 threads ----
// Pass A: stages 0-2 (address bits 12,11,10) straight from the loader into shared memory.
template <class Load>
HD void fwd_phase_a(int tid, u64 *smem, const LimbConst &lc, u32 hb, Load ld) {
    char *sm = reinterpret_cast<char *>(smem);
    const u64 nq = lc.nq, q3 = kBflyGrowth * lc.q;
#pragma unroll 1
    for (int r = 0; r < 4; r++) {
        u32 vt = tid + kNttThreads * r;
        u64 a[8];
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = ld((u32)(k << 10) | vt);
        ct_radix8(a, lc.tw, hb, nq, q3);
        if (lc.fwd_mask & 1u) quick_reduce8(a, lc);
        u32 base = tile_base<10>(vt);
#pragma unroll
        for (int k = 0; k < 8; k++) tile(sm, tile_off<10>(base, k)) = a[k];
    }
}
// Passes B, C, D: LO = 7, 4, 1 (lowest of the three address bits held in registers).
template <int LO>
HD void fwd_phase_mid(int tid, u64 *smem, const LimbConst &lc, u32 hb) {
    constexpr int PASS = (10 - LO) / 3;               // 1, 2, 3
    constexpr u32 S0 = 12 - (LO + 2);                 // first stage of the pass: 3, 6, 9
    char *sm = reinterpret_cast<char *>(smem);
    const u64 nq = lc.nq, q3 = kBflyGrowth * lc.q;
#pragma unroll 1
    for (int r = 0; r < 4; r++) {
        u32 vt = tid + kNttThreads * r;
        u32 base = tile_base<LO>(vt);
        u64 a[8];
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = tile(sm, tile_off<LO>(base, k));
        ct_radix8(a, lc.tw, (hb << S0) + (vt >> LO), nq, q3);
        if (lc.fwd_mask & (1u << PASS)) quick_reduce8(a, lc);
#pragma unroll
        for (int k = 0; k < 8; k++) tile(sm, tile_off<LO>(base, k)) = a[k];
    }
}
// Pass E: stage 12 (address bit 0); st(pair index, x, y) receives lazily reduced values in
// [0, (B+2)q) — the storer applies full_reduce or folds further work in first.
// Storer protocol: `typename Store::Aux`, `Store::kGroup` rounds in flight, `fetch(vt)` issues the epilogue's own
// global loads (parked data, key material) together with the twiddle loads, `operator()(vt, x, y, aux)` finishes.
struct NoAux {};
template <class Store>
HD void fwd_phase_e(int tid, const u64 *smem, const LimbConst &lc, u32 hb, Store st) {
    constexpr int G = Store::kGroup;
    const char *sm = reinterpret_cast<const char *>(smem);
    const u64 nq = lc.nq, q3 = kBflyGrowth * lc.q;
    const tw_t *__restrict__ tw = lc.tw + ((size_t)hb << (kLogN - 1));
#pragma unroll 1
    for (int r0 = 0; r0 < 16; r0 += G) {  // G rounds in flight: every long-latency load is issued before any math
        tw_t w[G];
        u64x2 v[G];
        typename Store::Aux aux[G];
#pragma unroll
        for (int i = 0; i < G; i++) w[i] = tw[tid + kNttThreads * (r0 + i)];
#pragma unroll
        for (int i = 0; i < G; i++) aux[i] = st.fetch((u32)(tid + kNttThreads * (r0 + i)));
#pragma unroll
        for (int i = 0; i < G; i++) v[i] = *reinterpret_cast<const u64x2 *>(sm + tile_pair(tid + kNttThreads * (r0 + i)));
#pragma unroll
        for (int i = 0; i < G; i++) {
            ct_bfly(v[i].x, v[i].y, w[i], nq, q3);
            st((u32)(tid + kNttThreads * (r0 + i)), v[i].x, v[i].y, aux[i]);
        }
    }
}

// unfused IEEE double operations (nvcc would otherwise contract a*b+c; the reference does not)
HD double fmul(double a, double b) {
#ifdef __CUDA_ARCH__
    return __dmul_rn(a, b);
#else
    return a * b;
#endif
}
HD double fadd(double a, double b) {
#ifdef __CUDA_ARCH__
    return __dadd_rn(a, b);
#else
    return a + b;
#endif
}
HD double fsub(double a, double b) {
#ifdef __CUDA_ARCH__
    return __dsub_rn(a, b);
#else
    return a - b;
#endif
}
// ------------------------------------------------------------------------------------------
// The same forward transform on the FP64 pipe, for limbs with q < 2^53 (PALISADE's scaling primes at <= 52 bits).
// B200 issues DFMA at 50 /clk/SM against 29 for IMAD.WIDE (measured: tools/ubench/pipes.cu), and a 53 x 53 ->
// 106-bit product is two instructions there (DMUL + DFMA) against four IMAD.WIDE: 56 ns per row against 70.
//
// Values are integers held exactly in doubles, centred: |v| <= 0.51 q between stages.
//   mulmod(y, w), |y| < q, |w| <= q/2:  h = RN(y w), l = y w - h (exact, FMA), c = rint(h / q) by the 1.5*2^52 magic
//   constant (|h / q| < 2^51), d = h - c q (exact: an integer below 2^53), r = d + l.  |y w / q - c| <= 1/2 + |c| 2^-52
//   so |r| <= 0.76 q for |y| <= 0.5 q.
//   butterfly: x +- r is at most 1.26 q < 2^53 (exact); renormalise with v - q rint(v / q) back to |v| <= q/2.
//   First-stage outputs of a radix-8 pass that the second stage only adds to skip that renormalisation
//   (d_bfly_lazy): they are at most q + 2^49, the sums that follow at most 1.5 q + 2^50.
//   (q <= 1.125 * 2^52 keeps every bound above below 2^53: tables.cpp admits only such limbs.)
// Every rounding that happens is of a value the derivation does not need exactly (h, c), so the residues — and after
// the final conversion to [0, q) the stored words — equal the integer path's bit for bit.
// ------------------------------------------------------------------------------------------
HD double d_fma(double a, double b, double c) {
#ifdef __CUDA_ARCH__
    return __fma_rn(a, b, c);
#else
    return fma(a, b, c);
#endif
}
HD u32 d_hi(double v) {
#ifdef __CUDA_ARCH__
    return (u32)__double2hiint(v);
#else
    u64 b; memcpy(&b, &v, 8); return (u32)(b >> 32);
#endif
}
HD double d_copysign_q(double q, double v) {   // q > 0 with the sign of v
#ifdef __CUDA_ARCH__
    return __hiloint2double((int)((u32)__double2hiint(q) | ((u32)__double2hiint(v) & 0x80000000u)), __double2loint(q));
#else
    return copysign(q, v);
#endif
}
HD double d_mulmod(double y, double w, double q, double qinv) {
    const double M = 6755399441055744.0;   // 1.5 * 2^52
    double h = fmul(y, w);
    double l = d_fma(y, w, -h);
    double c = fsub(d_fma(h, qinv, M), M);
    double d = d_fma(-c, q, h);
    return fadd(d, l);
}
// Renormalisation to |v| <= q/2 (+1): v - q rint(v / q), three FP64 operations and nothing on the ALU pipe.
// Measured on B200 (8140-row forward NTT, 52-bit prime): this form 58.3 ns/row; compare-on-high-word + two FSEL
// (1 FP64 + 5 ALU operations) 67.9; DSETP + predicated DADD (2 FP64 + 1 ALU) 64.8.  The others stay selectable.
#ifndef SHELFI_DRENORM
#define SHELFI_DRENORM 1
#endif
HD double d_renorm(double v, double q, double qinv, u32 thr) {
#if SHELFI_DRENORM == 0
    (void)qinv;
    return (d_hi(v) & 0x7fffffffu) > thr ? fsub(v, d_copysign_q(q, v)) : v;
#elif SHELFI_DRENORM == 1
    (void)thr;
    const double M = 6755399441055744.0;
    double c = fsub(d_fma(v, qinv, M), M);
    return d_fma(-c, q, v);
#else
    (void)qinv;
#ifdef __CUDA_ARCH__
    double t = __hiloint2double((int)thr, 0), s = d_copysign_q(q, v);
    asm("{\n\t"
        ".reg .pred p;\n\t"
        ".reg .f64 a;\n\t"
        "abs.f64 a, %0;\n\t"
        "setp.gt.f64 p, a, %1;\n\t"
        "@p sub.rn.f64 %0, %0, %2;\n\t"
        "}"
        : "+d"(v)
        : "d"(t), "d"(s));
    return v;
#else
    u64 tb = (u64)thr << 32;
    double t;
    memcpy(&t, &tb, 8);
    return fabs(v) > t ? v - copysign(q, v) : v;
#endif
#endif
}
HD void d_bfly(double &x, double &y, double w, double q, double qinv, u32 thr) {
    double t = d_mulmod(y, w, q, qinv);
    y = d_renorm(fsub(x, t), q, qinv, thr);
    x = d_renorm(fadd(x, t), q, qinv, thr);
}
// Both outputs left un-renormalised, |.| <= q + 2^49: allowed for a value that the NEXT stage uses only as the additive
// operand x of a full d_bfly (x +- t then stays below 1.5 q + 2^50 < 2^53 for q <= 1.125 * 2^52, still an exact
// integer, and the rint-style d_renorm folds any such value back).  Saves 6 of the 14 FP64 operations.
HD void d_bfly_lazy(double &x, double &y, double w, double q, double qinv) {
    double t = d_mulmod(y, w, q, qinv);
    y = fsub(x, t);
    x = fadd(x, t);
}
HD void d_radix8(double (&a)[8], const double *__restrict__ tw, u32 t0, double q, double qinv, u32 thr) {
    double w = tw[t0];
#if SHELFI_DRENORM == 1
    // stage 2 pairs (0,2) (1,3) (4,6) (5,7): a[0], a[1], a[4], a[5] are additive operands there
    d_bfly_lazy(a[0], a[4], w, q, qinv); d_bfly_lazy(a[1], a[5], w, q, qinv);
    d_bfly(a[2], a[6], w, q, qinv, thr); d_bfly(a[3], a[7], w, q, qinv, thr);
#else
#pragma unroll
    for (int k = 0; k < 4; k++) d_bfly(a[k], a[k + 4], w, q, qinv, thr);
#endif
    double w0 = tw[2 * t0], w1 = tw[2 * t0 + 1];
    d_bfly(a[0], a[2], w0, q, qinv, thr); d_bfly(a[1], a[3], w0, q, qinv, thr);
    d_bfly(a[4], a[6], w1, q, qinv, thr); d_bfly(a[5], a[7], w1, q, qinv, thr);
#pragma unroll
    for (int k = 0; k < 4; k++) d_bfly(a[2 * k], a[2 * k + 1], tw[4 * t0 + k], q, qinv, thr);
}
HD double &dtile(char *sm, u32 byte_off) { return *reinterpret_cast<double *>(sm + byte_off); }
// residue in [0, q) -> centred double; centred integer double (|v| < q) -> residue in [0, q)
HD double d_from_residue(u64 r, u64 q) { return (double)(r > (q >> 1) ? (i64)r - (i64)q : (i64)r); }
HD u64 d_to_residue(double v, u64 q) {
    i64 i = (i64)v;
    return i < 0 ? (u64)(i + (i64)q) : (u64)i;
}
template <class Load>   // ld(index) -> centred double, |v| <= q/2
HD void dfwd_phase_a(int tid, u64 *smem, const LimbConst &lc, u32 hb, Load ld) {
    char *sm = reinterpret_cast<char *>(smem);
    const double q = lc.qd, qinv = lc.qinv;
    const u32 thr = lc.dthr;
#pragma unroll 1
    for (int r = 0; r < 4; r++) {
        u32 vt = tid + kNttThreads * r;
        double a[8];
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = ld((u32)(k << 10) | vt);
        d_radix8(a, lc.dtw, hb, q, qinv, thr);
        u32 base = tile_base<10>(vt);
#pragma unroll
        for (int k = 0; k < 8; k++) dtile(sm, tile_off<10>(base, k)) = a[k];
    }
}
template <int LO>
HD void dfwd_phase_mid(int tid, u64 *smem, const LimbConst &lc, u32 hb) {
    constexpr u32 S0 = 12 - (LO + 2);
    char *sm = reinterpret_cast<char *>(smem);
    const double q = lc.qd, qinv = lc.qinv;
    const u32 thr = lc.dthr;
#pragma unroll 1
    for (int r = 0; r < 4; r++) {
        u32 vt = tid + kNttThreads * r;
        u32 base = tile_base<LO>(vt);
        double a[8];
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = dtile(sm, tile_off<LO>(base, k));
        d_radix8(a, lc.dtw, (hb << S0) + (vt >> LO), q, qinv, thr);
#pragma unroll
        for (int k = 0; k < 8; k++) dtile(sm, tile_off<LO>(base, k)) = a[k];
    }
}
// last stage + conversion back to residues: the storers of the integer path are used unchanged (they get x, y in [0, q))
template <class Store>
HD void dfwd_phase_e(int tid, const u64 *smem, const LimbConst &lc, u32 hb, Store st) {
    constexpr int G = Store::kGroup;
    const char *sm = reinterpret_cast<const char *>(smem);
    const double q = lc.qd, qinv = lc.qinv;
    const u32 thr = lc.dthr;
    const double *__restrict__ tw = lc.dtw + ((size_t)hb << (kLogN - 1));
#pragma unroll 1
    for (int r0 = 0; r0 < 16; r0 += G) {
        double w[G];
        double2 v[G];
        typename Store::Aux aux[G];
#pragma unroll
        for (int i = 0; i < G; i++) w[i] = tw[tid + kNttThreads * (r0 + i)];
#pragma unroll
        for (int i = 0; i < G; i++) aux[i] = st.fetch((u32)(tid + kNttThreads * (r0 + i)));
#pragma unroll
        for (int i = 0; i < G; i++) v[i] = *reinterpret_cast<const double2 *>(sm + tile_pair(tid + kNttThreads * (r0 + i)));
#pragma unroll
        for (int i = 0; i < G; i++) {
            d_bfly(v[i].x, v[i].y, w[i], q, qinv, thr);
            st((u32)(tid + kNttThreads * (r0 + i)), d_to_residue(v[i].x, lc.q), d_to_residue(v[i].y, lc.q), aux[i]);
        }
    }
}

// HeadLoad on doubles: the d outermost stages for tile blk, every intermediate renormalised (|result| <= q/2)
template <int D, class Inner>
struct DHeadLoad {
    Inner in; const LimbConst *lc; u32 blk;
    HD double operator()(u32 i) const {
        if (D == 0) return in(i);
        const double q = lc->qd, qinv = lc->qinv;
        const double *__restrict__ tw = lc->dtw;
        if (D == 1) {
            double x = in(i), t = d_mulmod(in(i + kN), tw[1], q, qinv);
            return d_renorm(blk ? fsub(x, t) : fadd(x, t), q, qinv, lc->dthr);
        }
        double x0 = in(i), x1 = in(i + kN);
        double t0 = d_mulmod(in(i + 2 * kN), tw[1], q, qinv), t1 = d_mulmod(in(i + 3 * kN), tw[1], q, qinv);
        double s0 = d_renorm(blk < 2 ? fadd(x0, t0) : fsub(x0, t0), q, qinv, lc->dthr);
        double s1 = d_renorm(blk < 2 ? fadd(x1, t1) : fsub(x1, t1), q, qinv, lc->dthr);
        double t = d_mulmod(s1, tw[2 + (blk >> 1)], q, qinv);
        return d_renorm((blk & 1u) ? fsub(s0, t) : fadd(s0, t), q, qinv, lc->dthr);
    }
};

// ---- the inverse transform on the FP64 pipe (same limbs as the forward one: q <= 1.125 * 2^52) ----
// Gentleman-Sande butterfly on centred integer-valued doubles: s = x + y, d = x - y, both renormalised to |.| <= q/2
// (|x|, |y| <= 0.76 q on entry, so the sums are exact integers below 1.52 q < 2^53), then r = d * w mod q, |r| <= 0.76 q,
// left as it is: the next stage's sums absorb it.  14 FP64 operations, none on the integer pipes; the integer
// butterfly costs 6 IMAD.WIDE + 4 IMAD + carries (108 ns per 8192-point row against 63 forward in round 1, which is
// what decrypt and rescale paid).  Every intermediate is an exact integer, so the residues equal the integer path's.
HD void d_gs_bfly(double &x, double &y, double w, double q, double qinv, u32 thr) {
    const double s = d_renorm(fadd(x, y), q, qinv, thr);
    const double d = d_renorm(fsub(x, y), q, qinv, thr);
    x = s;
    y = d_mulmod(d, w, q, qinv);
}
HD void d_gs_radix8(double (&a)[8], const double *__restrict__ itw, u32 t2, double q, double qinv, u32 thr) {
#pragma unroll
    for (int k = 0; k < 4; k++) d_gs_bfly(a[2 * k], a[2 * k + 1], itw[4 * t2 + k], q, qinv, thr);
    const double w0 = itw[2 * t2], w1 = itw[2 * t2 + 1];
    d_gs_bfly(a[0], a[2], w0, q, qinv, thr); d_gs_bfly(a[1], a[3], w0, q, qinv, thr);
    d_gs_bfly(a[4], a[6], w1, q, qinv, thr); d_gs_bfly(a[5], a[7], w1, q, qinv, thr);
    const double w = itw[t2];
#pragma unroll
    for (int k = 0; k < 4; k++) d_gs_bfly(a[k], a[k + 4], w, q, qinv, thr);
}
// first pass (address bit 0) from the loader; ld(pair index, &x, &y) delivers values below 4q (integer words)
template <class Load>
HD void dinv_phase_e(int tid, u64 *smem, const LimbConst &lc, u32 hb, Load ld) {
    char *sm = reinterpret_cast<char *>(smem);
    const double q = lc.qd, qinv = lc.qinv;
    const u32 thr = lc.dthr;
    const double *__restrict__ itw = lc.ditw + ((size_t)hb << (kLogN - 1));
#pragma unroll 2
    for (int r = 0; r < 16; r++) {
        const u32 vt = tid + kNttThreads * r;
        u64 xi, yi;
        ld(vt, xi, yi);
        double x = d_from_residue(csub(csub(xi, lc.q2), lc.q), lc.q), y = d_from_residue(csub(csub(yi, lc.q2), lc.q), lc.q);
        d_gs_bfly(x, y, itw[vt], q, qinv, thr);
        *reinterpret_cast<double2 *>(sm + tile_pair(vt)) = double2{x, y};
    }
}
template <int LO>  // LO = 1, 4, 7
HD void dinv_phase_mid(int tid, u64 *smem, const LimbConst &lc, u32 hb) {
    char *sm = reinterpret_cast<char *>(smem);
    const double q = lc.qd, qinv = lc.qinv;
    const u32 thr = lc.dthr;
#pragma unroll 1
    for (int r = 0; r < 4; r++) {
        const u32 vt = tid + kNttThreads * r;
        const u32 base = tile_base<LO>(vt);
        double a[8];
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = dtile(sm, tile_off<LO>(base, k));
        d_gs_radix8(a, lc.ditw, (hb << (10 - LO)) + (vt >> LO), q, qinv, thr);
#pragma unroll
        for (int k = 0; k < 8; k++) dtile(sm, tile_off<LO>(base, k)) = a[k];
    }
}
// last pass (address bits 10..12) with N^-1 folded into the final stage; st(index, residue in [0, q))
template <class Store>
HD void dinv_phase_last(int tid, const u64 *smem, const LimbConst &lc, Store st) {
    char *sm = const_cast<char *>(reinterpret_cast<const char *>(smem));
    const double q = lc.qd, qinv = lc.qinv;
    const u32 thr = lc.dthr;
    const double *__restrict__ itw = lc.ditw;
#pragma unroll 1
    for (int r = 0; r < 4; r++) {
        const u32 vt = tid + kNttThreads * r;
        const u32 base = tile_base<10>(vt);
        double a[8];
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = dtile(sm, tile_off<10>(base, k));
#pragma unroll
        for (int k = 0; k < 4; k++) d_gs_bfly(a[2 * k], a[2 * k + 1], itw[4 + k], q, qinv, thr);
        const double w0 = itw[2], w1 = itw[3];
        d_gs_bfly(a[0], a[2], w0, q, qinv, thr); d_gs_bfly(a[1], a[3], w0, q, qinv, thr);
        d_gs_bfly(a[4], a[6], w1, q, qinv, thr); d_gs_bfly(a[5], a[7], w1, q, qinv, thr);
#pragma unroll
        for (int k = 0; k < 4; k++) {   // x' = (x + y) / N, y' = (x - y) w / N
            const double s = d_renorm(fadd(a[k], a[k + 4]), q, qinv, thr), d = d_renorm(fsub(a[k], a[k + 4]), q, qinv, thr);
            a[k] = d_renorm(d_mulmod(s, lc.d_ninv, q, qinv), q, qinv, thr);
            a[k + 4] = d_renorm(d_mulmod(d, lc.d_wninv, q, qinv), q, qinv, thr);
        }
#pragma unroll
        for (int k = 0; k < 8; k++) st((u32)(k << 10) | vt, d_to_residue(a[k], lc.q));
    }
}
// rings above one tile: an ordinary radix-8 pass with twiddles offset by hb; canonical residues (the tail pass takes
// anything below 2q)
template <class Store>
HD void dinv_phase_last_lazy(int tid, const u64 *smem, const LimbConst &lc, u32 hb, Store st) {
    char *sm = const_cast<char *>(reinterpret_cast<const char *>(smem));
    const double q = lc.qd, qinv = lc.qinv;
    const u32 thr = lc.dthr;
#pragma unroll 1
    for (int r = 0; r < 4; r++) {
        const u32 vt = tid + kNttThreads * r;
        const u32 base = tile_base<10>(vt);
        double a[8];
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = dtile(sm, tile_off<10>(base, k));
        d_gs_radix8(a, lc.ditw, hb, q, qinv, thr);
#pragma unroll
        for (int k = 0; k < 8; k++) st((u32)(k << 10) | vt, d_to_residue(d_renorm(a[k], q, qinv, thr), lc.q));
    }
}

// ---- inverse NTT phases (bit-reversed in, natural out), Gentleman-Sande ----
// Inputs to a pass are < 2q; X' = X + Y doubles the bound per stage, Y' = (X - Y + c)*w with
// c = (bound of Y) so the difference stays non-negative.  Every pass ends with quick_reduce.
// gs_bfly: truncated quotient, y' < 3q (every use below keeps c >= the bound of y); gs_bfly_exact: y' < 2q.
HD void gs_bfly(u64 &x, u64 &y, const tw_t &w, u64 nq, u64 c) {
    u64 t = x + c - y;
    x = x + y;
    y = bfly_mul(t, w.w, w.ws, nq);
}
HD void gs_bfly_exact(u64 &x, u64 &y, const tw_t &w, u64 nq, u64 c) {
    u64 t = x + c - y;
    x = x + y;
    y = shoup_lazy_nq(t, w.w, w.ws, nq);
}
// radix-8 GS pass over address bits LO, LO+1, LO+2 (a[k]: k bit j <-> address bit LO+j).
// Stage for address bit b uses m = N >> (b+1) groups: twiddle index m + (x >> (b+1)).
HD void gs_radix8(u64 (&a)[8], const tw_t *__restrict__ itw, u32 t2, u64 nq, u64 q2) {
    // t2 = index for the LAST of the three stages (bit LO+2): (N >> (LO+3)) + high
    // first stage, bit LO: index 4*t2 + (k >> 1), pairs (k, k+1) for even k; bounds 2q -> 4q
#pragma unroll
    for (int k = 0; k < 4; k++) gs_bfly(a[2 * k], a[2 * k + 1], itw[4 * t2 + k], nq, q2);
    // second stage, bit LO+1: x elements < 4q, y elements < 2q (fresh products)
    tw_t w0 = itw[2 * t2], w1 = itw[2 * t2 + 1];
    gs_bfly(a[0], a[2], w0, nq, 2 * q2); gs_bfly(a[1], a[3], w0, nq, 2 * q2);
    gs_bfly(a[4], a[6], w1, nq, 2 * q2); gs_bfly(a[5], a[7], w1, nq, 2 * q2);
    // third stage, bit LO+2: bound 8q
    tw_t w = itw[t2];
#pragma unroll
    for (int k = 0; k < 4; k++) gs_bfly(a[k], a[k + 4], w, nq, 4 * q2);
}
// Pass E': address bit 0 from the loader (pairs), result quick-reduced into shared memory.
// ld(pair index, &x, &y) must deliver values < 4q.
template <class Load>
HD void inv_phase_e(int tid, u64 *smem, const LimbConst &lc, u32 hb, Load ld) {
    char *sm = reinterpret_cast<char *>(smem);
    const u64 nq = lc.nq, q2 = lc.q2;
#pragma unroll 2
    for (int r = 0; r < 16; r++) {
        u32 vt = tid + kNttThreads * r;
        u64 x, y;
        ld(vt, x, y);
        gs_bfly_exact(x, y, lc.itw[((size_t)hb << (kLogN - 1)) + vt], nq, 2 * q2);   // y is stored unreduced: < 2q
        *reinterpret_cast<u64x2 *>(sm + tile_pair(vt)) = u64x2{quick_reduce(x, lc), y};
    }
}
template <int LO>  // LO = 1, 4, 7
HD void inv_phase_mid(int tid, u64 *smem, const LimbConst &lc, u32 hb) {
    char *sm = reinterpret_cast<char *>(smem);
    const u64 nq = lc.nq, q2 = lc.q2;
#pragma unroll 1
    for (int r = 0; r < 4; r++) {
        u32 vt = tid + kNttThreads * r;
        u32 base = tile_base<LO>(vt);
        u64 a[8];
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = tile(sm, tile_off<LO>(base, k));
        gs_radix8(a, lc.itw, (hb << (10 - LO)) + (vt >> LO), nq, q2);
        quick_reduce8(a, lc);
#pragma unroll
        for (int k = 0; k < 8; k++) tile(sm, tile_off<LO>(base, k)) = a[k];
    }
}
// Last pass: address bits 10, 11, 12 with N^-1 folded into the final stage; st(index, value in [0,q)).
template <class Store>
HD void inv_phase_last(int tid, const u64 *smem, const LimbConst &lc, Store st) {
    const char *sm = reinterpret_cast<const char *>(smem);
    const u64 q = lc.q, q2 = lc.q2, nq = lc.nq;
#pragma unroll 1
    for (int r = 0; r < 4; r++) {
        u32 vt = tid + kNttThreads * r;
        u32 base = tile_base<10>(vt);
        u64 a[8];
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = tile(sm, tile_off<10>(base, k));
#pragma unroll
        for (int k = 0; k < 4; k++) gs_bfly(a[2 * k], a[2 * k + 1], lc.itw[4 + k], nq, q2);
        tw_t w0 = lc.itw[2], w1 = lc.itw[3];
        gs_bfly(a[0], a[2], w0, nq, 2 * q2); gs_bfly(a[1], a[3], w0, nq, 2 * q2);
        gs_bfly(a[4], a[6], w1, nq, 2 * q2); gs_bfly(a[5], a[7], w1, nq, 2 * q2);
#pragma unroll
        for (int k = 0; k < 4; k++) {  // x' = (x+y)/N, y' = (x-y)*w/N ; inputs < 8q
            u64 x = a[k], y = a[k + 4];
            u64 s = x + y, d = x + 4 * q2 - y;
            a[k] = csub(shoup_lazy_nq(s, lc.ninv, lc.ninv_s, nq), q);
            a[k + 4] = csub(shoup_lazy_nq(d, lc.wninv, lc.wninv_s, nq), q);
        }
#pragma unroll
        for (int k = 0; k < 8; k++) st((u32)(k << 10) | vt, a[k]);
    }
}

// Ring dimensions above one tile: the tile's last pass is an ordinary radix-8 pass (address bits 10-12 of the
// tile, twiddles offset by hb) whose lazily reduced results (< 2q) go back to global memory; inv_tail finishes.
template <class Store>
HD void inv_phase_last_lazy(int tid, const u64 *smem, const LimbConst &lc, u32 hb, Store st) {
    const char *sm = reinterpret_cast<const char *>(smem);
    const u64 q2 = lc.q2, nq = lc.nq;
#pragma unroll 1
    for (int r = 0; r < 4; r++) {
        u32 vt = tid + kNttThreads * r;
        u32 base = tile_base<10>(vt);
        u64 a[8];
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = tile(sm, tile_off<10>(base, k));
        gs_radix8(a, lc.itw, hb, nq, q2);
        quick_reduce8(a, lc);
#pragma unroll
        for (int k = 0; k < 8; k++) st((u32)(k << 10) | vt, a[k]);
    }
}
// The d outermost Gentleman-Sande stages (address bits 13..13+d-1) with N^-1 folded into the last one, on the
// 2^d values a[blk] that share one tile-local index.  Inputs < 2q, outputs in [0, q).
template <int D>
HD void inv_tail(u64 (&a)[1 << D], const LimbConst &lc) {
    const u64 q = lc.q, q2 = lc.q2, nq = lc.nq;
    if (D == 1) {
        u64 s = a[0] + a[1], d = a[0] + q2 - a[1];
        a[0] = csub(shoup_lazy_nq(s, lc.ninv, lc.ninv_s, nq), q);
        a[1] = csub(shoup_lazy_nq(d, lc.wninv, lc.wninv_s, nq), q);
    } else if (D == 2) {
        gs_bfly_exact(a[0], a[1], lc.itw[2], nq, q2);   // address bit 13: a[0], a[2] < 4q; a[1], a[3] < 2q
        gs_bfly_exact(a[2], a[3], lc.itw[3], nq, q2);
        u64 s0 = a[0] + a[2], d0 = a[0] + 2 * q2 - a[2];   // address bit 14 with N^-1
        u64 s1 = a[1] + a[3], d1 = a[1] + q2 - a[3];
        a[0] = csub(shoup_lazy_nq(s0, lc.ninv, lc.ninv_s, nq), q);
        a[2] = csub(shoup_lazy_nq(d0, lc.wninv, lc.wninv_s, nq), q);
        a[1] = csub(shoup_lazy_nq(s1, lc.ninv, lc.ninv_s, nq), q);
        a[3] = csub(shoup_lazy_nq(d1, lc.wninv, lc.wninv_s, nq), q);
    }
}
// The d outermost Cooley-Tukey stages seen from tile blk: value at tile-local index i after those stages, from
// the 2^d inputs in(i + k*8192) (each < q).  Redundant across the tiles of a polynomial (one extra modmul per
// coefficient for d = 1, three for d = 2) but keeps the whole transform in one kernel with no global round trip.
// Result < (1 + 2d) q: LimbConst::fwd_mask is scheduled from that bound (tables.cpp).
template <int D, class Inner>
struct HeadLoad {
    Inner in; const LimbConst *lc; u32 blk;
    HD u64 operator()(u32 i) const {
        if (D == 0) return in(i);
        const u64 nq = lc->nq, q2 = lc->q2;
        const tw_t *__restrict__ tw = lc->tw;
        if (D == 1) {
            u64 x = in(i), y = in(i + kN);
            u64 t = shoup_lazy_nq(y, tw[1].w, tw[1].ws, nq);
            return blk ? x + q2 - t : x + t;
        }
        u64 x0 = in(i), x1 = in(i + kN), x2 = in(i + 2 * kN), x3 = in(i + 3 * kN);
        u64 t0 = shoup_lazy_nq(x2, tw[1].w, tw[1].ws, nq), t1 = shoup_lazy_nq(x3, tw[1].w, tw[1].ws, nq);
        u64 s0, s1;
        if (blk < 2) { s0 = x0 + t0; s1 = x1 + t1; } else { s0 = x0 + q2 - t0; s1 = x1 + q2 - t1; }
        tw_t w = tw[2 + (blk >> 1)];
        u64 t = shoup_lazy_nq(s1, w.w, w.ws, nq);
        return (blk & 1u) ? s0 + q2 - t : s0 + t;
    }
};

// ------------------------------------------------------------------------------------------
// sampler: ChaCha20 in counter mode (the RFC 8439 block function, 20 rounds, 256-bit key).  Every draw of the
// path — ternary v and Gaussian e0, e1 of Encrypt, ternary s, Gaussian e and uniform a of KeyGen — is a fixed
// function of (key, nonce, block counter), so any thread can produce any coefficient's randomness on its own
// and the CPU oracle reproduces the stream word for word under a test key (DESIGN.md §4.2 has the layout).
// The reference draws the same quantities from PALISADE's Blake2-based PRNG seeded from OS entropy
// (ckks.cpp:46,81,95); keys here come from getrandom(2), one per call, secret and public material under
// independent keys (engine.cpp fresh_key).
// ------------------------------------------------------------------------------------------
HD u32 rotl32(u32 x, int n) {
#ifdef __CUDA_ARCH__
    return __funnelshift_l(x, x, n);
#else
    return (x << n) | (x >> (32 - n));
#endif
}
#define SHELFI_CHACHA_QR(a, b, c, d)                  \
    a += b; d ^= a; d = rotl32(d, 16);                \
    c += d; b ^= c; b = rotl32(b, 12);                \
    a += b; d ^= a; d = rotl32(d, 8);                 \
    c += d; b ^= c; b = rotl32(b, 7);
struct SamplerKey { u32 k[8]; };
#ifndef SHELFI_CHACHA_ROUNDS
#define SHELFI_CHACHA_ROUNDS 20   // RFC 8439.  (Reduced-round builds exist only to measure what the generator costs.)
#endif
// state = "expand 32-byte k" | key | counter | nonce[3]; out = state + 20 rounds(state)
HD void chacha20_block(const SamplerKey &key, u32 counter, u32 n0, u32 n1, u32 n2, u32 (&out)[16]) {
    const u32 in[16] = {0x61707865u, 0x3320646eu, 0x79622d32u, 0x6b206574u, key.k[0], key.k[1], key.k[2], key.k[3],
                        key.k[4], key.k[5], key.k[6], key.k[7], counter, n0, n1, n2};
    u32 x[16];
#pragma unroll
    for (int i = 0; i < 16; i++) x[i] = in[i];
#pragma unroll
    for (int r = 0; r < SHELFI_CHACHA_ROUNDS / 2; r++) {
        SHELFI_CHACHA_QR(x[0], x[4], x[8], x[12])
        SHELFI_CHACHA_QR(x[1], x[5], x[9], x[13])
        SHELFI_CHACHA_QR(x[2], x[6], x[10], x[14])
        SHELFI_CHACHA_QR(x[3], x[7], x[11], x[15])
        SHELFI_CHACHA_QR(x[0], x[5], x[10], x[15])
        SHELFI_CHACHA_QR(x[1], x[6], x[11], x[12])
        SHELFI_CHACHA_QR(x[2], x[7], x[8], x[13])
        SHELFI_CHACHA_QR(x[3], x[4], x[9], x[14])
    }
#pragma unroll
    for (int i = 0; i < 16; i++) out[i] = x[i] + in[i];
}
// Test keys: shelfi_set_sampler_seed(seed) makes the engine deterministic; the 64-bit seed is laid into a key per
// domain so that the CPU checker under oracle/ draws the same stream.  Never used for real data.
enum { kKeyDomainEncrypt = 0, kKeyDomainKeySecret = 1, kKeyDomainKeyPublic = 2, kKeyDomainDecode = 3 };
HD SamplerKey seed_key(u64 seed, u32 domain) {
    SamplerKey k;
    k.k[0] = (u32)seed; k.k[1] = (u32)(seed >> 32);
    k.k[2] = 0x4c454853u; k.k[3] = 0x422d4946u; k.k[4] = 0x20303032u; k.k[5] = 0x64656573u;   // "SHELFI-B200 seed"
    k.k[6] = domain; k.k[7] = 0;
    return k;
}
// nonce word 2 names the stream; nonce words 0, 1 hold the ciphertext index (Encrypt) or (limb, attempt) (KeyGen a)
enum { kStreamGauss = 0, kStreamTernary = 1, kStreamKeyS = 2, kStreamKeyE = 3, kStreamKeyA = 4, kStreamKeyTag = 5,
       kStreamDecodeNoise = 6, kStreamGaussRetry = 7, kStreamTernaryRetry = 8, kStreamKeySRetry = 9, kStreamKeyERetry = 10 };
// Both small distributions are sampled EXACTLY (to the precision of the CDF table) from as few random bits as that
// allows; the rare draws that need more take them from a retry stream of their own, so the common path stays
// branch-free and a ciphertext costs 1280 ChaCha20 blocks instead of the 2560 of one 64-bit word per Gaussian and one
// 32-bit word per ternary value.
//   Gaussian(sigma): CDF inversion.  cdf63[k] = floor(2^63 P(|X| <= k)), k < 32.  A draw is the 63-bit number
//     U = (h >> 1) * 2^32 + l with h, l uniform 32-bit words; magnitude = #{k : U >= cdf63[k]} (at most 31), sign = h & 1.
//     The top 31 bits decide unless they equal the top 31 bits of a table entry (probability < 2^-26 per draw); only
//     then is l drawn.  The result is the same as if l had always been drawn.
//   ternary: eight 2-bit attempts from a 16-bit half word, low bits first: 0, 1, 2 -> -1, 0, +1, 3 -> next attempt;
//     all eight rejected (probability 2^-16): the attempts continue in the retry stream.  Uniform on {-1, 0, 1}.
struct GaussTable {
    const u32 *hi31;   // cdf63[k] >> 32
    const u64 *cdf63;
};
HD int cdf_rank32(const u32 *t, u32 m) {      // #{k < 31 : m >= t[k]} for a sorted table
    int lo = 0;
#pragma unroll
    for (int step = 16; step >= 1; step >>= 1)
        if (m >= t[lo + step - 1]) lo += step;
    return lo;
}
HD int cdf_rank64(const u64 *t, u64 u) {
    int lo = 0;
#pragma unroll
    for (int step = 16; step >= 1; step >>= 1)
        if (u >= t[lo + step - 1]) lo += step;
    return lo;
}
template <class Retry>   // retry() -> the 32-bit word l of this draw
HD int gauss_lazy(const GaussTable &g, u32 h, Retry retry) {
    const u32 m = h >> 1;
    int lo = cdf_rank32(g.hi31, m);
    if (lo > 0 && g.hi31[lo - 1] == m) lo = cdf_rank64(g.cdf63, ((u64)m << 32) | retry());
    return (h & 1u) ? -lo : lo;
}
HD int first_set(u32 x) {   // index of the lowest set bit, x != 0
#ifdef __CUDA_ARCH__
    return __ffs((int)x) - 1;
#else
    return __builtin_ctz(x);
#endif
}
template <class Retry>   // retry() -> a value in {-1, 0, 1} from the retry stream
HD int ternary16(u32 h16, Retry retry) {
    const u32 ok = ~(h16 & (h16 >> 1)) & 0x5555u;   // bit 2a set iff attempt a is not 3
    if (ok == 0) return retry();
    return (int)((h16 >> first_set(ok)) & 3u) - 1;
}
// the retry streams: one block per coefficient, counter = coefficient index
HD_COLD u32 retry_word(const SamplerKey &key, u64 ct, u32 stream, u32 j, u32 which) {
    u32 o[16];
    chacha20_block(key, j, (u32)ct, (u32)(ct >> 32), stream, o);
    u32 r = 0;
#pragma unroll
    for (int i = 0; i < 16; i++)
        if ((u32)i == which) r = o[i];
    return r;
}
HD_COLD int retry_ternary(const SamplerKey &key, u64 ct, u32 stream, u32 j) {
    u32 o[16];
    chacha20_block(key, j, (u32)ct, (u32)(ct >> 32), stream, o);
    for (int i = 0; i < 16; i++)
        for (int a = 0; a < 16; a++) {
            const u32 c = (o[i] >> (2 * a)) & 3u;
            if (c != 3u) return (int)c - 1;
        }
    return 0;   // 256 rejections in a row: probability 2^-512
}
struct GaussRetry {
    const SamplerKey *key; u64 ct; u32 stream, j, which;
    HD u32 operator()() const { return retry_word(*key, ct, stream, j, which); }
};
struct TernaryRetry {
    const SamplerKey *key; u64 ct; u32 stream, j;
    HD int operator()() const { return retry_ternary(*key, ct, stream, j); }
};

// ------------------------------------------------------------------------------------------
// canonical-embedding FFT (FFTSpecial / FFTSpecialInv), FP64, no fused multiply-add: llround()
// results must match the reference's unfused butterflies bit for bit.
// ------------------------------------------------------------------------------------------
struct __attribute__((aligned(16))) cplx { double re, im; };
HD u32 bitrev32(u32 x, int bits) {
#ifdef __CUDA_ARCH__
    return bits ? __brev(x) >> (32 - bits) : 0u;
#else
    u32 r = 0;
    for (int i = 0; i < bits; i++) r |= ((x >> i) & 1u) << (bits - 1 - i);
    return r;
#endif
}
// FFT tile swizzle (16-byte complex elements): XOR index bits 2..0 with bits 5..3 so that both "8 consecutive
// lanes read 8 consecutive elements" and "8 consecutive lanes read elements 8 apart" are conflict-free.
HD u32 zsw(u32 i) { return i ^ ((i >> 3) & 7u); }
HD void dif_bfly(cplx &a, cplx &b, const cplx &w) {  // FFTSpecialInv: u = a + b, v = (a - b) * w
    double dr = fsub(a.re, b.re), di = fsub(a.im, b.im);
    a.re = fadd(a.re, b.re); a.im = fadd(a.im, b.im);
    b.re = fsub(fmul(dr, w.re), fmul(di, w.im));
    b.im = fadd(fmul(dr, w.im), fmul(di, w.re));
}
HD void dit_bfly(cplx &a, cplx &b, const cplx &w) {  // FFTSpecial: v = b * w, a' = a + v, b' = a - v
    double vr = fsub(fmul(b.re, w.re), fmul(b.im, w.im));
    double vi = fadd(fmul(b.re, w.im), fmul(b.im, w.re));
    b.re = fsub(a.re, vr); b.im = fsub(a.im, vi);
    a.re = fadd(a.re, vr); a.im = fadd(a.im, vi);
}
// One stage of FFTSpecialInv (decimation in frequency): butterflies j of block length `len`.
// twi = per-stage table, entry (len/2 - 1) + j  holds ksi[(lenq - rot[j] % lenq) * M/lenq].
HD void fft_inv_stage(int tid, int nthreads, cplx *z, u32 size, u32 len, const cplx *__restrict__ twi) {
    u32 lenh = len >> 1;
    for (u32 b = tid; b < (size >> 1); b += nthreads) {
        u32 j = b & (lenh - 1), i = (b / lenh) * len;
        cplx w = twi[lenh - 1 + j];
        cplx u = z[zsw(i + j)], v = z[zsw(i + j + lenh)];
        dif_bfly(u, v, w);
        z[zsw(i + j)] = u; z[zsw(i + j + lenh)] = v;
    }
}
// One stage of FFTSpecial (decimation in time, input already bit-reversed).
HD void fft_fwd_stage(int tid, int nthreads, cplx *z, u32 size, u32 len, const cplx *__restrict__ twf) {
    u32 lenh = len >> 1;
    for (u32 b = tid; b < (size >> 1); b += nthreads) {
        u32 j = b & (lenh - 1), i = (b / lenh) * len;
        cplx w = twf[lenh - 1 + j];
        cplx u = z[zsw(i + j)], v = z[zsw(i + j + lenh)];
        dit_bfly(u, v, w);
        z[zsw(i + j)] = u; z[zsw(i + j + lenh)] = v;
    }
}
// Full packing (4096 slots, the reference's default): the same butterflies grouped into four radix-8 register
// passes (3 stages each, 8 complex values per thread, 2 rounds of 256 threads) — 4 shared-memory round trips
// and barriers instead of 12.  LO = lowest of the three index bits held in registers.
constexpr u32 kFftFast = 4096;
template <int LO>  // LO = 9, 6, 3, 0 in that order
HD void fft_inv_pass8(int tid, cplx *z, const cplx *__restrict__ twi) {
#pragma unroll 1
    for (int r = 0; r < 2; r++) {
        u32 vt = tid + kFftThreads * r;
        u32 low = vt & ((1u << LO) - 1u), high = vt >> LO, base = (high << (LO + 3)) | low;
        cplx a[8];
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = z[zsw(base | ((u32)k << LO))];
        {   // stage on bit LO+2: lenh = 2^(LO+2), j = (k & 3) << LO | low
            const cplx *t = twi + ((1u << (LO + 2)) - 1u) + low;
#pragma unroll
            for (int k = 0; k < 4; k++) dif_bfly(a[k], a[k + 4], t[(u32)k << LO]);
        }
        {   // bit LO+1: lenh = 2^(LO+1), j = (k & 1) << LO | low
            const cplx *t = twi + ((1u << (LO + 1)) - 1u) + low;
            cplx w0 = t[0], w1 = t[1u << LO];
            dif_bfly(a[0], a[2], w0); dif_bfly(a[1], a[3], w1);
            dif_bfly(a[4], a[6], w0); dif_bfly(a[5], a[7], w1);
        }
        {   // bit LO: lenh = 2^LO, j = low
            cplx w = twi[((1u << LO) - 1u) + low];
#pragma unroll
            for (int k = 0; k < 4; k++) dif_bfly(a[2 * k], a[2 * k + 1], w);
        }
#pragma unroll
        for (int k = 0; k < 8; k++) z[zsw(base | ((u32)k << LO))] = a[k];
    }
}
template <int LO>  // LO = 0, 3, 6, 9 in that order
HD void fft_fwd_pass8(int tid, cplx *z, const cplx *__restrict__ twf) {
#pragma unroll 1
    for (int r = 0; r < 2; r++) {
        u32 vt = tid + kFftThreads * r;
        u32 low = vt & ((1u << LO) - 1u), high = vt >> LO, base = (high << (LO + 3)) | low;
        cplx a[8];
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = z[zsw(base | ((u32)k << LO))];
        {   // bit LO
            cplx w = twf[((1u << LO) - 1u) + low];
#pragma unroll
            for (int k = 0; k < 4; k++) dit_bfly(a[2 * k], a[2 * k + 1], w);
        }
        {   // bit LO+1
            const cplx *t = twf + ((1u << (LO + 1)) - 1u) + low;
            cplx w0 = t[0], w1 = t[1u << LO];
            dit_bfly(a[0], a[2], w0); dit_bfly(a[1], a[3], w1);
            dit_bfly(a[4], a[6], w0); dit_bfly(a[5], a[7], w1);
        }
        {   // bit LO+2
            const cplx *t = twf + ((1u << (LO + 2)) - 1u) + low;
#pragma unroll
            for (int k = 0; k < 4; k++) dit_bfly(a[k], a[k + 4], t[(u32)k << LO]);
        }
#pragma unroll
        for (int k = 0; k < 8; k++) z[zsw(base | ((u32)k << LO))] = a[k];
    }
}

// ------------------------------------------------------------------------------------------
// decode helpers: two-limb CRT to a centred 128-bit integer, correctly rounded to double
// ------------------------------------------------------------------------------------------
struct Crt2Const {
    u64 q0, q1;
    u64 q0inv, q0inv_s;    // q0^-1 mod q1 and Shoup companion
    u64 mu1;               // floor(2^64 / q1)
    u64 Q_hi, Q_lo;        // Q = q0*q1
    u64 Qh_hi, Qh_lo;      // floor(Q / 2)
};
HD void mul64x64(u64 a, u64 b, u64 &hi, u64 &lo) { lo = a * b; hi = mulhi(a, b); }
// sign/magnitude of the centred representative of x = a0 (mod q0), a1 (mod q1)
HD void crt2_centered(u64 a0, u64 a1, const Crt2Const &c, bool &neg, u64 &mag_hi, u64 &mag_lo) {
    u64 a0m = csub(a0 - mulhi(a0, c.mu1) * c.q1, c.q1);       // a0 mod q1
    u64 diff = a1 >= a0m ? a1 - a0m : a1 + c.q1 - a0m;
    u64 d = csub(shoup_lazy(diff, c.q0inv, c.q0inv_s, c.q1), c.q1);
    u64 hi, lo;
    mul64x64(c.q0, d, hi, lo);
    u64 xlo = lo + a0;
    u64 xhi = hi + (xlo < lo ? 1 : 0);
    bool gt = xhi > c.Qh_hi || (xhi == c.Qh_hi && xlo > c.Qh_lo);  // x > floor(Q/2) -> x - Q
    if (gt) {
        u64 rlo = c.Q_lo - xlo;
        u64 rhi = c.Q_hi - xhi - (c.Q_lo < xlo ? 1 : 0);
        neg = true; mag_hi = rhi; mag_lo = rlo;
    } else {
        neg = false; mag_hi = xhi; mag_lo = xlo;
    }
}
HD int clz64(u64 x) {
#ifdef __CUDA_ARCH__
    return __clzll((long long)x);
#else
    return x ? __builtin_clzll(x) : 64;
#endif
}
// (double) of a signed 128-bit integer, round to nearest even in ONE rounding (same as gcc's __floattidf)
HD double i128_to_double(bool neg, u64 hi, u64 lo) {
    double r;
    if (hi == 0) {
        r = (double)lo;
    } else {
        int s = clz64(hi);                         // 0..63
        u64 top = s ? (hi << s) | (lo >> (64 - s)) : hi;
        u64 rest = s ? (lo << s) : lo;
        if (rest) top |= 1ull;                     // sticky bit below the 64-bit window
        r = ldexp((double)top, 64 - s);
    }
    return neg ? -r : r;
}


// ---- chains deeper than two limbs: Garner mixed-radix digits + Horner into L 64-bit words ----
// ([PAL] DCRTPoly::CRTInterpolate yields the same integer in [0, Q); every q_i < 2^60 so Q fits L words)
struct CrtNConst {
    u32 limbs, pad;
    u64 q[kMaxLimbs], mu[kMaxLimbs];                    // mu = floor(2^64 / q)
    u64 inv[kMaxLimbs][kMaxLimbs], inv_s[kMaxLimbs][kMaxLimbs];   // inv[i][j] = q_j^-1 mod q_i (j < i) + Shoup
    u64 Q[kMaxLimbs], Qh[kMaxLimbs];                    // Q and floor(Q / 2), little-endian words
};
// sign / magnitude (L little-endian words) of the centred representative of x = a[i] (mod q_i)
template <int L>
HD void crtn_centered(const u64 (&a)[L], const CrtNConst &c, bool &neg, u64 (&mag)[L]) {
    u64 v[L];
#pragma unroll
    for (int i = 0; i < L; i++) {
        const u64 qi = c.q[i];
        u64 t = a[i];
#pragma unroll
        for (int j = 0; j < i; j++) {
            u64 vj = csub(v[j] - mulhi(v[j], c.mu[i]) * qi, qi);   // v_j mod q_i
            u64 diff = t >= vj ? t - vj : t + qi - vj;
            t = csub(shoup_lazy(diff, c.inv[i][j], c.inv_s[i][j], qi), qi);
        }
        v[i] = t;
    }
    u64 x[L];
#pragma unroll
    for (int i = 0; i < L; i++) x[i] = 0;
    x[0] = v[L - 1];
#pragma unroll
    for (int i = L - 2; i >= 0; i--) {   // x = x * q_i + v_i ; x has L-1-i words before the step
        const u64 qi = c.q[i];
        u64 carry = v[i];
#pragma unroll
        for (int k = 0; k < L - 1 - i; k++) {
            u64 hi, lo;
            mul64x64(x[k], qi, hi, lo);
            lo += carry;
            hi += lo < carry ? 1 : 0;
            x[k] = lo; carry = hi;
        }
        x[L - 1 - i] = carry;
    }
    bool gt = false, decided = false;   // x > floor(Q/2) ?
#pragma unroll
    for (int k = L - 1; k >= 0; k--) {
        if (!decided && x[k] != c.Qh[k]) { gt = x[k] > c.Qh[k]; decided = true; }
    }
    if (gt) {
        u64 borrow = 0;
#pragma unroll
        for (int k = 0; k < L; k++) {
            u64 qk = c.Q[k], xk = x[k], d = qk - xk - borrow;
            borrow = (qk < xk || (qk == xk && borrow)) ? 1 : 0;
            mag[k] = d;
        }
    } else {
#pragma unroll
        for (int k = 0; k < L; k++) mag[k] = x[k];
    }
    neg = gt;
}
// (double) of a sign/magnitude integer of L words, ONE round-to-nearest-even (i128_to_double for any width)
template <int L>
HD double words_to_double(bool neg, const u64 (&m)[L]) {
    u64 hi = 0, lo = 0;
    int top = 0;
    bool sticky = false;
#pragma unroll
    for (int i = L - 1; i >= 0; i--) {
        if (hi == 0) {
            if (m[i]) { hi = m[i]; top = i; lo = i > 0 ? m[i > 0 ? i - 1 : 0] : 0; }
        } else if (i < top - 1) {
            sticky = sticky || m[i] != 0;
        }
    }
    double r;
    if (top == 0) {
        r = (double)hi;
    } else {
        int s = clz64(hi);
        u64 win = s ? (hi << s) | (lo >> (64 - s)) : hi;
        u64 rest = s ? (lo << s) : lo;
        if (rest || sticky) win |= 1ull;
        r = ldexp((double)win, 64 * top - s);
    }
    return neg ? -r : r;
}

}  // namespace shelfi

// ==========================================================================================
// kernel argument blocks and the remaining per-thread phases
// ==========================================================================================
namespace shelfi {

constexpr int kMaxLearnersPerLaunch = 16;

struct SamplerConst {
    SamplerKey key;      // ChaCha20 key of this call (one learner's vector)
    u64 ct_index0;       // global index of the first ciphertext of this launch (shards draw the oracle's stream)
    u64 gauss_cdf[32];   // cdf63: floor(2^63 * P(|X| <= k)), k = 0..31
};

struct EncodeConst {
    u32 slots, gap, log_slots, batch;  // context slots, N/2/slots, log2(slots), ctor batchSize (chunk length)
    double pow_p;                      // scaling factor (q_last as double)
    double inv_size;                   // 1 / slots (exact power of two)
    const cplx *tw_inv;                // slots-1 per-stage twiddles of FFTSpecialInv
    const cplx *tw_fwd;                // ... of FFTSpecial
};

constexpr int kMaxPeerOuts = 7;   // further destinations of a decoded value: the other GPUs of one box
struct DecodeConst {
    Crt2Const crt;
    double pre;     // scalingFactor^-1 * 2^p
    double pow_p;   // 2^-p
    // decrypt-gather: every decoded value is also stored to peer[p][index] (peer-mapped memory of the other GPUs,
    // already offset to this rank's slice), so the decoded slices are assembled on every GPU by the decode kernel's
    // own stores over NVLink instead of by a collective that re-reads them.  Not combined with a selection list.
    double *peer[kMaxPeerOuts];
    u32 n_peer;
};

struct WavgArgs {
    const u64 *in[kMaxLearnersPerLaunch];
    u64 c[kMaxLearnersPerLaunch][kMaxLimbs];  // EvalMult constant reduced into every limb
    u32 k;
    u32 accumulate;  // add the values already in `out` (chained launches for > kMaxLearnersPerLaunch learners)
};

struct LimbScalars { u64 c[kMaxLimbs]; };

// plaintext remainder of a selectively encrypted model: out[i] = sum_k w_k * x_k[i] (unfused, in learner order)
struct PlainWavgArgs {
    const double *in[kMaxLearnersPerLaunch];
    double w[kMaxLearnersPerLaunch];
    u32 k;
    u32 accumulate;
};

// one launch may encrypt several learners' vectors (same length n): block b -> learner b / n_ct, ciphertext b % n_ct
constexpr int kMaxEncryptBatch = 16;
struct EncryptBatch {
    const u32 *sel;   // optional selection shared by the learners: learner l's vector is x[l][sel[i]]
    const double *x[kMaxEncryptBatch];
    u64 *ct[kMaxEncryptBatch];
    SamplerKey key[kMaxEncryptBatch];   // every learner's vector is encrypted under its own key
    u32 k, n_ct;
    u32 f32, pad;     // the x[] hold float32 values
};

// byte offsets of the 2L coefficient runs of a serialized vector<Ciphertext> (wire.h)
struct BlobRuns {
    u64 off0[2 * kMaxLimbs];   // ciphertext 0
    u64 off1[2 * kMaxLimbs];   // ciphertext 1 (absolute); ciphertext j >= 1: off1[r] + (j-1)*stride
    u64 stride;
    u32 runs;                  // 2L
    u32 n;                     // values per run (ring dimension)
};
HD u64 run_offset(const BlobRuns &br, u64 ct, u32 r) { return ct == 0 ? br.off0[r] : br.off1[r] + (ct - 1) * br.stride; }

// ---- encode: MakeCKKSPackedPlaintext up to the integer coefficients (ckks.cpp:80) ----
// phase 0: chunk of x -> z (zero padded)
// Selective encryption (benchmark_selection*.py encrypt a chosen subset of the parameters): with `sel` the chunk is
// x[sel[i]] instead of x[i], i.e. the gather is part of the load (and the scatter part of decode_store).
// `f32`: x holds float32 values (the reference's callers pass float32 state_dict arrays, code/benchmark.py:490-493,
// which pybind's array_t<double> caster widens on the host; here the widening — exact — is part of the load, so such
// an input crosses PCIe at 4 bytes per parameter).
HD double encode_value(const double *x, u64 idx, bool f32) {
    return f32 ? (double)reinterpret_cast<const float *>(x)[idx] : x[idx];
}
HD void encode_load(int tid, int nthreads, cplx *z, const EncodeConst &ec, const double *x, u64 n_valid,
                    const u32 *sel = nullptr, bool f32 = false) {
    if (ec.slots == kFftFast && !sel && nthreads == kFftThreads) {
        // full packing, the common case: all sixteen loads of a thread in flight before the first store
        double v[kFftFast / kFftThreads];
#pragma unroll
        for (int k = 0; k < (int)(kFftFast / kFftThreads); k++) {
            const u32 i = (u32)tid + (u32)kFftThreads * k;
            v[k] = i < n_valid ? encode_value(x, i, f32) : 0.0;
        }
#pragma unroll
        for (int k = 0; k < (int)(kFftFast / kFftThreads); k++) z[zsw((u32)tid + (u32)kFftThreads * k)] = cplx{v[k], 0.0};
        return;
    }
    for (u32 i = tid; i < ec.slots; i += nthreads) {
        z[zsw(i)].re = i < n_valid ? encode_value(x, sel ? sel[i] : i, f32) : 0.0;
        z[zsw(i)].im = 0.0;
    }
}
// phases 1..log_slots: fft_inv_stage(len = slots >> (p-1))
// final phase: bit reversal, 1/size, *Delta, llround — in place, z[i] reinterpreted as two int64
HD void encode_finalize(int tid, int nthreads, cplx *z, const EncodeConst &ec) {
    i64 *zi = reinterpret_cast<i64 *>(z);
    for (u32 i = tid; i < ec.slots; i += nthreads) {
        u32 j = bitrev32(i, (int)ec.log_slots);
        if (i > j) continue;
        u32 pi = zsw(i), pj = zsw(j);
        cplx a = z[pi], b = z[pj];
        i64 ar = llround(fmul(fmul(a.re, ec.inv_size), ec.pow_p)), ai = llround(fmul(fmul(a.im, ec.inv_size), ec.pow_p));
        i64 br = llround(fmul(fmul(b.re, ec.inv_size), ec.pow_p)), bi = llround(fmul(fmul(b.im, ec.inv_size), ec.pow_p));
        zi[2 * pj] = ar; zi[2 * pj + 1] = ai;
        zi[2 * pi] = br; zi[2 * pi + 1] = bi;
    }
}
// coefficient j of the plaintext polynomial from the finalized tile
HD i64 encode_coeff(const cplx *z, const EncodeConst &ec, u32 j) {
    const i64 *zi = reinterpret_cast<const i64 *>(z);
    u32 nh = kN >> 1;
    u32 half = j >= nh ? 1u : 0u, idx = j - half * nh;
    if (idx & (ec.gap - 1)) return 0;
    return zi[2 * zsw(idx / ec.gap) + half];
}
// two standard normals from four random words (Box-Muller on two 53-bit uniforms): the decode noise of [PAL] Decode
HD void normal_pair(u32 w0, u32 w1, u32 w2, u32 w3, double &z0, double &z1) {
    const double u1 = (double)(((((u64)w1 << 32) | w0) >> 11) + 1) * 1.1102230246251565e-16;   // (0, 1]
    const double u2 = (double)((((u64)w3 << 32) | w2) >> 11) * 1.1102230246251565e-16;          // [0, 1)
    const double r = sqrt(-2.0 * log(u1)), a = 6.283185307179586476925286766559 * u2;
    z0 = r * cos(a); z1 = r * sin(a);
}
// [PAL] Encrypt's draws for ciphertext `ct` (ckks.cpp:81,95), as counter-mode blocks:
//   ternary block b, nonce (ct_lo, ct_hi, kStreamTernary): v[32 b + 2 i] from the low half of word i, v[32 b + 2 i + 1]
//                     from its high half (ternary16), i < 16
//   Gaussian block b, nonce (ct_lo, ct_hi, kStreamGauss):  e0[8 b + i] from word 2 i, e1[8 b + i] from word 2 i + 1
//                     (gauss_lazy), i < 8
//   retries: block j (the coefficient), nonce (ct_lo, ct_hi, kStreamTernaryRetry / kStreamGaussRetry; the latter's
//            word 0 serves e0[j], word 1 e1[j])
// The small draws of one coefficient travel as one byte: bits 1..0 = v + 1, bits 7..2 = e1 + 32 (|e1| <= 31).
HD uint8_t pack_ve(int v, int e1) { return (uint8_t)((u32)(v + 1) | ((u32)(e1 + 32) << 2)); }
HD int ve_ternary(uint8_t b) { return (int)(b & 3u) - 1; }
HD int ve_gauss(uint8_t b) { return (int)(b >> 2) - 32; }
// 32 ternary values of block b, as v + 1 in eight words of four bytes each.  Draws whose eight attempts were all
// rejected are noted in a mask and settled afterwards, so the common path holds no call.
HD void sample_ternary_block(const SamplerKey &key, u64 ct, u32 b, u32 stream, u32 retry_stream, u32 (&packed)[8]) {
    u32 o[16];
    chacha20_block(key, b, (u32)ct, (u32)(ct >> 32), stream, o);
    u32 pending = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        u32 acc = 0;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const u32 idx = 4 * w + i;                       // coefficient 32 b + idx
            const u32 h16 = (o[idx >> 1] >> (16 * (idx & 1u))) & 0xFFFFu;
            const u32 ok = ~(h16 & (h16 >> 1)) & 0x5555u;    // ternary16 with the retry deferred
            if (ok == 0) pending |= 1u << idx;
            acc |= (ok ? ((h16 >> first_set(ok)) & 3u) : 1u) << (8 * i);
        }
        packed[w] = acc;
    }
    while (pending) {
        const u32 idx = (u32)first_set(pending);
        pending &= pending - 1;
        const u32 v1 = (u32)(retry_ternary(key, ct, retry_stream, 32 * b + idx) + 1);
#pragma unroll
        for (int w = 0; w < 8; w++)
            if ((u32)w == (idx >> 2)) packed[w] = (packed[w] & ~(0xFFu << (8 * (idx & 3u)))) | (v1 << (8 * (idx & 3u)));
    }
}
// one Gaussian draw from its first word; a tie of the top 31 bits with a table entry is reported, not settled
HD int gauss_fast(const GaussTable &g, u32 h, bool &tie) {
    const u32 m = h >> 1;
    const int lo = cdf_rank32(g.hi31, m);
    tie = lo > 0 && g.hi31[lo - 1] == m;
    return (h & 1u) ? -lo : lo;
}
HD_COLD int gauss_settle(const GaussTable &g, u32 h, const SamplerKey &key, u64 ct, u32 stream, u32 j, u32 which) {
    const int lo = cdf_rank64(g.cdf63, ((u64)(h >> 1) << 32) | retry_word(key, ct, stream, j, which));
    return (h & 1u) ? -lo : lo;
}
HD void sample_gauss_block(const SamplerKey &key, const GaussTable &g, u64 ct, u32 b, int (&e0)[8], int (&e1)[8]) {
    u32 o[16];
    chacha20_block(key, b, (u32)ct, (u32)(ct >> 32), kStreamGauss, o);
    u32 pending = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) {
        bool t0, t1;
        e0[i] = gauss_fast(g, o[2 * i], t0);
        e1[i] = gauss_fast(g, o[2 * i + 1], t1);
        pending |= (t0 ? 1u : 0u) << (2 * i) | (t1 ? 1u : 0u) << (2 * i + 1);
    }
    while (pending) {
        const u32 w = (u32)first_set(pending);
        pending &= pending - 1;
        u32 h = 0;
#pragma unroll
        for (int i = 0; i < 16; i++)
            if ((u32)i == w) h = o[i];
        const int v = gauss_settle(g, h, key, ct, kStreamGaussRetry, 8 * b + (w >> 1), w & 1u);
#pragma unroll
        for (int i = 0; i < 8; i++) {
            if ((u32)(2 * i) == w) e0[i] = v;
            if ((u32)(2 * i + 1) == w) e1[i] = v;
        }
    }
}
// [PAL] KeyGen's draws (ckks.cpp:46): s <- ternary and e <- Gaussian under the SECRET key-generation key, a <- uniform
// mod q_l per limb (EVALUATION form) and the key tag under the PUBLIC one.
//   s          : ternary blocks as above with nonce (0, 0, kStreamKeyS), retries (0, 0, kStreamKeySRetry)
//   e[16 b + i]: word i of block b, nonce (0, 0, kStreamKeyE), gauss_lazy with retries from word 0 of block 16 b + i,
//                nonce (0, 0, kStreamKeyERetry)
//   a_l[j]     : attempt t = 0, 1, ...: words (2 (j & 7), 2 (j & 7) + 1) of block j >> 3, nonce (l, t, kStreamKeyA),
//                top bits(q_l) bits of the 64-bit draw, accepted when below q_l
HD void keygen_error_block(const SamplerKey &key, const GaussTable &g, u32 b, int (&e)[16]) {
    u32 o[16];
    chacha20_block(key, b, 0, 0, kStreamKeyE, o);
    for (int i = 0; i < 16; i++) {
        bool tie;
        e[i] = gauss_fast(g, o[i], tie);
        if (tie) e[i] = gauss_settle(g, o[i], key, 0, kStreamKeyERetry, 16 * b + (u32)i, 0);
    }
}
HD u64 keygen_uniform(const SamplerKey &key, u32 l, u32 j, u64 q) {
    int qbits = 0;
    while (qbits < 64 && (q >> qbits)) qbits++;
    for (u32 t = 0;; t++) {
        u32 o[16];
        chacha20_block(key, j >> 3, l, t, kStreamKeyA, o);
        u64 r = 0;
#pragma unroll
        for (int i = 0; i < 8; i++)
            if ((u32)i == (j & 7u)) r = ((u64)o[2 * i + 1] << 32) | o[2 * i];
        const u64 a = r >> (64 - qbits);
        if (a < q) return a;
    }
}
// phase 1 of the sampler: ve[j] = v[j] + 1 for the n coefficients of one ciphertext (32 bytes per thread and block)
HD void sample_ternary_phase(int tid, int nthreads, const SamplerKey &key, u64 ct, u32 n, uint8_t *ve) {
    for (u32 b = tid; b < (n >> 5); b += nthreads) {
        u32 w[8];
        sample_ternary_block(key, ct, b, kStreamTernary, kStreamTernaryRetry, w);
        u32 *dst = reinterpret_cast<u32 *>(ve + 32 * (size_t)b);
#pragma unroll
        for (int i = 0; i < 8; i++) dst[i] = w[i];
    }
}
// phase 2 (after a barrier): m_e0[j] = coefficient j of the encoded message + e0[j]; e1[j] joins v[j] in ve[j].
// `z` = the finalized FFT tile (ring dimension 8192), or null when m_e0 already holds the message (larger rings).
HD void sample_gauss_phase(int tid, int nthreads, const cplx *z, const EncodeConst &ec, const SamplerKey &key,
                           const GaussTable &g, u64 ct, u32 n, i64 *m_e0, uint8_t *ve, u64 keep = 0) {
    for (u32 b = tid; b < (n >> 3); b += nthreads) {
        int e0[8], e1[8];
        sample_gauss_block(key, g, ct, b, e0, e1);
        u32 *vw = reinterpret_cast<u32 *>(ve + 8 * (size_t)b);
        u32 w0 = vw[0], w1 = vw[1];
        u64 m[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const u32 j = 8 * b + i;
            m[i] = (u64)((z ? encode_coeff(z, ec, j) : m_e0[j]) + (i64)e0[i]);
            if (i < 4) w0 |= (u32)(e1[i] + 32) << (8 * i + 2);
            else w1 |= (u32)(e1[i] + 32) << (8 * (i - 4) + 2);
        }
        vw[0] = w0; vw[1] = w1;
        u64x2 *dst = reinterpret_cast<u64x2 *>(m_e0 + 8 * (size_t)b);
#pragma unroll
        for (int i = 0; i < 4; i++) st_pair(dst + i, u64x2{m[2 * i], m[2 * i + 1]}, keep, false);
    }
}
HD u64 small_to_mod(int x, u64 q) { return x >= 0 ? (u64)x : q - (u64)(-x); }

// ---- decode: CRTInterpolate -> double -> FFTSpecial (ckks.cpp:189-199) ----
// phase 0: coefficients (limb-major, natural order) -> z in bit-reversed slot order
HD void decode_load(int tid, int nthreads, cplx *z, const EncodeConst &ec, const DecodeConst &dc, u32 limbs,
                    const u64 *coeff, i64 *tap_hi, u64 *tap_lo) {
    u32 nh = kN >> 1;
    for (u32 j = tid; j < (u32)kN; j += nthreads) {
        bool neg; u64 hi, lo;
        if (limbs == 1) {
            u64 a = coeff[j], q = dc.crt.q0;
            neg = a > (q >> 1); hi = 0; lo = neg ? q - a : a;
        } else {
            crt2_centered(coeff[j], coeff[kN + j], dc.crt, neg, hi, lo);
        }
        if (tap_hi) {  // two's complement of the centred integer, for the bit-exact integer tier
            u64 th = hi, tl = lo;
            if (neg) { tl = ~lo + 1; th = ~hi + (tl == 0 ? 1 : 0); }
            tap_hi[j] = (i64)th; tap_lo[j] = tl;
        }
        u32 half = j >= nh ? 1u : 0u, idx = j - half * nh;
        if (idx & (ec.gap - 1)) continue;
        u32 slot = zsw(bitrev32(idx / ec.gap, (int)ec.log_slots));
        double val = fmul(i128_to_double(neg, hi, lo), dc.pre);
        if (half) z[slot].im = val; else z[slot].re = val;
    }
}
// phases 1..log_slots: fft_fwd_stage(len = 2 << (p-1)); last phase: real parts * 2^-p
// `first` = index of out[0] in the whole decoded vector (what the peer destinations are indexed by)
HD void decode_store(int tid, int nthreads, const cplx *z, const DecodeConst &dc, double *out, u64 length,
                     const u32 *sel = nullptr, u64 first = 0) {
    if (dc.n_peer == 0) {
        for (u32 i = tid; i < length; i += nthreads) out[sel ? sel[i] : i] = fmul(z[zsw(i)].re, dc.pow_p);
        return;
    }
    for (u32 i = tid; i < length; i += nthreads) {
        const double v = fmul(z[zsw(i)].re, dc.pow_p);
        out[i] = v;
        for (u32 p = 0; p < dc.n_peer; p++) dc.peer[p][first + i] = v;
    }
}

// ---- slot counts above one FFT tile (4096 complex values, 64 KiB) and ring dimensions above one NTT tile ----
// FFTSpecialInv is decimation in frequency: after the df = log2(slots / 4096) outermost stages the array falls
// apart into 2^df independent 4096-point blocks.  Block fb's loader applies those stages to the 2^df inputs
// that share its tile-local index (same butterflies, same order as the reference => same doubles), the tile
// runs the 12 inner stages, and the bit reversal sends tile position r to slot (r << df) | bitrev(fb).
struct EncSrc {             // one chunk of encode input: x[j] or x[sel[j]], j < n_valid
    const double *x;
    u64 n_valid;
    const u32 *sel;
    bool f32;
};
HD cplx encode_input(const EncSrc &x, u64, u32 j) {
    return cplx{j < x.n_valid ? encode_value(x.x, x.sel ? x.sel[j] : j, x.f32) : 0.0, 0.0};
}
template <int DF>
HD cplx encode_head(const EncodeConst &ec, const EncSrc &x, u64 n_valid, u32 fb, u32 i) {
    if (DF == 0) return encode_input(x, n_valid, i);
    const cplx *__restrict__ twi = ec.tw_inv;
    if (DF == 1) {
        cplx a = encode_input(x, n_valid, i), b = encode_input(x, n_valid, i + kFftFast);
        dif_bfly(a, b, twi[kFftFast - 1 + i]);
        return fb ? b : a;
    }
    cplx a0 = encode_input(x, n_valid, i), a1 = encode_input(x, n_valid, i + kFftFast);
    cplx a2 = encode_input(x, n_valid, i + 2 * kFftFast), a3 = encode_input(x, n_valid, i + 3 * kFftFast);
    dif_bfly(a0, a2, twi[2 * kFftFast - 1 + i]);               // len = 16384: pairs (j, j + 8192)
    dif_bfly(a1, a3, twi[2 * kFftFast - 1 + i + kFftFast]);
    if (fb < 2) { dif_bfly(a0, a1, twi[kFftFast - 1 + i]); return fb ? a1 : a0; }   // len = 8192, first half
    dif_bfly(a2, a3, twi[kFftFast - 1 + i]);                                        // second half
    return (fb & 1u) ? a3 : a2;
}
template <int DF>
HD void encode_load_block(int tid, int nthreads, cplx *z, const EncodeConst &ec, const double *x, u64 n_valid, u32 fb,
                          const u32 *sel = nullptr, bool f32 = false) {
    const u32 tile_slots = ec.slots < kFftFast ? ec.slots : kFftFast;
    const EncSrc src{x, n_valid, sel, f32};
    for (u32 i = tid; i < tile_slots; i += nthreads) z[zsw(i)] = encode_head<DF>(ec, src, n_valid, fb, i);
}
// tile (finalized: bit-reversed within the tile, scaled, rounded) -> coefficients m[gap*slot], m[gap*slot + nh]
template <int DF>
HD void encode_emit_block(int tid, int nthreads, const cplx *z, const EncodeConst &ec, u32 fb, u32 nh, i64 *m) {
    const i64 *zi = reinterpret_cast<const i64 *>(z);
    const u32 tile_slots = ec.slots < kFftFast ? ec.slots : kFftFast;
    const u32 fbr = DF ? bitrev32(fb, DF) : 0u;
    for (u32 r = tid; r < tile_slots; r += nthreads) {
        u32 idx = (((r << DF) | fbr)) * ec.gap;
        m[idx] = zi[2 * zsw(r)];
        m[idx + nh] = zi[2 * zsw(r) + 1];
        for (u32 g = 1; g < ec.gap; g++) { m[idx + g] = 0; m[idx + nh + g] = 0; }
    }
}
// centred CRT value of coefficient j of a polynomial stored [L][n] -> double (and optionally its words)
template <int L>
HD double decode_value(const u64 *coeff, u32 n, u32 j, const Crt2Const &c2, const CrtNConst &cn, u64 *tap8) {
    bool neg;
    if (L <= 2) {
        u64 hi, lo;
        if (L == 1) {
            u64 a = coeff[j], q = c2.q0;
            neg = a > (q >> 1); hi = 0; lo = neg ? q - a : a;
        } else {
            crt2_centered(coeff[j], coeff[n + j], c2, neg, hi, lo);
        }
        if (tap8) {
            u64 th = hi, tl = lo;
            if (neg) { tl = ~lo + 1; th = ~hi + (tl == 0 ? 1 : 0); }
            u64 ext = neg ? ~0ull : 0ull;
            tap8[0] = tl; tap8[1] = th;
            for (int k = 2; k < 8; k++) tap8[k] = ext;
        }
        return i128_to_double(neg, hi, lo);
    } else {
        u64 a[L], mag[L];
#pragma unroll
        for (int l = 0; l < L; l++) a[l] = coeff[(size_t)l * n + j];
        crtn_centered<L>(a, cn, neg, mag);
        if (tap8) {   // two's complement, sign-extended to 8 words
            u64 carry = 1;
#pragma unroll
            for (int k = 0; k < 8; k++) {
                u64 mk = k < L ? mag[k < L ? k : 0] : 0ull;
                if (neg) { u64 t = ~mk + carry; carry = (carry && t == 0) ? 1 : 0; tap8[k] = t; }
                else tap8[k] = mk;
            }
        }
        return words_to_double<L>(neg, mag);
    }
}
// FFTSpecial is decimation in time on the bit-reversed array: tile fb holds positions [fb*4096, (fb+1)*4096) of it,
// i.e. slots (bitrev(p) << df) | bitrev(fb); the df outermost stages run in decode_tail.
template <int DF>
HD void decode_load_block(int tid, int nthreads, cplx *z, const EncodeConst &ec, u32 fb, u32 n, const double *vals) {
    const u32 nh = n >> 1;
    const u32 tile_slots = ec.slots < kFftFast ? ec.slots : kFftFast;
    const int tile_log = (int)ec.log_slots - DF;
    const u32 fbr = DF ? bitrev32(fb, DF) : 0u;
    for (u32 p = tid; p < tile_slots; p += nthreads) {
        u32 idx = ((bitrev32(p, tile_log) << DF) | fbr) * ec.gap;
        z[zsw(p)] = cplx{vals[idx], vals[idx + nh]};
    }
}
// the df outermost FFTSpecial stages on the 2^df values sharing tile-local index i, then real parts * 2^-p
template <int DF>
HD void decode_tail(u32 i, const cplx *zs, const EncodeConst &ec, const DecodeConst &dc, double *out, u64 length,
                    const u32 *sel = nullptr, u64 first = 0) {
    const cplx *__restrict__ twf = ec.tw_fwd;
    cplx a[1 << DF];
#pragma unroll
    for (int k = 0; k < (1 << DF); k++) a[k] = zs[i + (u32)k * kFftFast];
    if (DF == 1) {
        dit_bfly(a[0], a[1], twf[kFftFast - 1 + i]);
    } else if (DF == 2) {
        dit_bfly(a[0], a[1], twf[kFftFast - 1 + i]);               // len = 8192 on both halves
        dit_bfly(a[2], a[3], twf[kFftFast - 1 + i]);
        dit_bfly(a[0], a[2], twf[2 * kFftFast - 1 + i]);           // len = 16384
        dit_bfly(a[1], a[3], twf[2 * kFftFast - 1 + i + kFftFast]);
    }
#pragma unroll
    for (int k = 0; k < (1 << DF); k++) {
        u64 slot = i + (u64)k * kFftFast;
        if (slot < length) {
            const double v = fmul(a[k].re, dc.pow_p);
            out[sel ? sel[slot] : slot] = v;
            for (u32 p = 0; p < dc.n_peer; p++) dc.peer[p][first + slot] = v;
        }
    }
}

// ---- weighted sum: EvalMult(ct, const) + EvalAdd over learners (ckks.cpp:286-297) ----
struct acc128 { u64 lo, hi; };
HD void mac128(acc128 &a, u64 x, u64 c) {
    u64 lo = x * c, hi = mulhi(x, c);
    a.lo += lo;
    a.hi += hi + (a.lo < lo ? 1 : 0);
}
HD u64 reduce128(const acc128 &a, const LimbConst &lc) {
    u64 t1 = shoup_lazy(a.hi, lc.r64, lc.r64_s, lc.q);       // hi * 2^64 mod q, < 2q
    u64 t2 = a.lo - mulhi(a.lo, lc.mu) * lc.q;               // < 2q
    return csub(csub(t1 + t2, lc.q2), lc.q);
}

// ---- serialized image <-> dense: coefficient runs start at arbitrary byte offsets ----
HD u64 load_unaligned(const u64 *image, u64 byte_off) {
    u64 w = byte_off >> 3;
    u32 sh = (u32)(byte_off & 7u) * 8u;
    u64 a = image[w];
    if (sh == 0) return a;
    return (a >> sh) | (image[w + 1] << (64u - sh));
}
// word `widx` of the image, overlapping the run that starts at byte a0 and holds n values
HD void pack_word(u64 *image, u64 a0, u64 widx, const u64 *run, u32 n) {
    i64 o = (i64)(widx * 8) - (i64)a0;                 // run byte offset of the word's first byte: -7 .. 8N-1
    i64 i = o >= 0 ? (o >> 3) : -1;
    u32 t = (u32)(o - i * 8);                          // 0..7
    u64 lo = i >= 0 ? run[i] : 0ull;
    u64 hi = (t && i + 1 < (i64)n) ? run[i + 1] : 0ull;
    u64 val = t ? (lo >> (8 * t)) | (hi << (64 - 8 * t)) : lo;
    if (o >= 0 && o + 8 <= (i64)n * 8) { image[widx] = val; return; }
    unsigned char *bytes = reinterpret_cast<unsigned char *>(image + widx);
    for (int b = 0; b < 8; b++) {
        i64 ob = o + b;
        if (ob >= 0 && ob < (i64)n * 8) bytes[b] = (unsigned char)(val >> (8 * b));
    }
}

// ------------------------------------------------------------------------------------------
// loaders / storers plugged into the NTT phases by the kernels (and by the CPU replay harness)
// ------------------------------------------------------------------------------------------
HD u64x2 ld_readonly(const u64x2 *p) {
#ifdef __CUDA_ARCH__
    ulonglong2 v = __ldg(reinterpret_cast<const ulonglong2 *>(p));
    return u64x2{v.x, v.y};
#else
    return *p;
#endif
}
struct LoadGlobal {
    const u64 *p;
    HD u64 operator()(u32 i) const { return p[i]; }
};
// FP64-path loaders: centred doubles, |v| <= q/2
struct DLoadGlobal {
    const u64 *p; u64 q;
    HD double operator()(u32 i) const { return d_from_residue(p[i], q); }
};
struct DLoadTernary {
    const uint8_t *ve;
    HD double operator()(u32 i) const { return (double)ve_ternary(ve[i]); }
};
struct DLoadGaussHi {
    const uint8_t *ve;
    HD double operator()(u32 i) const { return (double)ve_gauss(ve[i]); }
};
struct DLoadSigned {
    const i64 *m; const LimbConst *lc;
    HD double operator()(u32 i) const { return d_from_residue(signed_to_mod(m[i], *lc), lc->q); }
};
struct StoreFullReducePair {
    typedef NoAux Aux;
    static constexpr int kGroup = 4;
    u64 *p; const LimbConst *lc;
    HD Aux fetch(u32) const { return Aux{}; }
    HD void operator()(u32 vt, u64 x, u64 y, const Aux &) const {
        reinterpret_cast<u64x2 *>(p)[vt] = u64x2{full_reduce(x, *lc), full_reduce(y, *lc)};
    }
};
struct LoadPairGlobal {
    const u64 *p;
    HD void operator()(u32 vt, u64 &x, u64 &y) const {
        u64x2 v = reinterpret_cast<const u64x2 *>(p)[vt];
        x = v.x; y = v.y;
    }
};
struct StoreGlobal {
    u64 *p;
    HD void operator()(u32 i, u64 v) const { p[i] = v; }
};
// encrypt: ve[j] = packed (v, e1) of coefficient j (pack_ve); m[j] = m + e0 (signed)
struct LoadTernary {
    const uint8_t *ve; u64 q;
    HD u64 operator()(u32 i) const { return small_to_mod(ve_ternary(ve[i]), q); }
};
struct LoadGaussHi {
    const uint8_t *ve; u64 q;
    HD u64 operator()(u32 i) const { return small_to_mod(ve_gauss(ve[i]), q); }
};
struct LoadSigned {
    const i64 *m; const LimbConst *lc;
    HD u64 operator()(u32 i) const { return signed_to_mod(m[i], *lc); }
};
// encrypt epilogues.  NTT(v) parks v^ itself (lazy, any 64-bit value) in the c1 row; NTT(m + e0) then forms
// c0 = b*v^ + . and NTT(e1) overwrites the parked row with c1 = a*v^ + . — each thread re-reads only what it
// wrote itself, so no fence is needed, and the key material is fetched 32 bytes per pair per pass.
struct StoreLazyPair {
    typedef NoAux Aux;
    static constexpr int kGroup = 4;
    u64 *park; u64 keep;
    HD Aux fetch(u32) const { return Aux{}; }
    HD void operator()(u32 vt, u64 x, u64 y, const Aux &) const {
        st_pair(reinterpret_cast<u64x2 *>(park) + vt, u64x2{x, y}, keep, false);
    }
};
struct PkAux { u64x2 v, k0, k1; };
struct StoreMulAdd {
    typedef PkAux Aux;
    static constexpr int kGroup = 2;
    const tw_t *key; const u64 *park; u64 *out; const LimbConst *lc; bool stream;
    HD Aux fetch(u32 vt) const {
        const u64x2 *kp = reinterpret_cast<const u64x2 *>(key) + 2 * vt;
        return Aux{reinterpret_cast<const u64x2 *>(park)[vt], ld_readonly(kp), ld_readonly(kp + 1)};
    }
    HD void operator()(u32 vt, u64 x, u64 y, const Aux &a) const {
        const u64 q = lc->q, q2 = lc->q2, nq = lc->nq;
        u64 r0 = shoup_lazy_nq(a.v.x, a.k0.x, a.k0.y, nq) + quick_reduce(x, *lc);   // < 4q
        u64 r1 = shoup_lazy_nq(a.v.y, a.k1.x, a.k1.y, nq) + quick_reduce(y, *lc);
        st_pair(reinterpret_cast<u64x2 *>(out) + vt, u64x2{csub(csub(r0, q2), q), csub(csub(r1, q2), q)}, 0, stream);
    }
};
// rescale (PALISADE DCRTPoly::DropLastElementAndScale): the dropped tower's coefficients t (canonical, mod q_last) enter
// the forward transform of every remaining tower as their centred representative reduced mod q_i (NativeVector::
// SwitchModulus), and the epilogue forms (c_i - NTT_i(t)) * q_last^-1 mod q_i.
struct RescaleConst { u64 q_last, inv[kMaxLimbs], inv_s[kMaxLimbs]; };   // inv[i] = q_last^-1 mod q_i + Shoup companion
HD i64 centre_mod(u64 t, u64 q) { return t > (q >> 1) ? (i64)t - (i64)q : (i64)t; }
struct LoadSwitch {
    const u64 *t; u64 q_last; const LimbConst *lc;
    HD u64 operator()(u32 i) const { return signed_to_mod(centre_mod(t[i], q_last), *lc); }
};
struct DLoadSwitch {
    const u64 *t; u64 q_last; const LimbConst *lc;
    HD double operator()(u32 i) const { return d_from_residue(signed_to_mod(centre_mod(t[i], q_last), *lc), lc->q); }
};
struct StoreRescale {
    typedef u64x2 Aux;
    static constexpr int kGroup = 4;
    const u64 *c; u64 *out; const LimbConst *lc; u64 inv, inv_s;
    HD Aux fetch(u32 vt) const { return ld_readonly(reinterpret_cast<const u64x2 *>(c) + vt); }
    HD void operator()(u32 vt, u64 x, u64 y, const Aux &a) const {
        const u64 q = lc->q, nq = lc->nq;
        u64 d0 = a.x + q - full_reduce(x, *lc), d1 = a.y + q - full_reduce(y, *lc);       // (0, 2q)
        reinterpret_cast<u64x2 *>(out)[vt] = u64x2{csub(shoup_lazy_nq(d0, inv, inv_s, nq), q),
                                                   csub(shoup_lazy_nq(d1, inv, inv_s, nq), q)};
    }
};
// decrypt: t = c0 + c1*s on the way into the inverse transform (< 3q for reduced input)
struct LoadDecryptPair {
    const u64 *c0, *c1; const tw_t *s; u64 q;
    HD void operator()(u32 vt, u64 &x, u64 &y) const {
        u64x2 u = ld_readonly(reinterpret_cast<const u64x2 *>(c0) + vt);
        u64x2 v = ld_readonly(reinterpret_cast<const u64x2 *>(c1) + vt);
        const u64x2 *sp = reinterpret_cast<const u64x2 *>(s) + 2 * vt;
        u64x2 s0 = ld_readonly(sp), s1 = ld_readonly(sp + 1);
        x = u.x + shoup_lazy_nq(v.x, s0.x, s0.y, (u64)0 - q);
        y = u.y + shoup_lazy_nq(v.y, s1.x, s1.y, (u64)0 - q);
    }
};

}  // namespace shelfi
