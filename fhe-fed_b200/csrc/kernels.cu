// kernels.cu — the sm_100a kernels of the SHELFI CKKS path.  Build: nvcc -gencode arch=compute_100a,code=sm_100a.
//
// Data layout in HBM: a ciphertext batch is dense limb-major SoA u64 [n_ct][2][L][N];
// every kernel maps one CTA to one 64 KiB limb polynomial (or a slice of one) so all global traffic is
// contiguous 8/16-byte-per-lane runs.  NTT working set lives in a swizzled 64 KiB shared-memory tile,
// three CTAs per SM; butterflies are Shoup/Harvey lazy 64-bit modmuls on the IMAD/ALU pipes.
#include "kernels.h"

#include <cooperative_groups.h>

#include <mutex>

namespace shelfi {
namespace cg = cooperative_groups;

#define NTT_SMEM_BYTES (kN * 8)

// ------------------------------------------------------------------------------------------
// Rings above one tile, N = 2^(13+D): the 2^D CTAs that hold the tiles of one polynomial form a thread-block
// cluster.  The D outermost Cooley-Tukey stages pair element i of every tile; CTA r does them for its share
// i in [r, r+1) * 8192 / 2^D of the index range — reading the 2^D strided inputs, one butterfly per output —
// and writes output b straight into tile b, i.e. into the shared memory of CTA b (distributed shared memory).
// Same arithmetic, operand for operand, as HeadLoad / DHeadLoad (which recompute the head per tile: 3 butterflies per
// output at D = 2 instead of 1, and which the CPU replay harness keeps using), so results are bit-identical.
// Protocol per transform: cluster barrier (every tile is free), scatter, cluster barrier (every tile is full).
// Measured (profiles/README.md): at D = 2 the cluster form wins (config 4 encrypt 208 -> 172 ms); at D = 1 the
// recomputed head costs only +7 % multiplications and the two cluster barriers cost as much (standalone NTT
// 0.216 vs 0.240 ms), so D = 1 keeps HeadLoad.
// ------------------------------------------------------------------------------------------
__host__ __device__ constexpr bool cluster_head_for(int d) { return d >= 2; }
struct LoadTile {
    const u64 *sm;
    __device__ __forceinline__ u64 operator()(u32 i) const { return sm[swz(i)]; }
};
struct DLoadTile {
    const u64 *sm;
    __device__ __forceinline__ double operator()(u32 i) const { return reinterpret_cast<const double *>(sm)[swz(i)]; }
};
template <int D, class Inner>
__device__ __forceinline__ void head_scatter(int tid, u64 *sm, const LimbConst &lc, Inner in) {
    cg::cluster_group cl = cg::this_cluster();
    const u32 rank = cl.block_rank();
    u64 *peer[1 << D];
#pragma unroll
    for (int b = 0; b < (1 << D); b++) peer[b] = cl.map_shared_rank(sm, b);
    const u64 nq = lc.nq, q2 = lc.q2;
    const tw_t *__restrict__ tw = lc.tw;
    constexpr u32 span = (u32)kN >> D;
    const tw_t w1 = tw[1];
    for (u32 i = rank * span + tid; i < (rank + 1) * span; i += kNttThreads) {
        const u32 pos = swz(i);
        if (D == 1) {
            u64 x = in(i), y = in(i + kN);
            u64 t = shoup_lazy_nq(y, w1.w, w1.ws, nq);
            peer[0][pos] = x + t;
            peer[1][pos] = x + q2 - t;
        } else {
            u64 x0 = in(i), x1 = in(i + kN), x2 = in(i + 2 * kN), x3 = in(i + 3 * kN);
            u64 t0 = shoup_lazy_nq(x2, w1.w, w1.ws, nq), t1 = shoup_lazy_nq(x3, w1.w, w1.ws, nq);
            u64 a0 = x0 + t0, a1 = x1 + t1, b0 = x0 + q2 - t0, b1 = x1 + q2 - t1;
            const tw_t w2 = tw[2], w3 = tw[3];
            u64 ta = shoup_lazy_nq(a1, w2.w, w2.ws, nq), tb = shoup_lazy_nq(b1, w3.w, w3.ws, nq);
            peer[0][pos] = a0 + ta;
            peer[1][pos] = a0 + q2 - ta;
            peer[2 % (1 << D)][pos] = b0 + tb;
            peer[3 % (1 << D)][pos] = b0 + q2 - tb;
        }
    }
}
template <int D, class Inner>
__device__ __forceinline__ void dhead_scatter(int tid, u64 *sm, const LimbConst &lc, Inner in) {
    cg::cluster_group cl = cg::this_cluster();
    const u32 rank = cl.block_rank();
    double *peer[1 << D];
#pragma unroll
    for (int b = 0; b < (1 << D); b++) peer[b] = reinterpret_cast<double *>(cl.map_shared_rank(sm, b));
    const double q = lc.qd, qinv = lc.qinv;
    const u32 thr = lc.dthr;
    const double *__restrict__ tw = lc.dtw;
    constexpr u32 span = (u32)kN >> D;
    const double w1 = tw[1];
    for (u32 i = rank * span + tid; i < (rank + 1) * span; i += kNttThreads) {
        const u32 pos = swz(i);
        if (D == 1) {
            double x = in(i), t = d_mulmod(in(i + kN), w1, q, qinv);
            peer[0][pos] = d_renorm(fadd(x, t), q, qinv, thr);
            peer[1][pos] = d_renorm(fsub(x, t), q, qinv, thr);
        } else {
            double x0 = in(i), x1 = in(i + kN);
            double t0 = d_mulmod(in(i + 2 * kN), w1, q, qinv), t1 = d_mulmod(in(i + 3 * kN), w1, q, qinv);
            double a0 = d_renorm(fadd(x0, t0), q, qinv, thr), a1 = d_renorm(fadd(x1, t1), q, qinv, thr);
            double b0 = d_renorm(fsub(x0, t0), q, qinv, thr), b1 = d_renorm(fsub(x1, t1), q, qinv, thr);
            double ta = d_mulmod(a1, tw[2], q, qinv), tb = d_mulmod(b1, tw[3], q, qinv);
            peer[0][pos] = d_renorm(fadd(a0, ta), q, qinv, thr);
            peer[1][pos] = d_renorm(fsub(a0, ta), q, qinv, thr);
            peer[2 % (1 << D)][pos] = d_renorm(fadd(b0, tb), q, qinv, thr);
            peer[3 % (1 << D)][pos] = d_renorm(fsub(b0, tb), q, qinv, thr);
        }
    }
}
// cluster barrier, scatter, cluster barrier: afterwards this CTA's tile holds its head-stage outputs
template <int D, class Inner>
__device__ __forceinline__ void cluster_head(int tid, u64 *sm, const LimbConst &lc, Inner in) {
    cg::cluster_group cl = cg::this_cluster();
    cl.sync();
    head_scatter<D>(tid, sm, lc, in);
    cl.sync();
}
template <int D, class Inner>
__device__ __forceinline__ void cluster_head_d(int tid, u64 *sm, const LimbConst &lc, Inner in) {
    cg::cluster_group cl = cg::this_cluster();
    cl.sync();
    dhead_scatter<D>(tid, sm, lc, in);
    cl.sync();
}

// ------------------------------------------------------------------------------------------
// standalone forward / inverse NTT (parity taps, keygen, and the "NTT HBM GB/s" metric)
// ------------------------------------------------------------------------------------------
template <class Load, class Store>
__device__ __forceinline__ void ntt_forward_tile_d(int tid, u64 *sm, const LimbConst &lc, u32 hb, Load ld, Store st) {
    dfwd_phase_a(tid, sm, lc, hb, ld);
    __syncthreads();
    dfwd_phase_mid<7>(tid, sm, lc, hb);
    __syncthreads();
    dfwd_phase_mid<4>(tid, sm, lc, hb);
    __syncthreads();
    dfwd_phase_mid<1>(tid, sm, lc, hb);
    __syncthreads();
    dfwd_phase_e(tid, sm, lc, hb, st);
}

// One CTA per 8192-point tile: a polynomial of ring dimension 2^(13+D) is 2^D tiles (ckks_core.cuh).  For D > 0 the
// forward transform reads all tiles of its row (HeadLoad) and so must not run in place: `in` != `out`.
template <int D>
__global__ void __launch_bounds__(kNttThreads, 3) ntt_forward_kernel(const __grid_constant__ RingConst rc, const u64 *in,
                                                                     u64 *out) {
    extern __shared__ __align__(16) u64 sm[];
    const int tid = threadIdx.x;
    const u32 blk = D ? blockIdx.x & ((1u << D) - 1u) : 0u;
    const u32 row = blockIdx.x >> D;
    const u32 hb = D ? head_bits(D, blk) : 1u;
    const LimbConst lc = rc.limb[row % rc.limbs];
    const u64 *src = in + ((size_t)row << (kLogN + D));
    u64 *dst = out + ((size_t)row << (kLogN + D)) + (size_t)blk * kN;
    if (lc.dtw) {   // q < 2^53: the same transform on the FP64 pipe
        if (!cluster_head_for(D)) {
            ntt_forward_tile_d(tid, sm, lc, hb, DHeadLoad<D, DLoadGlobal>{DLoadGlobal{src, lc.q}, &lc, blk},
                               StoreFullReducePair{dst, &lc});
        } else {
            cluster_head_d<D>(tid, sm, lc, DLoadGlobal{src, lc.q});
            ntt_forward_tile_d(tid, sm, lc, hb, DLoadTile{sm}, StoreFullReducePair{dst, &lc});
        }
        return;
    }
    if (!cluster_head_for(D)) {
        fwd_phase_a(tid, sm, lc, hb, HeadLoad<D, LoadGlobal>{LoadGlobal{src}, &lc, blk});
    } else {
        cluster_head<D>(tid, sm, lc, LoadGlobal{src});
        fwd_phase_a(tid, sm, lc, hb, LoadTile{sm});
    }
    __syncthreads();
    fwd_phase_mid<7>(tid, sm, lc, hb);
    __syncthreads();
    fwd_phase_mid<4>(tid, sm, lc, hb);
    __syncthreads();
    fwd_phase_mid<1>(tid, sm, lc, hb);
    __syncthreads();
    fwd_phase_e(tid, sm, lc, hb, StoreFullReducePair{dst, &lc});
}

// EXPERIMENT (north_star: "warp-shuffle butterflies" for the last stages; measured in profiles/README.md, not used by
// the product): the forward transform with the last FOUR stages in registers.  After the radix-8 pass over address bits
// 3..1, the partner of every value for the last stage (address bit 0) sits in the neighbouring lane: lanes exchange
// half of their values (__shfl_xor 1), so each ends up with four complete pairs, does their butterflies and stores them.
// One shared-memory round trip and one barrier fewer than the five-pass form; in exchange 8 SHFL + 24 SEL per 8 values
// and 16-byte stores that are 64 bytes apart within a warp instead of contiguous.
template <class Store>
__device__ __forceinline__ void fwd_phase_de_shfl(int tid, u64 *smem, const LimbConst &lc, u32 hb, Store st) {
    char *sm = reinterpret_cast<char *>(smem);
    const u64 nq = lc.nq, q3 = kBflyGrowth * lc.q;
    const tw_t *__restrict__ twe = lc.tw + ((size_t)hb << (kLogN - 1));
#pragma unroll 1
    for (int r = 0; r < 4; r++) {
        const u32 vt = tid + kNttThreads * r;
        const u32 base = tile_base<1>(vt);
        u64 a[8];
#pragma unroll
        for (int k = 0; k < 8; k++) a[k] = tile(sm, tile_off<1>(base, k));
        ct_radix8(a, lc.tw, (hb << 9) + (vt >> 1), nq, q3);
        if (lc.fwd_mask & (1u << 3)) quick_reduce8(a, lc);
        const bool odd = vt & 1u;
        const u32 high = vt >> 1;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const u64 send = odd ? a[j] : a[j + 4];
            const u64 recv = __shfl_xor_sync(0xffffffffu, send, 1);
            u64 x = odd ? recv : a[j], y = odd ? a[j + 4] : recv;
            const u32 p = (high << 3) | (u32)(j + (odd ? 4 : 0));        // pair index = element index >> 1
            const tw_t w = twe[p];
            typename Store::Aux aux = st.fetch(p);
            ct_bfly(x, y, w, nq, q3);
            st(p, x, y, aux);
        }
    }
}
__global__ void __launch_bounds__(kNttThreads, 3) ntt_forward_shfl_kernel(const __grid_constant__ RingConst rc, const u64 *in,
                                                                          u64 *out) {
    extern __shared__ __align__(16) u64 sm[];
    const int tid = threadIdx.x;
    const u32 row = blockIdx.x;
    const LimbConst lc = rc.limb[row % rc.limbs];
    const u64 *src = in + ((size_t)row << kLogN);
    u64 *dst = out + ((size_t)row << kLogN);
    if (lc.dtw) {   // FP64 limbs keep the regular form (the exchange idea is the same; the integer limb is the slow one)
        ntt_forward_tile_d(tid, sm, lc, 1u, DLoadGlobal{src, lc.q}, StoreFullReducePair{dst, &lc});
        return;
    }
    fwd_phase_a(tid, sm, lc, 1u, LoadGlobal{src});
    __syncthreads();
    fwd_phase_mid<7>(tid, sm, lc, 1u);
    __syncthreads();
    fwd_phase_mid<4>(tid, sm, lc, 1u);
    __syncthreads();
    fwd_phase_de_shfl(tid, sm, lc, 1u, StoreFullReducePair{dst, &lc});
}

// Plaintext polynomial of an encoded weight vector -> evaluation form in every limb (EvalMult(ct, plaintext)): the
// signed coefficients m[ct][N] are shared by the limbs, one CTA per (ciphertext, limb, tile).
template <int D>
__global__ void __launch_bounds__(kNttThreads, 3) plain_ntt_kernel(const __grid_constant__ RingConst rc, const i64 *m,
                                                                   u64 *pt_hat) {
    extern __shared__ __align__(16) u64 sm[];
    const int tid = threadIdx.x;
    const u32 L = rc.limbs;
    const u32 blk = D ? blockIdx.x & ((1u << D) - 1u) : 0u;
    const u32 hb = D ? head_bits(D, blk) : 1u;
    const u32 poly = blockIdx.x >> D;          // ct * L + l
    const u32 l = poly % L;
    const u64 ct = poly / L;
    const u64 n = (u64)kN << D;
    const LimbConst lc = rc.limb[l];
    const i64 *m_row = m + ct * n;
    u64 *dst = pt_hat + (size_t)poly * n + (size_t)blk * kN;
    if (lc.dtw) {
        if (!cluster_head_for(D)) {
            ntt_forward_tile_d(tid, sm, lc, hb, DHeadLoad<D, DLoadSigned>{DLoadSigned{m_row, &lc}, &lc, blk},
                               StoreFullReducePair{dst, &lc});
        } else {
            cluster_head_d<D>(tid, sm, lc, DLoadSigned{m_row, &lc});
            ntt_forward_tile_d(tid, sm, lc, hb, DLoadTile{sm}, StoreFullReducePair{dst, &lc});
        }
        return;
    }
    if (!cluster_head_for(D)) {
        fwd_phase_a(tid, sm, lc, hb, HeadLoad<D, LoadSigned>{LoadSigned{m_row, &lc}, &lc, blk});
    } else {
        cluster_head<D>(tid, sm, lc, LoadSigned{m_row, &lc});
        fwd_phase_a(tid, sm, lc, hb, LoadTile{sm});
    }
    __syncthreads();
    fwd_phase_mid<7>(tid, sm, lc, hb);
    __syncthreads();
    fwd_phase_mid<4>(tid, sm, lc, hb);
    __syncthreads();
    fwd_phase_mid<1>(tid, sm, lc, hb);
    __syncthreads();
    fwd_phase_e(tid, sm, lc, hb, StoreFullReducePair{dst, &lc});
}

// inverse: the 13 inner stages per tile in place; for D > 0 the results stay lazy (< 2q) and intt_tail_kernel
// applies the D outermost stages and N^-1 as a streaming pass
template <int D, class Load>
__device__ __forceinline__ void intt_tile(int tid, u64 *sm, const LimbConst &lc, u32 hb, Load ld, u64 *dst) {
    if (lc.ditw) {   // q < 2^53: the same transform on the FP64 pipe
        dinv_phase_e(tid, sm, lc, hb, ld);
        __syncthreads();
        dinv_phase_mid<1>(tid, sm, lc, hb);
        __syncthreads();
        dinv_phase_mid<4>(tid, sm, lc, hb);
        __syncthreads();
        dinv_phase_mid<7>(tid, sm, lc, hb);
        __syncthreads();
        if (D == 0) dinv_phase_last(tid, sm, lc, StoreGlobal{dst});
        else dinv_phase_last_lazy(tid, sm, lc, hb, StoreGlobal{dst});
        return;
    }
    inv_phase_e(tid, sm, lc, hb, ld);
    __syncthreads();
    inv_phase_mid<1>(tid, sm, lc, hb);
    __syncthreads();
    inv_phase_mid<4>(tid, sm, lc, hb);
    __syncthreads();
    inv_phase_mid<7>(tid, sm, lc, hb);
    __syncthreads();
    if (D == 0) inv_phase_last(tid, sm, lc, StoreGlobal{dst});
    else inv_phase_last_lazy(tid, sm, lc, hb, StoreGlobal{dst});
}
template <int D>
__global__ void __launch_bounds__(kNttThreads, 3) ntt_inverse_kernel(const __grid_constant__ RingConst rc, u64 *data) {
    extern __shared__ __align__(16) u64 sm[];
    const u32 blk = D ? blockIdx.x & ((1u << D) - 1u) : 0u;
    const u32 row = blockIdx.x >> D;
    const LimbConst lc = rc.limb[row % rc.limbs];
    u64 *tile_ptr = data + (size_t)blockIdx.x * kN;   // tiles of a row are contiguous
    intt_tile<D>(threadIdx.x, sm, lc, D ? head_bits(D, blk) : 1u, LoadPairGlobal{tile_ptr}, tile_ptr);
}
// ------------------------------------------------------------------------------------------
// rescale: drop the last tower of every polynomial (PALISADE ModReduce under EXACTRESCALE)
//   pass 1: inverse transform of the last tower of each polynomial, read in place from the ciphertext batch,
//           coefficients to t[n_ct * 2][N]   (rings above one tile: + intt_tail_kernel on t)
//   pass 2: one CTA per (polynomial, remaining tower i, tile): forward transform of t switched into q_i, with
//           (c_i - .) * q_last^-1 folded into the store -> out[n_ct][2][L-1][N]
// ------------------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(kNttThreads, 3) rescale_intt_kernel(const __grid_constant__ RingConst rc, const u64 *ct,
                                                                      u64 *t) {
    extern __shared__ __align__(16) u64 sm[];
    const u32 L = rc.limbs;                       // towers of the input
    const u32 blk = D ? blockIdx.x & ((1u << D) - 1u) : 0u;
    const u64 ctp = blockIdx.x >> D;              // ct * 2 + poly
    const u64 n = (u64)kN << D;
    const LimbConst lc = rc.limb[L - 1];
    const u64 *src = ct + (ctp * L + (L - 1)) * n + (size_t)blk * kN;
    intt_tile<D>(threadIdx.x, sm, lc, D ? head_bits(D, blk) : 1u, LoadPairGlobal{src}, t + ctp * n + (size_t)blk * kN);
}
template <int D>
__global__ void __launch_bounds__(kNttThreads, 3) rescale_ntt_kernel(const __grid_constant__ RingConst rc,
                                                                     const __grid_constant__ RescaleConst rs,
                                                                     const u64 *ct, const u64 *t, u64 *out) {
    extern __shared__ __align__(16) u64 sm[];
    const int tid = threadIdx.x;
    const u32 L = rc.limbs, Lo = L - 1;
    const u32 blk = D ? blockIdx.x & ((1u << D) - 1u) : 0u;
    const u32 hb = D ? head_bits(D, blk) : 1u;
    const u32 row = blockIdx.x >> D;              // ctp * (L - 1) + i
    const u32 i = row % Lo;
    const u64 ctp = row / Lo;
    const u64 n = (u64)kN << D;
    const LimbConst lc = rc.limb[i];
    const u64 *t_row = t + ctp * n;
    const StoreRescale st{ct + (ctp * L + i) * n + (size_t)blk * kN, out + (size_t)row * n + (size_t)blk * kN, &lc,
                          rs.inv[i], rs.inv_s[i]};
    if (lc.dtw) {
        if (!cluster_head_for(D)) {
            ntt_forward_tile_d(tid, sm, lc, hb, DHeadLoad<D, DLoadSwitch>{DLoadSwitch{t_row, rs.q_last, &lc}, &lc, blk}, st);
        } else {
            cluster_head_d<D>(tid, sm, lc, DLoadSwitch{t_row, rs.q_last, &lc});
            ntt_forward_tile_d(tid, sm, lc, hb, DLoadTile{sm}, st);
        }
        return;
    }
    if (!cluster_head_for(D)) {
        fwd_phase_a(tid, sm, lc, hb, HeadLoad<D, LoadSwitch>{LoadSwitch{t_row, rs.q_last, &lc}, &lc, blk});
    } else {
        cluster_head<D>(tid, sm, lc, LoadSwitch{t_row, rs.q_last, &lc});
        fwd_phase_a(tid, sm, lc, hb, LoadTile{sm});
    }
    __syncthreads();
    fwd_phase_mid<7>(tid, sm, lc, hb);
    __syncthreads();
    fwd_phase_mid<4>(tid, sm, lc, hb);
    __syncthreads();
    fwd_phase_mid<1>(tid, sm, lc, hb);
    __syncthreads();
    fwd_phase_e(tid, sm, lc, hb, st);
}
constexpr int kTailThreads = 256;
template <int D>
__global__ void __launch_bounds__(kTailThreads) intt_tail_kernel(const __grid_constant__ RingConst rc, u64 *data) {
    const u32 row = blockIdx.y;
    const u32 i = blockIdx.x * kTailThreads + threadIdx.x;      // tile-local index
    const LimbConst &lc = rc.limb[row % rc.limbs];
    u64 *p = data + ((size_t)row << (kLogN + D)) + i;
    u64 a[1 << D];
#pragma unroll
    for (int k = 0; k < (1 << D); k++) a[k] = p[(size_t)k * kN];
    inv_tail<D>(a, lc);
#pragma unroll
    for (int k = 0; k < (1 << D); k++) p[(size_t)k * kN] = a[k];
}

// ------------------------------------------------------------------------------------------
// encode (+ sampler): one CTA per ciphertext, FFT tile in shared memory
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ const double *chunk_ptr(const double *x, u64 first, bool f32) {
    return f32 ? reinterpret_cast<const double *>(reinterpret_cast<const float *>(x) + first) : x + first;
}
__global__ void __launch_bounds__(kFftThreads) encode_kernel(const __grid_constant__ EncodeConst ec, const double *x,
                                                             u64 n, i64 *m, const u32 *sel, u32 f32) {
    extern __shared__ __align__(16) unsigned char smraw[];
    cplx *z = reinterpret_cast<cplx *>(smraw);
    const int tid = threadIdx.x;
    const u64 ct = blockIdx.x;
    u64 first = ct * ec.batch;
    u64 n_valid = n - first < ec.batch ? n - first : ec.batch;
    encode_load(tid, kFftThreads, z, ec, sel ? x : chunk_ptr(x, first, f32), n_valid, sel ? sel + first : nullptr, f32);
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
    i64 *mrow = m + ct * (u64)kN;
    for (u32 j = tid; j < (u32)kN; j += kFftThreads) mrow[j] = encode_coeff(z, ec, j);
}

// Ring dimensions / slot counts above one tile: one CTA per (ciphertext, FFT block), message coefficients only;
// sample_add_kernel then adds e0 and writes (v, e1) as a pass of its own (the sampler draws blocks of consecutive
// coefficients, the FFT blocks own strided ones).
template <int DF>
__global__ void __launch_bounds__(kFftThreads) encode_block_kernel(const __grid_constant__ EncodeConst ec, u32 n_ring,
                                                                   const double *x, u64 n, i64 *m, const u32 *sel, u32 f32) {
    extern __shared__ __align__(16) unsigned char smraw[];
    cplx *z = reinterpret_cast<cplx *>(smraw);
    const int tid = threadIdx.x;
    const u32 fb = DF ? blockIdx.x & ((1u << DF) - 1u) : 0u;
    const u64 ct = blockIdx.x >> DF;
    const u64 first = ct * ec.batch;
    const u64 n_valid = n - first < ec.batch ? n - first : ec.batch;
    encode_load_block<DF>(tid, kFftThreads, z, ec, sel ? x : chunk_ptr(x, first, f32), n_valid, fb, sel ? sel + first : nullptr, f32);
    __syncthreads();
    EncodeConst et = ec;                       // the tile's own geometry for the bit reversal
    if (ec.slots >= kFftFast) {
        et.slots = kFftFast; et.log_slots = 12;
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
    encode_finalize(tid, kFftThreads, z, et);
    __syncthreads();
    encode_emit_block<DF>(tid, kFftThreads, z, ec, fb, n_ring >> 1, m + ct * (u64)n_ring);
}
constexpr int kSampleThreads = 256;
__global__ void __launch_bounds__(kSampleThreads) sample_add_kernel(const __grid_constant__ SamplerConst sc, u32 n_ring,
                                                                    i64 *m_e0, uint8_t *ve) {
    __shared__ u64 cdf[32];
    __shared__ u32 cdf_hi[32];
    if (threadIdx.x < 32) {
        cdf[threadIdx.x] = sc.gauss_cdf[threadIdx.x];
        cdf_hi[threadIdx.x] = (u32)(sc.gauss_cdf[threadIdx.x] >> 32);
    }
    const u64 ct = blockIdx.x;
    uint8_t *ve_row = ve + ct * (u64)n_ring;
    sample_ternary_phase(threadIdx.x, kSampleThreads, sc.key, sc.ct_index0 + ct, n_ring, ve_row);
    __syncthreads();   // the bytes written above are completed by other threads of this block below
    sample_gauss_phase(threadIdx.x, kSampleThreads, nullptr, EncodeConst{}, sc.key, GaussTable{cdf_hi, cdf}, sc.ct_index0 + ct, n_ring,
                       m_e0 + ct * (u64)n_ring, ve_row);
}

// ------------------------------------------------------------------------------------------
// encrypt: per (ciphertext, limb) CTA, three forward NTTs through one shared-memory tile.
//   NTT(v)      -> v^ parked in the c1 row (re-read by the same thread, L2-resident)
//   NTT(m + e0) -> c0 = b*v^ + .     NTT(e1) -> c1 = a*v^ + .  (overwrites the parked row)
// ------------------------------------------------------------------------------------------
template <class Load, class Store>
__device__ __forceinline__ void ntt_forward_tile(int tid, u64 *sm, const LimbConst &lc, u32 hb, Load ld, Store st) {
    fwd_phase_a(tid, sm, lc, hb, ld);
    __syncthreads();
    fwd_phase_mid<7>(tid, sm, lc, hb);
    __syncthreads();
    fwd_phase_mid<4>(tid, sm, lc, hb);
    __syncthreads();
    fwd_phase_mid<1>(tid, sm, lc, hb);
    __syncthreads();
    fwd_phase_e(tid, sm, lc, hb, st);
}

template <int D>
__global__ void __launch_bounds__(kNttThreads, 3) encrypt_kernel(const __grid_constant__ RingConst rc, const i64 *m_e0,
                                                                 const uint8_t *ve, const tw_t *pk_b, const tw_t *pk_a,
                                                                 u64 *ct_out) {
    extern __shared__ __align__(16) u64 sm[];
    const int tid = threadIdx.x;
    const u32 L = rc.limbs;
    const u32 blk = D ? blockIdx.x & ((1u << D) - 1u) : 0u;
    const u32 hb = D ? head_bits(D, blk) : 1u;
    const u32 poly = blockIdx.x >> D;          // ct * L + l
    const u32 l = poly % L;
    const u64 ct = poly / L;
    const u64 n = (u64)kN << D;
    const LimbConst lc = rc.limb[l];
    const uint8_t *ve_row = ve + ct * n;
    const i64 *m_row = m_e0 + ct * n;
    const size_t tile_off = (size_t)blk * kN;
    u64 *c0 = ct_out + ((ct * 2 + 0) * L + l) * n + tile_off;
    u64 *c1 = ct_out + ((ct * 2 + 1) * L + l) * n + tile_off;
    const tw_t *kb = pk_b + (size_t)l * n + tile_off, *ka = pk_a + (size_t)l * n + tile_off;
    if (cluster_head_for(D)) {   // the 2^D tiles of this (ciphertext, limb) are one cluster: head stages through DSMEM
        if (lc.dtw) {
            cluster_head_d<D>(tid, sm, lc, DLoadTernary{ve_row});
            ntt_forward_tile_d(tid, sm, lc, hb, DLoadTile{sm}, StoreLazyPair{c1, 0});
            cluster_head_d<D>(tid, sm, lc, DLoadSigned{m_row, &lc});
            ntt_forward_tile_d(tid, sm, lc, hb, DLoadTile{sm}, StoreMulAdd{kb, c1, c0, &lc, false});
            cluster_head_d<D>(tid, sm, lc, DLoadGaussHi{ve_row});
            ntt_forward_tile_d(tid, sm, lc, hb, DLoadTile{sm}, StoreMulAdd{ka, c1, c1, &lc, false});
            return;
        }
        cluster_head<D>(tid, sm, lc, LoadTernary{ve_row, lc.q});
        ntt_forward_tile(tid, sm, lc, hb, LoadTile{sm}, StoreLazyPair{c1, 0});
        cluster_head<D>(tid, sm, lc, LoadSigned{m_row, &lc});
        ntt_forward_tile(tid, sm, lc, hb, LoadTile{sm}, StoreMulAdd{kb, c1, c0, &lc, false});
        cluster_head<D>(tid, sm, lc, LoadGaussHi{ve_row, lc.q});
        ntt_forward_tile(tid, sm, lc, hb, LoadTile{sm}, StoreMulAdd{ka, c1, c1, &lc, false});
        return;
    }
    if (lc.dtw) {
        ntt_forward_tile_d(tid, sm, lc, hb, DHeadLoad<D, DLoadTernary>{DLoadTernary{ve_row}, &lc, blk}, StoreLazyPair{c1, 0});
        __syncthreads();
        ntt_forward_tile_d(tid, sm, lc, hb, DHeadLoad<D, DLoadSigned>{DLoadSigned{m_row, &lc}, &lc, blk},
                           StoreMulAdd{kb, c1, c0, &lc, false});
        __syncthreads();
        ntt_forward_tile_d(tid, sm, lc, hb, DHeadLoad<D, DLoadGaussHi>{DLoadGaussHi{ve_row}, &lc, blk},
                           StoreMulAdd{ka, c1, c1, &lc, false});
        return;
    }
    ntt_forward_tile(tid, sm, lc, hb, HeadLoad<D, LoadTernary>{LoadTernary{ve_row, lc.q}, &lc, blk}, StoreLazyPair{c1, 0});
    __syncthreads();
    ntt_forward_tile(tid, sm, lc, hb, HeadLoad<D, LoadSigned>{LoadSigned{m_row, &lc}, &lc, blk},
                     StoreMulAdd{kb, c1, c0, &lc, false});
    __syncthreads();
    ntt_forward_tile(tid, sm, lc, hb, HeadLoad<D, LoadGaussHi>{LoadGaussHi{ve_row, lc.q}, &lc, blk},
                     StoreMulAdd{ka, c1, c1, &lc, false});
}

// Fused encrypt (ring dimension 8192): one CTA per ciphertext does encode + sampling in the FFT tile, then reuses the
// same 64 KiB tile for the 3 NTTs of every limb.  Nothing of a ciphertext's intermediate state has a buffer of its
// own: the packed (v, e1) draws (1 byte per coefficient) stay in 8 KiB of shared memory, and m + e0 (int64) is parked
// in the c0 row of the limb this CTA transforms LAST — an output row that is not written until that limb's own
// NTT(m + e0), which reads all of it (pass A) before any thread stores to it (pass E).  Compared with
// encode_kernel + sample_add_kernel + encrypt_kernel this samples once per ciphertext, is one launch, lets the
// latency-bound FFT / sampler phases of one CTA hide under the transforms of its two co-resident CTAs, and (round 2)
// moves no scratch bytes to DRAM.
#ifndef SHELFI_FUSED_CTAS_PER_SM
#define SHELFI_FUSED_CTAS_PER_SM 3
#endif
// The CTAs that share an SM should not march through the same limb together: the 60-bit limb's transforms live on the
// multiplier pipe, the scaling limbs' on the FP64 pipe, and two CTAs on different pipes overlap where two on the same
// one queue.  blockIdx parity does not tell co-resident CTAs apart (a wave puts blocks b, b + 148, b + 296 on one SM:
// all of one parity), so every CTA takes a turn number from a counter of the SM it landed on and the odd turns walk
// the limbs backwards.  The counters only ever grow; their parity is all that is used.
__device__ unsigned int g_sm_turn[512];
__device__ __forceinline__ unsigned int sm_turn() {
    unsigned int smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    return atomicAdd(&g_sm_turn[smid & 511u], 1u);
}
__global__ void __launch_bounds__(kNttThreads, SHELFI_FUSED_CTAS_PER_SM) encrypt_fused_kernel(const __grid_constant__ RingConst rc,
                                                                       const __grid_constant__ EncodeConst ec,
                                                                       const __grid_constant__ SamplerConst sc,
                                                                       const __grid_constant__ EncryptBatch eb,
                                                                       u64 n, const tw_t *pk_b, const tw_t *pk_a) {
    extern __shared__ __align__(16) u64 sm[];
    cplx *z = reinterpret_cast<cplx *>(sm);
    __shared__ u64 cdf[32];
    __shared__ u32 cdf_hi[32];
    __shared__ u32 turn;
    __shared__ __align__(16) uint8_t ve[kN];
    const int tid = threadIdx.x;
    if (tid == 0) turn = sm_turn();
    const u32 learner = blockIdx.x / eb.n_ct;
    const u64 ct = blockIdx.x % eb.n_ct;
    const double *x = eb.x[learner];
    u64 *ct_out = eb.ct[learner];
    const u64 ct_index = sc.ct_index0 + ct;
    if (tid < 32) {
        cdf[tid] = sc.gauss_cdf[tid];
        cdf_hi[tid] = (u32)(sc.gauss_cdf[tid] >> 32);
    }
    const u64 first = ct * ec.batch;
    const u64 n_valid = n - first < ec.batch ? n - first : ec.batch;
    const u32 L = rc.limbs;
    const u64 keep = make_keep_policy();
    if (!eb.sel) {   // this ciphertext's inputs on their way into L2 while the ternary draws are made (0.4 %)
        const char *xp = reinterpret_cast<const char *>(chunk_ptr(x, first, eb.f32));
        const size_t bytes = n_valid * (eb.f32 ? 4 : 8), off = (size_t)tid * 128;
        if (off < bytes) asm volatile("prefetch.global.L2 [%0];" ::"l"(xp + off));
    }
    sample_ternary_phase(tid, kNttThreads, eb.key[learner], ct_index, kN, ve);   // touches only `ve`
    encode_load(tid, kNttThreads, z, ec, eb.sel ? x : chunk_ptr(x, first, eb.f32), n_valid, eb.sel ? eb.sel + first : nullptr, eb.f32);
    __syncthreads();
    // odd turns on this SM walk the limbs backwards (see sm_turn)
    const u32 flip = (turn & 1u) ? L - 1 : 0u;
    const u32 l_last = flip ? 0u : L - 1;
    i64 *m_row = reinterpret_cast<i64 *>(ct_out + ((ct * 2 + 0) * L + l_last) * (u64)kN);
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
    sample_gauss_phase(tid, kNttThreads, z, ec, eb.key[learner], GaussTable{cdf_hi, cdf}, ct_index, kN, m_row, ve, keep);
    __syncthreads();  // the parked row and `ve` are visible to all threads of the block; the tile is free
    for (u32 li = 0; li < L; li++) {
        const u32 l = flip ? flip - li : li;
        const LimbConst lc = rc.limb[l];
        u64 *c0 = ct_out + ((ct * 2 + 0) * L + l) * (u64)kN;
        u64 *c1 = ct_out + ((ct * 2 + 1) * L + l) * (u64)kN;
        const StoreMulAdd st0{pk_b + (size_t)l * kN, c1, c0, &lc, true}, st1{pk_a + (size_t)l * kN, c1, c1, &lc, true};
        if (lc.dtw) {
            ntt_forward_tile_d(tid, sm, lc, 1u, DLoadTernary{ve}, StoreLazyPair{c1, keep});
            __syncthreads();
            ntt_forward_tile_d(tid, sm, lc, 1u, DLoadSigned{m_row, &lc}, st0);
            __syncthreads();
            ntt_forward_tile_d(tid, sm, lc, 1u, DLoadGaussHi{ve}, st1);
            __syncthreads();
            continue;
        }
        ntt_forward_tile(tid, sm, lc, 1u, LoadTernary{ve, lc.q}, StoreLazyPair{c1, keep});
        __syncthreads();
        ntt_forward_tile(tid, sm, lc, 1u, LoadSigned{m_row, &lc}, st0);
        __syncthreads();
        ntt_forward_tile(tid, sm, lc, 1u, LoadGaussHi{ve, lc.q}, st1);
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------
// key generation ([PAL] LPAlgorithmCKKS::KeyGen, ckks.cpp:46) on the device:
//   keygen_sample_kernel : rows[0..L) = s, rows[L..2L) = e reduced into every limb (coefficient form), a[L][N] uniform
//   (ntt_forward_kernel over the 2L rows)
//   keygen_combine_kernel: b = e^ - a * s^ per limb
// ------------------------------------------------------------------------------------------
constexpr int kKeygenThreads = 256;
__global__ void __launch_bounds__(kKeygenThreads) keygen_sample_kernel(const __grid_constant__ RingConst rc,
                                                                       const __grid_constant__ SamplerConst secret,
                                                                       const __grid_constant__ SamplerKey pub, u64 *rows,
                                                                       u64 *a) {
    const u32 n = rc.n, L = rc.limbs;
    const u32 t = blockIdx.x * kKeygenThreads + threadIdx.x, stride = gridDim.x * kKeygenThreads;
    __shared__ u64 cdf[32];
    __shared__ u32 cdf_hi[32];
    if (threadIdx.x < 32) {
        cdf[threadIdx.x] = secret.gauss_cdf[threadIdx.x];
        cdf_hi[threadIdx.x] = (u32)(secret.gauss_cdf[threadIdx.x] >> 32);
    }
    __syncthreads();
    for (u32 b = t; b < (n >> 5); b += stride) {
        u32 packed[8];
        sample_ternary_block(secret.key, 0, b, kStreamKeyS, kStreamKeySRetry, packed);
        for (u32 l = 0; l < L; l++)
#pragma unroll
            for (int i = 0; i < 32; i++)
                rows[(size_t)l * n + 32 * b + i] = small_to_mod((int)((packed[i >> 2] >> (8 * (i & 3))) & 3u) - 1, rc.limb[l].q);
    }
    for (u32 b = t; b < (n >> 4); b += stride) {
        int ev[16];
        keygen_error_block(secret.key, GaussTable{cdf_hi, cdf}, b, ev);
        for (u32 l = 0; l < L; l++)
#pragma unroll
            for (int i = 0; i < 16; i++) rows[(size_t)(L + l) * n + 16 * b + i] = small_to_mod(ev[i], rc.limb[l].q);
    }
    for (u32 i = t; i < L * n; i += stride) {
        const u32 l = i / n, j = i - l * n;
        a[i] = keygen_uniform(pub, l, j, rc.limb[l].q);
    }
}
__global__ void __launch_bounds__(kKeygenThreads) keygen_combine_kernel(const __grid_constant__ RingConst rc,
                                                                        const u64 *rows_hat, const u64 *a, u64 *b) {
    const u32 n = rc.n, L = rc.limbs;
    const u32 i = blockIdx.x * kKeygenThreads + threadIdx.x;
    if (i >= L * n) return;
    const LimbConst &lc = rc.limb[i / n];
    acc128 p{0, 0};
    mac128(p, a[i], rows_hat[i]);                       // a * s^ as a 128-bit integer
    const u64 as = reduce128(p, lc), e = rows_hat[(size_t)L * n + i];
    b[i] = csub(e + lc.q - as, lc.q);
}
void launch_keygen_sample(const RingConst &rc, const SamplerConst &secret, const SamplerKey &pub, u64 *rows, u64 *a,
                          cudaStream_t s) {
    keygen_sample_kernel<<<(rc.limbs * rc.n + kKeygenThreads - 1) / kKeygenThreads, kKeygenThreads, 0, s>>>(rc, secret, pub,
                                                                                                         rows, a);
}
void launch_keygen_combine(const RingConst &rc, const u64 *rows_hat, const u64 *a, u64 *b, cudaStream_t s) {
    keygen_combine_kernel<<<(rc.limbs * rc.n + kKeygenThreads - 1) / kKeygenThreads, kKeygenThreads, 0, s>>>(rc, rows_hat,
                                                                                                          a, b);
}
// known-answer tap of the device sampler: one ChaCha20 block (RFC 8439 2.3.2 test vector in tests/)
__global__ void chacha20_block_kernel(const __grid_constant__ SamplerKey key, u32 counter, u32 n0, u32 n1, u32 n2, u32 *out) {
    u32 o[16];
    chacha20_block(key, counter, n0, n1, n2, o);
    for (int i = 0; i < 16; i++) out[i] = o[i];
}
void launch_chacha20_block(const SamplerKey &key, u32 counter, u32 n0, u32 n1, u32 n2, u32 *out, cudaStream_t s) {
    chacha20_block_kernel<<<1, 1, 0, s>>>(key, counter, n0, n1, n2, out);
}

// ------------------------------------------------------------------------------------------
// decrypt stage 1: t = c0 + c1*s per limb, inverse NTT to coefficients
// ------------------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(kNttThreads, 3) decrypt_intt_kernel(const __grid_constant__ RingConst rc,
                                                                      const u64 *ct_in, const tw_t *sk, u64 *coeff) {
    extern __shared__ __align__(16) u64 sm[];
    const u32 L = rc.limbs;
    const u32 blk = D ? blockIdx.x & ((1u << D) - 1u) : 0u;
    const u32 poly = blockIdx.x >> D;
    const u32 l = poly % L;
    const u64 ct = poly / L;
    const u64 n = (u64)kN << D;
    const LimbConst lc = rc.limb[l];
    const size_t tile_off = (size_t)blk * kN;
    const u64 *c0 = ct_in + ((ct * 2 + 0) * L + l) * n + tile_off;
    const u64 *c1 = ct_in + ((ct * 2 + 1) * L + l) * n + tile_off;
    intt_tile<D>(threadIdx.x, sm, lc, D ? head_bits(D, blk) : 1u,
                 LoadDecryptPair{c0, c1, sk + (size_t)l * n + tile_off, lc.q}, coeff + (ct * L + l) * n + tile_off);
}

// ------------------------------------------------------------------------------------------
// decrypt stage 2: CRT -> double -> FFTSpecial -> real parts
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kFftThreads) decode_kernel(const __grid_constant__ EncodeConst ec,
                                                             const __grid_constant__ DecodeConst dc, u32 limbs,
                                                             const u64 *coeff, double *out, u64 n, i64 *tap_hi,
                                                             u64 *tap_lo, const u32 *sel) {
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
    decode_store(tid, kFftThreads, z, dc, sel ? out : out + first, length, sel ? sel + first : nullptr, first);
}

// Both stages in one launch (ring dimension 8192, at most two limbs): the CTA of a ciphertext runs the inverse
// transforms of its limbs one after the other in the tile, leaves the coefficients in `coeff` (they stay in L2) and
// decodes them in the same tile.  Same phases as decrypt_intt_kernel<0> + decode_kernel, hence the same doubles; one
// launch latency instead of two, and the FFT phases of one CTA run under the transforms of its neighbours.
__global__ void __launch_bounds__(kNttThreads, 3) decrypt_fused_kernel(const __grid_constant__ RingConst rc,
                                                                       const __grid_constant__ EncodeConst ec,
                                                                       const __grid_constant__ DecodeConst dc,
                                                                       const u64 *ct_in, const tw_t *sk, u64 *coeff,
                                                                       double *out, u64 n, const u32 *sel) {
    extern __shared__ __align__(16) u64 sm[];
    const int tid = threadIdx.x;
    const u64 ct = blockIdx.x;
    const u32 L = rc.limbs;
    for (u32 l = 0; l < L; l++) {
        const LimbConst lc = rc.limb[l];
        const u64 *c0 = ct_in + ((ct * 2 + 0) * L + l) * (u64)kN;
        const u64 *c1 = ct_in + ((ct * 2 + 1) * L + l) * (u64)kN;
        intt_tile<0>(tid, sm, lc, 1u, LoadDecryptPair{c0, c1, sk + (size_t)l * kN, lc.q}, coeff + (ct * L + l) * (u64)kN);
        __syncthreads();   // the tile is free again, and this CTA's coefficient stores are visible to all its threads
    }
    cplx *z = reinterpret_cast<cplx *>(sm);
    decode_load(tid, kFftThreads, z, ec, dc, L, coeff + ct * L * (u64)kN, nullptr, nullptr);
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
    const u64 first = ct * ec.batch;
    if (first >= n) return;
    const u64 length = n - first < ec.batch ? n - first : ec.batch;
    decode_store(tid, kFftThreads, z, dc, sel ? out : out + first, length, sel ? sel + first : nullptr, first);
}

// General decode (any ring dimension, limb count, slot count): crt_kernel turns the L residues of every coefficient
// into the centred integer's double (x scalingFactor^-1 * 2^p), decode_fft_kernel runs FFTSpecial per 4096-slot
// tile, decode_tail_kernel finishes slot counts above one tile.
constexpr int kCrtThreads = 128;
template <int L>
__global__ void __launch_bounds__(kCrtThreads) crt_kernel(const __grid_constant__ DecodeConst dc,
                                                           const __grid_constant__ CrtNConst cn, u32 n_ring,
                                                           const u64 *coeff, double *vals, u64 *tap8, i64 *tap_hi,
                                                           u64 *tap_lo) {
    const u64 ct = blockIdx.y;
    const u32 j = blockIdx.x * kCrtThreads + threadIdx.x;
    if (j >= n_ring) return;
    u64 t[8];
    const bool tap = tap8 || tap_hi;
    double v = decode_value<L>(coeff + ct * L * (u64)n_ring, n_ring, j, dc.crt, cn, tap ? t : nullptr);
    if (vals) vals[ct * (u64)n_ring + j] = fmul(v, dc.pre);
    if (tap8) {
#pragma unroll
        for (int k = 0; k < 8; k++) tap8[(ct * (u64)n_ring + j) * 8 + k] = t[k];
    }
    if (tap_hi) { tap_hi[ct * (u64)n_ring + j] = (i64)t[1]; tap_lo[ct * (u64)n_ring + j] = t[0]; }
}
// [PAL >= 1.10.6] CKKSPackedEncoding::Decode's noise flooding (opt-in: shelfi_set_option(DECODE_NOISE)), between
// crt_kernel and decode_fft_kernel: one CTA per ciphertext estimates the standard deviation of the decryption error
// from the part of the polynomial an exact real message cannot have (m(X) - m(1/X) = 2 Im z), floors it at sqrt(N)/8,
// and adds N(0, 2 sd^2) to the real and imaginary part of every slot.  vals = doubles of the CRT integers at scale 2^p.
constexpr int kNoiseThreads = 256;
__device__ __forceinline__ double block_sum(double v, double *red) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();                       // `red` may still be read from the previous reduction
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    double t = 0.0;
    for (int w = 0; w < kNoiseThreads / 32; w++) t += red[w];
    return t;
}
__global__ void __launch_bounds__(kNoiseThreads) decode_noise_kernel(const __grid_constant__ EncodeConst ec,
                                                                     const __grid_constant__ SamplerKey key, u64 ct_index0,
                                                                     u32 n_ring, double *vals) {
    __shared__ double red[kNoiseThreads / 32];
    const u64 ct = blockIdx.x;
    double *v = vals + ct * (u64)n_ring;
    const u32 n = ec.slots, gap = ec.gap, nh = n_ring >> 1;
    auto re = [&](u32 i) { return v[gap * i]; };
    auto im = [&](u32 i) { return v[gap * i + nh]; };
    double sd;
    if (n == 1) {
        sd = fabs(im(0));
    } else {
        // cv[i] = vec[i] - Conjugate(vec)[i]:  cv[0] = (0, 2 Im vec[0]),  cv[i] = (Re vec[i] + Im vec[n-i], Im vec[i] + Re vec[n-i])
        const double cv0 = 2.0 * im(0), half_re = re(n / 2) + im(n - n / 2);
        double part = 0.0;
        for (u32 i = 1 + threadIdx.x; i < n / 2; i += kNoiseThreads) part += (re(i) + im(n - i)) + (im(i) + re(n - i));
        const double mean = (2.0 * block_sum(part, red) + cv0 + 2.0 * half_re) / ((double)(2 * n) - 1.0);
        part = 0.0;
        for (u32 i = 1 + threadIdx.x; i < n / 2; i += kNoiseThreads) {
            const double a = re(i) + im(n - i) - mean, b = im(i) + re(n - i) - mean;
            part += a * a + b * b;
        }
        const double var = (2.0 * block_sum(part, red) + (cv0 - mean) * (cv0 - mean) + 2.0 * (half_re - mean) * (half_re - mean)) /
                           ((double)(2 * n) - 2.0);
        sd = 0.5 * sqrt(var);
    }
    const double floor_sd = 0.125 * sqrt((double)n_ring);
    sd = 1.4142135623730951 * (sd < floor_sd ? floor_sd : sd);        // sqrt(M + 1), M = 1
    __syncthreads();                                                  // every thread has read what it needs of v
    const u64 cti = ct_index0 + ct;
    for (u32 b = threadIdx.x; b < (n + 3) / 4; b += kNoiseThreads) {  // block b: slots 4 b .. 4 b + 3
        u32 o[16];
        chacha20_block(key, b, (u32)cti, (u32)(cti >> 32), kStreamDecodeNoise, o);
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const u32 i = 4 * b + k;
            if (i >= n) break;
            double z0, z1;
            normal_pair(o[4 * k], o[4 * k + 1], o[4 * k + 2], o[4 * k + 3], z0, z1);
            v[gap * i] += sd * z0;
            v[gap * i + nh] += sd * z1;
        }
    }
}
void launch_decode_noise(const RingConst &rc, const EncodeConst &ec, const SamplerKey &key, u64 ct_index0, double *vals,
                         u32 n_ct, cudaStream_t s) {
    if (n_ct) decode_noise_kernel<<<n_ct, kNoiseThreads, 0, s>>>(ec, key, ct_index0, rc.n, vals);
}

template <int DF>
__global__ void __launch_bounds__(kFftThreads) decode_fft_kernel(const __grid_constant__ EncodeConst ec,
                                                                 const __grid_constant__ DecodeConst dc, u32 n_ring,
                                                                 const double *vals, cplx *zs, double *out, u64 n,
                                                                 const u32 *sel) {
    extern __shared__ __align__(16) unsigned char smraw[];
    cplx *z = reinterpret_cast<cplx *>(smraw);
    const int tid = threadIdx.x;
    const u32 fb = DF ? blockIdx.x & ((1u << DF) - 1u) : 0u;
    const u64 ct = blockIdx.x >> DF;
    decode_load_block<DF>(tid, kFftThreads, z, ec, fb, n_ring, vals + ct * (u64)n_ring);
    __syncthreads();
    if (ec.slots >= kFftFast) {
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
    if (DF == 0) {
        u64 first = ct * ec.batch;
        if (first >= n) return;
        u64 length = n - first < ec.batch ? n - first : ec.batch;
        decode_store(tid, kFftThreads, z, dc, sel ? out : out + first, length, sel ? sel + first : nullptr, first);
    } else {
        cplx *dst = zs + ct * (u64)ec.slots + (u64)fb * kFftFast;
        for (u32 i = tid; i < kFftFast; i += kFftThreads) dst[i] = z[zsw(i)];
    }
}
template <int DF>
__global__ void __launch_bounds__(kTailThreads) decode_tail_kernel(const __grid_constant__ EncodeConst ec,
                                                                   const __grid_constant__ DecodeConst dc,
                                                                   const cplx *zs, double *out, u64 n, const u32 *sel) {
    const u64 ct = blockIdx.y;
    const u32 i = blockIdx.x * kTailThreads + threadIdx.x;
    u64 first = ct * ec.batch;
    if (first >= n) return;
    u64 length = n - first < ec.batch ? n - first : ec.batch;
    decode_tail<DF>(i, zs + ct * (u64)ec.slots, ec, dc, sel ? out : out + first, length, sel ? sel + first : nullptr, first);
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
    const u32 l = (u32)((base >> rc.logn) % rc.limbs);
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

// out (+)= ct (*) pt_hat: EvalMult(ct, plaintext) + EvalAdd.  pt_hat[ct][L][N] multiplies both polynomials of the
// ciphertext; the 128-bit product (plus the running value) is reduced once, so `out` always holds canonical residues.
constexpr int kMacThreads = 256;
__global__ void __launch_bounds__(kMacThreads) ct_pt_mac_kernel(const __grid_constant__ RingConst rc, const u64 *ct,
                                                                const u64 *pt_hat, u64 *out, u32 accumulate) {
    const u64 e = ((u64)blockIdx.x * kMacThreads + threadIdx.x) * 2;        // two adjacent coefficients per thread
    const u32 L = rc.limbs;
    const u64 row = e >> rc.logn;                                           // (ct * 2 + poly) * L + l
    const u32 l = (u32)(row % L);
    const u64 ctp = row / L;                                                // ct * 2 + poly
    const u64 pt_off = (((ctp >> 1) * L + l) << rc.logn) + (e & ((u64)rc.n - 1));
    const LimbConst &lc = rc.limb[l];
    const ulonglong2 a = *reinterpret_cast<const ulonglong2 *>(ct + e);
    const ulonglong2 b = *reinterpret_cast<const ulonglong2 *>(pt_hat + pt_off);
    acc128 s0{0, 0}, s1{0, 0};
    if (accumulate) {
        const ulonglong2 o = *reinterpret_cast<const ulonglong2 *>(out + e);
        s0.lo = o.x; s1.lo = o.y;
    }
    mac128(s0, a.x, b.x);
    mac128(s1, a.y, b.y);
    *reinterpret_cast<ulonglong2 *>(out + e) = make_ulonglong2(reduce128(s0, lc), reduce128(s1, lc));
}

// plaintext weighted sum (selective encryption: the parameters that are not encrypted); streaming, 8 B per lane
constexpr int kPlainThreads = 256;
constexpr int kPlainPerThread = 4;
__global__ void __launch_bounds__(kPlainThreads) plain_weighted_sum_kernel(const __grid_constant__ PlainWavgArgs pa, u64 n,
                                                                           double *out) {
    const u64 base = (u64)blockIdx.x * (kPlainThreads * kPlainPerThread) + threadIdx.x;
    double acc[kPlainPerThread];
#pragma unroll
    for (int p = 0; p < kPlainPerThread; p++) {
        const u64 i = base + (u64)p * kPlainThreads;
        acc[p] = (pa.accumulate && i < n) ? out[i] : 0.0;
    }
    for (u32 k = 0; k < pa.k; k++) {
        const double *__restrict__ x = pa.in[k];
        const double w = pa.w[k];
        double v[kPlainPerThread];
#pragma unroll
        for (int p = 0; p < kPlainPerThread; p++) {
            const u64 i = base + (u64)p * kPlainThreads;
            v[p] = i < n ? x[i] : 0.0;
        }
#pragma unroll
        for (int p = 0; p < kPlainPerThread; p++) acc[p] = fadd(acc[p], fmul(w, v[p]));
    }
#pragma unroll
    for (int p = 0; p < kPlainPerThread; p++) {
        const u64 i = base + (u64)p * kPlainThreads;
        if (i < n) out[i] = acc[p];
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
    u64 *dst = dense + run * (u64)br.n;
    for (u32 i = blockIdx.y * kPackThreads + threadIdx.x; i < br.n; i += gridDim.y * kPackThreads)
        dst[i] = load_unaligned(image, a0 + 8ull * i);
}

__global__ void __launch_bounds__(kPackThreads) pack_kernel(const __grid_constant__ BlobRuns br, const u64 *dense,
                                                            u64 *image) {
    const u64 run = blockIdx.x;
    const u64 ct = run / br.runs;
    const u32 r = (u32)(run % br.runs);
    const u64 a0 = run_offset(br, ct, r);
    const u64 w0 = a0 >> 3, w1 = (a0 + 8ull * br.n - 1) >> 3;
    const u64 *src = dense + run * (u64)br.n;
    for (u64 w = w0 + blockIdx.y * kPackThreads + threadIdx.x; w <= w1; w += (u64)gridDim.y * kPackThreads)
        pack_word(image, a0, w, src, br.n);
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
    u64 *dst = acc + run * (u64)br.n;
    for (u32 i = blockIdx.y * kPackThreads + threadIdx.x; i < br.n; i += gridDim.y * kPackThreads) {
        acc128 a{first ? 0ull : dst[i], 0ull};
        mac128(a, load_unaligned(image, a0 + 8ull * i), c);
        dst[i] = reduce128(a, lc);
    }
}

// ------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------
// D (tiles per polynomial = 1 << D) and DF (FFT tiles per ciphertext = 1 << DF) are compile-time in the kernels
#define DISPATCH_D(d, ...)                                   \
    switch (d) {                                             \
        case 0: { constexpr int D = 0; __VA_ARGS__; break; } \
        case 1: { constexpr int D = 1; __VA_ARGS__; break; } \
        default: { constexpr int D = 2; __VA_ARGS__; break; } \
    }
#define DISPATCH_L(l, ...)                                   \
    switch (l) {                                             \
        case 1: { constexpr int L = 1; __VA_ARGS__; break; } \
        case 2: { constexpr int L = 2; __VA_ARGS__; break; } \
        case 3: { constexpr int L = 3; __VA_ARGS__; break; } \
        case 4: { constexpr int L = 4; __VA_ARGS__; break; } \
        case 5: { constexpr int L = 5; __VA_ARGS__; break; } \
        case 6: { constexpr int L = 6; __VA_ARGS__; break; } \
        case 7: { constexpr int L = 7; __VA_ARGS__; break; } \
        default: { constexpr int L = 8; __VA_ARGS__; break; } \
    }
static u32 fft_log_blocks(const EncodeConst &ec) { return ec.log_slots > 12 ? ec.log_slots - 12 : 0; }
static size_t fft_smem(const EncodeConst &ec) { return (ec.slots < kFftFast ? ec.slots : kFftFast) * sizeof(cplx); }

cudaError_t kernels_init() {
    cudaError_t e;
#define OPT_IN(k, bytes)                                                                      \
    e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes));  \
    if (e != cudaSuccess) return e;
#define OPT_IN_D(k, bytes) OPT_IN(k<0>, bytes) OPT_IN(k<1>, bytes) OPT_IN(k<2>, bytes)
    OPT_IN_D(ntt_forward_kernel, NTT_SMEM_BYTES)
    OPT_IN_D(ntt_inverse_kernel, NTT_SMEM_BYTES)
    OPT_IN(ntt_forward_shfl_kernel, NTT_SMEM_BYTES)
    OPT_IN_D(encrypt_kernel, NTT_SMEM_BYTES)
    OPT_IN_D(plain_ntt_kernel, NTT_SMEM_BYTES)
    OPT_IN_D(decrypt_intt_kernel, NTT_SMEM_BYTES)
    OPT_IN_D(rescale_intt_kernel, NTT_SMEM_BYTES)
    OPT_IN_D(rescale_ntt_kernel, NTT_SMEM_BYTES)
    OPT_IN(encrypt_fused_kernel, NTT_SMEM_BYTES)
    OPT_IN(encode_kernel, kFftFast * sizeof(cplx))
    OPT_IN(decode_kernel, kFftFast * sizeof(cplx))
    OPT_IN(decrypt_fused_kernel, NTT_SMEM_BYTES)
    OPT_IN_D(encode_block_kernel, kFftFast * sizeof(cplx))
    OPT_IN_D(decode_fft_kernel, kFftFast * sizeof(cplx))
#undef OPT_IN_D
#undef OPT_IN
    return cudaSuccess;
}

// launch with the 2^D tiles of a polynomial as one thread-block cluster (consecutive blockIdx.x)
template <class... KArgs, class... Args>
static void launch_tiles(void (*kernel)(KArgs...), unsigned polys, unsigned log_blocks, cudaStream_t s, Args &&...args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(polys << log_blocks);
    cfg.blockDim = dim3(kNttThreads);
    cfg.dynamicSmemBytes = NTT_SMEM_BYTES;
    cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 1u << log_blocks;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = cluster_head_for((int)log_blocks) ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
void launch_ntt_forward(const RingConst &rc, const u64 *in, u64 *out, u32 rows, cudaStream_t s) {
    if (!rows) return;
    DISPATCH_D(rc.log_blocks, launch_tiles(ntt_forward_kernel<D>, rows, D, s, rc, in, out))
}
void launch_ntt_forward_shfl(const RingConst &rc, const u64 *in, u64 *out, u32 rows, cudaStream_t s) {
    if (rows && rc.log_blocks == 0) ntt_forward_shfl_kernel<<<rows, kNttThreads, NTT_SMEM_BYTES, s>>>(rc, in, out);
}
void launch_ntt_inverse(const RingConst &rc, u64 *data, u32 rows, cudaStream_t s) {
    if (!rows) return;
    DISPATCH_D(rc.log_blocks, ntt_inverse_kernel<D><<<rows << D, kNttThreads, NTT_SMEM_BYTES, s>>>(rc, data))
    if (rc.log_blocks == 1) intt_tail_kernel<1><<<dim3(kN / kTailThreads, rows), kTailThreads, 0, s>>>(rc, data);
    if (rc.log_blocks == 2) intt_tail_kernel<2><<<dim3(kN / kTailThreads, rows), kTailThreads, 0, s>>>(rc, data);
}
// rings above one tile: message coefficients, then the sampler pass (the fused kernel covers ring dimension 8192)
void launch_encode_sample(const RingConst &rc, const EncodeConst &ec, const SamplerConst &sc, const double *x, u64 n,
                          u32 n_ct, i64 *m_e0, uint8_t *ve, cudaStream_t s, const u32 *sel, bool f32) {
    if (!n_ct) return;
    launch_encode_only(rc, ec, x, n, n_ct, m_e0, s, sel, f32);
    sample_add_kernel<<<n_ct, kSampleThreads, 0, s>>>(sc, rc.n, m_e0, ve);
}
void launch_encode_only(const RingConst &rc, const EncodeConst &ec, const double *x, u64 n, u32 n_ct, i64 *m,
                        cudaStream_t s, const u32 *sel, bool f32) {
    if (!n_ct) return;
    if (rc.log_blocks == 0) {
        encode_kernel<<<n_ct, kFftThreads, ec.slots * sizeof(cplx), s>>>(ec, x, n, m, sel, f32 ? 1u : 0u);
        return;
    }
    DISPATCH_D(fft_log_blocks(ec), encode_block_kernel<D><<<n_ct << D, kFftThreads, fft_smem(ec), s>>>(ec, rc.n, x, n, m, sel, f32 ? 1u : 0u))
}
void launch_encrypt(const RingConst &rc, const i64 *m_e0, const uint8_t *ve, const tw_t *pk_b, const tw_t *pk_a,
                    u64 *ct, u32 n_ct, cudaStream_t s) {
    if (!n_ct) return;
    DISPATCH_D(rc.log_blocks, launch_tiles(encrypt_kernel<D>, n_ct * rc.limbs, D, s,
                                  rc, m_e0, ve, pk_b, pk_a, ct))
}
void launch_encrypt_fused(const RingConst &rc, const EncodeConst &ec, const SamplerConst &sc, const EncryptBatch &eb,
                          u64 n, const tw_t *pk_b, const tw_t *pk_a, cudaStream_t s) {
    if (eb.n_ct && eb.k)   // ring dimension 8192 only (one tile per polynomial); callers check rc.log_blocks == 0
        encrypt_fused_kernel<<<eb.n_ct * eb.k, kNttThreads, NTT_SMEM_BYTES, s>>>(rc, ec, sc, eb, n, pk_b, pk_a);
}
void launch_decrypt_intt(const RingConst &rc, const u64 *ct, const tw_t *sk, u64 *coeff, u32 n_ct, cudaStream_t s) {
    if (!n_ct) return;
    const u32 rows = n_ct * rc.limbs;
    DISPATCH_D(rc.log_blocks, decrypt_intt_kernel<D><<<rows << D, kNttThreads, NTT_SMEM_BYTES, s>>>(rc, ct, sk, coeff))
    if (rc.log_blocks == 1) intt_tail_kernel<1><<<dim3(kN / kTailThreads, rows), kTailThreads, 0, s>>>(rc, coeff);
    if (rc.log_blocks == 2) intt_tail_kernel<2><<<dim3(kN / kTailThreads, rows), kTailThreads, 0, s>>>(rc, coeff);
}
void launch_rescale(const RingConst &rc, const RescaleConst &rs, const u64 *ct, u64 *t, u64 *out, u32 n_ct,
                    cudaStream_t s) {
    if (!n_ct || rc.limbs < 2) return;
    const u32 polys = n_ct * 2;
    DISPATCH_D(rc.log_blocks, rescale_intt_kernel<D><<<polys << D, kNttThreads, NTT_SMEM_BYTES, s>>>(rc, ct, t))
    if (rc.log_blocks) {   // outer inverse stages + N^-1 over t, whose rows all belong to the dropped tower
        RingConst last = rc;
        last.limbs = 1;
        last.limb[0] = rc.limb[rc.limbs - 1];
        if (rc.log_blocks == 1) intt_tail_kernel<1><<<dim3(kN / kTailThreads, polys), kTailThreads, 0, s>>>(last, t);
        if (rc.log_blocks == 2) intt_tail_kernel<2><<<dim3(kN / kTailThreads, polys), kTailThreads, 0, s>>>(last, t);
    }
    DISPATCH_D(rc.log_blocks, launch_tiles(rescale_ntt_kernel<D>, polys * (rc.limbs - 1), D, s, rc, rs, ct, t, out))
}
void launch_decode(const RingConst &rc, const EncodeConst &ec, const DecodeConst &dc, const u64 *coeff, u32 n_ct,
                   double *out, u64 n, cudaStream_t s, const u32 *sel) {
    if (n_ct)   // ring dimension 8192, at most two limbs: CRT and FFT in one kernel
        decode_kernel<<<n_ct, kFftThreads, ec.slots * sizeof(cplx), s>>>(ec, dc, rc.limbs, coeff, out, n, nullptr,
                                                                        nullptr, sel);
}
void launch_decrypt_fused(const RingConst &rc, const EncodeConst &ec, const DecodeConst &dc, const u64 *ct,
                          const tw_t *sk, u64 *coeff, u32 n_ct, double *out, u64 n, cudaStream_t s, const u32 *sel) {
    static_assert(kNttThreads == kFftThreads && NTT_SMEM_BYTES >= kFftFast * sizeof(cplx), "one tile serves both stages");
    if (n_ct)   // ring dimension 8192, at most two limbs (callers check)
        decrypt_fused_kernel<<<n_ct, kNttThreads, NTT_SMEM_BYTES, s>>>(rc, ec, dc, ct, sk, coeff, out, n, sel);
}
void launch_crt(const RingConst &rc, const DecodeConst &dc, const CrtNConst &cn, const u64 *coeff, u32 n_ct,
                double *vals, u64 *tap8, i64 *tap_hi, u64 *tap_lo, cudaStream_t s) {
    if (!n_ct) return;
    dim3 grid((rc.n + kCrtThreads - 1) / kCrtThreads, n_ct);
    DISPATCH_L(rc.limbs, crt_kernel<L><<<grid, kCrtThreads, 0, s>>>(dc, cn, rc.n, coeff, vals, tap8, tap_hi, tap_lo))
}
void launch_decode_fft(const RingConst &rc, const EncodeConst &ec, const DecodeConst &dc, const double *vals, cplx *zs,
                       u32 n_ct, double *out, u64 n, cudaStream_t s, const u32 *sel) {
    if (!n_ct) return;
    const u32 df = fft_log_blocks(ec);
    DISPATCH_D(df, decode_fft_kernel<D><<<n_ct << D, kFftThreads, fft_smem(ec), s>>>(ec, dc, rc.n, vals, zs, out, n, sel))
    if (df == 1) decode_tail_kernel<1><<<dim3(kFftFast / kTailThreads, n_ct), kTailThreads, 0, s>>>(ec, dc, zs, out, n, sel);
    if (df == 2) decode_tail_kernel<2><<<dim3(kFftFast / kTailThreads, n_ct), kTailThreads, 0, s>>>(ec, dc, zs, out, n, sel);
}
void launch_crt_only(const RingConst &rc, const DecodeConst &dc, const u64 *coeff, u32 n_ct, i64 *hi, u64 *lo,
                     cudaStream_t s) {
    EncodeConst ec{};
    ec.slots = kN / 2; ec.gap = 1; ec.log_slots = kLogN - 1; ec.batch = kN / 2;
    if (n_ct)
        decode_kernel<<<n_ct, kFftThreads, ec.slots * sizeof(cplx), s>>>(ec, dc, rc.limbs, coeff, nullptr, 0, hi, lo, nullptr);
}
void launch_weighted_sum(const RingConst &rc, const WavgArgs &wa, u64 *out, u64 n_ct, cudaStream_t s) {
    u64 elems = n_ct * 2 * rc.limbs * (u64)rc.n;
    if (elems) weighted_sum_kernel<<<(unsigned)(elems / kWavgElemsPerCta), kWavgThreads, 0, s>>>(rc, wa, out);
}
// One warp per segment: cereal's bookkeeping between the coefficient runs (8 to ~150 bytes per segment, one of 2.7 KB
// holding the crypto context).  Byte stores only: the neighbouring run bytes belong to pack_kernel.
__global__ void __launch_bounds__(256) write_segments_kernel(const SegDesc *__restrict__ desc,
                                                             const unsigned char *__restrict__ bytes,
                                                             unsigned char *image, u32 n_seg) {
    const u32 seg = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (seg >= n_seg) return;
    const SegDesc d = desc[seg];
    for (u32 i = threadIdx.x & 31; i < d.len; i += 32) image[d.dst + i] = bytes[d.src + i];
}
void launch_write_segments(const SegDesc *desc, const unsigned char *bytes, unsigned char *image, u32 n_seg, cudaStream_t s) {
    if (n_seg) write_segments_kernel<<<(n_seg + 7) / 8, 256, 0, s>>>(desc, bytes, image, n_seg);
}

// ------------------------------------------------------------------------------------------
// decrypt-gather: stream-ordered barrier between the GPUs of one box (flags in peer-mapped memory)
// ------------------------------------------------------------------------------------------
// Lane r publishes this rank's epoch to rank r and waits for rank r's epoch in the local array.  The fence orders the
// decode kernel's peer stores (earlier in the stream) before the flag; release/acquire at system scope pair the flag
// with the reader's later loads.  Epochs only grow, so a flag never has to be reset.
__global__ void __launch_bounds__(32) peer_barrier_kernel(const __grid_constant__ PeerFlags pf, u32 n_ranks, u32 me,
                                                          u64 epoch, u64 timeout_ns, double *poison, u64 poison_stride) {
    const u32 r = threadIdx.x;
    if (r >= n_ranks) return;
    __threadfence_system();
    unsigned long long *theirs = pf.flags[r] + me;
    const unsigned long long *mine = pf.flags[me] + r;
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(theirs), "l"(epoch) : "memory");
    unsigned long long t0, t1, seen;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(mine) : "memory");
        if (seen >= epoch) break;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (t1 - t0 > timeout_ns) {   // a peer never arrived: record it and let the stream go on ...
            pf.flags[me][8] = epoch;
            // ... but not with a half-gathered model that looks whole: the slice rank r owed us starts with a NaN
            if (poison) poison[(u64)r * poison_stride] = __longlong_as_double(0x7ff8000000000000ll);
            break;
        }
        __nanosleep(100);
    }
    __threadfence_system();
}
void launch_peer_barrier(const PeerFlags &pf, u32 n_ranks, u32 me, u64 epoch, u64 timeout_ns, double *poison,
                         u64 poison_stride, cudaStream_t s) {
    peer_barrier_kernel<<<1, 32, 0, s>>>(pf, n_ranks, me, epoch, timeout_ns, poison, poison_stride);
}

static dim3 run_grid(const BlobRuns &br, u64 n_ct) { return dim3((unsigned)(n_ct * br.runs), br.n / 2048); }
void launch_plain_ntt(const RingConst &rc, const i64 *m, u64 *pt_hat, u32 n_ct, cudaStream_t s) {
    if (!n_ct) return;
    DISPATCH_D(rc.log_blocks, launch_tiles(plain_ntt_kernel<D>, n_ct * rc.limbs, D, s, rc, m, pt_hat))
}
void launch_ct_pt_mac(const RingConst &rc, const u64 *ct, const u64 *pt_hat, u64 *out, u64 n_ct, bool accumulate,
                      cudaStream_t s) {
    const u64 elems = n_ct * 2 * rc.limbs * (u64)rc.n;
    if (elems) ct_pt_mac_kernel<<<(unsigned)(elems / (2 * kMacThreads)), kMacThreads, 0, s>>>(rc, ct, pt_hat, out,
                                                                                              accumulate ? 1u : 0u);
}
void launch_plain_weighted_sum(const PlainWavgArgs &pa, u64 n, double *out, cudaStream_t s) {
    const u64 per = (u64)kPlainThreads * kPlainPerThread;
    if (n) plain_weighted_sum_kernel<<<(unsigned)((n + per - 1) / per), kPlainThreads, 0, s>>>(pa, n, out);
}
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

// ------------------------------------------------------------------------------------------
// decrypt with its scratch in the module image (shelfi_decrypt_static_device): no allocation at all
// ------------------------------------------------------------------------------------------
// No counterpart in the reference; added for callers that may not allocate
// (stream capture, serving loops).  The array lives in every device that loads the library until the process exits.
// Layout of one call, in u64 words: coeff[n_ct][limbs][N] | vals[n_ct][N] (general decode only) | zs[n_ct][slots]
// complex (general decode with slots > kFftFast only).  Every word a call reads it has written first (the inverse
// transforms fill coeff, crt_kernel fills vals, decode_fft_kernel fills zs), so nothing has to start at zero; the
// array keeps the previous call's contents, and calls that overlap on the GPU race on it.
static __device__ __align__(256) unsigned long long g_static_scratch[kStaticScratchWords];

// In host code g_static_scratch is only a stand-in for the symbol: the device address is looked up once per device,
// with that device current, and kept.
static cudaError_t static_scratch_ptr(u64 **out) {
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static u64 *ptr_of[kMaxDevices] = {nullptr};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr_of[dev]) {
        u64 *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, g_static_scratch);
        if (e != cudaSuccess) return e;
        if (!p) return cudaErrorInvalidSymbol;
        ptr_of[dev] = p;
    }
    *out = ptr_of[dev];
    return cudaSuccess;
}

u64 static_scratch_words(const RingConst &rc, const EncodeConst &ec, u64 n_ct, bool general) {
    u64 per = (u64)rc.limbs * rc.n;
    if (general) {
        per += rc.n;
        if (ec.slots > kFftFast) per += 2ull * ec.slots;
    }
    return n_ct * per;
}

cudaError_t launch_decrypt_static(const RingConst &rc, const EncodeConst &ec, const DecodeConst &dc, const CrtNConst &cn,
                                  bool general, const u64 *ct, const tw_t *sk, u32 n_ct, double *out, u64 n,
                                  cudaStream_t s) {
    if (!n_ct) return cudaSuccess;
    if (static_scratch_words(rc, ec, n_ct, general) > kStaticScratchWords) return cudaErrorInvalidValue;
    u64 *base = nullptr;
    cudaError_t e = static_scratch_ptr(&base);
    if (e != cudaSuccess) return e;
    u64 *coeff = base;
    launch_decrypt_intt(rc, ct, sk, coeff, n_ct, s);
    if (!general) {
        launch_decode(rc, ec, dc, coeff, n_ct, out, n, s, nullptr);
    } else {
        double *vals = reinterpret_cast<double *>(base + (u64)n_ct * rc.limbs * rc.n);
        cplx *zs = reinterpret_cast<cplx *>(base + (u64)n_ct * (rc.limbs + 1) * rc.n);   // 16-byte aligned: N is
        launch_crt(rc, dc, cn, coeff, n_ct, vals, nullptr, nullptr, nullptr, s);         // a multiple of 8192 words
        launch_decode_fft(rc, ec, dc, vals, zs, n_ct, out, n, s, nullptr);
    }
    return cudaGetLastError();
}

}  // namespace shelfi
