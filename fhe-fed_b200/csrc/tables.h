// tables.h — host-side precomputation shared by the engine and the CPU replay harness (tests/emu):
// per-limb constants with twiddle tables and the lazy-reduction schedule, FFT stage tables,
// the sampler's CDF, CRT constants.  No CUDA here.
#pragma once
#include <string>
#include <vector>

#include "ckks_core.cuh"

namespace shelfi {

// Fills *lc (tw / itw pointers left null) and the two tables; returns "" or why the modulus is unusable.
// logn = log2(ring dimension), 13..15; the tables hold 2^logn entries.
// dtw (optional): centred forward twiddles as doubles for the FP64-pipe transform; left empty (and lc->dtw null) when
// the limb does not qualify (q > 1.25 * 2^52).
std::string make_limb_const(uint64_t q, uint64_t psi, uint32_t logn, LimbConst *lc, std::vector<tw_t> *tw,
                            std::vector<tw_t> *itw, std::vector<double> *dtw = nullptr);
// [PAL] DiscreteFourierTransform::PreComputeTable(M) flattened per stage: entry (len/2 - 1) + j
void make_fft_tables(uint32_t cyclotomic_order, uint32_t slots, std::vector<cplx> *inv, std::vector<cplx> *fwd);
// floor(2^64 * P(|X| <= k)), k = 0..31, for the discrete Gaussian of parameter sigma
void build_gauss_cdf(double sigma, uint64_t cdf[32]);
Crt2Const make_crt2(uint64_t q0, uint64_t q1);  // q1 == 0: single limb
CrtNConst make_crtn(const uint64_t *q, uint32_t limbs);  // Garner constants for any limb count

}  // namespace shelfi
