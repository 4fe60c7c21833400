// numtheory.h — host-side modular arithmetic and parameter generation for the CKKS context:
// the rules PALISADE's genCryptoContextCKKS(multDepth=1, scaleFactorBits, batchSize) follows
// (palisade_pybind/SHELFI_FHE/src/ckks.cpp:26-28), verified against both fixture contexts.
#pragma once
#include <cstdint>

namespace shelfi {
namespace nt {

typedef unsigned __int128 u128;

inline uint64_t mul_mod(uint64_t a, uint64_t b, uint64_t q) { return (uint64_t)((u128)a * b % q); }
inline uint64_t add_mod(uint64_t a, uint64_t b, uint64_t q) { uint64_t s = a + b; return s >= q ? s - q : s; }
inline uint64_t sub_mod(uint64_t a, uint64_t b, uint64_t q) { return a >= b ? a - b : a + q - b; }
uint64_t pow_mod(uint64_t a, uint64_t e, uint64_t q);
inline uint64_t inv_mod(uint64_t a, uint64_t q) { return pow_mod(a, q - 2, q); }
inline uint64_t shoup(uint64_t w, uint64_t q) { return (uint64_t)(((u128)w << 64) / q); }
inline uint64_t signed_mod(int64_t x, uint64_t q) {
    if (x >= 0) return (uint64_t)x % q;
    uint64_t m = ((uint64_t)0 - (uint64_t)x) % q;
    return m ? q - m : 0;
}
bool is_prime(uint64_t n);
uint64_t first_prime(uint32_t bits, uint32_t m);      // smallest prime >= 2^bits congruent 1 mod m
uint64_t previous_prime(uint64_t q, uint32_t m);      // next smaller prime congruent 1 mod m
uint64_t min_primitive_root(uint32_t m, uint64_t q);  // numerically smallest primitive m-th root of unity (m = 2^k)
uint64_t next_prime(uint64_t q, uint32_t m);          // next larger prime congruent 1 mod m
uint32_t num_large_digits(uint32_t mult_depth);       // PALISADE's default dnum for the HYBRID key switch
// HEStd_128_classic table, HYBRID key switching; ckks.cpp:26 passes mult_depth = 1
uint32_t select_ring_dim(uint32_t scale_bits, uint32_t mult_depth = 1);
// q[0..mult_depth] of genCryptoContextCKKS(multDepth, scaleFactorBits) under EXACTRESCALE; returns the limb count
int modulus_chain(uint32_t scale_bits, uint32_t mult_depth, uint32_t ring_dim, uint64_t *q);

}  // namespace nt
}  // namespace shelfi
