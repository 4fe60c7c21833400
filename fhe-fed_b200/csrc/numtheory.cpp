#include "numtheory.h"

#include <cmath>

namespace shelfi {
namespace nt {

uint64_t pow_mod(uint64_t a, uint64_t e, uint64_t q) {
    uint64_t r = 1 % q;
    a %= q;
    for (; e; e >>= 1) {
        if (e & 1) r = mul_mod(r, a, q);
        a = mul_mod(a, a, q);
    }
    return r;
}

bool is_prime(uint64_t n) {
    static const uint64_t bases[] = {2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 37};  // deterministic below 2^64
    if (n < 2) return false;
    for (uint64_t p : bases) {
        if (n == p) return true;
        if (n % p == 0) return false;
    }
    uint64_t d = n - 1;
    int s = 0;
    while ((d & 1) == 0) { d >>= 1; s++; }
    for (uint64_t a : bases) {
        uint64_t x = pow_mod(a, d, n);
        if (x == 1 || x == n - 1) continue;
        bool witness = true;
        for (int r = 1; r < s && witness; r++) {
            x = mul_mod(x, x, n);
            if (x == n - 1) witness = false;
        }
        if (witness) return false;
    }
    return true;
}

uint64_t first_prime(uint32_t bits, uint32_t m) {
    uint64_t lower = (uint64_t)1 << bits;
    uint64_t q = lower + (m - lower % m) + 1;   // PALISADE: 2^bits + (m - 2^bits mod m) + 1, then steps of m
    while (!is_prime(q)) q += m;
    return q;
}

uint64_t previous_prime(uint64_t q, uint32_t m) {
    do q -= m; while (!is_prime(q));
    return q;
}

uint64_t min_primitive_root(uint32_t m, uint64_t q) {
    // any generator of the order-m subgroup, then scan its odd powers for the smallest
    uint64_t g = 0;
    for (uint64_t x = 2; x < q && !g; x++) {
        uint64_t c = pow_mod(x, (q - 1) / m, q);
        if (pow_mod(c, m / 2, q) == q - 1) g = c;
    }
    uint64_t g2 = mul_mod(g, g, q), cur = g, best = g;
    for (uint32_t i = 1; i < m / 2; i++) {
        cur = mul_mod(cur, g2, q);
        if (cur < best) best = cur;
    }
    return best;
}

uint64_t next_prime(uint64_t q, uint32_t m) {
    do q += m; while (!is_prime(q));
    return q;
}

uint32_t num_large_digits(uint32_t mult_depth) { return mult_depth > 3 ? 3 : mult_depth > 0 ? 2 : 1; }

uint32_t select_ring_dim(uint32_t scale_bits, uint32_t mult_depth) {
    // log2(Q*P) budget per ring dimension at 128-bit classical security, ternary secret
    static const uint32_t dims[] = {1024, 2048, 4096, 8192, 16384, 32768};
    static const uint32_t max_bits[] = {27, 54, 109, 218, 438, 881};
    uint32_t dnum = num_large_digits(mult_depth);
    uint32_t q_bits = 60 + mult_depth * scale_bits;    // first modulus 60 bits + one scaling modulus per level
    uint32_t part = (q_bits + dnum - 1) / dnum;        // bits per digit of the hybrid key switch
    uint32_t total = q_bits + ((part + 59) / 60) * 60; // its auxiliary moduli
    for (int i = 0; i < 6; i++)
        if (total <= max_bits[i]) return dims[i];
    return 65536;
}

int modulus_chain(uint32_t scale_bits, uint32_t mult_depth, uint32_t ring_dim, uint64_t *q) {
    // EXACTRESCALE: the inner moduli alternate just below / just above the running scaling factor so that the
    // scale stays near 2^bits after each rescale; two limbs (multDepth 1) reduce to the rule the fixtures pin.
    const uint32_t m = 2 * ring_dim;
    const int L = (int)mult_depth + 1;
    if (L < 1 || L > 8) return -1;
    q[L - 1] = first_prime(scale_bits, m);
    if (L == 1) return L;
    double sf = (double)q[L - 1];
    uint32_t cnt = 0;
    for (int i = L - 2; i >= 1; i--, cnt++) {
        sf = std::pow(sf, 2) / (double)q[i + 1];
        uint64_t sf_int = (uint64_t)std::llround(sf), rem = sf_int % m;
        uint64_t cand = (cnt % 2) == 0 ? sf_int - m - rem + 1 : sf_int + m - rem + 1;
        for (bool clash = true; clash;) {
            cand = (cnt % 2) == 0 ? previous_prime(cand, m) : next_prime(cand, m);
            clash = false;
            for (int j = i + 1; j < L; j++) clash = clash || cand == q[j];
        }
        q[i] = cand;
    }
    q[0] = previous_prime(first_prime(60, m), m);
    return L;
}

}  // namespace nt
}  // namespace shelfi
