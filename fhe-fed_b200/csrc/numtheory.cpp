#include "numtheory.h"

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

uint32_t select_ring_dim(uint32_t scale_bits) {
    // log2(Q*P) budget per ring dimension at 128-bit classical security, ternary secret
    static const uint32_t dims[] = {1024, 2048, 4096, 8192, 16384, 32768};
    static const uint32_t max_bits[] = {27, 54, 109, 218, 438, 881};
    uint32_t q_bits = 60 + scale_bits;                 // first modulus 60 bits + one scaling modulus
    uint32_t part = (q_bits + 1) / 2;                  // dnum = 2 digits
    uint32_t total = q_bits + ((part + 59) / 60) * 60; // auxiliary moduli of the hybrid key switch
    for (int i = 0; i < 6; i++)
        if (total <= max_bits[i]) return dims[i];
    return 65536;
}

}  // namespace nt
}  // namespace shelfi
