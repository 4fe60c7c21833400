// wire.h — PALISADE / cereal PortableBinary wire format for the objects on the SHELFI path:
// CryptoContext, LPPublicKey, LPPrivateKey and vector<Ciphertext<DCRTPoly>>.
//
// The reference moves these through Serial::Serialize / Deserialize
// (palisade_pybind/SHELFI_FHE/src/ckks.cpp:12,16,20,41,48,53,98-101,172-176,278-281,308-310); the
// grammar is the one decoded from the reference's binary fixtures.  Coefficient
// payload is never copied here: readers report where the 2L runs of N u64 live, writers emit the
// bytes around them, and the GPU pack/unpack kernels move the runs.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace shelfi {
namespace wire {

constexpr int kMaxLimbs = 8;

struct ContextParams {
    uint32_t cyclotomic_order = 0, ring_dim = 0;
    uint32_t limbs = 0;
    uint64_t modulus[kMaxLimbs] = {0}, root[kMaxLimbs] = {0};
    uint32_t big_modulus[110] = {0};      // most-significant word first
    uint16_t big_modulus_msb = 0;
    uint64_t plaintext_modulus = 0;       // scaleFactorBits
    uint32_t batch_size = 0;              // slots
    float sigma = 0, assurance = 0, hermite = 0;
    uint32_t rlwe_ints[9] = {0};          // relinWindow, depth, maxDepth, [extra], mode, stdLevel, rsTech, ksTech, dnum
    int rlwe_int_count = 8;               // 8 (older PALISADE) or 9
    uint32_t enabled_features = 0;
};

struct KeyFile {
    ContextParams ctx;
    std::string tag;                      // 32 ASCII hex chars
    std::vector<uint64_t> polys;          // public key: b then a (2*L*N); private key: s (L*N); EVALUATION form
};

struct CiphertextMeta {
    uint64_t depth = 1, level = 0;
    double scaling_factor = 0;
    uint32_t encoding_type = 4;           // CKKSPacked
    char tag[33] = {0};
    uint32_t limbs = 0;                   // towers per polynomial; 0 when writing = the context's count (level 0)
};

// Where things are inside a serialized vector<Ciphertext>.
struct BlobMap {
    ContextParams ctx;
    uint64_t n_ct = 0;
    uint32_t ct_limbs = 0;                // towers per polynomial (< ctx.limbs after a rescale)
    uint64_t size = 0;
    uint64_t run0[2 * kMaxLimbs] = {0};   // byte offsets of ciphertext 0's coefficient runs
    uint64_t run1[2 * kMaxLimbs] = {0};   // ... of ciphertext 1 (absolute)
    uint64_t stride = 0;                  // ciphertext j >= 1 starts stride bytes after j-1
    std::vector<CiphertextMeta> meta;     // one per ciphertext
};

// Each parser returns an empty string on success, else a description of the first mismatch.
std::string parse_context(const uint8_t *data, size_t len, ContextParams *out);
std::string parse_public_key(const uint8_t *data, size_t len, KeyFile *out);
std::string parse_private_key(const uint8_t *data, size_t len, KeyFile *out);
// single != false: a lone Ciphertext (no vector length word), as in the reference's CT1.txt
std::string map_ciphertext_blob(const uint8_t *data, size_t len, bool single, BlobMap *out);

std::vector<uint8_t> emit_context(const ContextParams &p);
std::vector<uint8_t> emit_public_key(const KeyFile &k);
std::vector<uint8_t> emit_private_key(const KeyFile &k);

// A header segment of an outgoing blob: `bytes` go at byte `offset`; the gaps between segments are
// the coefficient runs.
struct Segment { uint64_t offset; std::vector<uint8_t> bytes; };
struct BlobPlan {
    uint64_t size = 0;
    uint64_t run0[2 * kMaxLimbs] = {0}, run1[2 * kMaxLimbs] = {0}, stride = 0;
    std::vector<Segment> segments;
};
// meta: one entry (applied to every ciphertext) or n_ct entries; meta[].limbs < p.limbs writes rescaled ciphertexts
BlobPlan plan_ciphertext_blob(const ContextParams &p, uint64_t n_ct, const std::vector<CiphertextMeta> &meta);
uint64_t ciphertext_blob_size(const ContextParams &p, uint64_t n_ct, uint32_t ct_limbs = 0);

void set_big_modulus(ContextParams *p);  // fills big_modulus / msb from the limb moduli

}  // namespace wire
}  // namespace shelfi
