// wire.cpp — see wire.h.  One bidirectional description of each object (the way cereal itself is
// written): the same visitor runs over an input cursor, an output buffer, or an output buffer that
// leaves holes where the GPU will drop the coefficient runs.
#include "wire.h"

#include <cstring>

namespace shelfi {
namespace wire {

namespace {

constexpr uint32_t kSame = 0x40000000u;  // cereal: static type == dynamic type
constexpr uint32_t kNew = 0x80000000u;   // cereal: first occurrence of this shared pointer / type name
const char kParamsName[] = "lbcrypto::LPCryptoParametersCKKS<lbcrypto::DCRTPoly>";
const char kSchemeName[] = "lbcrypto::LPPublicKeyEncryptionSchemeCKKS<lbcrypto::DCRTPoly>";

// C++ types whose cereal class-version word appears once per archive, at first use
enum Versioned {
    kVCryptoContext, kVCkksParams, kVRlweParams, kVLpParams, kVDcrtParams, kVElemParamsBig, kVBigInteger,
    kVNativeParams, kVElemParamsNative, kVNativeInteger, kVEncodingParams, kVCkksScheme, kVSchemeBase,
    kVKey, kVKeyBase, kVCryptoObject, kVCiphertext, kVCiphertextObject, kVDcrtPoly, kVNativePoly, kVNativeVector,
    kVersionedCount
};

struct Ids {                       // shared-pointer ids handed out so far in this archive
    uint32_t next = 1;
    uint32_t context = 0, dcrt = 0, native[kMaxLimbs] = {0};
};

class Archive {
  public:
    // reading
    Archive(const uint8_t *data, size_t len) : rd_(data), len_(len), reading_(true) {}
    // writing; with holes == true coefficient runs are skipped and segments recorded
    explicit Archive(bool holes) : reading_(false), holes_(holes) {}

    bool reading() const { return reading_; }
    bool ok() const { return err_.empty(); }
    const std::string &error() const { return err_; }
    size_t position() const { return reading_ ? pos_ : base_ + buf_.size(); }
    void fail(const std::string &what) {
        if (err_.empty()) err_ = what + " at byte " + std::to_string(position());
    }

    void raw(void *p, size_t n) {
        if (reading_) {
            if (!ok() || pos_ + n > len_) { fail("truncated input"); std::memset(p, 0, n); return; }
            std::memcpy(p, rd_ + pos_, n);
            pos_ += n;
        } else {
            const uint8_t *b = static_cast<const uint8_t *>(p);
            buf_.insert(buf_.end(), b, b + n);
        }
    }
    template <class T> void value(T &v) { raw(&v, sizeof(T)); }
    template <class T> void literal(T want, const char *what) {
        T v = want;
        value(v);
        if (reading_ && v != want) fail(std::string("unexpected ") + what);
    }
    void same() { literal<uint32_t>(kSame, "polymorphic marker"); }
    void version(Versioned t, uint32_t v) {
        if (seen_ & (1u << t)) return;
        seen_ |= 1u << t;
        literal<uint32_t>(v, "class version");
    }
    uint32_t new_pointer() {
        uint32_t id = ids.next++;
        literal<uint32_t>(kNew | id, "shared pointer id");
        return id;
    }
    void type_name(uint32_t poly_id, const char *name) {
        literal<uint32_t>(kNew | poly_id, "polymorphic type id");
        uint64_t n = std::strlen(name);
        literal<uint64_t>(n, "type name length");
        char tmp[96];
        std::memcpy(tmp, name, n);
        raw(tmp, n);
        if (reading_ && std::memcmp(tmp, name, n)) fail("unexpected type name");
    }
    uint32_t peek32() {
        uint32_t v = 0;
        if (reading_ && pos_ + 4 <= len_) std::memcpy(&v, rd_ + pos_, 4);
        return v;
    }
    uint32_t peek32_at(size_t off) {
        uint32_t v = 0;
        if (reading_ && pos_ + off + 4 <= len_) std::memcpy(&v, rd_ + pos_ + off, 4);
        return v;
    }
    // a run of n coefficient words: the reader notes where it is (and copies if dst != nullptr), the
    // writer either appends src or leaves a hole
    uint64_t run(uint64_t *dst, const uint64_t *src, size_t n) {
        uint64_t at = position();
        if (reading_) {
            if (!ok() || pos_ + n * 8 > len_) { fail("truncated coefficient run"); return at; }
            if (dst) std::memcpy(dst, rd_ + pos_, n * 8);
            pos_ += n * 8;
        } else if (holes_) {
            flush_segment();
            base_ = at + n * 8;
        } else {
            const uint8_t *b = reinterpret_cast<const uint8_t *>(src);
            buf_.insert(buf_.end(), b, b + n * 8);
        }
        return at;
    }
    void flush_segment() {
        if (!buf_.empty()) {
            segments.push_back(Segment{base_, std::move(buf_)});
            base_ += segments.back().bytes.size();
            buf_.clear();
        }
    }
    std::vector<uint8_t> take() { return std::move(buf_); }

    Ids ids;
    std::vector<Segment> segments;

  private:
    const uint8_t *rd_ = nullptr;
    size_t len_ = 0, pos_ = 0;
    bool reading_, holes_ = false;
    std::vector<uint8_t> buf_;
    uint64_t base_ = 0;  // blob offset of buf_[0] (hole mode)
    uint32_t seen_ = 0;
    std::string err_;
};

void big_integer(Archive &ar, uint32_t (&words)[110], uint16_t &msb) {
    for (auto &w : words) ar.value(w);
    ar.value(msb);
}
void big_zero(Archive &ar) {
    for (int i = 0; i < 110; i++) ar.literal<uint32_t>(0, "big integer word");
    ar.literal<uint16_t>(0, "big integer msb");
}

void context_body(Archive &ar, ContextParams &p) {
    ar.version(kVCryptoContext, 1);
    ar.type_name(1, kParamsName);
    ar.new_pointer();
    ar.version(kVCkksParams, 0); ar.version(kVRlweParams, 0); ar.version(kVLpParams, 0);
    // ILDCRTParams<BigInteger>
    ar.same();
    ar.ids.dcrt = ar.new_pointer();
    ar.version(kVDcrtParams, 1); ar.version(kVElemParamsBig, 1);
    ar.value(p.cyclotomic_order); ar.value(p.ring_dim);
    ar.literal<uint8_t>(1, "isPowerOfTwo");
    ar.version(kVBigInteger, 1);
    big_integer(ar, p.big_modulus, p.big_modulus_msb);
    big_zero(ar); big_zero(ar); big_zero(ar);
    uint64_t limbs = p.limbs;
    ar.value(limbs);
    if (limbs < 1 || limbs > (uint64_t)kMaxLimbs) { ar.fail("unsupported tower count"); return; }
    p.limbs = (uint32_t)limbs;
    for (uint32_t l = 0; l < p.limbs; l++) {  // ILNativeParams per tower
        ar.same();
        ar.ids.native[l] = ar.new_pointer();
        ar.version(kVNativeParams, 1); ar.version(kVElemParamsNative, 1);
        ar.literal<uint32_t>(p.cyclotomic_order, "tower cyclotomic order");
        ar.literal<uint32_t>(p.ring_dim, "tower ring dimension");
        ar.literal<uint8_t>(1, "isPowerOfTwo");
        ar.version(kVNativeInteger, 1);
        ar.value(p.modulus[l]); ar.value(p.root[l]);
        ar.literal<uint64_t>(0, "bigModulus"); ar.literal<uint64_t>(0, "bigRootOfUnity");
    }
    big_zero(ar);  // originalModulus
    // EncodingParams
    ar.same();
    ar.new_pointer();
    ar.version(kVEncodingParams, 1);
    ar.value(p.plaintext_modulus);
    ar.literal<uint64_t>(0, "plaintextRootOfUnity"); ar.literal<uint64_t>(0, "plaintextBigModulus");
    ar.literal<uint64_t>(0, "plaintextBigRootOfUnity"); ar.literal<uint32_t>(0, "plaintextGenerator");
    ar.value(p.batch_size);
    ar.value(p.sigma); ar.value(p.assurance); ar.value(p.hermite);
    if (ar.reading()) {  // 8 or 9 integers precede the scheme's type id
        p.rlwe_int_count = 0;
        for (int k = 8; k <= 9; k++)
            if (ar.peek32_at(4u * k) == (kNew | 2u)) { p.rlwe_int_count = k; break; }
        if (!p.rlwe_int_count) { ar.fail("unrecognised RLWE parameter block"); return; }
    }
    for (int i = 0; i < p.rlwe_int_count; i++) ar.value(p.rlwe_ints[i]);
    ar.type_name(2, kSchemeName);
    ar.new_pointer();
    ar.version(kVCkksScheme, 0); ar.version(kVSchemeBase, 0);
    ar.value(p.enabled_features);
    ar.literal<uint64_t>(4, "scheme id length");
    char id[4] = {'C', 'K', 'K', 'S'};
    ar.raw(id, 4);
    if (ar.reading() && std::memcmp(id, "CKKS", 4)) ar.fail("scheme id is not CKKS");
}

// shared_ptr<CryptoContextImpl>: full body the first time, a back reference afterwards
void context(Archive &ar, ContextParams &p) {
    ar.same();
    bool first = ar.reading() ? (ar.peek32() & kNew) != 0 : ar.ids.context == 0;
    if (first) {
        ar.ids.context = ar.new_pointer();
        context_body(ar, p);
    } else {
        ar.literal<uint32_t>(ar.ids.context, "crypto context back reference");
    }
}

// DCRTPoly in EVALUATION format; offsets[l] receives where tower l's coefficients are.
// *limbs = towers of this polynomial: the context's count at level 0, fewer after a rescale (0 on input when reading =
// accept what the blob says).  A rescaled polynomial no longer shares the context's parameter object: PALISADE's
// DCRTPolyImpl::DropLastElement gives it a copy with the last tower popped and the modulus divided, and cereal writes
// that object in line — a new pointer id, the ElemParams fields, and back references to the ILNativeParams the context
// carried.  No fixture pins this form.
void dcrt_poly(Archive &ar, const ContextParams &p, uint32_t *limbs, uint64_t *dst, const uint64_t *src,
               uint64_t *offsets) {
    ar.version(kVDcrtPoly, 1);
    uint64_t count = *limbs;
    ar.value(count);
    if (count < 1 || count > p.limbs || (*limbs && count != *limbs)) { ar.fail("unexpected tower count"); return; }
    *limbs = (uint32_t)count;
    const uint32_t lc = *limbs;
    for (uint32_t l = 0; l < lc && ar.ok(); l++) {
        ar.version(kVNativePoly, 1);
        ar.same();
        ar.literal<uint8_t>(1, "values present flag");
        ar.version(kVNativeVector, 1);
        ar.literal<uint64_t>(p.ring_dim, "vector length");
        uint64_t at = ar.run(dst ? dst + (size_t)l * p.ring_dim : nullptr, src ? src + (size_t)l * p.ring_dim : nullptr,
                             p.ring_dim);
        if (offsets) offsets[l] = at;
        ar.literal<uint64_t>(p.modulus[l], "tower modulus");
        ar.literal<uint32_t>(0, "tower format (EVALUATION)");
        ar.same();
        ar.literal<uint32_t>(ar.ids.native[l], "tower parameter reference");
    }
    ar.literal<uint32_t>(0, "polynomial format (EVALUATION)");
    ar.same();
    if (lc == p.limbs) { ar.literal<uint32_t>(ar.ids.dcrt, "element parameter reference"); return; }
    ar.new_pointer();
    ar.version(kVDcrtParams, 1); ar.version(kVElemParamsBig, 1);
    ar.literal<uint32_t>(p.cyclotomic_order, "reduced parameters: cyclotomic order");
    ar.literal<uint32_t>(p.ring_dim, "reduced parameters: ring dimension");
    ar.literal<uint8_t>(1, "isPowerOfTwo");
    ar.version(kVBigInteger, 1);
    ContextParams reduced = p;
    reduced.limbs = lc;
    set_big_modulus(&reduced);
    for (uint32_t w : reduced.big_modulus) ar.literal<uint32_t>(w, "reduced modulus word");
    ar.literal<uint16_t>(reduced.big_modulus_msb, "reduced modulus msb");
    big_zero(ar); big_zero(ar); big_zero(ar);
    ar.literal<uint64_t>(lc, "reduced tower count");
    for (uint32_t l = 0; l < lc; l++) {
        ar.same();
        ar.literal<uint32_t>(ar.ids.native[l], "reduced tower parameter reference");
    }
    big_zero(ar);  // originalModulus
}

void key_tag(Archive &ar, char (&tag)[33]) {
    ar.literal<uint64_t>(32, "key tag length");
    ar.raw(tag, 32);
    tag[32] = 0;
}

void key_header(Archive &ar, KeyFile &k, char (&tag)[33]) {
    ar.literal<uint8_t>(1, "archive endianness flag");
    ar.same();
    ar.new_pointer();
    ar.version(kVKey, 0); ar.version(kVKeyBase, 0); ar.version(kVCryptoObject, 0);
    context(ar, k.ctx);
    if (!ar.reading()) std::memcpy(tag, k.tag.c_str(), k.tag.size() < 32 ? k.tag.size() : 32);
    key_tag(ar, tag);
    if (ar.reading()) k.tag = tag;
}

void key_file(Archive &ar, KeyFile &k, int n_polys) {
    char tag[33] = {0};
    key_header(ar, k, tag);
    if (!ar.ok()) return;
    if (n_polys == 2) ar.literal<uint64_t>(2, "public key element count");
    size_t words = (size_t)k.ctx.limbs * k.ctx.ring_dim;
    if (ar.reading()) k.polys.assign(words * n_polys, 0);
    else if (k.polys.size() != words * n_polys) { ar.fail("key polynomial size mismatch"); return; }
    for (int i = 0; i < n_polys && ar.ok(); i++) {
        uint32_t limbs = k.ctx.limbs;  // keys always carry the full chain
        dcrt_poly(ar, k.ctx, &limbs, k.polys.data() + i * words, k.polys.data() + i * words, nullptr);
    }
}

void ciphertext(Archive &ar, ContextParams &p, CiphertextMeta &m, uint64_t *runs) {
    ar.same();
    ar.new_pointer();
    ar.version(kVCiphertext, 1); ar.version(kVCiphertextObject, 0);
    context(ar, p);
    if (!ar.ok()) return;
    key_tag(ar, m.tag);
    ar.literal<uint64_t>(2, "ciphertext element count");
    if (!ar.reading() && !m.limbs) m.limbs = p.limbs;
    if (ar.reading()) m.limbs = 0;
    dcrt_poly(ar, p, &m.limbs, nullptr, nullptr, runs);
    if (ar.ok()) dcrt_poly(ar, p, &m.limbs, nullptr, nullptr, runs + m.limbs);
    ar.value(m.depth); ar.value(m.level); ar.value(m.scaling_factor); ar.value(m.encoding_type);
    ar.new_pointer();                                   // shared_ptr<map<string, Metadata>>, always empty here
    ar.literal<uint64_t>(0, "metadata map size");
}

std::string finish(Archive &ar, size_t len) {
    if (ar.ok() && ar.position() != len) ar.fail("trailing bytes");
    return ar.error();
}

}  // namespace

std::string parse_context(const uint8_t *data, size_t len, ContextParams *out) {
    Archive ar(data, len);
    ar.literal<uint8_t>(1, "archive endianness flag");
    context(ar, *out);
    return finish(ar, len);
}
std::string parse_public_key(const uint8_t *data, size_t len, KeyFile *out) {
    Archive ar(data, len);
    key_file(ar, *out, 2);
    return finish(ar, len);
}
std::string parse_private_key(const uint8_t *data, size_t len, KeyFile *out) {
    Archive ar(data, len);
    key_file(ar, *out, 1);
    return finish(ar, len);
}

std::string map_ciphertext_blob(const uint8_t *data, size_t len, bool single, BlobMap *out) {
    Archive ar(data, len);
    ar.literal<uint8_t>(1, "archive endianness flag");
    uint64_t n_ct = 1;
    if (!single) ar.value(n_ct);
    if (!ar.ok() || n_ct == 0 || n_ct > len / 64) return "implausible ciphertext count";
    out->n_ct = n_ct;
    out->size = len;
    out->meta.assign(n_ct, CiphertextMeta());
    uint64_t runs[2 * kMaxLimbs];
    size_t start1 = 0;
    // ciphertexts 0 and 1 are walked field by field; from 2 on the layout repeats with a fixed stride,
    // so only the per-ciphertext scalars (ids, metadata) are read at their known offsets
    for (uint64_t j = 0; j < n_ct && ar.ok(); j++) {
        size_t start = ar.position();
        ciphertext(ar, out->ctx, out->meta[j], runs);
        if (!ar.ok()) break;
        if (j == 0) out->ct_limbs = out->meta[0].limbs;
        else if (out->meta[j].limbs != out->ct_limbs) { ar.fail("ciphertexts of one blob at different levels"); break; }
        if (j == 0) std::memcpy(out->run0, runs, sizeof runs);
        if (j == 1) { std::memcpy(out->run1, runs, sizeof runs); start1 = start; }
        if (j == 2) out->stride = start - start1;
    }
    if (ar.ok() && n_ct == 2) out->stride = ar.position() - start1;
    return finish(ar, len);
}

std::vector<uint8_t> emit_context(const ContextParams &p) {
    Archive ar(false);
    ContextParams c = p;
    ar.literal<uint8_t>(1, "");
    context(ar, c);
    return ar.take();
}
std::vector<uint8_t> emit_public_key(const KeyFile &k) {
    Archive ar(false);
    KeyFile c = k;
    key_file(ar, c, 2);
    return ar.ok() ? ar.take() : std::vector<uint8_t>();
}
std::vector<uint8_t> emit_private_key(const KeyFile &k) {
    Archive ar(false);
    KeyFile c = k;
    key_file(ar, c, 1);
    return ar.ok() ? ar.take() : std::vector<uint8_t>();
}

BlobPlan plan_ciphertext_blob(const ContextParams &p, uint64_t n_ct, const std::vector<CiphertextMeta> &meta) {
    BlobPlan plan;
    Archive ar(true);
    ContextParams c = p;
    ar.literal<uint8_t>(1, "");
    ar.value(n_ct);
    uint64_t runs[2 * kMaxLimbs];
    size_t start1 = 0;
    for (uint64_t j = 0; j < n_ct; j++) {
        size_t start = ar.position();
        CiphertextMeta m = meta.size() == 1 ? meta[0] : meta[j];
        ciphertext(ar, c, m, runs);
        if (j == 0) std::memcpy(plan.run0, runs, sizeof runs);
        if (j == 1) { std::memcpy(plan.run1, runs, sizeof runs); start1 = start; }
        if (j == 2) plan.stride = start - start1;
    }
    if (n_ct == 2) plan.stride = ar.position() - start1;
    plan.size = ar.position();
    ar.flush_segment();
    plan.segments = std::move(ar.segments);
    return plan;
}

uint64_t ciphertext_blob_size(const ContextParams &p, uint64_t n_ct, uint32_t ct_limbs) {
    std::vector<CiphertextMeta> m(1);
    m[0].limbs = ct_limbs;
    if (n_ct <= 2) return plan_ciphertext_blob(p, n_ct, m).size;
    uint64_t s1 = plan_ciphertext_blob(p, 1, m).size, s2 = plan_ciphertext_blob(p, 2, m).size;
    return s1 + (n_ct - 1) * (s2 - s1);
}

void set_big_modulus(ContextParams *p) {
    uint32_t le[110] = {0};
    le[0] = 1;
    for (uint32_t l = 0; l < p->limbs; l++) {
        unsigned __int128 carry = 0;
        for (int i = 0; i < 110; i++) {
            unsigned __int128 cur = (unsigned __int128)le[i] * p->modulus[l] + carry;
            le[i] = (uint32_t)cur;
            carry = cur >> 32;
        }
    }
    int top = 109;
    while (top > 0 && !le[top]) top--;
    int bits = 0;
    for (uint32_t v = le[top]; v; v >>= 1) bits++;
    p->big_modulus_msb = (uint16_t)(top * 32 + bits);
    for (int i = 0; i < 110; i++) p->big_modulus[i] = le[109 - i];
}

}  // namespace wire
}  // namespace shelfi
