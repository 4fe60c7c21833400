// shelfi_ckks.hpp — header-only C++ mirror of the reference's CKKS class for C++ callers.
//
// Same constructor and the same three *_cpp methods as palisade_pybind/SHELFI_FHE/include/ckks.h:27-54
// (used by src/main.cpp:48,67,73), implemented on the C ABI of shelfi_b200.h.  Link with -lshelfi_b200.
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "shelfi_b200.h"

class CKKS {
  public:
    // include/ckks.h:40 — CKKS(string scheme, uint batchSize, uint scaleFactorBits, string cryptodir)
    CKKS(std::string scheme, unsigned batchSize, unsigned scaleFactorBits, std::string cryptodir, int device = 0)
        : cryptodir_(cryptodir) {
        if (shelfi_ckks_create(scheme.c_str(), batchSize, scaleFactorBits, cryptodir.c_str(), device, &h_))
            throw std::runtime_error("shelfi_ckks_create failed");
    }
    ~CKKS() { shelfi_ckks_destroy(h_); }
    CKKS(const CKKS &) = delete;
    CKKS &operator=(const CKKS &) = delete;

    int genCryptoContextAndKeyGen() { return shelfi_gen_crypto_context_and_keygen(h_); }  // ckks.cpp:25-59
    void loadCryptoParams() {                                                              // ckks.cpp:11-23
        int missing = shelfi_load_crypto_params(h_);
        if (missing < 0) throw std::runtime_error(shelfi_last_error(h_));
        load_status = missing;
    }
    // ckks.cpp:107-167
    std::string encrypt_cpp(const std::vector<double> &learner_Data) {
        uint8_t *blob = nullptr;
        uint64_t len = 0;
        check(shelfi_encrypt(h_, learner_Data.data(), learner_Data.size(), &blob, &len));
        return take(blob, len);
    }
    // ckks.cpp:323-371 — a size mismatch returns "" like the reference (:324-327)
    std::string computeWeightedAverage_cpp(const std::vector<std::string> &learners_Data,
                                           const std::vector<float> &scalingFactors) {
        if (learners_Data.size() != scalingFactors.size()) return "";
        std::vector<const uint8_t *> ptr;
        std::vector<uint64_t> len;
        for (const auto &s : learners_Data) {
            ptr.push_back(reinterpret_cast<const uint8_t *>(s.data()));
            len.push_back(s.size());
        }
        uint8_t *blob = nullptr;
        uint64_t n = 0;
        check(shelfi_compute_weighted_average(h_, ptr.data(), len.data(), scalingFactors.data(),
                                              (uint32_t)scalingFactors.size(), &blob, &n));
        return take(blob, n);
    }
    // ckks.cpp:217-260
    std::vector<double> decrypt_cpp(const std::string &learner_Data, unsigned long data_dimensions) {
        std::vector<double> out(data_dimensions);
        check(shelfi_decrypt(h_, reinterpret_cast<const uint8_t *>(learner_Data.data()), learner_Data.size(),
                             data_dimensions, out.data()));
        return out;
    }

    shelfi_ckks *handle() { return h_; }
    int load_status = -1;

  private:
    void check(int rc) {
        if (rc) throw std::runtime_error(shelfi_last_error(h_));
    }
    static std::string take(uint8_t *blob, uint64_t len) {
        std::string s(reinterpret_cast<const char *>(blob), len);
        shelfi_free(blob);
        return s;
    }
    shelfi_ckks *h_ = nullptr;
    std::string cryptodir_;
};
