// The reference's C++ smoke driver (palisade_pybind/SHELFI_FHE/src/main.cpp:8-23,28-78) as an asserting test:
// 100 uniform [0,100) values, three identical learners, weights .5/.3/.5, result must equal 1.3 * x within 1e-6.
#include <cmath>
#include <cstdio>
#include <random>

#include "../../include/shelfi_ckks.hpp"

int main(int argc, char **argv) {
    if (argc < 2) { std::fprintf(stderr, "usage: %s <cryptodir/>\n", argv[0]); return 2; }
    CKKS ckks("ckks", 4096, 52, argv[1]);
    ckks.loadCryptoParams();
    if (ckks.load_status != 0) { std::fprintf(stderr, "load status %d\n", ckks.load_status); return 3; }
    std::mt19937_64 rng(1);
    std::uniform_real_distribution<double> dist(0.0, 100.0);
    std::vector<double> x(100);
    for (auto &v : x) v = dist(rng);
    std::string enc = ckks.encrypt_cpp(x);
    std::vector<std::string> learners{enc, enc, enc};
    std::vector<float> w{0.5f, 0.3f, 0.5f};
    std::string avg = ckks.computeWeightedAverage_cpp(learners, w);
    if (!ckks.computeWeightedAverage_cpp(learners, {0.5f}).empty()) return 4;   // size mismatch -> ""
    std::vector<double> out = ckks.decrypt_cpp(avg, x.size());
    double scale = 0.5 + (double)0.3f + 0.5, worst = 0;
    for (size_t i = 0; i < x.size(); i++) worst = std::fmax(worst, std::fabs(out[i] - scale * x[i]) / 130.0);
    std::printf("max relative error %.3e over %zu values\n", worst, x.size());
    return worst < 1e-6 ? 0 : 1;
}
