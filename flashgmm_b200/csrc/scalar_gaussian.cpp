// scalar_gaussian.cpp -- the three scalar Gaussian-CDF approximations of the reference's scalar
// (non-SIMD) path, /root/reference/compressai/cpp_exts/rans/rans_interface.cpp:148-150 (Polya),
// :188-204 (Abramowitz-Stegun 26.2.17), :216-219 (logistic), used only by the single-Gaussian search
// coder (a14, abi_table.cu). Evaluated with libm like the reference does. Built by g++ with the
// reference's own flags (-O3 -std=c++17 -ffp-contract=fast, FMA available: setup.py:72-76), because
// what the reference computes is whatever GCC makes of these expressions.
// Parity: unpinned across hosts (glibc expf / compiler version); verified against the reference
// compiled on the same machine (tests/test_host.py).
#include <cmath>

namespace {

const float kPi = 3.14159265358979323846f;
const float kInvSqrt2Pi = 0.3989422804014327f;

float polya(float x) {
    return 0.5f * (1.0f + std::copysign(std::sqrt(1.0f - std::exp(-2.0f * x * x / kPi)), x));
}

float abramowitz_stegun(float x) {
    const float p = 0.2316419f;
    const float b1 = 0.319381530f, b2 = -0.356563782f, b3 = 1.781477937f, b4 = -1.821255978f, b5 = 1.330274429f;
    const float ax = std::abs(x);
    const float zx = kInvSqrt2Pi * std::exp(-0.5f * x * x);
    const float t = 1.0f / (1.0f + p * ax);
    const float poly = t * (b1 + t * (b2 + t * (b3 + t * (b4 + t * b5))));
    const float upper = 1.0f - zx * poly;
    return (x >= 0.0f) ? upper : 1.0f - upper;
}

float logistic(float x) {
    const float k = 1.702f;
    return 1.0f / (1.0f + std::exp(-k * x));
}

}  // namespace

extern "C" float flashgmm_scalar_gaussian_cdf(float t, int mode) {
    switch (mode) {
    case 1: return abramowitz_stegun(t);
    case 2: return logistic(t);
    default: return polya(t);
    }
}
