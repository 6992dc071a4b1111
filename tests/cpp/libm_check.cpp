// TEST INFRASTRUCTURE. Pins elevenrender_b200/csrc/er_libm.cuh (compiled here for the HOST, the very source
// the device code is built from) against the host's libm — the libm the reference's sycl::sin/cos/acos/
// atan2/pow/log resolve to when its sources are compiled as the oracle (oracle/shim, src/sycl.h:8).
//
//   libm_check [stride]      prints one JSON object: per function, arguments compared and mismatching bits
//
// stride 1 walks every binary32 argument of the unary functions (minutes); the test suite uses a stride
// that finishes in seconds. Built by tests/test_libm_cpu.py with g++ -O2 -ffp-contract=off.
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "../../elevenrender_b200/csrc/er_libm.cuh"

static float F(uint32_t u) {
    float f;
    memcpy(&f, &u, 4);
    return f;
}
static bool same(float a, float b) { return memcmp(&a, &b, 4) == 0 || (a != a && b != b); }
static uint32_t rs = 0x2545F491u;
static float rnd() {
    rs ^= rs << 13;
    rs ^= rs >> 17;
    rs ^= rs << 5;
    return (float)rs / 4294967296.0f;
}

int main(int argc, char** argv) {
    const uint64_t stride = argc > 1 ? strtoull(argv[1], nullptr, 10) : 97;
    long n_sin = 0, b_sin = 0, b_cos = 0, n_acos = 0, b_acos = 0, n_log = 0, b_log = 0, n_at = 0, b_at = 0, n_pw = 0, b_pw = 0,
         n_p22 = 0, b_p22 = 0;
    // sin / cos: every stride-th |x| < 120, both signs
    for (uint64_t u = 0; u < 0x42f00000u; u += stride)
        for (int g = 0; g < 2; g++) {
            const float x = F((uint32_t)u | (g ? 0x80000000u : 0u));
            n_sin++;
            if (!same(sinf(x), erlm_sinf(x))) b_sin++;
            if (!same(cosf(x), erlm_cosf(x))) b_cos++;
        }
    // acos: [-1, 1] and a little beyond (NaN)
    for (uint64_t u = 0; u <= 0x3f800010u; u += stride)
        for (int g = 0; g < 2; g++) {
            const float x = F((uint32_t)u | (g ? 0x80000000u : 0u));
            n_acos++;
            if (!same(acosf(x), erlm_acosf(x))) b_acos++;
        }
    // log: every positive finite argument, zero and a negative
    for (uint64_t u = 0; u <= 0x7f800000u; u += stride) {
        const float x = F((uint32_t)u);
        n_log++;
        if (!same(logf(x), erlm_logf(x))) b_log++;
    }
    if (!same(logf(-1.f), erlm_logf(-1.f))) b_log++;
    // pow(x, 2.2): what generateHitData evaluates (src/kernel.cpp:152-153), x in [0, 4]
    for (uint64_t u = 0; u <= 0x40800000u; u += stride) {
        const float x = F((uint32_t)u);
        n_p22++;
        if (!same(powf(x, 2.2f), erlm_powf(x, 2.2f))) b_p22++;
    }
    // atan2 / pow on random pairs (incl. tiny, huge ratios and zeros)
    const long N2 = (long)(4000000000ull / (stride < 8 ? 8 : stride));
    for (long i = 0; i < N2; i++) {
        float y = rnd() * 2 - 1, x = rnd() * 2 - 1;
        if (i & 1) y *= rnd() * 1e-3f;
        if ((i & 7) == 3) x *= 1e-4f;
        if ((i & 15) == 5) x *= 1e-30f;
        if ((i & 31) == 9) y = 0.f;
        if ((i & 63) == 17) x = 0.f;
        if ((i & 127) == 33) x = 1.f;
        n_at++;
        if (!same(atan2f(y, x), erlm_atan2f(y, x))) b_at++;
        float px = rnd() * 4, py = (rnd() * 2 - 1) * 8;
        if (i & 2) px = F((rs >> 1) % 0x7f800000u), py = (rnd() * 2 - 1) * 2;
        if (py == 0.f) continue;
        n_pw++;
        if (!same(powf(px, py), erlm_powf(px, py))) b_pw++;
    }
    printf("{\"stride\": %llu, \"sin\": [%ld, %ld], \"cos\": [%ld, %ld], \"acos\": [%ld, %ld], \"log\": [%ld, %ld], "
           "\"pow22\": [%ld, %ld], \"atan2\": [%ld, %ld], \"pow\": [%ld, %ld]}\n",
           (unsigned long long)stride, n_sin, b_sin, n_sin, b_cos, n_acos, b_acos, n_log, b_log, n_p22, b_p22, n_at, b_at, n_pw, b_pw);
    return 0;
}
