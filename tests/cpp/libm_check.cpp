// libm_check.cpp -- CPU suite: csrc/pt_libm.h against the installed C library, bit for bit.
// Every float is tried for logf; sinf / cosf over every float up to |x| = 2^20 plus a stride through the rest
// (the shading path only ever passes |x| <= 2 pi + 1).  Prints "OK <count>" or the first mismatches.
// Build: g++ -O2 -ffp-contract=off -fopenmp -I<repo>/monte-carlo-raytracer_b200/csrc libm_check.cpp -o libm_check
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <atomic>

#include "pt_libm.h"

static inline uint32_t bits(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    return u;
}
static inline bool same(float a, float b) { return bits(a) == bits(b) || (a != a && b != b); }

int main(int argc, char** argv) {
    const bool quick = argc > 1 && argv[1][0] == 'q';
    std::atomic<long long> bad{0}, n{0};
    const uint32_t step = quick ? 37u : 1u;
#pragma omp parallel for schedule(static)
    for (long long i = 0; i < (1LL << 32); i += step) {
        const uint32_t u = (uint32_t)i;
        float x;
        memcpy(&x, &u, 4);
        const uint32_t mag = u & 0x7fffffffu;
        long long local_bad = 0;
        if (!same(mcrt::libm::logf_glibc(x), logf(x))) ++local_bad;
        // |x| < 2^20 exhaustively; above that every 1021st value (reduce_large path, inf, NaN)
        if (mag < 0x49800000u || (u % 1021u) == 0u) {
            if (!same(mcrt::libm::sinf_glibc(x), sinf(x))) ++local_bad;
            if (!same(mcrt::libm::cosf_glibc(x), cosf(x))) ++local_bad;
        }
        if (local_bad) {
            if (bad.fetch_add(local_bad) < 8)
                printf("mismatch at %08x (%g): sin %08x/%08x cos %08x/%08x log %08x/%08x\n", u, (double)x,
                       bits(mcrt::libm::sinf_glibc(x)), bits(sinf(x)), bits(mcrt::libm::cosf_glibc(x)), bits(cosf(x)),
                       bits(mcrt::libm::logf_glibc(x)), bits(logf(x)));
        }
        n.fetch_add(1, std::memory_order_relaxed);
    }
    if (bad.load() == 0) {
        printf("OK %lld\n", n.load());
        return 0;
    }
    printf("FAILED %lld mismatches\n", bad.load());
    return 1;
}
