// Exactness of the division by invariant multipliers the int32 emit pass uses (kernels_narrow.cuh: narrow_magic;
// eval_narrow.cu: qmagic = umulhi(2x, m) >> l): floor(x / d) for every dividend 0 <= x < 2^31 and divisor 1 <= d < 2^31.
// Host-only: the same header function computes the pairs for the device tables.
#include <cstdint>
#include <cstdio>
#include <random>

#include "kernels_narrow.cuh"

static inline uint32_t umulhi(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }

int main() {
    long long checks = 0, bad = 0;
    auto check = [&](uint32_t x, uint32_t d, uint32_t m, int l) {
        if (x >= (1u << 31)) return;
        const uint32_t q = umulhi(2u * x, m) >> l;
        ++checks;
        if (q != x / d) {
            if (bad < 10) std::printf("x=%u d=%u m=%u l=%d: %u != %u\n", x, d, m, l, q, x / d);
            ++bad;
        }
    };
    auto divisor = [&](uint32_t d, std::mt19937_64& rng) {
        int l = 0;
        const uint32_t m = fgs::narrow_magic(d, &l);
        const uint32_t top = 0x7fffffffu;
        for (uint32_t x : {0u, 1u, d - 1, d, d + 1, 2 * d - 1, 2 * d, top, top - 1, top - d, top / d * d, top / d * d - 1}) check(x, d, m, l);
        for (int k = 0; k < 64; ++k) {
            const uint32_t x = (uint32_t)(rng() & top);
            check(x, d, m, l);
            const uint32_t y = x / d * d; // multiples and their neighbours: where a wrong multiplier shows
            check(y, d, m, l);
            if (y) check(y - 1, d, m, l);
        }
    };
    std::mt19937_64 rng(0xF65);
    for (uint32_t d = 1; d <= 200000; ++d) divisor(d, rng);                       // every small divisor (capacities, maxima)
    for (int e = 1; e < 31; ++e)                                                    // around every power of two
        for (int o = -3; o <= 3; ++o) {
            const int64_t d = ((int64_t)1 << e) + o;
            if (d >= 1 && d < ((int64_t)1 << 31)) divisor((uint32_t)d, rng);
        }
    for (int k = 0; k < 200000; ++k) divisor((uint32_t)(rng() % 0x7fffffffu) + 1, rng); // random divisors up to 2^31 - 1
    int l = 7;
    if (fgs::narrow_magic(0, &l) != 0 || l != 0) ++bad; // capacity 0 → quotient 0 (least_allocated.go:48-50)
    // narrow_magic_fast == narrow_magic
    long long mchecks = 0;
    auto mcheck = [&](uint32_t d) {
        int l0 = -1, l1 = -1;
        const uint32_t m0 = fgs::narrow_magic(d, &l0), m1 = fgs::narrow_magic_fast(d, &l1);
        ++mchecks;
        if (m0 != m1 || l0 != l1) {
            if (bad < 10) std::printf("narrow_magic_fast(%u): %u/%d != %u/%d\n", d, m1, l1, m0, l0);
            ++bad;
        }
    };
    for (uint32_t d = 0; d <= (1u << 24) + 16; ++d) mcheck(d);
    for (int e = 24; e <= 31; ++e)
        for (int o = -3; o <= 3; ++o) {
            const int64_t d = ((int64_t)1 << e) + o;
            if (d >= 1 && d < ((int64_t)1 << 31)) mcheck((uint32_t)d);
        }
    for (int k = 0; k < 2000000; ++k) mcheck((uint32_t)(rng() % 0x7fffffffu) + 1);
    std::printf("%lld magic pairs compared\n", mchecks);
    // alloc_div100: (x * 100) / range as Go computes it (wrapping product, truncation toward zero), for every kind of input
    long long dchecks = 0;
    auto dcheck = [&](long long x, long long range) {
        if (range == 0) return;
        const long long prod = (long long)((unsigned long long)x * 100ull);
        if (prod == INT64_MIN && range == -1) return; // the one quotient int64 cannot hold
        const int want = (int)(prod / range);
        const int got = fgs::alloc_div100(x, range, fgs::alloc_recip(range));
        if (fgs::alloc_div100(x, range, 0) != want) ++bad; // no reciprocal: the division itself
        ++dchecks;
        if (want != got) {
            if (bad < 10) std::printf("alloc_div100(%lld, %lld): %d != %d\n", x, range, got, want);
            ++bad;
        }
    };
    for (int k = 0; k < 4000000; ++k) {
        const int rb = 1 + (int)(rng() % 62), xb = 1 + (int)(rng() % 63);
        long long range = (long long)(rng() >> (64 - rb));
        long long x = (long long)(rng() >> (64 - xb));
        if (rng() & 1) x = -x;
        if ((rng() & 15) == 0) range = -range;
        dcheck(x, range);
        if (range > 0) { // exact multiples and their neighbours, x inside [0, range] (the case the scores come from)
            const long long y = (long long)(rng() % (unsigned long long)range);
            dcheck(y, range);
            const long long q = (long long)(rng() % 101);
            if (range < (1ll << 55)) {
                const long long m = (q * range) / 100;
                dcheck(m, range), dcheck(m + 1, range), dcheck(m - 1, range);
            }
            dcheck(range, range), dcheck(range - 1, range), dcheck(-range, range);
        }
    }
    for (long long range : {1ll, 2ll, 3ll, (1ll << 24), (1ll << 24) + 1, (1ll << 53) - 1, (1ll << 53) + 1, (1ll << 56), (long long)INT64_MAX})
        for (long long x : {0ll, 1ll, -1ll, (1ll << 56) - 1, -(1ll << 56) + 1, (1ll << 56), -(1ll << 56), (long long)INT64_MAX, (long long)INT64_MIN, (long long)INT64_MIN + 1})
            dcheck(x, range);
    std::printf("%lld checks, %lld division checks, %lld failed\n", checks, dchecks, bad);
    return bad ? 1 : 0;
}
