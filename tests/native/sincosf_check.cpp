// Host check of csrc/sincosf_glibc.h against the running glibc's sinf/cosf over every float in [0, 6.3].
// Build: g++ -O2 -ffp-contract=off -pthread.  Prints "<inputs> <sin mismatches> <cos mismatches>".
#include "sincosf_glibc.h"
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <thread>
#include <vector>

int main(int argc, char** argv)
{
    const uint64_t stride = argc > 1 ? (uint64_t)atoll(argv[1]) : 1;   // 1 = exhaustive
    const float lim = 6.3f;
    uint32_t top;
    memcpy(&top, &lim, 4);
    const unsigned nt = std::max(1u, std::thread::hardware_concurrency());
    std::vector<long> bs(nt, 0), bc(nt, 0);
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nt; t++)
        th.emplace_back([&, t] {
            long ls = 0, lc = 0;
            for (uint64_t u = t * stride; u <= top; u += nt * stride) {
                uint32_t v = (uint32_t)u;
                float f, s, c;
                memcpy(&f, &v, 4);
                orbx_sc::sincosf_glibc(f, &s, &c);
                const float rs = sinf(f), rc = cosf(f);
                ls += memcmp(&s, &rs, 4) != 0;
                lc += memcmp(&c, &rc, 4) != 0;
            }
            bs[t] = ls; bc[t] = lc;
        });
    for (auto& x : th) x.join();
    long a = 0, b = 0;
    for (unsigned t = 0; t < nt; t++) { a += bs[t]; b += bc[t]; }
    printf("%llu %ld %ld\n", (unsigned long long)((top + stride) / stride), a, b);
    return 0;
}
