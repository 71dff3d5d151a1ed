// Bit-equality of gdtrig::sinf_glibc / cosf_glibc with this machine's libm (sinf, cosf, sincosf).
// Build: g++ -O2 -mfma -ffp-contract=off check_trig.cpp -lm
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include "../../gpudrive_lab_b200/csrc/glibc_trig.h"

static float from_bits(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }
static uint32_t bits(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }

int main(int argc, char **argv) {
    uint32_t stride = argc > 1 ? (uint32_t)atoi(argv[1]) : 7;
    uint64_t n = 0, bad_s = 0, bad_c = 0, bad_sc = 0;
    for (int sign = 0; sign < 2; sign++) {
        for (uint32_t u = 0; u < 0x42f00000u; u += stride) {   // 0 <= |x| < 120
            float x = from_bits(u | (sign ? 0x80000000u : 0u));
            volatile float vx = x;     // keep the compiler from folding the libm calls
            float s = sinf(vx), c = cosf(vx), s2, c2;
            sincosf(vx, &s2, &c2);
            float ms = gdtrig::sinf_glibc(x), mc = gdtrig::cosf_glibc(x), qs, qc;
            gdtrig::sincosf_glibc(x, qs, qc);      // the one-pass form the lidar ray directions use
            if (bits(qs) != bits(s) || bits(qc) != bits(c)) { if (bad_sc++ < 5) printf("sincosf_glibc(%a): %a %a vs libm %a %a\n", x, qs, qc, s, c); }
            n++;
            if (bits(s) != bits(ms)) { if (bad_s++ < 5) printf("sin(%a): libm %a mine %a\n", x, s, ms); }
            if (bits(c) != bits(mc)) { if (bad_c++ < 5) printf("cos(%a): libm %a mine %a\n", x, c, mc); }
            if (bits(s2) != bits(s) || bits(c2) != bits(c)) { if (bad_sc++ < 5) printf("sincosf(%a) != sinf/cosf\n", x); }
        }
    }
    printf("checked %llu inputs: sin mismatches %llu, cos mismatches %llu, sincosf-vs-separate mismatches %llu\n",
           (unsigned long long)n, (unsigned long long)bad_s, (unsigned long long)bad_c, (unsigned long long)bad_sc);
    return (bad_s || bad_c || bad_sc) ? 1 : 0;
}
