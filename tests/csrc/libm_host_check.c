/* Exhaustive host check of magritte_b200/csrc/mgb_libm.h against the box's own libm (test infrastructure).
 * usage: libm_host_check <fn> [stride]   fn in expf logf erff erff_u expm1f powf; prints mismatch count. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <omp.h>
#include "mgb_libm.h"
static inline int same(float a, float b) {
    uint32_t x, y; memcpy(&x, &a, 4); memcpy(&y, &b, 4);
    if (x == y) return 1;
    return (a != a) && (b != b); /* any NaN == any NaN */
}
int main(int argc, char** argv) {
    const char* fn = argc > 1 ? argv[1] : "expf";
    const uint64_t stride = argc > 2 ? strtoull(argv[2], 0, 10) : 1;
    uint64_t bad = 0, n = 0;
    uint32_t first_bad = 0; int have = 0;
    if (strcmp(fn, "powf") == 0) {
        /* x over a stride-lattice of all floats, y over interpolation factors 1 - i/n (n <= 12) and a few others */
        float ys[96]; int ny = 0;
        for (int nn = 2; nn <= 12; nn++) for (int i = 1; i < nn; i++) ys[ny++] = (float)(1.0 - ((double)i) / nn);
        ys[ny++] = 0.123456f; ys[ny++] = 0.987654f; ys[ny++] = 1e-7f; ys[ny++] = 3.5f; ys[ny++] = -0.75f; ys[ny++] = 2.0f; ys[ny++]=3.0f;
#pragma omp parallel for reduction(+ : bad, n) schedule(static)
        for (uint64_t u = 0; u < (1ull << 32); u += stride) {
            float x; uint32_t b = (uint32_t)u; memcpy(&x, &b, 4);
            for (int j = 0; j < ny; j++) { n++; if (!same(mgb_powf(x, ys[j]), powf(x, ys[j]))) { bad++; if (!have) { have = 1; first_bad = b; } } }
        }
    } else {
        int which = !strcmp(fn, "expf") ? 0 : !strcmp(fn, "logf") ? 1 : !strcmp(fn, "erff") ? 2 : !strcmp(fn, "erff_u") ? 4 : 3;
#pragma omp parallel for reduction(+ : bad, n) schedule(static)
        for (uint64_t u = 0; u < (1ull << 32); u += stride) {
            float x; uint32_t b = (uint32_t)u; memcpy(&x, &b, 4);
            float a, r;
            switch (which) {
            case 0: a = mgb_expf(x); r = expf(x); break;
            case 1: a = mgb_logf(x); r = logf(x); break;
            case 2: a = mgb_erff(x); r = erff(x); break;
            case 4: a = mgb_erff_u(x); r = erff(x); break;
            default: a = mgb_expm1f(x); r = expm1f(x); break;
            }
            n++;
            if (!same(a, r)) { bad++; if (!have) { have = 1; first_bad = b; } }
        }
    }
    printf("%s checked=%llu mismatches=%llu first_bad=0x%08x\n", fn, (unsigned long long)n, (unsigned long long)bad, first_bad);
    return bad ? 1 : 0;
}
