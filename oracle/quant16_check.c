// Exhaustive proof behind the integer form of the 16-bit reference quantiser (frx_common.cuh frx_quant_ref, fast path).
// The reference computes, in separately rounded float64 steps (converter.py:56-86 then astype(int16)):
//     q = trunc(RN(RN(RN(2 v / R) - 1) * 32767))      v = x - min in [0, R],  R = max - min in [1, 65535]
// The kernel computes N = (2 v - R) * 32767 and t = N / R truncated toward zero in integers, and keeps t whenever the
// division leaves a remainder; when N / R is an integer it runs the float64 chain (a quotient that is an exact integer is
// the only place where the rounding of the chain can land on the other side of the truncation).  This program checks
// every (R, v) pair: remainder != 0  =>  t == q; the exact quotients 0 and +-32767 (v = R/2, 0, R) also agree, the other
// exact pairs are counted (they take the float64 chain in the kernel).  Second, the kernel's way of dividing is checked:
// qh = RZ(RZ(float(|N|)) * RZ(1 / float(R))) is floor(|N| / R) or one less, for every pair (so one upward correction by the
// remainder suffices).
// Build and run: gcc -O2 -fopenmp -frounding-math -ffp-contract=off oracle/quant16_check.c -o oracle/_ref/quant16_check -lm && oracle/_ref/quant16_check
#include <fenv.h>
#include <math.h>
#include <stdlib.h>
#include <stdio.h>
int main(void) {
    long bad = 0, total = 0, exact = 0, exact_differs = 0, trivial_bad = 0, div_bad = 0;
#pragma omp parallel for reduction(+:bad,total,exact,exact_differs,trivial_bad,div_bad) schedule(dynamic, 64)
    for (int R = 1; R <= 65535; R++) {
        const double dR = (double)R;
        fesetround(FE_TOWARDZERO);
        volatile float one = 1.0f, fR = (float)R;
        volatile float rcpf = one / fR;                   // RZ(1 / R)
        fesetround(FE_TONEAREST);
        for (int v = 0; v <= R; v++) {
            volatile double a = (2.0 * (double)v) / dR;
            volatile double b = a - 1.0;
            volatile double c = b * 32767.0;
            const int q = (int)c;                         // C conversion truncates toward zero
            const long N = (long)(2 * v - R) * 32767;
            const int t = (int)(N / R);
            const int rem = (int)(N % R);
            total++;
            if (rem != 0) { if (t != q) bad++; }
            else if (t == 0 || t == 32767 || t == -32767) { if (t != q) trivial_bad++; }
            else { exact++; if (t != q) exact_differs++; }
            {   // the kernel's division
                const unsigned M = (unsigned)labs(N);
                fesetround(FE_TOWARDZERO);
                volatile float fm = (float)M;             // RZ conversion
                volatile float pr = fm * rcpf;            // RZ product
                const unsigned qh = (unsigned)pr;
                fesetround(FE_TONEAREST);
                const unsigned qt = M / (unsigned)R;
                if (!(qh == qt || qh + 1 == qt)) div_bad++;
            }
        }
    }
    printf("pairs %ld, with remainder: mismatches %ld; trivial exact quotients (0, +-32767) mismatching %ld; other exact quotients %ld "
           "(of which the float64 chain differs from N/R: %ld); float division outside {floor - 1, floor}: %ld\n",
           total, bad, trivial_bad, exact, exact_differs, div_bad);
    return bad != 0 || trivial_bad != 0 || div_bad != 0;
}
