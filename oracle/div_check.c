// exhaustive check: for integer a in [0, 131070], b in [1, 65535]: RN(a/b) == fma(r, rcp, y0) with rcp = RN(1/b), y0 = RN(a*rcp), r = fma(-y0, b, a)
#include <math.h>
#include <stdio.h>
#include <omp.h>
int main(void) {
    long bad = 0, total = 0;
#pragma omp parallel for reduction(+:bad,total) schedule(dynamic, 64)
    for (int b = 1; b <= 65535; b++) {
        const double db = (double)b, rcp = 1.0 / db;
        for (int a = 0; a <= 131070; a++) {
            const double da = (double)a;
            const double y0 = da * rcp;
            const double r = fma(-y0, db, da);
            const double y = fma(r, rcp, y0);
            if (y != da / db) bad++;
            total++;
        }
    }
    printf("pairs %ld mismatches %ld\n", total, bad);
    return bad != 0;
}
