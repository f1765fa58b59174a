// ORACLE (test infrastructure only).  C entry points so tests/ can drive the
// CPU restatement through ctypes.  Nothing here is linked into the product.
#include "om_affine.h"

extern "C" {

// kind: 0 calibrated (3 pts), 1 shared focal (4 pts), 2 two focal (4 pts)
// x1,x2: npts x 3 row-major homogeneous points.  out: up to 8 rows of
// (b1, a2, b2*a2, f1, f2)  -- same convention as solver.cpp:117-121 (1,b1,a2,b2*a2[,f..]).
int om_solve_affine(int kind, const double *x1, const double *x2, const double *d1, const double *d2, double *out) {
    std::vector<om::AffineSol> s;
    if (kind == 0) {
        double a[3][3], b[3][3];
        std::memcpy(a, x1, sizeof(a)); std::memcpy(b, x2, sizeof(b));
        s = om::solve_scale_and_shift(a, b, d1, d2);
    } else {
        double a[4][3], b[4][3];
        std::memcpy(a, x1, sizeof(a)); std::memcpy(b, x2, sizeof(b));
        s = kind == 1 ? om::solve_scale_and_shift_shared_focal(a, b, d1, d2)
                      : om::solve_scale_and_shift_two_focal(a, b, d1, d2);
    }
    for (size_t i = 0; i < s.size(); i++) {
        out[5 * i + 0] = s[i].b1; out[5 * i + 1] = s[i].a2; out[5 * i + 2] = s[i].b2 * s[i].a2;
        out[5 * i + 3] = s[i].f1; out[5 * i + 4] = s[i].f2;
    }
    return (int)s.size();
}

} // extern "C"
