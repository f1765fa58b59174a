// Real roots of a real polynomial of degree <= MAXD on fixed-size arrays: Sturm chain
// isolation, bisection, Newton polish.  One thread = one polynomial.  This is the
// register/local-memory replacement of the eigenvalue solves the reference uses to
// extract roots (Eigen::EigenSolver on action matrices, solver.cpp:104,269,455) and of
// PoseLib's sturm.h inside relpose_5pt.  Coefficients ascending: p[0] + p[1] x + ...
#pragma once
#include "mp_common.cuh"

namespace mp {

MP_HD double poly_val(const double *p, int n, double x) {
    double r = p[n];
    for (int i = n - 1; i >= 0; i--) r = fma(r, x, p[i]);
    return r;
}
MP_HD double poly_val_d(const double *p, int n, double x, double *dp) {
    double r = 0, d = 0;
    for (int i = n; i >= 0; i--) {
        d = fma(d, x, r);
        r = fma(r, x, p[i]);
    }
    *dp = d;
    return r;
}

// The same two evaluations on a coefficient array padded with zeros up to degree MAXD, fully unrolled: the
// coefficients stay in registers (static indexing) and the leading zeros change nothing (fma(0, x, c) == c).
template <int MAXD> MP_HD double poly_val_pad(const double (&p)[MAXD + 1], double x) {
    double r = p[MAXD];
#pragma unroll
    for (int i = MAXD - 1; i >= 0; i--) r = fma(r, x, p[i]);
    return r;
}
template <int MAXD> MP_HD double poly_val_d_pad(const double (&p)[MAXD + 1], double x, double *dp) {
    double r = 0, d = 0;
#pragma unroll
    for (int i = MAXD; i >= 0; i--) {
        d = fma(d, x, r);
        r = fma(r, x, p[i]);
    }
    *dp = d;
    return r;
}

template <int MAXD> struct Sturm {
    double c[MAXD + 1][MAXD + 1]; // chain polynomials
    int deg[MAXD + 1];
    int len;

    // p has exact degree n >= 2 (leading coefficient non-negligible)
    MP_HD void build(const double *p, int n) {
#pragma unroll
        for (int k = 0; k <= MAXD; k++) // rows stay zero above their degree: changes() evaluates them at full length
#pragma unroll
            for (int i = 0; i <= MAXD; i++) c[k][i] = 0.0;
        for (int i = 0; i <= n; i++) c[0][i] = p[i];
        deg[0] = n;
        for (int i = 1; i <= n; i++) c[1][i - 1] = i * p[i];
        deg[1] = n - 1;
        len = 2;
        while (deg[len - 1] > 0 && len <= MAXD) {
            const double *a = c[len - 2], *b = c[len - 1];
            const int na = deg[len - 2], nb = deg[len - 1];
            double *r = c[len];
            for (int i = 0; i <= na; i++) r[i] = a[i];
            const double ilead = 1.0 / b[nb];
            for (int k = na; k >= nb; k--) { // long division, remainder left in r[0..nb-1]
                double f = r[k] * ilead;
                for (int i = 0; i <= nb; i++) r[k - nb + i] = fma(-f, b[i], r[k - nb + i]);
            }
            int nr = nb - 1;
            double mx = 0;
            for (int i = 0; i <= nr; i++) { r[i] = -r[i]; mx = fmax(mx, fabs(r[i])); }
            if (!(mx > 0.0)) break;
            while (nr > 0 && fabs(r[nr]) <= 1e-13 * mx) nr--;
            for (int i = nr + 1; i <= MAXD; i++) r[i] = 0.0; // (what the division left above the degree, and trimmed terms)
            deg[len] = nr;
            len++;
        }
    }
    // Every chain polynomial is evaluated at the full length MAXD with its zero padding (fma(0, x, c) == c: the values are
    // those of the exact-degree Horner form): fixed trip counts and static offsets let the rows' chains run side by side
    // instead of one data-dependent loop after the other.
    MP_HD int changes(double x) const {
        double vals[MAXD + 1];
#pragma unroll
        for (int k = 0; k <= MAXD; k++) {
            double r = c[k][MAXD];
#pragma unroll
            for (int i = MAXD - 1; i >= 0; i--) r = fma(r, x, c[k][i]);
            vals[k] = r;
        }
        int n = 0, last = 0;
#pragma unroll
        for (int k = 0; k <= MAXD; k++) {
            if (k >= len) break;
            const double v = vals[k];
            int s = (v > 0) - (v < 0);
            if (s != 0) {
                if (last != 0 && s != last) n++;
                last = s;
            }
        }
        return n;
    }
};

// distinct real roots (ascending) of p (degree n after the caller's own trimming is NOT
// assumed: leading coefficients below 1e-14 of the largest are dropped here).
template <int MAXD> MP_HD int real_roots_unscaled(const double *pin, int n, double *roots) {
    double p[MAXD + 1], mx = 0;
    for (int i = 0; i <= n; i++) { p[i] = pin[i]; mx = fmax(mx, fabs(p[i])); }
    while (n > 0 && fabs(p[n]) <= 1e-14 * mx) n--;
    if (n <= 0) return 0;
    if (n == 1) { roots[0] = -p[0] / p[1]; return 1; }
    double B = 0;
    for (int i = 0; i < n; i++) B = fmax(B, fabs(p[i] / p[n]));
    B += 1.0; // Cauchy bound
    Sturm<MAXD> sc;
    sc.build(p, n);
    double lo_s[MAXD + 2], hi_s[MAXD + 2];
    int cl_s[MAXD + 2], ch_s[MAXD + 2], sp = 0, nroots = 0;
    {
        int c0 = sc.changes(-B), c1 = sc.changes(B);
        if (c0 - c1 <= 0) return 0;
        lo_s[0] = -B; hi_s[0] = B; cl_s[0] = c0; ch_s[0] = c1; sp = 1;
    }
    // phase 1: isolation only (every lane of a warp is in the same loop: bisect on the sign-change count);
    // isolated intervals are queued in the order the one-pass version would have refined them
    double qlo[MAXD + 1], qhi[MAXD + 1];
    int nq = 0;
    while (sp > 0 && nq < MAXD) {
        sp--;
        const double ilo = lo_s[sp], ihi = hi_s[sp];
        const int clo = cl_s[sp], chi = ch_s[sp], nr = clo - chi;
        if (nr <= 0) continue;
        if (nr == 1 || (ihi - ilo) < 1e-14 * (1.0 + fabs(ilo))) {
            qlo[nq] = ilo; qhi[nq] = ihi; nq++;
            continue;
        }
        const double mid = 0.5 * (ilo + ihi);
        const int cm = sc.changes(mid);
        if (cm - chi > 0 && sp < MAXD + 2) { lo_s[sp] = mid; hi_s[sp] = ihi; cl_s[sp] = cm; ch_s[sp] = chi; sp++; }
        if (clo - cm > 0 && sp < MAXD + 2) { lo_s[sp] = ilo; hi_s[sp] = mid; cl_s[sp] = clo; ch_s[sp] = cm; sp++; }
    }
    // phase 2: refinement of every isolated interval (bisection, then Newton) on a register copy of p
    double pp[MAXD + 1];
#pragma unroll
    for (int i = 0; i <= MAXD; i++) pp[i] = i <= n ? p[i] : 0.0;
    for (int r = 0; r < nq; r++) {
        const double ilo = qlo[r], ihi = qhi[r];
        double lo = ilo, hi = ihi, flo = poly_val_pad<MAXD>(pp, lo), fhi = poly_val_pad<MAXD>(pp, hi);
        double x = 0.5 * (lo + hi);
        if (flo * fhi <= 0) {
            for (int it = 0; it < 200 && hi - lo > 1e-15 * (1.0 + fabs(x)); it++) {
                x = 0.5 * (lo + hi);
                double fx = poly_val_pad<MAXD>(pp, x);
                if (fx == 0.0) break;
                if ((fx < 0) == (flo < 0)) { lo = x; flo = fx; } else hi = x;
            }
            x = 0.5 * (lo + hi);
        }
        for (int it = 0; it < 8; it++) {
            double d, f = poly_val_d_pad<MAXD>(pp, x, &d);
            if (d == 0.0) break;
            double xn = x - f / d;
            if (!(xn >= ilo && xn <= ihi)) break;
            if (xn == x) break;
            x = xn;
        }
        roots[nroots++] = x;
    }
    for (int i = 1; i < nroots; i++) { // insertion sort (already ascending except for clustered roots)
        double v = roots[i];
        int j = i - 1;
        while (j >= 0 && roots[j] > v) { roots[j + 1] = roots[j]; j--; }
        roots[j + 1] = v;
    }
    return nroots;
}

// Balanced front-end: x = s*y with s = |p0/pn|^(1/n) brings the roots to unit size,
// the polynomial is made monic, roots are found on the balanced form and polished on
// the original coefficients.
template <int MAXD> MP_HD int real_roots(const double *pin, int n, double *roots) {
    while (n > 0 && pin[n] == 0.0) n--;
    if (n >= 2 && pin[0] != 0.0 && finite_d(pin[0]) && finite_d(pin[n])) {
        double s = pow(fabs(pin[0] / pin[n]), 1.0 / n);
        if (s > 0 && finite_d(s)) {
            double q[MAXD + 1], sp = 1.0;
            for (int i = 0; i <= n; i++) { q[i] = pin[i] * sp; sp *= s; }
            double qn = q[n];
            for (int i = 0; i <= n; i++) q[i] /= qn;
            int nr = real_roots_unscaled<MAXD>(q, n, roots);
            for (int k = 0; k < nr; k++) {
                double x = roots[k] * s;
                for (int it = 0; it < 4; it++) {
                    double d, f = poly_val_d(pin, n, x, &d);
                    if (d == 0.0 || !finite_d(f / d)) break;
                    double xn = x - f / d;
                    if (xn == x) break;
                    x = xn;
                }
                roots[k] = x;
            }
            return nr;
        }
    }
    return real_roots_unscaled<MAXD>(pin, n, roots);
}

} // namespace mp
