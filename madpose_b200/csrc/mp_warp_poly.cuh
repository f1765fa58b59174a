// Real roots of a real polynomial of degree <= MAXD by ONE WARP (the warp-per-sample minimal solvers, mp_warp6.cuh ...):
// the same method as real_roots<MAXD> of mp_poly.cuh -- balancing, Sturm chain, isolation on the sign-change count,
// refinement, polish on the original coefficients -- spread over the lanes:
//   chain build   lane = coefficient (each long-division step is one FMA per lane, leading terms by shuffle)
//   isolation     lane = evaluation point: an interval holding several roots is cut into 32 parts at once
//   refinement    lane = isolated root (safeguarded Newton inside the isolating interval)
// Everything lives in registers and in the caller's shared-memory workspace; no local memory.
// All 32 lanes must call these functions together.
#pragma once
#include "mp_poly.cuh"

namespace mp {

constexpr unsigned WP_FULL = 0xffffffffu;

template <int MAXD> struct WarpPolyWork {
    double q[MAXD + 1];                  // balanced monic polynomial
    double chain[MAXD + 1][MAXD + 1];    // Sturm chain, ascending coefficients
    int deg[MAXD + 1];
    int len;
    double ilo[MAXD + 2], ihi[MAXD + 2]; // interval queue (several roots inside)
    int clo[MAXD + 2], chi[MAXD + 2];
    double qlo[MAXD + 1], qhi[MAXD + 1]; // isolated intervals
    double r[MAXD + 1];                  // refined roots (unsorted)
};

__device__ __forceinline__ double warp_max_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(WP_FULL, v, o));
    return v;
}

// number of sign changes of the chain at x (every lane its own x; coefficient reads are broadcasts)
template <int MAXD> __device__ __forceinline__ int sturm_changes(const WarpPolyWork<MAXD> &w, double x) {
    int n = 0, last = 0;
    const int len = w.len;
    for (int k = 0; k < len; k++) {
        const int d = w.deg[k];
        double v = w.chain[k][d];
        for (int i = d - 1; i >= 0; i--) v = fma(v, x, w.chain[k][i]);
        const int s = (v > 0) - (v < 0);
        if (s != 0) {
            if (last != 0 && s != last) n++;
            last = s;
        }
    }
    return n;
}

// distinct real roots of w.q (degree n, already trimmed) in [lo0, hi0]; refined roots are left in w.r (unsorted);
// returns their number (uniform over the warp)
template <int MAXD> __device__ inline int warp_roots_unscaled(WarpPolyWork<MAXD> &w, int n, bool positive_only, int lane) {
    // ---- Sturm chain: c0 = p, c1 = p', c_{k+1} = -(c_{k-1} mod c_k); lane i holds coefficient i
    double a = lane <= n ? w.q[lane] : 0.0;                       // c_{k-1}
    double b = lane < n ? (double)(lane + 1) * w.q[lane + 1] : 0.0; // c_k  (needs q[lane+1]: lane < n <= MAXD)
    int na = n, nb = n - 1;
    if (lane <= MAXD) { w.chain[0][lane] = a; w.chain[1][lane] = b; }
    if (lane == 0) { w.deg[0] = na; w.deg[1] = nb; }
    int len = 2;
    while (nb > 0 && len <= MAXD) {
        const double ilead = 1.0 / __shfl_sync(WP_FULL, b, nb);
        double r = a;
        for (int k = na; k >= nb; k--) { // long division: r[k - nb + i] -= f * b[i]
            const double f = __shfl_sync(WP_FULL, r, k) * ilead;
            const int src = lane - (k - nb);
            const double bs = __shfl_sync(WP_FULL, b, src < 0 ? 0 : src);
            if (src >= 0 && src <= nb) r = fma(-f, bs, r);
        }
        int nr = nb - 1;
        r = lane <= nr ? -r : 0.0;
        const double mx = warp_max_d(fabs(r));
        if (!(mx > 0.0)) break;
        const unsigned big = __ballot_sync(WP_FULL, lane <= nr && fabs(r) > 1e-13 * mx);
        // degree = highest coefficient above the relative floor (at least 0)
        nr = big ? 31 - __clz((int)big) : 0;
        if (lane > nr) r = 0.0;
        if (lane <= MAXD) w.chain[len][lane] = r;
        if (lane == 0) w.deg[len] = nr;
        a = b; na = nb; b = r; nb = nr;
        len++;
    }
    if (lane == 0) w.len = len;
    __syncwarp();
    // ---- Cauchy bound
    const double lead = w.q[n];
    const double B = warp_max_d(lane < n ? fabs(w.q[lane] / lead) : 0.0) + 1.0;
    const double lo0 = positive_only ? 0.0 : -B;
    int c0 = 0, c1 = 0;
    {
        const int c = sturm_changes<MAXD>(w, lane == 0 ? lo0 : B);
        c0 = __shfl_sync(WP_FULL, c, 0); c1 = __shfl_sync(WP_FULL, c, 1);
    }
    if (c0 - c1 <= 0) return 0;
    int sp = 0, nq = 0;
    if (lane == 0) { w.ilo[0] = lo0; w.ihi[0] = B; w.clo[0] = c0; w.chi[0] = c1; }
    sp = 1;
    __syncwarp();
    // ---- isolation: an interval with several roots is cut at 31 interior points (one per lane)
    while (sp > 0 && nq < MAXD) {
        sp--;
        const double lo = w.ilo[sp], hi = w.ihi[sp];
        const int clo = w.clo[sp], chi = w.chi[sp], nr = clo - chi;
        __syncwarp();
        if (nr <= 0) continue;
        if (nr == 1 || (hi - lo) < 1e-14 * (1.0 + fabs(lo))) {
            if (lane == 0) { w.qlo[nq] = lo; w.qhi[nq] = hi; }
            nq++;
            continue;
        }
        const double step = (hi - lo) * (1.0 / 32.0);
        const double xj = lane < 31 ? lo + step * (double)(lane + 1) : hi; // upper end of sub-interval `lane`
        const int cj = lane < 31 ? sturm_changes<MAXD>(w, xj) : chi;
        int cprev = __shfl_up_sync(WP_FULL, cj, 1);
        if (lane == 0) cprev = clo;
        const int nj = cprev - cj;
        const double xl = lane == 0 ? lo : lo + step * (double)lane;
        const unsigned one = __ballot_sync(WP_FULL, nj == 1), many = __ballot_sync(WP_FULL, nj > 1);
        const unsigned below = (1u << lane) - 1u;
        if (nj == 1) { const int k = nq + __popc(one & below); if (k < MAXD + 1) { w.qlo[k] = xl; w.qhi[k] = xj; } }
        if (nj > 1) { const int k = sp + __popc(many & below); if (k < MAXD + 2) { w.ilo[k] = xl; w.ihi[k] = xj; w.clo[k] = cprev; w.chi[k] = cj; } }
        nq = min(nq + __popc(one), MAXD + 1);
        sp = min(sp + __popc(many), MAXD + 2);
        __syncwarp();
    }
    if (nq > MAXD) nq = MAXD;
    __syncwarp();
    // ---- refinement.  First every isolating interval with a sign change is narrowed by two rounds of a 32-way
    // sign scan (lane = interior point: the bracket shrinks ~1000 x whatever the Cauchy bound was), then lane = root:
    // Newton safeguarded by bisection (a step that leaves the bracket or does not halve the previous one is replaced).
    auto val = [&](double x) {
        double r = w.q[n];
        for (int i = n - 1; i >= 0; i--) r = fma(r, x, w.q[i]);
        return r;
    };
    for (int j = 0; j < nq; j++) {
        double lo = w.qlo[j], hi = w.qhi[j];
        const double flo = val(lo), fhi = val(hi);
        if (!(flo * fhi < 0)) continue; // uniform: every lane evaluated the same two points
        for (int round = 0; round < 2; round++) {
            const double step = (hi - lo) * (1.0 / 32.0);
            const double x = lo + step * (double)(lane + 1);
            const double fx = lane < 31 ? val(x) : fhi;
            const unsigned flip = __ballot_sync(WP_FULL, (fx < 0) != (flo < 0) || fx == 0.0);
            const int l = flip ? __ffs((int)flip) - 1 : 31; // first point on the other side
            const double nlo = lo + step * (double)l, nhi = l < 31 ? lo + step * (double)(l + 1) : hi;
            lo = nlo; hi = nhi;
        }
        if (lane == 0) { w.qlo[j] = lo; w.qhi[j] = hi; }
    }
    __syncwarp();
    if (lane < nq) {
        const double ilo = w.qlo[lane], ihi = w.qhi[lane];
        auto val_d = [&](double x, double *dp) {
            double r = 0, d = 0;
            for (int i = n; i >= 0; i--) { d = fma(d, x, r); r = fma(r, x, w.q[i]); }
            *dp = d;
            return r;
        };
        double lo = ilo, hi = ihi, dd;
        const double flo = val_d(lo, &dd), fhi = val_d(hi, &dd);
        double x = 0.5 * (lo + hi);
        if (flo == 0.0) x = lo;
        else if (fhi == 0.0) x = hi;
        else if (flo * fhi < 0) {
            double dxold = hi - lo, dx = dxold, d;
            double f = val_d(x, &d);
            for (int it = 0; it < 120; it++) {
                if (f == 0.0) break;
                if ((f < 0) == (flo < 0)) lo = x; else hi = x;
                const double xn_newton = x - f / d;
                const bool bisect = !(xn_newton > lo && xn_newton < hi) || fabs(2.0 * f) > fabs(dxold * d);
                dxold = dx;
                double xn;
                if (bisect) { dx = 0.5 * (hi - lo); xn = lo + dx; }
                else { dx = xn_newton - x; xn = xn_newton; }
                if (xn == x || fabs(dx) <= 1e-16 * (1.0 + fabs(xn))) { x = xn; break; }
                x = xn;
                f = val_d(x, &d);
            }
        } else { // no sign change over the isolating interval (clustered roots): Newton from the middle, kept inside
            for (int it = 0; it < 8; it++) {
                double d;
                const double f = val_d(x, &d);
                if (d == 0.0) break;
                const double xn = x - f / d;
                if (!(xn >= ilo && xn <= ihi)) break;
                if (xn == x) break;
                x = xn;
            }
        }
        w.r[lane] = x;
    }
    __syncwarp();
    return nq;
}

// Real roots of pin[0..n] (ascending coefficients, in shared memory), ascending, into roots[] (shared memory).
// positive_only: only the roots > 0 are wanted (the search is restricted to (0, B]).  Returns the number of roots.
template <int MAXD> __device__ inline int warp_real_roots(const double *pin, int n, double *roots, WarpPolyWork<MAXD> &w,
                                                         bool positive_only, int lane) {
    // exact-zero leading coefficients
    {
        const unsigned nz = __ballot_sync(WP_FULL, lane <= n && pin[lane <= n ? lane : 0] != 0.0);
        n = nz ? 31 - __clz((int)nz) : 0;
    }
    if (n <= 0) return 0;
    double s = 1.0;
    bool balanced = false;
    if (n >= 2 && pin[0] != 0.0 && finite_d(pin[0]) && finite_d(pin[n])) {
        const double sv = pow(fabs(pin[0] / pin[n]), 1.0 / n);
        if (sv > 0 && finite_d(sv)) { s = sv; balanced = true; }
    }
    {
        double v = 0.0;
        if (lane <= n) {
            double spw = 1.0;
            for (int i = 0; i < lane; i++) spw *= s;
            v = pin[lane] * spw;
        }
        const double qn = __shfl_sync(WP_FULL, v, n);
        if (balanced) v /= qn;
        if (lane <= MAXD) w.q[lane] = lane <= n ? v : 0.0;
    }
    __syncwarp();
    // relative trimming of the leading coefficients (real_roots_unscaled)
    {
        const double v = lane <= n ? fabs(w.q[lane]) : 0.0;
        const double mx = warp_max_d(v);
        const unsigned big = __ballot_sync(WP_FULL, lane <= n && v > 1e-14 * mx);
        n = big ? 31 - __clz((int)big) : 0;
    }
    if (n <= 0) return 0;
    int nr;
    if (n == 1) {
        if (lane == 0) w.r[0] = -w.q[0] / w.q[1];
        nr = 1;
        __syncwarp();
    } else nr = warp_roots_unscaled<MAXD>(w, n, positive_only, lane);
    // back to the original variable + polish on the original coefficients
    double x = 0.0;
    bool have = lane < nr;
    if (have) x = w.r[lane] * s;
    // the polish runs on pin with its exact-zero-trimmed degree (uniform)
    int np_;
    {
        const unsigned nz = __ballot_sync(WP_FULL, lane <= MAXD && pin[lane <= MAXD ? lane : 0] != 0.0);
        np_ = nz ? 31 - __clz((int)nz) : 0;
    }
    if (have && balanced) {
        for (int it = 0; it < 4; it++) {
            double r = 0, d = 0;
            for (int i = np_; i >= 0; i--) { d = fma(d, x, r); r = fma(r, x, pin[i]); }
            if (d == 0.0 || !finite_d(r / d)) break;
            const double xn = x - r / d;
            if (xn == x) break;
            x = xn;
        }
    }
    if (positive_only) have = have && x > 0.0;
    // ascending order: rank = number of kept roots that are smaller (ties by lane)
    const unsigned keep = __ballot_sync(WP_FULL, have);
    int rank = 0;
    for (int j = 0; j < nr; j++) {
        const double xj = __shfl_sync(WP_FULL, x, j);
        if (((keep >> j) & 1u) && (xj < x || (xj == x && j < lane))) rank++;
    }
    if (have) roots[rank] = x;
    __syncwarp();
    return __popc(keep);
}

} // namespace mp
