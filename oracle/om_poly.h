// ORACLE (test infrastructure only).  Real roots of real polynomials by Sturm
// isolation + bisection + Newton polish.  Stands in for the eigenvalue solves
// the reference uses to extract roots (Eigen::EigenSolver on action/companion
// matrices, /root/reference/src/solver.cpp:104,269,455,590) and for PoseLib
// v2.0.4's sturm.h used by relpose_5pt (not vendored in /root/reference).
#pragma once
#include <algorithm>
#include <cmath>
#include <vector>

namespace om {

// p[0] + p[1] x + ... + p[n] x^n
inline double poly_eval(const std::vector<double> &p, double x) {
    double r = 0;
    for (int i = (int)p.size() - 1; i >= 0; i--) r = r * x + p[i];
    return r;
}
inline double poly_eval_d(const std::vector<double> &p, double x, double *dp) {
    double r = 0, d = 0;
    for (int i = (int)p.size() - 1; i >= 0; i--) {
        d = d * x + r;
        r = r * x + p[i];
    }
    *dp = d;
    return r;
}

inline void poly_trim(std::vector<double> &p) {
    double mx = 0;
    for (double c : p) mx = std::fmax(mx, std::fabs(c));
    while (p.size() > 1 && std::fabs(p.back()) <= 1e-14 * mx) p.pop_back();
}

// remainder of a / b (negated -> Sturm chain element)
inline std::vector<double> poly_neg_rem(std::vector<double> a, const std::vector<double> &b) {
    int nb = (int)b.size() - 1;
    while ((int)a.size() - 1 >= nb && a.size() > 0) {
        int na = (int)a.size() - 1;
        double f = a[na] / b[nb];
        for (int i = 0; i <= nb; i++) a[na - nb + i] -= f * b[i];
        a.pop_back();
    }
    if (a.empty()) a.push_back(0.0);
    for (double &c : a) c = -c;
    return a;
}

struct SturmChain {
    std::vector<std::vector<double>> seq;
    explicit SturmChain(const std::vector<double> &p) {
        seq.push_back(p);
        std::vector<double> d(p.size() > 1 ? p.size() - 1 : 1, 0.0);
        for (size_t i = 1; i < p.size(); i++) d[i - 1] = i * p[i];
        seq.push_back(d);
        while (seq.back().size() > 1) {
            std::vector<double> r = poly_neg_rem(seq[seq.size() - 2], seq.back());
            // trim numerically-zero leading terms relative to the chain scale
            double mx = 0;
            for (double c : r) mx = std::fmax(mx, std::fabs(c));
            while (r.size() > 1 && std::fabs(r.back()) <= 1e-13 * mx) r.pop_back();
            if (mx == 0) break;
            seq.push_back(r);
        }
    }
    int changes(double x) const {
        int n = 0, last = 0;
        for (auto &q : seq) {
            double v = poly_eval(q, x);
            int s = (v > 0) - (v < 0);
            if (s != 0) {
                if (last != 0 && s != last) n++;
                last = s;
            }
        }
        return n;
    }
};

// all distinct real roots, ascending
inline std::vector<double> poly_real_roots_unscaled(std::vector<double> p);
inline std::vector<double> poly_real_roots(std::vector<double> p) {
    // balance coefficients with x = s*y, s = |p0/pn|^(1/n) (roots of size ~1)
    while (p.size() > 1 && p.back() == 0.0) p.pop_back();
    int n = (int)p.size() - 1;
    if (n >= 2 && p[0] != 0.0 && std::isfinite(p[0]) && std::isfinite(p[n])) {
        double s = std::pow(std::fabs(p[0] / p[n]), 1.0 / n);
        if (s > 0 && std::isfinite(s)) {
            std::vector<double> q(p);
            double sp = 1.0;
            for (int i = 0; i <= n; i++) { q[i] = p[i] * sp; sp *= s; }
            double qn = q[n];
            for (double &c : q) c /= qn;
            std::vector<double> r = poly_real_roots_unscaled(q);
            for (double &x : r) {
                x *= s;
                for (int it = 0; it < 4; it++) { // polish on the original coefficients
                    double d, f = poly_eval_d(p, x, &d);
                    if (d == 0 || !std::isfinite(f / d)) break;
                    double xn = x - f / d;
                    if (xn == x) break;
                    x = xn;
                }
            }
            return r;
        }
    }
    return poly_real_roots_unscaled(p);
}
inline std::vector<double> poly_real_roots_unscaled(std::vector<double> p) {
    std::vector<double> roots;
    poly_trim(p);
    int n = (int)p.size() - 1;
    if (n <= 0) return roots;
    if (n == 1) { roots.push_back(-p[0] / p[1]); return roots; }
    double lead = p[n], B = 0;
    for (int i = 0; i < n; i++) B = std::fmax(B, std::fabs(p[i] / lead));
    B = 1.0 + B; // Cauchy bound
    SturmChain sc(p);
    struct Iv { double lo, hi; int clo, chi; };
    std::vector<Iv> stack;
    int c0 = sc.changes(-B), c1 = sc.changes(B);
    if (c0 - c1 <= 0) return roots;
    stack.push_back({-B, B, c0, c1});
    while (!stack.empty()) {
        Iv iv = stack.back();
        stack.pop_back();
        int nr = iv.clo - iv.chi;
        if (nr <= 0) continue;
        double w = iv.hi - iv.lo;
        if (nr == 1 || w < 1e-14 * (1.0 + std::fabs(iv.lo))) {
            // bisection on sign of p, then Newton polish
            double lo = iv.lo, hi = iv.hi, flo = poly_eval(p, lo), fhi = poly_eval(p, hi);
            double x = 0.5 * (lo + hi);
            if (flo * fhi <= 0) {
                for (int it = 0; it < 200 && hi - lo > 1e-15 * (1 + std::fabs(x)); it++) {
                    x = 0.5 * (lo + hi);
                    double fx = poly_eval(p, x);
                    if (fx == 0) break;
                    if ((fx < 0) == (flo < 0)) { lo = x; flo = fx; } else { hi = x; }
                }
                x = 0.5 * (lo + hi);
            }
            for (int it = 0; it < 8; it++) {
                double d, f = poly_eval_d(p, x, &d);
                if (d == 0) break;
                double xn = x - f / d;
                if (!(xn >= iv.lo && xn <= iv.hi)) break;
                if (xn == x) break;
                x = xn;
            }
            roots.push_back(x);
            continue;
        }
        double mid = 0.5 * (iv.lo + iv.hi);
        int cm = sc.changes(mid);
        stack.push_back({mid, iv.hi, cm, iv.chi});
        stack.push_back({iv.lo, mid, iv.clo, cm});
    }
    std::sort(roots.begin(), roots.end());
    return roots;
}

} // namespace om
