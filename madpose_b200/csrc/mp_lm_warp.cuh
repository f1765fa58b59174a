// K5 (part 3) -- the LM driver of mp_lm2.cuh executed by ONE WARP instead of one thread.
// Same state (an LmDriver in shared memory) and the same algorithm as the scalar driver; what
// changes is who does it and the form of the elimination: the 11 x 11 systems are spread over
// the lanes (lane = row), the damped normal equations are factored as L D L^T (right-looking,
// rows in the driver's shared-memory workspace, pivot column broadcast by shuffles: no square roots, one reciprocal per
// pivot, substitutions without divisions) and the model cost change uses A y = -g - D y.  The
// scalar driver (L L^T, mp_lm2.cuh) is what the host check compiles against the oracle; the two
// agree to rounding (the GPU LM tests hold both to the oracle within 1e-8).  Scalar decisions are
// taken by lane 0 and broadcast.  All 32 lanes of the warp must call these functions together.
#pragma once

namespace mp {

constexpr unsigned LMW_FULL = 0xffffffffu;
constexpr int LM_TRI = LM_NT * (LM_NT + 1) / 2;

// reduced full-slot sums (LmAcc as a flat array) -> compact nt x nt system; = lm_compress
__device__ __forceinline__ void warp_compress(const LmSetup &L, const double *tot, double *A, double *g, int lane) {
    const int nt = L.nt;
    for (int k = lane; k < LM_TRI; k += 32) {
        int a = 0, r = k;
        while (r >= LM_NT - a) { r -= LM_NT - a; a++; }
        const int b = a + r, ia = L.slot[a], ib = L.slot[b];
        if (ia >= 0 && ib >= 0) {
            const double v = tot[k];
            A[ia * nt + ib] = v; A[ib * nt + ia] = v;
        }
    }
    if (lane < LM_NT && L.slot[lane] >= 0) g[L.slot[lane]] = tot[LM_TRI + lane];
    __syncwarp();
}

// lm_plus, one ambient component per lane: returns component `lane` (< LM_NA) of Plus(x, sign * dsrc) projected onto
// the bounds; dsrc is the compact tangent vector (shared memory).  Same expressions per component as lm_plus.
__device__ __forceinline__ double warp_plus_component(const LmSetup &L, const double *x, const double *dsrc, double sign, int lane) {
    const int k = lane < LM_NA ? lane : 0;
    auto dv = [&](int s_) { const int c = L.slot[s_]; return c >= 0 ? sign * dsrc[c] : 0.0; };
    double v;
    if (k == 0 || k >= 5) v = x[k] + dv(k == 0 ? 0 : k - 1);
    else {
        const double d1 = dv(1), d2 = dv(2), d3 = dv(3);
        const double n = sqrt(d1 * d1 + d2 * d2 + d3 * d3);
        v = x[k];
        if (n > 0) {
            double sv, e0;
            sincos(n, &sv, &e0);
            const double sn = sv / n, e1 = sn * d1, e2 = sn * d2, e3 = sn * d3;
            const double *q = x + 1;
            if (k == 1) v = e0 * q[0] - e1 * q[1] - e2 * q[2] - e3 * q[3];
            else if (k == 2) v = e0 * q[1] + e1 * q[0] + e2 * q[3] - e3 * q[2];
            else if (k == 3) v = e0 * q[2] - e1 * q[3] + e2 * q[0] + e3 * q[1];
            else v = e0 * q[3] + e1 * q[2] - e2 * q[1] + e3 * q[0];
        }
    }
    if (L.has_lower[k] && v < L.lower[k]) v = L.lower[k];
    return v;
}
// LmDriver::grad_max on the warp: max_k |Plus(x, -g)_k - x_k|, valid on every lane
__device__ __forceinline__ double warp_grad_max(const LmSetup &L, const double *x, const double *g, int lane) {
    const double v = warp_plus_component(L, x, g, -1.0, lane);
    double m = lane < LM_NA ? fabs(v - x[lane < LM_NA ? lane : 0]) : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(LMW_FULL, m, o));
    return m;
}

// LmDriver::first
#define LMW_T0() long long lmw_t_ = clock64()
#define LMW_T(slot_) do { if (lane == 0) { const long long n_ = clock64(); d.prof[slot_] += (unsigned long long)(n_ - lmw_t_); lmw_t_ = n_; } } while (0)
static __device__ __noinline__ bool warp_lm_first(LmDriver &d, const double *tot, int lane) {
    const LmSetup &L = *d.L;
    LMW_T0();
    const int nt = L.nt;
    warp_compress(L, tot, d.A, d.g, lane);
    if (lane < nt) d.scale[lane] = 1.0 / (1.0 + sqrt(d.A[lane * nt + lane]));
    __syncwarp();
    int go = 0;
    const double gm0 = warp_grad_max(L, d.x, d.g, lane);
    if (lane == 0) {
        d.cost = tot[LM_TRI + LM_NT];
        d.info.cost0 = d.cost; d.info.cost = d.cost;
        if (!finite_d(d.cost)) d.info.termination = 3;
        else if (gm0 <= L.gtol) d.info.termination = 1;
        else {
            d.radius = 1e4; d.decrease = 2.0;
            d.min_cost = d.cur_cost = d.ref_cost = d.cand_cost = d.cost; d.acc_ref = 0; d.acc_cand = 0;
            d.n_nonmono = 0; d.n_invalid = 0; d.it = 0;
            d.info.termination = 2;
            go = 1;
        }
    }
    go = __shfl_sync(LMW_FULL, go, 0);
    __syncwarp();
    LMW_T(7);
    return go != 0;
}

// LmDriver::propose.  Compact form: ROLLED loops over a working matrix in shared memory (d.Lc, lane = row), one copy
// of the code for every number of active slots.  The first warp form kept the rows in registers, which needs the
// elimination fully unrolled: with two instantiations (compile-time / run-time nt) at two call sites block_lm was
// 17 040 instructions = 272 KB of straight-line code that ONE warp walks once per LM iteration, and lo_kernel 609 KB:
// the instruction cache thrashed for every phase of the kernel (profiles/r2_whole_path.md).  The arithmetic (operands,
// order, fused multiply-adds) is the register form's.
static __device__ __noinline__ bool warp_lm_propose(LmDriver &d, int lane) {
    const LmSetup &L = *d.L;
    const int nt = L.nt;
    double *M = d.Lc;
    while (true) {
        int stop = 0;
        if (lane == 0) {
            if (++d.it > L.max_it) { d.info.cost = d.cost; stop = 1; }
            else {
                d.info.iterations = d.it;
                if (d.radius < 1e-32) { d.info.termination = 1; d.info.cost = d.cost; stop = 1; }
            }
        }
        stop = __shfl_sync(LMW_FULL, stop, 0);
        if (stop) return false;
        __syncwarp();
        LMW_T0();
        const double radius = d.radius;
        const bool rowl = lane < nt;
        double *row = M + (rowl ? lane : 0) * nt; // idle lanes never store
        // my row of the scaled, damped system
        double gsl = 0.0;
        if (rowl) {
            const double sa = d.scale[lane];
            const double *arow = d.A + lane * nt;
#pragma unroll 2
            for (int b = 0; b < nt; b++) row[b] = arow[b] * sa * d.scale[b];
            const double diag = row[lane];
            const double add = fmin(fmax(diag, 1e-6), 1e32) / radius;
            row[lane] = diag + add;
            d.Ad[lane] = add; // the damping D_aa, for the model cost change below
            gsl = d.g[lane] * sa;
            d.gs[lane] = gsl;
        }
        __syncwarp();
        LMW_T(1);
        // right-looking LDL^T, no square roots and ONE division per pivot (the reciprocal, computed by every lane):
        // after step k, M[i][k] = l_ik for i > k; pivots d_k > 0 <=> the scalar driver's Cholesky succeeds.
        // (The scalar driver of mp_lm2.cuh uses L L^T; the two agree to rounding, ~1e-16 relative in the step.)
        bool bad = false;
        for (int k = 0; k < nt; k++) {
            const double dk = M[k * nt + k];
            bad = bad || !(dk > 0.0) || !finite_d(dk);
            const double rk = 1.0 / dk;
            const double aik = row[k], lik = aik * rk; // a_ik = l_ik d_k
            if (lane == k) d.wr[k] = rk;
#pragma unroll 2
            for (int j = k + 1; j < nt; j++) {
                const double ajk = __shfl_sync(LMW_FULL, aik, j);
                if (rowl && j <= lane) row[j] -= lik * ajk; // a_ij -= l_ik d_k l_jk
            }
            if (rowl && lane > k) row[k] = lik;
            __syncwarp();
        }
        LMW_T(2);
        // forward substitution with the unit lower factor (column-oriented), then the diagonal
        {
            double s = gsl;
            for (int k = 0; k < nt; k++) {
                const double yk = __shfl_sync(LMW_FULL, s, k);
                if (lane == k) d.yw[k] = yk * d.wr[k];
                if (rowl && lane > k) s -= row[k] * yk;
            }
        }
        __syncwarp();
        LMW_T(3);
        // back substitution with the unit upper factor: every lane computes the same values redundantly (FMAs only) and
        // stores them to the same slots of d.xs -- a lane reads back what it (and every other lane, identically) wrote
        bool ok = !bad;
        for (int i = nt - 1; i >= 0; i--) {
            double v = d.yw[i];
#pragma unroll 2
            for (int k = i + 1; k < nt; k++) v -= M[k * nt + i] * d.xs[k];
            ok = ok && finite_d(v);
            d.xs[i] = v;
        }
        __syncwarp();
        LMW_T(4);
        // model cost change -(g.y + y.A y / 2) with A y = -g - D y (y solves the damped system): 2 x nt terms
        double mcc = 0;
        if (ok) {
            double dg = 0, dDd = 0;
            for (int a = 0; a < nt; a++) {
                const double xa = -d.xs[a];
                dg += xa * d.gs[a];
                dDd += xa * xa * d.Ad[a];
            }
            mcc = 0.5 * (dDd - dg);
        }
        LMW_T(5);
        if (!ok || !(mcc > 0)) {
            int fail_ = 0;
            if (lane == 0) {
                if (++d.n_invalid >= 5) { d.info.termination = 3; d.info.cost = d.cost; fail_ = 1; }
                else { d.radius /= d.decrease; d.decrease *= 2; }
            }
            fail_ = __shfl_sync(LMW_FULL, fail_, 0);
            if (fail_) return false;
            __syncwarp();
            continue;
        }
        if (lane == 0) { d.mcc = mcc; d.n_invalid = 0; }
        if (rowl) d.y[lane] = -d.xs[lane] * d.scale[lane]; // the step in the unscaled tangent space, one entry per lane
        __syncwarp();
        const double xcv = warp_plus_component(L, d.x, d.y, 1.0, lane);
        if (lane < LM_NA) d.xc[lane] = xcv;
        __syncwarp();
        LMW_T(6);
        return true;
    }
}

// LmDriver::feed
static __device__ __noinline__ bool warp_lm_feed(LmDriver &d, const double *tot, int lane) {
    const LmSetup &L = *d.L;
    const int nt = L.nt;
    LMW_T0();
    warp_compress(L, tot, d.An, d.gn, lane);
    int verdict = 0; // 0 finished, 1 rejected step (continue), 2 accepted step
    double cc = 0, rel = 0;
    double sn = 0, xn = 0; // |xc - x|^2 and |x|^2: one component per lane, butterfly sums
    if (lane < LM_NA) { const double xv = d.x[lane], dx = d.xc[lane] - xv; sn = dx * dx; xn = xv * xv; }
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) { sn += __shfl_xor_sync(LMW_FULL, sn, o); xn += __shfl_xor_sync(LMW_FULL, xn, o); }
    if (lane == 0) {
        cc = tot[LM_TRI + LM_NT];
        if (!finite_d(cc)) cc = DBL_MAX;
        if (sqrt(sn) <= L.ptol * (sqrt(xn) + L.ptol)) { d.info.termination = 1; d.info.cost = d.cost; }
        else if (fabs(d.cost - cc) <= L.ftol * d.cost) { d.info.termination = 1; d.info.cost = d.cost; }
        else {
            rel = cc >= DBL_MAX ? -DBL_MAX : fmax((d.cur_cost - cc) / d.mcc, (d.ref_cost - cc) / (d.acc_ref + d.mcc));
            if (rel > 1e-3) verdict = 2;
            else { d.radius /= d.decrease; d.decrease *= 2; verdict = 1; }
        }
    }
    verdict = __shfl_sync(LMW_FULL, verdict, 0);
    if (verdict == 2) {
        for (int e = lane; e < nt * nt; e += 32) d.A[e] = d.An[e];
        if (lane < nt) d.g[lane] = d.gn[lane];
        __syncwarp();
        if (lane < LM_NA) d.x[lane] = d.xc[lane];
        __syncwarp();
        const double gm = warp_grad_max(L, d.x, d.g, lane);
        if (lane == 0) {
            d.cost = cc;
            const double tt = 2.0 * rel - 1.0;
            d.radius = fmin(1e16, d.radius / fmax(1.0 / 3.0, 1.0 - tt * tt * tt));
            d.decrease = 2.0;
            d.cur_cost = cc; d.acc_cand += d.mcc; d.acc_ref += d.mcc;
            if (d.cur_cost < d.min_cost) { d.min_cost = d.cur_cost; d.n_nonmono = 0; d.cand_cost = d.cur_cost; d.acc_cand = 0; }
            else { ++d.n_nonmono; if (d.cur_cost > d.cand_cost) { d.cand_cost = d.cur_cost; d.acc_cand = 0; } }
            if (d.n_nonmono == L.max_nonmono) { d.ref_cost = d.cand_cost; d.acc_ref = d.acc_cand; }
            if (gm <= L.gtol) { d.info.termination = 1; d.info.cost = d.cost; verdict = 0; }
        }
        verdict = __shfl_sync(LMW_FULL, verdict, 0);
    }
    __syncwarp();
    LMW_T(0);
    return verdict != 0;
}

} // namespace mp
