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
// The driver state, its set-up and the evaluation's totals live in the CTA's shared memory (BlockScratch): telling the
// compiler turns the generic loads / stores of these out-of-line functions (each with its descriptor moves: a fifth of
// their instructions) into plain shared-memory accesses.
#define LMW_SHARED(p_) __builtin_assume(__isShared(p_))
constexpr int LM_TRI = LM_NT * (LM_NT + 1) / 2;

// reduced full-slot sums (LmAcc as a flat array) -> compact nt x nt system; = lm_compress.  The map from the 66 upper-
// triangle slots to compact positions depends on the set-up only: warp_lm_first decodes it once per solve into
// d.cpos / d.cposT, every evaluation after that is a table-driven scatter.
__device__ __forceinline__ void warp_compress_table(LmDriver &d, int lane) {
    const LmSetup &L = *d.L;
    const int nt = L.nt;
    for (int k = lane; k < LM_TRI; k += 32) {
        int a = 0, r = k;
        while (r >= LM_NT - a) { r -= LM_NT - a; a++; }
        const int b = a + r, ia = L.slot[a], ib = L.slot[b];
        const bool on = ia >= 0 && ib >= 0;
        d.cpos[k] = (short)(on ? ia * nt + ib : -1);
        d.cposT[k] = (short)(on ? ib * nt + ia : -1);
    }
    __syncwarp();
}
__device__ __forceinline__ void warp_compress(const LmDriver &d, const double *tot, double *A, double *g, int lane) {
    const LmSetup &L = *d.L;
    for (int k = lane; k < LM_TRI; k += 32) {
        const int p = d.cpos[k];
        if (p >= 0) {
            const double v = tot[k];
            A[p] = v; A[d.cposT[k]] = v;
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
// cycles per driver section (printed under MP_TRACE=1): a build with -DMP_LM_PROF only, the reads cost ~2 % of the driver
#ifdef MP_LM_PROF
#define LMW_T0() long long lmw_t_ = clock64()
#define LMW_T(slot_) do { if (lane == 0) { const long long n_ = clock64(); d.prof[slot_] += (unsigned long long)(n_ - lmw_t_); lmw_t_ = n_; } } while (0)
#else
#define LMW_T0() do { } while (0)
#define LMW_T(slot_) do { } while (0)
#endif
static __device__ __noinline__ bool warp_lm_first(LmDriver &d, const double *tot, int lane) {
    LMW_SHARED(&d); LMW_SHARED(tot);
    const LmSetup &L = *d.L;
    LMW_SHARED(&L);
    LMW_T0();
    const int nt = L.nt;
    warp_compress_table(d, lane);
    warp_compress(d, tot, d.A, d.g, lane);
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

// LmDriver::propose.  One copy of the code for every number of active slots, working matrices in shared memory at the
// fixed stride LM_NT (d.Lc: the matrix being eliminated, d.As: the unit-lower factor).  The first warp form kept the
// rows in registers, fully unrolled, in two instantiations at two call sites: block_lm was 17 040 instructions = 272 KB
// of straight-line code (lo_kernel 609 KB) and the instruction cache thrashed for every phase of the kernel
// (profiles/r2_whole_path.md).  A lone warp is latency-bound (~8 cycles per instruction), so what counts here is the
// length of the dependent chain: the elimination is spread over the ENTRIES of the lower triangle (<= 66 entries, <= 3
// per lane: one step = load, reciprocal, multiply-add, store for every lane at once) instead of over the rows (a
// shuffle + read-modify-write loop of up to 10 trips per step).  The arithmetic -- operands, order of the sums, fused
// multiply-adds -- is the row form's.
static __device__ __noinline__ bool warp_lm_propose(LmDriver &d, int lane) {
    LMW_SHARED(&d);
    const LmSetup &L = *d.L;
    LMW_SHARED(&L);
    const int nt = L.nt;
    constexpr int S = LM_NT;
    double *M = d.Lc, *Lf = d.As;
    // the lower-triangle entries (i >= j) this lane eliminates: e = i (i + 1) / 2 + j for e = lane, lane + 32, lane + 64
    int ei[3], ej[3];
#pragma unroll
    for (int q = 0; q < 3; q++) {
        const int e = lane + 32 * q;
        const int i = (int)((sqrtf((float)(8 * e + 1)) - 1.0f) * 0.5f); // exact: 8 e + 1 is a square only at j = 0
        ei[q] = i; ej[q] = e - i * (i + 1) / 2;
    }
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
        // the scaled, damped system: one row per lane
        double gsl = 0.0;
        if (rowl) {
            double *row = M + lane * S;
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
        // step k takes a_ij -= (a_ik / d_k) a_jk on every entry i >= j > k and leaves l_ik = a_ik / d_k in the factor;
        // pivots d_k > 0 <=> the scalar driver's Cholesky succeeds.  (The scalar driver of mp_lm2.cuh uses L L^T; the
        // two agree to rounding, ~1e-16 relative in the step.)  Column k is read-only during step k.
        bool bad = false;
        for (int k = 0; k < nt; k++) {
            const double dk = M[k * S + k];
            bad = bad || !(dk > 0.0) || !finite_d(dk);
            const double rk = 1.0 / dk;
            if (lane == 0) d.wr[k] = rk;
#pragma unroll
            for (int q = 0; q < 3; q++) {
                if (q == 2 && nt < LM_NT) break; // entries 64, 65 belong to row 10: the two-focal kind only (uniform)
                const int i = ei[q], j = ej[q];
                if (i < nt && i > k && j >= k) {
                    const double lik = M[i * S + k] * rk; // a_ik = l_ik d_k
                    if (j == k) Lf[i * S + k] = lik;
                    else M[i * S + j] -= lik * M[j * S + k];
                }
            }
            __syncwarp();
        }
        LMW_T(2);
        // forward substitution with the unit lower factor (column-oriented), then the diagonal
        {
            double s = gsl;
            for (int k = 0; k < nt; k++) {
                const double yk = __shfl_sync(LMW_FULL, s, k);
                if (lane == k) d.yw[k] = yk * d.wr[k];
                if (rowl && lane > k) s -= Lf[lane * S + k] * yk;
            }
        }
        __syncwarp();
        LMW_T(3);
        // back substitution with the unit upper factor (every lane, redundantly, in registers): FMAs only
        double xs[LM_NT];
        bool ok = !bad;
#pragma unroll
        for (int i = LM_NT - 1; i >= 0; i--) {
            double v = 0.0;
            if (i < nt) {
                v = d.yw[i];
#pragma unroll
                for (int k = i + 1; k < LM_NT; k++)
                    if (k < nt) v -= Lf[k * S + i] * xs[k];
                ok = ok && finite_d(v);
            }
            xs[i] = v;
        }
        LMW_T(4);
        // model cost change -(g.y + y.A y / 2) with A y = -g - D y (y solves the damped system): 2 x nt terms
        double mcc = 0;
        if (ok) {
            double dg = 0, dDd = 0;
#pragma unroll
            for (int a = 0; a < LM_NT; a++) xs[a] = -xs[a];
#pragma unroll
            for (int a = 0; a < LM_NT; a++) {
                if (a < nt) {
                    dg += xs[a] * d.gs[a];
                    dDd += xs[a] * xs[a] * d.Ad[a];
                }
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
        if (rowl) { // the step in the unscaled tangent space, one entry per lane
            double mine = 0.0;
#pragma unroll
            for (int a = 0; a < LM_NT; a++)
                if (a == lane) mine = xs[a];
            d.y[lane] = mine * d.scale[lane];
        }
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
    LMW_SHARED(&d); LMW_SHARED(tot);
    const LmSetup &L = *d.L;
    LMW_SHARED(&L);
    LMW_T0();
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
    if (verdict == 2) { // accepted: the system at xc replaces the current one in place (a rejected one is never needed)
        warp_compress(d, tot, d.A, d.g, lane);
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
