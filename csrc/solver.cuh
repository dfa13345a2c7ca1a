// In-kernel local solver: L-BFGS (two-loop, memory PSB_HIST) with the strong-Wolfe cubic line search,
// or one SGD step.  One CTA runs the whole solve; every thread executes the same scalar control flow
// (all scalars come out of block-wide reductions), so there is no divergence between warps.
//
// Algorithm statement: psb200/oracle/lbfgs.py (Breeze L-BFGS as used by Spark's LogisticRegression,
// which the reference calls at LogisticRegressionTaskSpark.java:179-184).  The evaluator is a template
// parameter so the SIMT and the tcgen05 kernels share this file.
#pragma once
#include "common.cuh"

#define PSB_LS_C1 1e-4
#define PSB_LS_C2 0.9
#define PSB_LS_MAX_ITER 10
#define PSB_LS_MAX_ZOOM 10

struct SolveVecs {
    float* x;   // current point (scaled space)
    float* g;   // gradient at x
    float* d;   // search direction
    float* xt;  // trial point
    float* gt;  // gradient at the trial point
    float* tmp;
    float* s[PSB_HIST];
    float* y[PSB_HIST];
    int n4;
};

struct SolveResult {
    double loss;
    int evals, iterations, status;   // status: 0 max_iter, 1 gradient converged, 2 line search failed
};

__device__ __forceinline__ void vec_axpy_to(float* out, const float* x, double a, const float* d, int n4) {
    vec_prefetch(x, n4);
    vec_prefetch(d, n4);
    const float fa = (float)a;
    const float4* x4 = (const float4*)x;
    const float4* d4 = (const float4*)d;
    float4* o4 = (float4*)out;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) {
        float4 xv = x4[i], dv = d4[i];
        o4[i] = make_float4(fmaf(fa, dv.x, xv.x), fmaf(fa, dv.y, xv.y), fmaf(fa, dv.z, xv.z), fmaf(fa, dv.w, xv.w));
    }
}
__device__ __forceinline__ void vec_scale_copy(float* out, const float* a, double s, int n4) {
    vec_prefetch(a, n4);
    const float fs = (float)s;
    const float4* a4 = (const float4*)a;
    float4* o4 = (float4*)out;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) {
        float4 u = a4[i];
        o4[i] = make_float4(u.x * fs, u.y * fs, u.z * fs, u.w * fs);
    }
}

struct Bracket {
    double t, dd, f;
};

__device__ __forceinline__ double ls_interp(const Bracket& l, const Bracket& r) {
    const double d1 = l.dd + r.dd - 3.0 * (l.f - r.f) / (l.t - r.t);
    const double disc = d1 * d1 - l.dd * r.dd;
    const double d2 = disc > 0.0 ? sqrt(disc) : 0.0;
    const double den = r.dd - l.dd + 2.0 * d2;
    double t = den != 0.0 ? r.t - (r.t - l.t) * (r.dd + d2 - d1) / den : 0.5 * (l.t + r.t);
    const double lb = l.t + 0.1 * (r.t - l.t), ub = l.t + 0.9 * (r.t - l.t);
    if (!(t >= lb)) return lb;
    if (t > ub) return ub;
    return t;
}

// phi(t): xt = x + t d, evaluate, return value and directional derivative.
template <class Eval>
__device__ __forceinline__ Bracket ls_phi(Eval& ev, SolveVecs& v, double t, double* red, int& evals) {
    vec_axpy_to(v.xt, v.x, t, v.d, v.n4);
    __syncthreads();
    Bracket b;
    b.t = t;
    b.f = ev.eval(v.xt, v.gt);
    ++evals;
    b.dd = block_dot(v.gt, v.d, v.n4, red);
    return b;
}

// Strong-Wolfe line search on a scalar function phi(t) -> Bracket{t, phi'(t), phi(t)}.
// Returns true and the accepted step (value in f_out) or false.
template <class Phi>
__device__ bool strong_wolfe_phi(Phi& phi, double f0, double dd0, double init, double& alpha, double& f_out) {
    if (dd0 > 0.0) return false;
    double t = init;
    Bracket low = {0.0, dd0, f0};
    Bracket lo, hi;
    bool zoom = false;
    for (int i = 0; i < PSB_LS_MAX_ITER; ++i) {
        Bracket c = phi(t);
        if (isinf(c.f) || isnan(c.f)) {
            t *= 0.5;
            continue;
        }
        if (c.f > f0 + PSB_LS_C1 * t * dd0 || (c.f >= low.f && i > 0)) {
            lo = low, hi = c, zoom = true;
            break;
        }
        if (fabs(c.dd) <= PSB_LS_C2 * fabs(dd0)) {
            alpha = c.t, f_out = c.f;
            return true;
        }
        if (c.dd >= 0.0) {
            lo = c, hi = low, zoom = true;
            break;
        }
        low = c;
        t *= 1.5;
    }
    if (!zoom) return false;
    for (int i = 0; i < PSB_LS_MAX_ZOOM; ++i) {
        const double tz = lo.t > hi.t ? ls_interp(hi, lo) : ls_interp(lo, hi);
        Bracket c = phi(tz);
        if (c.f > f0 + PSB_LS_C1 * c.t * dd0 || c.f >= lo.f) {
            hi = c;
        } else {
            if (fabs(c.dd) <= PSB_LS_C2 * fabs(dd0)) {
                alpha = c.t, f_out = c.f;
                return true;
            }
            if (c.dd * (hi.t - lo.t) >= 0.0) hi = lo;
            lo = c;
        }
    }
    return false;
}

// The classic form: every trial is a full loss/gradient evaluation at x + t d (gradient left in v.gt).
template <class Eval>
__device__ bool strong_wolfe(Eval& ev, SolveVecs& v, double f0, double dd0, double init, double* red, int& evals,
                             double& alpha, double& f_out) {
    auto phi = [&](double t) { return ls_phi(ev, v, t, red, evals); };
    return strong_wolfe_phi(phi, f0, dd0, init, alpha, f_out);
}

// d = -H g by the two-loop recursion over the `hist` newest pairs (newest at index 0).  False on a bad pair.
__device__ inline bool lbfgs_direction(SolveVecs& v, int hist, double* red) {
    double a[PSB_HIST], rho[PSB_HIST];
    vec_scale_copy(v.d, v.g, 1.0, v.n4);
    __syncthreads();
    double diag = 1.0;
    if (hist > 0) {
        const double sy = block_dot(v.s[0], v.y[0], v.n4, red);
        const double yy = block_dot(v.y[0], v.y[0], v.n4, red);
        if (sy < 0.0 || isnan(sy)) return false;
        diag = sy / yy;
    }
    for (int i = 0; i < hist; ++i) {
        rho[i] = block_dot(v.s[i], v.y[i], v.n4, red);
        a[i] = block_dot(v.s[i], v.d, v.n4, red) / rho[i];
        vec_axpy_to(v.d, v.d, -a[i], v.y[i], v.n4);
        __syncthreads();
    }
    vec_scale_copy(v.d, v.d, diag, v.n4);
    __syncthreads();
    for (int i = hist - 1; i >= 0; --i) {
        const double beta = block_dot(v.y[i], v.d, v.n4, red) / rho[i];
        vec_axpy_to(v.d, v.d, a[i] - beta, v.s[i], v.n4);
        __syncthreads();
    }
    vec_scale_copy(v.d, v.d, -1.0, v.n4);
    __syncthreads();
    return true;
}

// Several block-wide sums at once (all threads get all N results; deterministic order).  `red`: >= 33 doubles.
template <int N>
__device__ __forceinline__ void block_sum_multi(double (&v)[N], double* red) {
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int k = 0; k < N; ++k)
        for (int o = 16; o; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
#pragma unroll
    for (int k = 0; k < N; ++k) {
        __syncthreads();
        if (l == 0) red[w] = v[k];
        __syncthreads();
        double t = 0.0;
        for (int i = 0; i < nw; ++i) t += red[i];
        v[k] = t;
    }
}

// Steepest-descent start and the one-pair two-loop recursion written as fused sweeps: the solver's vectors live in
// global memory, so what a CTA pays for is the number of passes and block-wide reductions, not the flops.
//   hist == 0: one pass      — d = -g, gg = g.g                               (dd0 = -gg, |d| = sqrt(gg))
//   hist == 1: two passes    — (s.y, y.y, s.g, y.g, g.g), then d = -(diag (g - a y) + (a - beta) s) with
//                              a = s.g / s.y, diag = s.y / y.y, beta = diag (y.g - a y.y) / s.y, and g.d, d.d
// Same algebra as lbfgs_direction.  Returns false on a bad pair (s.y < 0 or NaN); gg is valid either way.
__device__ inline bool lbfgs_direction_fused(SolveVecs& v, int hist, double* red, double& gg, double& dd0, double& dn) {
    const float4* g4 = (const float4*)v.g;
    float4* d4 = (float4*)v.d;
    vec_prefetch(v.g, v.n4);
    if (hist == 0) {
        double acc[1] = {0.0};
        float s = 0.f;
        for (int i = threadIdx.x; i < v.n4; i += blockDim.x) {
            const float4 g = g4[i];
            d4[i] = make_float4(-g.x, -g.y, -g.z, -g.w);
            s += g.x * g.x + g.y * g.y + g.z * g.z + g.w * g.w;
        }
        acc[0] = (double)s;
        block_sum_multi(acc, red);
        gg = acc[0], dd0 = -gg, dn = sqrt(gg);
        return true;
    }
    const float4* s4 = (const float4*)v.s[0];
    const float4* y4 = (const float4*)v.y[0];
    vec_prefetch(v.s[0], v.n4);
    vec_prefetch(v.y[0], v.n4);
    float sy = 0.f, yy = 0.f, sg = 0.f, yg = 0.f, g2 = 0.f;
    for (int i = threadIdx.x; i < v.n4; i += blockDim.x) {
        const float4 g = g4[i], s = s4[i], y = y4[i];
        sy += s.x * y.x + s.y * y.y + s.z * y.z + s.w * y.w;
        yy += y.x * y.x + y.y * y.y + y.z * y.z + y.w * y.w;
        sg += s.x * g.x + s.y * g.y + s.z * g.z + s.w * g.w;
        yg += y.x * g.x + y.y * g.y + y.z * g.z + y.w * g.w;
        g2 += g.x * g.x + g.y * g.y + g.z * g.z + g.w * g.w;
    }
    double acc[5] = {(double)sy, (double)yy, (double)sg, (double)yg, (double)g2};
    block_sum_multi(acc, red);
    gg = acc[4];
    if (acc[0] < 0.0 || isnan(acc[0])) return false;
    const double a = acc[2] / acc[0], diag = acc[0] / acc[1];
    const double beta = diag * (acc[3] - a * acc[1]) / acc[0];
    const float fa = (float)a, fdiag = (float)diag, fc = (float)(a - beta);
    float gd = 0.f, d2 = 0.f;
    for (int i = threadIdx.x; i < v.n4; i += blockDim.x) {
        const float4 g = g4[i], s = s4[i], y = y4[i];
        float4 d;
        d.x = -fmaf(fc, s.x, fdiag * fmaf(-fa, y.x, g.x));
        d.y = -fmaf(fc, s.y, fdiag * fmaf(-fa, y.y, g.y));
        d.z = -fmaf(fc, s.z, fdiag * fmaf(-fa, y.z, g.z));
        d.w = -fmaf(fc, s.w, fdiag * fmaf(-fa, y.w, g.w));
        d4[i] = d;
        gd += g.x * d.x + g.y * d.y + g.z * d.z + g.w * d.w;
        d2 += d.x * d.x + d.y * d.y + d.z * d.z + d.w * d.w;
    }
    double acc2[2] = {(double)gd, (double)d2};
    block_sum_multi(acc2, red);
    dd0 = acc2[0], dn = sqrt(acc2[1]);
    return true;
}

// s = xt - x and y = gt - g in one sweep (the new curvature pair)
__device__ __forceinline__ void vec_pair_update(float* s, float* y, const float* xt, const float* x, const float* gt,
                                                const float* g, int n4) {
    vec_prefetch(xt, n4), vec_prefetch(x, n4), vec_prefetch(gt, n4), vec_prefetch(g, n4);
    for (int i = threadIdx.x; i < n4; i += blockDim.x) {
        const float4 a = ((const float4*)xt)[i], b = ((const float4*)x)[i];
        const float4 c = ((const float4*)gt)[i], e = ((const float4*)g)[i];
        ((float4*)s)[i] = make_float4(a.x - b.x, a.y - b.y, a.z - b.z, a.w - b.w);
        ((float4*)y)[i] = make_float4(c.x - e.x, c.y - e.y, c.z - e.z, c.w - e.w);
    }
}

// direction + the three scalars the line search needs; fused sweeps for hist <= 1, the general recursion otherwise
__device__ inline bool lbfgs_step_setup(SolveVecs& v, int hist, double* red, double& gg, double& dd0, double& dn) {
    if (hist <= 1) {
        const bool ok = lbfgs_direction_fused(v, hist, red, gg, dd0, dn);
        __syncthreads();
        return ok;
    }
    gg = block_dot(v.g, v.g, v.n4, red);
    if (!lbfgs_direction(v, hist, red)) return false;
    dd0 = block_dot(v.g, v.d, v.n4, red);
    dn = sqrt(block_dot(v.d, v.d, v.n4, red));
    return true;
}

// L-BFGS for evaluators that cache the logits.  The logits are linear in the parameters,
// z(x + t d) = z(x) + t z(d), which makes every line-search trial after the first one free:
//   * the first trial of an iteration is a full evaluation at x + t1 d (loss, gradient, and its logits Zt);
//     when the line search accepts it — the common case — the iteration costs exactly what the classic
//     solver pays, and Zt simply becomes the cached logits of the new point;
//   * when it is rejected, z(d) = (Zt - Zx) / t1 comes out of the two cached logit sets without touching the
//     data, every further trial is an O(rows x classes) update (Eval::trial), and only the accepted point
//     pays one more pass for its gradient (Eval::gradient_at) — none at all in the last iteration.
// Same line search on the same phi as local_solve, so the same iterates up to rounding.
// On entry: v.x = start point, v.g = gradient there, f0 = loss there, logits of v.x cached by the evaluator.
template <class Eval>
__device__ SolveResult local_solve_cached(Eval& ev, SolveVecs& v, int max_iter, double* red, double f0) {
    SolveResult r;
    r.evals = 1, r.iterations = 0, r.status = 0;
    double f = f0;
    int hist = 0;
    const double tol = 1e-6;
    for (int it = 0; it < max_iter; ++it) {
        double gg = 0.0, dd0 = 0.0, dn = 0.0;
        PSB_TRACE(10);
        bool ok = lbfgs_step_setup(v, hist, red, gg, dd0, dn);
        PSB_TRACE(11);
        const double gnorm = sqrt(gg);
        if (gnorm <= fmax(tol * fabs(f), 1e-8)) {
            r.status = 1;
            break;
        }
        double alpha = 0.0, f_new = f, t_first = 0.0;
        bool derived = false;
        int calls = 0;
        if (ok) {
            auto phi = [&](double t) {
                ++r.evals;
                Bracket b;
                b.t = t;
                if (calls++ == 0) {                                       // full evaluation, logits parked in Zd
                    vec_axpy_to(v.xt, v.x, t, v.d, v.n4);
                    __syncthreads();
                    PSB_TRACE(12);
                    b.f = ev.eval_store(v.xt, v.gt, ev.Zd);
                    PSB_TRACE(13);
                    b.dd = block_dot(v.gt, v.d, v.n4, red);
                    PSB_TRACE(14);
                    t_first = t;
                    if (isinf(b.f) || isnan(b.f)) calls = 0;              // unusable logits: evaluate fully again
                } else {
                    if (!derived) {
                        ev.derive_direction_logits(t_first);              // Zd <- (Zt - Zx) / t_first
                        derived = true;
                    }
                    ev.trial(t, v.x, v.d, b.f, b.dd);                     // no pass over the data
                }
                return b;
            };
            ok = strong_wolfe_phi(phi, f, dd0, it == 0 ? 1.0 / dn : 1.0, alpha, f_new);
            if (ok && alpha * gnorm < 1e-10) ok = false;
        }
        if (!ok) {
            r.status = 2;
            break;
        }
        const bool need_grad = it + 1 < max_iter;
        if (!derived) {                                                   // the fully evaluated trial was accepted
            float* z = ev.Zx;
            ev.Zx = ev.Zd, ev.Zd = z;                                     // its logits are the new point's logits
        } else {
            vec_axpy_to(v.xt, v.x, alpha, v.d, v.n4);
            ev.advance(alpha);                                            // cached logits move to x + alpha d
            __syncthreads();
            if (need_grad) ev.gradient_at(v.xt, v.gt);                    // one pass: gradient at the new point
        }
        if (need_grad) {
            for (int i = (hist < PSB_HIST - 2 ? hist : PSB_HIST - 2); i > 0; --i) {   // shift pairs, newest first
                float* ps = v.s[i];
                float* py = v.y[i];
                v.s[i] = v.s[i - 1], v.y[i] = v.y[i - 1], v.s[i - 1] = ps, v.y[i - 1] = py;
            }
            vec_pair_update(v.s[0], v.y[0], v.xt, v.x, v.gt, v.g, v.n4);
            __syncthreads();
            PSB_TRACE(15);
            if (hist < PSB_HIST - 1) ++hist;
            float* p = v.g;
            v.g = v.gt, v.gt = p;
        }
        float* p = v.x;
        v.x = v.xt, v.xt = p;
        f = f_new;
        r.iterations = it + 1;
    }
    r.loss = f;
    return r;
}

// On entry v.x holds the start point.  On exit v.x holds the result.  hist_cap: pairs kept (<= PSB_HIST); the pair
// buffers at index >= hist_cap are never touched (a kernel may park other data there).
template <class Eval>
__device__ SolveResult local_solve(Eval& ev, SolveVecs& v, int solver, int max_iter, float sgd_lr, double* red,
                                   bool have_f0 = false, double f0 = 0.0, int hist_cap = PSB_HIST) {
    SolveResult r;
    r.evals = 0, r.iterations = 0, r.status = 0;
    // have_f0: the caller already evaluated loss (f0) and gradient (in v.g) at the start point v.x
    double f = have_f0 ? f0 : ev.eval(v.x, v.g);
    r.evals = 1;
    if (solver == PSB_SOLVER_SGD) {
        vec_axpy_to(v.x, v.x, -(double)sgd_lr, v.g, v.n4);
        __syncthreads();
        r.loss = f, r.iterations = 1;
        return r;
    }
    int hist = 0;            // pairs stored; newest at index 0
    const double tol = 1e-6;
    for (int it = 0; it < max_iter; ++it) {
        double gg = 0.0, dd0 = 0.0, dn = 0.0;
        bool ok = lbfgs_step_setup(v, hist, red, gg, dd0, dn);
        const double gnorm = sqrt(gg);
        if (gnorm <= fmax(tol * fabs(f), 1e-8)) {
            r.status = 1;
            break;
        }
        double alpha = 0.0, f_new = f;
        if (ok) {
            ok = strong_wolfe(ev, v, f, dd0, it == 0 ? 1.0 / dn : 1.0, red, r.evals, alpha, f_new);
            if (ok && alpha * gnorm < 1e-10) ok = false;
        }
        if (!ok) {
            r.status = 2;
            break;
        }
        // ---- accept: history (rotate, newest first), x <- xt, g <- gt
        float* s_last = v.s[hist_cap - 1];
        float* y_last = v.y[hist_cap - 1];
        for (int i = hist_cap - 1; i > 0; --i) v.s[i] = v.s[i - 1], v.y[i] = v.y[i - 1];
        v.s[0] = s_last, v.y[0] = y_last;
        vec_pair_update(v.s[0], v.y[0], v.xt, v.x, v.gt, v.g, v.n4);
        __syncthreads();
        if (hist < hist_cap) ++hist;
        float* p;
        p = v.x, v.x = v.xt, v.xt = p;
        p = v.g, v.g = v.gt, v.gt = p;
        f = f_new;
        r.iterations = it + 1;
    }
    r.loss = f;
    return r;
}
