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
    const float fa = (float)a;
    const float4* x4 = (const float4*)x;
    const float4* d4 = (const float4*)d;
    float4* o4 = (float4*)out;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) {
        float4 xv = x4[i], dv = d4[i];
        o4[i] = make_float4(fmaf(fa, dv.x, xv.x), fmaf(fa, dv.y, xv.y), fmaf(fa, dv.z, xv.z), fmaf(fa, dv.w, xv.w));
    }
}
__device__ __forceinline__ void vec_sub_to(float* out, const float* a, const float* b, int n4) {
    const float4* a4 = (const float4*)a;
    const float4* b4 = (const float4*)b;
    float4* o4 = (float4*)out;
    for (int i = threadIdx.x; i < n4; i += blockDim.x) {
        float4 u = a4[i], v = b4[i];
        o4[i] = make_float4(u.x - v.x, u.y - v.y, u.z - v.z, u.w - v.w);
    }
}
__device__ __forceinline__ void vec_scale_copy(float* out, const float* a, double s, int n4) {
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

// Returns true and the accepted step (whose gradient is in v.gt, value in f_out) or false.
template <class Eval>
__device__ bool strong_wolfe(Eval& ev, SolveVecs& v, double f0, double dd0, double init, double* red, int& evals,
                             double& alpha, double& f_out) {
    if (dd0 > 0.0) return false;
    double t = init;
    Bracket low = {0.0, dd0, f0};
    Bracket lo, hi;
    bool zoom = false;
    for (int i = 0; i < PSB_LS_MAX_ITER; ++i) {
        Bracket c = ls_phi(ev, v, t, red, evals);
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
        Bracket c = ls_phi(ev, v, tz, red, evals);
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

// On entry v.x holds the start point.  On exit v.x holds the result.
template <class Eval>
__device__ SolveResult local_solve(Eval& ev, SolveVecs& v, int solver, int max_iter, float sgd_lr, double* red,
                                   bool have_f0 = false, double f0 = 0.0) {
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
        const double gg = block_dot(v.g, v.g, v.n4, red);
        const double gnorm = sqrt(gg);
        if (gnorm <= fmax(tol * fabs(f), 1e-8)) {
            r.status = 1;
            break;
        }
        // ---- two-loop recursion: d = -H g
        double a[PSB_HIST], rho[PSB_HIST];
        vec_scale_copy(v.d, v.g, 1.0, v.n4);
        __syncthreads();
        bool bad = false;
        double diag = 1.0;
        if (hist > 0) {
            const double sy = block_dot(v.s[0], v.y[0], v.n4, red);
            const double yy = block_dot(v.y[0], v.y[0], v.n4, red);
            if (sy < 0.0 || isnan(sy)) bad = true;
            diag = sy / yy;
        }
        if (!bad) {
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
        }
        double alpha = 0.0, f_new = f;
        bool ok = !bad;
        if (ok) {
            const double dd0 = block_dot(v.g, v.d, v.n4, red);
            const double dn = sqrt(block_dot(v.d, v.d, v.n4, red));
            ok = strong_wolfe(ev, v, f, dd0, it == 0 ? 1.0 / dn : 1.0, red, r.evals, alpha, f_new);
            if (ok && alpha * gnorm < 1e-10) ok = false;
        }
        if (!ok) {
            r.status = 2;
            break;
        }
        // ---- accept: history (rotate, newest first), x <- xt, g <- gt
        float* s_last = v.s[PSB_HIST - 1];
        float* y_last = v.y[PSB_HIST - 1];
        for (int i = PSB_HIST - 1; i > 0; --i) v.s[i] = v.s[i - 1], v.y[i] = v.y[i - 1];
        v.s[0] = s_last, v.y[0] = y_last;
        vec_sub_to(v.s[0], v.xt, v.x, v.n4);
        vec_sub_to(v.y[0], v.gt, v.g, v.n4);
        __syncthreads();
        if (hist < PSB_HIST) ++hist;
        float* p;
        p = v.x, v.x = v.xt, v.xt = p;
        p = v.g, v.g = v.gt, v.gt = p;
        f = f_new;
        r.iterations = it + 1;
    }
    r.loss = f;
    return r;
}
