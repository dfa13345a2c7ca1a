// Pieces shared by the SIMT and the tcgen05 worker kernels: scratch carving, softmax + residual of one
// row, standardisation statistics, scaled-space conversion, centring and delta.
#pragma once
#include "common.cuh"
#include "solver.cuh"

// Per-worker global scratch: PSB_SCRATCH_VECS vectors of nv floats (L2-resident).
struct WorkerScratch {
    float* vec[PSB_SCRATCH_VECS];
    float* w0;
    __device__ WorkerScratch(float* base, int nv) {
        for (int i = 0; i < PSB_SCRATCH_VECS; ++i) vec[i] = base + (size_t)i * nv;
        w0 = vec[6];
    }
    __device__ void bind(SolveVecs& v) const {
        v.x = vec[0], v.g = vec[1], v.d = vec[2], v.xt = vec[3], v.gt = vec[4], v.tmp = vec[5];
        for (int i = 0; i < PSB_HIST; ++i) v.s[i] = vec[7 + i], v.y[i] = vec[7 + PSB_HIST + i];
    }
};

// z: raw logits without intercept (8 values, pads ignored); b: intercepts; returns -log p[label] and the
// residual r = p - onehot(label) (pads 0).  Softmax over the R class rows only (class 0 included: Q4).
__device__ __forceinline__ float softmax_residual(const float* z, const float* b, int label, int R, float* r) {
    float v[8], m = -INFINITY, vy = 0.f;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        v[c] = c < R ? z[c] + b[c] : -INFINITY;
        m = fmaxf(m, v[c]);
        if (c == label) vy = v[c];
    }
    float sum = 0.f;
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        v[c] = c < R ? __expf(v[c] - m) : 0.f;
        sum += v[c];
    }
    const float is = 1.f / sum;
#pragma unroll
    for (int c = 0; c < 8; ++c) r[c] = v[c] * is - (c == label ? 1.f : 0.f);
    return __logf(sum) - (vy - m);          // log-sum-exp form: exact even when p[label] underflows
}

// inv[f] = 1/std_f over the window (unbiased, zeros included) or 0 when the variance is 0 (Q5);
// all ones when standardisation is off.
__device__ __forceinline__ void feature_inv_std(const __half* X, int start, int count, int cap, int F, int standardize,
                                                float* inv) {
    if (!standardize) {
        for (int f = threadIdx.x; f < F; f += blockDim.x) inv[f] = 1.f;
        return;
    }
    for (int fp = threadIdx.x; fp < (F >> 1); fp += blockDim.x) {
        const __half2* xc = (const __half2*)X + fp;
        float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll 4
        for (int i = 0; i < count; ++i) {
            int s = start + i;
            if (s >= cap) s -= cap;
            const float2 xv = __half22float2(xc[(size_t)s * (F >> 1)]);
            s0 += xv.x, q0 = fmaf(xv.x, xv.x, q0);
            s1 += xv.y, q1 = fmaf(xv.y, xv.y, q1);
        }
        const float n = (float)count;
        const float v0 = count > 1 ? (q0 - s0 * s0 / n) / (n - 1.f) : 0.f;
        const float v1 = count > 1 ? (q1 - s1 * s1 / n) / (n - 1.f) : 0.f;
        inv[2 * fp] = v0 > 1e-12f ? rsqrtf(v0) : 0.f;
        inv[2 * fp + 1] = v1 > 1e-12f ? rsqrtf(v1) : 0.f;
    }
}

// x = w0 * std  (std = 1/inv, 0 when inv == 0); intercept row copied
__device__ __forceinline__ void to_scaled_space(float* x, const float* w0, const float* inv, int F) {
    vec_prefetch(w0, (F + 1) * 2);
    for (int i = threadIdx.x; i < (F + 1) * 2; i += blockDim.x) {
        const int f = i >> 1;
        float4 v = ((const float4*)w0)[i];
        float sc = 1.f;
        if (f < F) sc = inv[f] > 0.f ? 1.f / inv[f] : 0.f;
        ((float4*)x)[i] = make_float4(v.x * sc, v.y * sc, v.z * sc, v.w * sc);
    }
}

// delta = (x * inv, centred across the R class rows when asked) - w0
__device__ __forceinline__ void finish_delta(float* delta, const float* x, const float* w0, const float* inv, int F,
                                             int R, int center) {
    vec_prefetch(x, (F + 1) * 2), vec_prefetch(w0, (F + 1) * 2);
    for (int f = threadIdx.x; f <= F; f += blockDim.x) {
        const float sc = f < F ? inv[f] : 1.f;
        float v[8];
        const float4 a = ((const float4*)x)[2 * f], b = ((const float4*)x)[2 * f + 1];
        v[0] = a.x * sc, v[1] = a.y * sc, v[2] = a.z * sc, v[3] = a.w * sc;
        v[4] = b.x * sc, v[5] = b.y * sc, v[6] = b.z * sc, v[7] = b.w * sc;
        if (center) {
            float m = 0.f;
#pragma unroll
            for (int c = 0; c < 8; ++c) m += c < R ? v[c] : 0.f;
            m /= (float)R;
#pragma unroll
            for (int c = 0; c < 8; ++c) v[c] = c < R ? v[c] - m : 0.f;
        }
        const float4 o0 = ((const float4*)w0)[2 * f], o1 = ((const float4*)w0)[2 * f + 1];
        ((float4*)delta)[2 * f] = make_float4(v[0] - o0.x, v[1] - o0.y, v[2] - o0.z, v[3] - o0.w);
        ((float4*)delta)[2 * f + 1] = make_float4(v[4] - o1.x, v[5] - o1.y, v[6] - o1.z, v[7] - o1.w);
    }
}
