// mha_mp.cu -- `--gnn_type MultiHeadAttention`: RelationalMultiHeadAttentionMP (fs_mol/modules/gnn.py:268-387) on the
// type-sorted edge layout of the tower kernels (edges sorted by (type, target, message index); tptr[t*V + v] .. tptr[t*V + v + 1]
// = positions of the type-t edges arriving at v; rev[e] = position of the reversed copy of edge e).
//
//   XG [E,256]    = [x_src | x_tgt]                                  adkf_mha_gather_inputs   (message MLP input, :327-332)
//   pre [E,256]   = per-type Linear(256 -> 256) on the tensor cores  adkf_gemm_tf32x3
//   QK [V,1536]   = x [Wq_0 Wq_1 Wq_2 Wk_0 Wk_1 Wk_2]^T              adkf_gemm_tf32x3: the reference projects x[tgt] / x[src]
//                   per EDGE (:334-341); projecting every node once gives the same rows
//   score[e,h]    = 64^-1/2 <Q_type(e)[tgt,h,:], K_type(e)[src,h,:]>  (:343-348)
//   p[e,h]        = exp(scatter_log_softmax over the edges arriving at tgt, all types)  (:365-371)
//   out[v,h,:]    = sum_e p[e,h] relu(pre[e,h,:])                    (:373-382)            adkf_mha_attention_fwd
// One warp per (node, head); lane owns channels lane and lane + 32 of the 64-wide head.  Deterministic: no atomics, the
// key gradient is gathered through rev[] instead of scattered.
#include "common.cuh"
#include "../../include/adkf_b200.h"

namespace adkf {

constexpr int mH = 4, mD = 64, mHid = 128, mMsg = mH * mD, mTypes = 3, mQK = 2 * mTypes * mMsg;   // 256, 1536
constexpr float mScale = 0.125f;   // per_head_dim ** -0.5

__global__ void mha_gather_inputs_kernel(const float* __restrict__ x, const int32_t* __restrict__ src,
                                         const int32_t* __restrict__ tgt, int64_t E, float* __restrict__ XG) {
    const int64_t total = E * (2 * mHid / 4);
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < total; p += (int64_t)gridDim.x * blockDim.x) {
        const int64_t e = p >> 6;
        const int q = (int)(p & 63);                     // float4 index inside the 256-float row
        const int node = (q < 32) ? src[e] : tgt[e];
        *reinterpret_cast<float4*>(XG + e * (2 * mHid) + q * 4) =
            *reinterpret_cast<const float4*>(x + (size_t)node * mHid + (q & 31) * 4);
    }
}

// d_x[v] = sum over edges e arriving at v of ( dXG[e, 128:256]  (v is the target of e)  +  dXG[rev[e], 0:128]  (v is the source of rev[e]) )
__global__ void mha_dx_gather_kernel(const float* __restrict__ dXG, const int32_t* __restrict__ tptr,
                                     const int32_t* __restrict__ rev, int V, float* __restrict__ d_x) {
    const int64_t total = (int64_t)V * (mHid / 4);
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < total; p += (int64_t)gridDim.x * blockDim.x) {
        const int v = (int)(p >> 5), c4 = (int)(p & 31);
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int t = 0; t < mTypes; ++t) {
            const int beg = tptr[(size_t)t * V + v], end = tptr[(size_t)t * V + v + 1];
            for (int r = beg; r < end; ++r) {
                const float4 a = *reinterpret_cast<const float4*>(dXG + (size_t)r * (2 * mHid) + mHid + c4 * 4);
                const float4 b = *reinterpret_cast<const float4*>(dXG + (size_t)rev[r] * (2 * mHid) + c4 * 4);
                acc.x += a.x + b.x; acc.y += a.y + b.y; acc.z += a.z + b.z; acc.w += a.w + b.w;
            }
        }
        *reinterpret_cast<float4*>(d_x + (size_t)v * mHid + c4 * 4) = acc;
    }
}

__global__ void __launch_bounds__(256)
mha_attention_fwd_kernel(const float* __restrict__ pre, const float* __restrict__ QK, const int32_t* __restrict__ src,
                         const int32_t* __restrict__ tptr, int V, float* __restrict__ out, float* __restrict__ probs) {
    const int lane = threadIdx.x & 31;
    const int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (w >= (int64_t)V * mH) return;
    const int v = (int)(w >> 2), h = (int)(w & 3);
    int beg[mTypes], end[mTypes];
#pragma unroll
    for (int t = 0; t < mTypes; ++t) { beg[t] = tptr[(size_t)t * V + v]; end[t] = tptr[(size_t)t * V + v + 1]; }
    // pass 1: scores (kept in probs[]) and their maximum
    float mx = -INFINITY;
#pragma unroll
    for (int t = 0; t < mTypes; ++t) {
        const float* q = QK + (size_t)v * mQK + t * mMsg + h * mD + lane;
        const float q0 = q[0], q1 = q[32];
        for (int r = beg[t]; r < end[t]; ++r) {
            const float* k = QK + (size_t)src[r] * mQK + (mTypes + t) * mMsg + h * mD + lane;
            const float s = warp_sum((mScale * q0) * k[0] + (mScale * q1) * k[32]);
            if (lane == 0) probs[(size_t)r * mH + h] = s;
            mx = fmaxf(mx, s);
        }
    }
    __syncwarp();
    float den = 0.f;
#pragma unroll
    for (int t = 0; t < mTypes; ++t)
        for (int r = beg[t]; r < end[t]; ++r) den += expf(probs[(size_t)r * mH + h] - mx);
    const float lse = logf(den);
    float a0 = 0.f, a1 = 0.f;
#pragma unroll
    for (int t = 0; t < mTypes; ++t)
        for (int r = beg[t]; r < end[t]; ++r) {
            const float p = expf(probs[(size_t)r * mH + h] - mx - lse);     // exp(log_softmax), as the reference
            const float* m = pre + (size_t)r * mMsg + h * mD + lane;
            a0 = fmaf(p, fmaxf(m[0], 0.f), a0);
            a1 = fmaf(p, fmaxf(m[32], 0.f), a1);
            __syncwarp();
            if (lane == 0) probs[(size_t)r * mH + h] = p;
        }
    float* o = out + (size_t)v * mMsg + h * mD + lane;
    o[0] = a0;
    o[32] = a1;
}

// per (target v, head h): d_pre, the score gradients ds (softmax backward) and the query gradient
__global__ void __launch_bounds__(256)
mha_attention_bwd_kernel(const float* __restrict__ pre, const float* __restrict__ QK, const int32_t* __restrict__ src,
                         const int32_t* __restrict__ tptr, int V, const float* __restrict__ probs,
                         const float* __restrict__ d_out, float* __restrict__ d_pre, float* __restrict__ ds,
                         float* __restrict__ dQK) {
    const int lane = threadIdx.x & 31;
    const int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (w >= (int64_t)V * mH) return;
    const int v = (int)(w >> 2), h = (int)(w & 3);
    int beg[mTypes], end[mTypes];
#pragma unroll
    for (int t = 0; t < mTypes; ++t) { beg[t] = tptr[(size_t)t * V + v]; end[t] = tptr[(size_t)t * V + v + 1]; }
    const float* g = d_out + (size_t)v * mMsg + h * mD + lane;
    const float g0 = g[0], g1 = g[32];
    float S = 0.f;                                            // sum_e p_e <g, m_e>
#pragma unroll
    for (int t = 0; t < mTypes; ++t)
        for (int r = beg[t]; r < end[t]; ++r) {
            const float* m = pre + (size_t)r * mMsg + h * mD + lane;
            const float dp = warp_sum(g0 * fmaxf(m[0], 0.f) + g1 * fmaxf(m[32], 0.f));
            if (lane == 0) ds[(size_t)r * mH + h] = dp;
            S = fmaf(probs[(size_t)r * mH + h], dp, S);
        }
    __syncwarp();
#pragma unroll
    for (int t = 0; t < mTypes; ++t) {
        float dq0 = 0.f, dq1 = 0.f;
        for (int r = beg[t]; r < end[t]; ++r) {
            const float p = probs[(size_t)r * mH + h];
            const float dsv = p * (ds[(size_t)r * mH + h] - S);
            const float* m = pre + (size_t)r * mMsg + h * mD + lane;
            float* dm = d_pre + (size_t)r * mMsg + h * mD + lane;
            dm[0] = (m[0] > 0.f) ? p * g0 : 0.f;
            dm[32] = (m[32] > 0.f) ? p * g1 : 0.f;
            const float* k = QK + (size_t)src[r] * mQK + (mTypes + t) * mMsg + h * mD + lane;
            dq0 = fmaf(mScale * dsv, k[0], dq0);
            dq1 = fmaf(mScale * dsv, k[32], dq1);
            __syncwarp();
            if (lane == 0) ds[(size_t)r * mH + h] = dsv;
        }
        float* dq = dQK + (size_t)v * mQK + t * mMsg + h * mD + lane;
        dq[0] = dq0;
        dq[32] = dq1;
    }
}

// per (source u, head h): dK_t[u] = sum over the edges u -> w of type t of scale ds[e] Q_t[w]; the edges out of u are the
// reversed copies of the edges arriving at u
__global__ void __launch_bounds__(256)
mha_attention_bwd_key_kernel(const float* __restrict__ QK, const int32_t* __restrict__ src, const int32_t* __restrict__ rev,
                             const int32_t* __restrict__ tptr, int V, const float* __restrict__ ds, float* __restrict__ dQK) {
    const int lane = threadIdx.x & 31;
    const int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (w >= (int64_t)V * mH) return;
    const int u = (int)(w >> 2), h = (int)(w & 3);
#pragma unroll
    for (int t = 0; t < mTypes; ++t) {
        const int beg = tptr[(size_t)t * V + u], end = tptr[(size_t)t * V + u + 1];
        float dk0 = 0.f, dk1 = 0.f;
        for (int r = beg; r < end; ++r) {
            const int tgt_of_rev = src[r];                    // rev[r] runs u -> src[r]
            const float dsv = ds[(size_t)rev[r] * mH + h];
            const float* q = QK + (size_t)tgt_of_rev * mQK + t * mMsg + h * mD + lane;
            dk0 = fmaf(mScale * dsv, q[0], dk0);
            dk1 = fmaf(mScale * dsv, q[32], dk1);
        }
        float* dk = dQK + (size_t)u * mQK + (mTypes + t) * mMsg + h * mD + lane;
        dk[0] = dk0;
        dk[32] = dk1;
    }
}

}  // namespace adkf

using namespace adkf;

static inline int mha_grid_cap(int64_t n, int block, int cap = 148 * 32) {
    int64_t g = (n + block - 1) / block;
    if (g > cap) g = cap;
    return g < 1 ? 1 : (int)g;
}

extern "C" int adkf_mha_gather_inputs(const float* x, const int32_t* src, const int32_t* tgt, int64_t E, float* XG, void* stream) {
    if (E <= 0) return 0;
    if (!x || !src || !tgt || !XG) return ADKF_ERR_BADARG;
    mha_gather_inputs_kernel<<<mha_grid_cap(E * 64, 256), 256, 0, (cudaStream_t)stream>>>(x, src, tgt, E, XG);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_mha_dx_gather(const float* dXG, const int32_t* tptr, const int32_t* rev, int32_t V, float* d_x, void* stream) {
    if (V <= 0) return 0;
    if (!tptr || !d_x) return ADKF_ERR_BADARG;
    mha_dx_gather_kernel<<<mha_grid_cap((int64_t)V * 32, 256), 256, 0, (cudaStream_t)stream>>>(dXG, tptr, rev, V, d_x);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_mha_attention_fwd(const float* pre, const float* QK, const int32_t* src, const int32_t* tptr, int32_t V,
                                      float* out, float* probs, void* stream) {
    if (V <= 0) return 0;
    if (!QK || !tptr || !out) return ADKF_ERR_BADARG;
    const int64_t threads = (int64_t)V * mH * 32;
    mha_attention_fwd_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(pre, QK, src, tptr, V, out, probs);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_mha_attention_bwd(const float* pre, const float* QK, const int32_t* src, const int32_t* rev,
                                      const int32_t* tptr, int32_t V, const float* probs, const float* d_out, float* d_pre,
                                      float* ds_ws, float* dQK, void* stream) {
    if (V <= 0) return 0;
    if (!QK || !tptr || !d_out || !dQK) return ADKF_ERR_BADARG;
    const int64_t threads = (int64_t)V * mH * 32;
    const unsigned grid = (unsigned)((threads + 255) / 256);
    cudaStream_t s = (cudaStream_t)stream;
    mha_attention_bwd_kernel<<<grid, 256, 0, s>>>(pre, QK, src, tptr, V, probs, d_out, d_pre, ds_ws, dQK);
    ADKF_LAUNCH_CHECK();
    mha_attention_bwd_key_kernel<<<grid, 256, 0, s>>>(QK, src, rev, tptr, V, ds_ws, dQK);
    adkf_count_launches(2);
    ADKF_LAUNCH_CHECK();
    return 0;
}
