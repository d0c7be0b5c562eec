// pna_gemm.cu -- PNA relational message passing as "gather -> batched tensor-core GEMM -> segment kernels" (sm_100a).
//
// Same maths as graph.cu's fused SIMT kernels (RelationalMP.forward + RelationalMultiAggrMP._aggregate_messages,
// fs_mol/modules/gnn.py:129-150, 197-265), reorganised so that the three dense products of a block -- the message MLP
// relu(W[k][type] [x_src ; x_tgt] + b), its pull-back to the node states and its weight gradient -- run on the tcgen05
// 3xTF32 GEMM (adkf_gemm_tf32x3_batched: 4 towers per launch, one launch per edge type), and only memory-bound work is
// left to the kernels here:
//   pna_gather_inputs    XG[r] = [h[src_r] | h[tgt_r]] per tower, edges in (type, target, message index) order
//   pna_aggregate_fwd    pre-activations [E, 4*192] -> relu -> sum | mean | pseudo-std | max x 3 degree scalers -> [V, 3072]
//   pna_dpre             d out -> d pre-activation of every edge (relu masks, mean / std / arg-max chain rule)
//   pna_dh_gather        d h[v] = sum over incoming edges of dXG[., tgt half] + over outgoing edges (their reverses) of dXG[., src half]
// Edge order: r = position in the (type, target node, reference message index) sort; tptr[type * V + v] .. tptr[.. + 1] are
// the type-`type` edges arriving at v.  Walking types 0,1,2 at a node visits its edges in the reference's message order,
// which fixes the arg-max tie-breaking exactly as graph.cu does.  No float atomics: every output element has one owner.
#include "common.cuh"
#include "../../include/adkf_b200.h"

namespace adkf {

constexpr int pTin = 32, pMsg = 192, pPmd = 64, pTypes = 3, pTowers = 4, pHid = 128, pOutTower = 768;
constexpr int pPre = pTowers * pMsg;   // 768 pre-activations per edge
constexpr int pXg = pTowers * 64;      // 256 gathered inputs per edge
constexpr float pSmall = 1e-7f, pDelta = 1.1515f;

__global__ void pna_gather_inputs_kernel(const float* __restrict__ h, const int32_t* __restrict__ src,
                                         const int32_t* __restrict__ tgt, int64_t E, float* __restrict__ XG) {
    const int64_t total = E * (pXg / 4);
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < total; p += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = p / (pXg / 4);
        const int q = (int)(p - r * (pXg / 4));          // float4 index inside the 256-float row
        const int k = q >> 4, j4 = q & 15;               // tower, float4 inside [x_src(8 float4) | x_tgt(8 float4)]
        const int node = (j4 < 8) ? src[r] : tgt[r];
        const float4 v = *reinterpret_cast<const float4*>(h + (size_t)node * pHid + k * pTin + (j4 & 7) * 4);
        *reinterpret_cast<float4*>(XG + (size_t)r * pXg + q * 4) = v;
    }
}

// one warp per (node, tower); lane owns channels lane + 32 c, c = 0..5  (0,1 sum part | 2,3 mean/std part | 4,5 max part)
template <bool SCALERS>
__global__ void __launch_bounds__(256)
pna_aggregate_fwd_kernel(const float* __restrict__ pre, const int32_t* __restrict__ tptr, int V, float* __restrict__ out,
                         uint8_t* __restrict__ aux_cnt, uint8_t* __restrict__ aux_arg) {
    const int lane = threadIdx.x & 31;
    const int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (w >= (int64_t)V * pTowers) return;
    const int node = (int)(w >> 2), k = (int)(w & 3);
    float sum[2] = {0.f, 0.f}, macc[2] = {0.f, 0.f}, mx[2] = {0.f, 0.f};
    int arg[2] = {0, 0}, deg = 0;
    int beg[pTypes], end[pTypes];
#pragma unroll
    for (int t = 0; t < pTypes; ++t) { beg[t] = tptr[(size_t)t * V + node]; end[t] = tptr[(size_t)t * V + node + 1]; }
#pragma unroll
    for (int t = 0; t < pTypes; ++t)
        for (int r = beg[t]; r < end[t]; ++r) {
            const float* p = pre + (size_t)r * pPre + k * pMsg + lane;
            float m[6];
#pragma unroll
            for (int c = 0; c < 6; ++c) m[c] = fmaxf(p[32 * c], 0.f);
            sum[0] += m[0]; sum[1] += m[1];
            macc[0] += m[2]; macc[1] += m[3];
            if (deg == 0 || m[4] > mx[0]) { mx[0] = m[4]; arg[0] = deg; }
            if (deg == 0 || m[5] > mx[1]) { mx[1] = m[5]; arg[1] = deg; }
            ++deg;
        }
    const float fdeg = (float)max(deg, 1);
    const float mean[2] = {macc[0] / fdeg, macc[1] / fdeg};     // true division, like scatter_mean
    float ssum[2] = {0.f, 0.f};
    int cpos[2] = {0, 0};
#pragma unroll
    for (int t = 0; t < pTypes; ++t)
        for (int r = beg[t]; r < end[t]; ++r) {
            const float* p = pre + (size_t)r * pPre + k * pMsg + lane;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const float mm = fmaxf(p[32 * (2 + j)], 0.f);
                // no FMA contraction: m^2 - mean^2 must be exactly 0 when m == mean (degree-1 rows), as in torch
                const float v = __fsub_rn(__fmul_rn(mm, mm), __fmul_rn(mean[j], mean[j]));
                if (v > 0.f) { ssum[j] += v; ++cpos[j]; }
                ssum[j] += pSmall;
            }
        }
    const float logd = logf((float)deg + 1.f);
    const float amp = logd / pDelta, att = pDelta / (logd + pSmall);
    constexpr int OT = SCALERS ? pOutTower : 4 * pPmd;      // 768 with the three PNA degree scalers, 256 for "MultiAggr"
    float* o = out + (size_t)node * (pTowers * OT) + (size_t)k * OT;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const float vals[4] = {sum[j], mean[j], sqrtf(ssum[j]), (deg > 0) ? mx[j] : 0.f};
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            const int col = a * pPmd + j * 32 + lane;
            o[col] = vals[a];
            if (SCALERS) {
                o[256 + col] = amp * vals[a];
                o[512 + col] = att * vals[a];
            }
        }
        aux_cnt[((size_t)node * pTowers + k) * pPmd + j * 32 + lane] = (uint8_t)min(cpos[j], 255);
        aux_arg[((size_t)node * pTowers + k) * pPmd + j * 32 + lane] = (uint8_t)min(arg[j], 255);
    }
}

// d out [V,3072] -> d pre [E,768]; reads the saved pre-activations, the forward output (mean, std) and the aux bytes
template <bool SCALERS>
__global__ void __launch_bounds__(256)
pna_dpre_kernel(const float* __restrict__ pre, const int32_t* __restrict__ tptr, int V, const float* __restrict__ fwd_out,
                const uint8_t* __restrict__ aux_cnt, const uint8_t* __restrict__ aux_arg, const float* __restrict__ d_out,
                float* __restrict__ d_pre) {
    const int lane = threadIdx.x & 31;
    const int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (w >= (int64_t)V * pTowers) return;
    const int node = (int)(w >> 2), k = (int)(w & 3);
    int beg[pTypes], end[pTypes], deg = 0;
#pragma unroll
    for (int t = 0; t < pTypes; ++t) {
        beg[t] = tptr[(size_t)t * V + node]; end[t] = tptr[(size_t)t * V + node + 1];
        deg += end[t] - beg[t];
    }
    if (deg == 0) return;
    const float logd = logf((float)deg + 1.f);
    const float amp = logd / pDelta, att = pDelta / (logd + pSmall);
    constexpr int OT = SCALERS ? pOutTower : 4 * pPmd;
    const float* go = d_out + (size_t)node * (pTowers * OT) + (size_t)k * OT;
    const float* fo = fwd_out + (size_t)node * (pTowers * OT) + (size_t)k * OT;
    float g_sum[2], g_mean[2], g_std[2], g_max[2], mean[2], stdv[2], cpos[2];
    int arg[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
        const int col = j * 32 + lane;
        g_sum[j] = go[col];
        g_mean[j] = go[pPmd + col];
        g_std[j] = go[2 * pPmd + col];
        g_max[j] = go[3 * pPmd + col];
        if (SCALERS) {
            g_sum[j] += amp * go[256 + col] + att * go[512 + col];
            g_mean[j] += amp * go[256 + pPmd + col] + att * go[512 + pPmd + col];
            g_std[j] += amp * go[256 + 2 * pPmd + col] + att * go[512 + 2 * pPmd + col];
            g_max[j] += amp * go[256 + 3 * pPmd + col] + att * go[512 + 3 * pPmd + col];
        }
        mean[j] = fo[pPmd + col];
        stdv[j] = fo[2 * pPmd + col];
        cpos[j] = (float)aux_cnt[((size_t)node * pTowers + k) * pPmd + col];
        arg[j] = aux_arg[((size_t)node * pTowers + k) * pPmd + col];
    }
    const float invdeg = 1.f / (float)deg;
    int idx = 0;
#pragma unroll
    for (int t = 0; t < pTypes; ++t)
        for (int r = beg[t]; r < end[t]; ++r, ++idx) {
            const float* p = pre + (size_t)r * pPre + k * pMsg + lane;
            float* q = d_pre + (size_t)r * pPre + k * pMsg + lane;
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const float a_sum = p[32 * j], a_mean = p[32 * (2 + j)], a_max = p[32 * (4 + j)];
                const float m_mean = fmaxf(a_mean, 0.f);
                const float cc = 0.5f * g_std[j] / stdv[j];
                const float ind = (__fsub_rn(__fmul_rn(m_mean, m_mean), __fmul_rn(mean[j], mean[j])) > 0.f) ? 1.f : 0.f;
                const float dm_mean = g_mean[j] * invdeg + cc * (2.f * m_mean * ind - 2.f * mean[j] * cpos[j] * invdeg);
                q[32 * j] = (a_sum > 0.f) ? g_sum[j] : 0.f;
                q[32 * (2 + j)] = (a_mean > 0.f) ? dm_mean : 0.f;
                q[32 * (4 + j)] = (a_max > 0.f && (idx == arg[j] || (idx >= 255 && arg[j] == 255))) ? g_max[j] : 0.f;
            }
        }
}

// "Plain" relational message passing (gnn.py:152-165): out[v, 64k + c] = sum over incoming edges of relu(pre[e, 64k + c]);
// pre [E, 256] (4 towers x 64), one warp per (node, tower), lane owns channels lane and lane + 32.
__global__ void __launch_bounds__(256)
plain_aggregate_fwd_kernel(const float* __restrict__ pre, const int32_t* __restrict__ tptr, int V, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (w >= (int64_t)V * pTowers) return;
    const int node = (int)(w >> 2), k = (int)(w & 3);
    float s0 = 0.f, s1 = 0.f;
#pragma unroll
    for (int t = 0; t < pTypes; ++t) {
        const int beg = tptr[(size_t)t * V + node], end = tptr[(size_t)t * V + node + 1];
        for (int r = beg; r < end; ++r) {
            const float* p = pre + (size_t)r * (pTowers * pPmd) + k * pPmd + lane;
            s0 += fmaxf(p[0], 0.f);
            s1 += fmaxf(p[32], 0.f);
        }
    }
    float* o = out + (size_t)node * (pTowers * pPmd) + k * pPmd + lane;
    o[0] = s0;
    o[32] = s1;
}

__global__ void __launch_bounds__(256)
plain_dpre_kernel(const float* __restrict__ pre, const int32_t* __restrict__ tptr, int V, const float* __restrict__ d_out,
                  float* __restrict__ d_pre) {
    const int lane = threadIdx.x & 31;
    const int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (w >= (int64_t)V * pTowers) return;
    const int node = (int)(w >> 2), k = (int)(w & 3);
    const float* go = d_out + (size_t)node * (pTowers * pPmd) + k * pPmd + lane;
    const float g0 = go[0], g1 = go[32];
#pragma unroll
    for (int t = 0; t < pTypes; ++t) {
        const int beg = tptr[(size_t)t * V + node], end = tptr[(size_t)t * V + node + 1];
        for (int r = beg; r < end; ++r) {
            const float* p = pre + (size_t)r * (pTowers * pPmd) + k * pPmd + lane;
            float* q = d_pre + (size_t)r * (pTowers * pPmd) + k * pPmd + lane;
            q[0] = (p[0] > 0.f) ? g0 : 0.f;
            q[32] = (p[32] > 0.f) ? g1 : 0.f;
        }
    }
}

__global__ void pna_dh_gather_kernel(const float* __restrict__ dXG, const int32_t* __restrict__ tptr,
                                     const int32_t* __restrict__ rev, int V, float* __restrict__ d_h) {
    const int64_t total = (int64_t)V * (pHid / 4);
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < total; p += (int64_t)gridDim.x * blockDim.x) {
        const int v = (int)(p / (pHid / 4)), c4 = (int)(p - (int64_t)v * (pHid / 4));   // float4 column of h
        const int k = c4 >> 3, j4 = c4 & 7;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int t = 0; t < pTypes; ++t) {
            const int beg = tptr[(size_t)t * V + v], end = tptr[(size_t)t * V + v + 1];
            for (int r = beg; r < end; ++r) {
                const float4 a = *reinterpret_cast<const float4*>(dXG + (size_t)r * pXg + k * 64 + 32 + j4 * 4);        // v is the target
                const float4 b = *reinterpret_cast<const float4*>(dXG + (size_t)rev[r] * pXg + k * 64 + j4 * 4);        // v is the source
                acc.x += a.x + b.x; acc.y += a.y + b.y; acc.z += a.z + b.z; acc.w += a.w + b.w;
            }
        }
        *reinterpret_cast<float4*>(d_h + (size_t)v * pHid + c4 * 4) = acc;
    }
}

}  // namespace adkf

using namespace adkf;

static inline int grid_cap(int64_t n, int block, int cap = 148 * 32) {
    int64_t g = (n + block - 1) / block;
    if (g > cap) g = cap;
    return g < 1 ? 1 : (int)g;
}

extern "C" int adkf_pna_gather_inputs(const float* h, const int32_t* src, const int32_t* tgt, int64_t E, float* XG, void* stream) {
    if (E <= 0) return 0;
    if (!h || !src || !tgt || !XG) return ADKF_ERR_BADARG;
    pna_gather_inputs_kernel<<<grid_cap(E * (pXg / 4), 256), 256, 0, (cudaStream_t)stream>>>(h, src, tgt, E, XG);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

// mode: ADKF_MP_PNA (sum|mean|std|max x 3 degree scalers, out [V,3072]), ADKF_MP_MULTIAGGR (no scalers, out [V,1024]),
// ADKF_MP_PLAIN (relu + sum of 64-wide messages, pre [E,256], out [V,256]; aux arrays unused)
extern "C" int adkf_mp_aggregate_fwd(const float* pre, const int32_t* tptr, int32_t V, int32_t mode, float* out, uint8_t* aux_cnt,
                                     uint8_t* aux_arg, void* stream) {
    if (V <= 0) return 0;
    if (!tptr || !out || mode < ADKF_MP_PNA || mode > ADKF_MP_PLAIN) return ADKF_ERR_BADARG;
    if (mode != ADKF_MP_PLAIN && (!aux_cnt || !aux_arg)) return ADKF_ERR_BADARG;
    const int64_t threads = (int64_t)V * pTowers * 32;
    const unsigned grid = (unsigned)((threads + 255) / 256);
    cudaStream_t s = (cudaStream_t)stream;
    if (mode == ADKF_MP_PNA) pna_aggregate_fwd_kernel<true><<<grid, 256, 0, s>>>(pre, tptr, V, out, aux_cnt, aux_arg);
    else if (mode == ADKF_MP_MULTIAGGR) pna_aggregate_fwd_kernel<false><<<grid, 256, 0, s>>>(pre, tptr, V, out, aux_cnt, aux_arg);
    else plain_aggregate_fwd_kernel<<<grid, 256, 0, s>>>(pre, tptr, V, out);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_mp_dpre(const float* pre, const int32_t* tptr, int32_t V, int32_t mode, const float* fwd_out,
                            const uint8_t* aux_cnt, const uint8_t* aux_arg, const float* d_out, float* d_pre, void* stream) {
    if (V <= 0) return 0;
    if (!tptr || !d_out || mode < ADKF_MP_PNA || mode > ADKF_MP_PLAIN) return ADKF_ERR_BADARG;
    if (mode != ADKF_MP_PLAIN && (!fwd_out || !aux_cnt || !aux_arg)) return ADKF_ERR_BADARG;
    const int64_t threads = (int64_t)V * pTowers * 32;
    const unsigned grid = (unsigned)((threads + 255) / 256);
    cudaStream_t s = (cudaStream_t)stream;
    if (mode == ADKF_MP_PNA) pna_dpre_kernel<true><<<grid, 256, 0, s>>>(pre, tptr, V, fwd_out, aux_cnt, aux_arg, d_out, d_pre);
    else if (mode == ADKF_MP_MULTIAGGR) pna_dpre_kernel<false><<<grid, 256, 0, s>>>(pre, tptr, V, fwd_out, aux_cnt, aux_arg, d_out, d_pre);
    else plain_dpre_kernel<<<grid, 256, 0, s>>>(pre, tptr, V, d_out, d_pre);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_pna_aggregate_fwd(const float* pre, const int32_t* tptr, int32_t V, float* out, uint8_t* aux_cnt,
                                      uint8_t* aux_arg, void* stream) {
    return adkf_mp_aggregate_fwd(pre, tptr, V, ADKF_MP_PNA, out, aux_cnt, aux_arg, stream);
}

extern "C" int adkf_pna_dpre(const float* pre, const int32_t* tptr, int32_t V, const float* fwd_out, const uint8_t* aux_cnt,
                             const uint8_t* aux_arg, const float* d_out, float* d_pre, void* stream) {
    return adkf_mp_dpre(pre, tptr, V, ADKF_MP_PNA, fwd_out, aux_cnt, aux_arg, d_out, d_pre, stream);
}

extern "C" int adkf_pna_dh_gather(const float* dXG, const int32_t* tptr, const int32_t* rev, int32_t V, float* d_h, void* stream) {
    if (V <= 0) return 0;
    if (!tptr || !d_h) return ADKF_ERR_BADARG;
    pna_dh_gather_kernel<<<grid_cap((int64_t)V * (pHid / 4), 256), 256, 0, (cudaStream_t)stream>>>(dXG, tptr, rev, V, d_h);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}
