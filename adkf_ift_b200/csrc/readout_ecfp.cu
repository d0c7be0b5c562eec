// readout_ecfp.cu -- graph readout segment kernels and the sparse ECFP branch of the first FC layer (sm_100a).
//
// Readout replaces torch_scatter's scatter_softmax / index_add_ / scatter(max) in
// MultiHeadWeightedGraphReadout.forward and UnweightedGraphReadout.forward
// (fs_mol/modules/graph_readout.py:224-259, 286-296).  Nodes of a graph are contiguous in the collated
// batch (fs_mol/data/fsmol_batcher.py:200-202), so every graph is one CTA and every reduction is an
// ordered loop over its node range: no atomics, bit-stable.
//
// ECFP replaces `support_features.fingerprints.float()` + the fingerprint columns of fc[0]
// (fs_mol/models/adaptive_dkt.py:145-146,158-160): the count fingerprint has ~2% non-zeros, so the
// [G,2048]x[2048,2048] GEMM becomes a gather-accumulate over the non-zero bits (HBM/L2 bound).
#include "common.cuh"
#include "../../include/adkf_b200.h"

namespace adkf {

// ---------------------------------------------------------------------------------------------
// multi-head weighted pooling
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
readout_weighted_fwd_kernel(const float* __restrict__ scores, const float* __restrict__ values,
                            const int32_t* __restrict__ graph_ptr, int heads, int head_dim, int mode,
                            float* __restrict__ weights, float* __restrict__ out) {
    const int g = blockIdx.x;
    const int v0 = graph_ptr[g], v1 = graph_ptr[g + 1];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    // step 1: weights
    for (int hd = wid; hd < heads; hd += nw) {
        if (mode == 0) {
            float mx = -INFINITY;
            for (int v = v0 + lane; v < v1; v += 32) mx = fmaxf(mx, scores[(size_t)v * heads + hd]);
            mx = warp_max(mx);
            float sm = 0.f;
            for (int v = v0 + lane; v < v1; v += 32) sm += expf(scores[(size_t)v * heads + hd] - mx);
            sm = warp_sum(sm);
            for (int v = v0 + lane; v < v1; v += 32)
                weights[(size_t)v * heads + hd] = expf(scores[(size_t)v * heads + hd] - mx) / sm;
        } else {
            for (int v = v0 + lane; v < v1; v += 32)
                weights[(size_t)v * heads + hd] = sigmoidf_(scores[(size_t)v * heads + hd]);
        }
    }
    __syncthreads();
    // step 2: weighted sum per output column
    const int D = heads * head_dim;
    for (int c = threadIdx.x; c < D; c += blockDim.x) {
        const int hd = c / head_dim;
        float acc = 0.f;
        for (int v = v0; v < v1; ++v) acc = fmaf(weights[(size_t)v * heads + hd], values[(size_t)v * D + c], acc);
        out[(size_t)g * D + c] = acc;
    }
}

__global__ void __launch_bounds__(256)
readout_weighted_bwd_kernel(const float* __restrict__ weights, const float* __restrict__ values,
                            const int32_t* __restrict__ graph_ptr, int heads, int head_dim, int mode,
                            const float* __restrict__ d_out, float* __restrict__ d_scores,
                            float* __restrict__ d_values) {
    const int g = blockIdx.x;
    const int v0 = graph_ptr[g], v1 = graph_ptr[g + 1];
    const int D = heads * head_dim;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    // d_values and d_w (stored temporarily in d_scores)
    for (int c = threadIdx.x; c < D; c += blockDim.x) {
        const int hd = c / head_dim;
        const float go = d_out[(size_t)g * D + c];
        for (int v = v0; v < v1; ++v) d_values[(size_t)v * D + c] = weights[(size_t)v * heads + hd] * go;
    }
    const int pairs = (v1 - v0) * heads;
    for (int p = wid; p < pairs; p += nw) {
        const int v = v0 + p / heads, hd = p % heads;
        float acc = 0.f;
        for (int c = lane; c < head_dim; c += 32)
            acc = fmaf(values[(size_t)v * D + hd * head_dim + c], d_out[(size_t)g * D + hd * head_dim + c], acc);
        acc = warp_sum(acc);
        if (lane == 0) d_scores[(size_t)v * heads + hd] = acc;
    }
    __syncthreads();
    for (int hd = wid; hd < heads; hd += nw) {
        if (mode == 0) {
            float dot = 0.f;
            for (int v = v0 + lane; v < v1; v += 32) dot += weights[(size_t)v * heads + hd] * d_scores[(size_t)v * heads + hd];
            dot = warp_sum(dot);
            for (int v = v0 + lane; v < v1; v += 32) {
                const float w = weights[(size_t)v * heads + hd];
                d_scores[(size_t)v * heads + hd] = w * (d_scores[(size_t)v * heads + hd] - dot);
            }
        } else {
            for (int v = v0 + lane; v < v1; v += 32) {
                const float w = weights[(size_t)v * heads + hd];
                d_scores[(size_t)v * heads + hd] *= w * (1.f - w);
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// max pooling
// ---------------------------------------------------------------------------------------------
__global__ void readout_max_fwd_kernel(const float* __restrict__ x, const int32_t* __restrict__ graph_ptr, int dim,
                                       float* __restrict__ out, int32_t* __restrict__ argmax) {
    const int g = blockIdx.x;
    const int v0 = graph_ptr[g], v1 = graph_ptr[g + 1];
    for (int c = threadIdx.x; c < dim; c += blockDim.x) {
        float best = 0.f;
        int arg = -1;
        for (int v = v0; v < v1; ++v) {
            const float val = x[(size_t)v * dim + c];
            if (arg < 0 || val > best) { best = val; arg = v; }
        }
        out[(size_t)g * dim + c] = best;   // empty graph -> 0 (torch_scatter semantics)
        argmax[(size_t)g * dim + c] = arg;
    }
}

__global__ void readout_max_bwd_kernel(const float* __restrict__ d_out, const int32_t* __restrict__ argmax,
                                       const int32_t* __restrict__ graph_ptr, int dim, float* __restrict__ d_x) {
    const int g = blockIdx.x;
    const int v0 = graph_ptr[g], v1 = graph_ptr[g + 1];
    for (int c = threadIdx.x; c < dim; c += blockDim.x) {
        const int arg = argmax[(size_t)g * dim + c];
        const float go = d_out[(size_t)g * dim + c];
        for (int v = v0; v < v1; ++v) d_x[(size_t)v * dim + c] = (v == arg) ? go : 0.f;
    }
}

// ---------------------------------------------------------------------------------------------
// ECFP sparse linear
// ---------------------------------------------------------------------------------------------
constexpr int kEcfpList = 512;

// block-wide ordered compaction of the non-zero fingerprint bits of one molecule (index order preserved)
__device__ int compact_nonzeros(const int32_t* __restrict__ fp, int base, int count, int* s_idx, float* s_val,
                                int* s_warp, int* s_total) {
    // processes fp[base .. base+count) with count <= blockDim.x; appends to the list after *s_total
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nw = blockDim.x >> 5;
    const int v = (tid < count) ? fp[base + tid] : 0;
    const unsigned bal = __ballot_sync(0xffffffffu, v != 0);
    if (lane == 0) s_warp[wid] = __popc(bal);
    __syncthreads();
    int prior = *s_total;
    for (int w = 0; w < wid; ++w) prior += s_warp[w];
    if (v != 0) {
        const int pos = prior + __popc(bal & ((1u << lane) - 1u));
        if (pos < kEcfpList) { s_idx[pos] = base + tid; s_val[pos] = (float)v; }
    }
    __syncthreads();
    if (tid == 0) {
        int tot = *s_total;
        for (int w = 0; w < nw; ++w) tot += s_warp[w];
        *s_total = tot;
    }
    __syncthreads();
    return *s_total;
}

__global__ void __launch_bounds__(256)
ecfp_linear_fwd_kernel(const int32_t* __restrict__ fp, int fp_dim, const float* __restrict__ Wt, int out_dim,
                       float* __restrict__ out) {
    __shared__ int s_idx[kEcfpList];
    __shared__ float s_val[kEcfpList];
    __shared__ int s_warp[8];
    __shared__ int s_total;
    const int g = blockIdx.x;
    const int32_t* f = fp + (size_t)g * fp_dim;
    float* o = out + (size_t)g * out_dim;
    int base = 0;
    while (base < fp_dim) {
        if (threadIdx.x == 0) s_total = 0;
        __syncthreads();
        // fill the list (up to kEcfpList entries; a window stops early only if the list could overflow)
        int n = 0;
        while (base < fp_dim && n + (int)blockDim.x <= kEcfpList) {
            const int cnt = min((int)blockDim.x, fp_dim - base);
            n = compact_nonzeros(f, base, cnt, s_idx, s_val, s_warp, &s_total);
            base += cnt;
        }
        for (int c = threadIdx.x * 4; c < out_dim; c += blockDim.x * 4) {
            float4 acc = *reinterpret_cast<float4*>(o + c);
            for (int i = 0; i < n; ++i) {
                const float4 w = *reinterpret_cast<const float4*>(Wt + (size_t)s_idx[i] * out_dim + c);
                const float v = s_val[i];
                acc.x = fmaf(v, w.x, acc.x); acc.y = fmaf(v, w.y, acc.y);
                acc.z = fmaf(v, w.z, acc.z); acc.w = fmaf(v, w.w, acc.w);
            }
            *reinterpret_cast<float4*>(o + c) = acc;
        }
        __syncthreads();
    }
}

// d_Wt[j, c] += sum_g fp[g, j] * d_out[g, c].  CTA = (32 fingerprint bits) x (256 output columns); the 32x256
// accumulator tile lives in shared memory and graphs are visited in order -> deterministic.
constexpr int kEbJ = 32, kEbC = 256;
__global__ void __launch_bounds__(256)
ecfp_linear_bwd_kernel(const int32_t* __restrict__ fp, int G, int fp_dim, const float* __restrict__ d_out,
                       int out_dim, float* __restrict__ d_Wt) {
    __shared__ float acc[kEbJ][kEbC];
    __shared__ int s_g[256 * 4];
    __shared__ int s_j[256 * 4];
    __shared__ float s_v[256 * 4];
    __shared__ int s_cnt[256];
    const int j0 = blockIdx.x * kEbJ, c0 = blockIdx.y * kEbC;
    const int tid = threadIdx.x;
    for (int idx = tid; idx < kEbJ * kEbC; idx += blockDim.x) (&acc[0][0])[idx] = 0.f;
    __syncthreads();
    for (int g0 = 0; g0 < G; g0 += 256) {
        // each thread scans the 32-bit block of one molecule; up to 4 non-zeros per pass, repeat if more
        const int g = g0 + tid;
        int consumed = 0;   // bits already handled for this molecule
        bool more = true;
        while (true) {
            int cnt = 0;
            int lj[4];
            float lv[4];
            if (g < G && more) {
                int j = consumed;
                for (; j < kEbJ && cnt < 4; ++j) {
                    const int jj = j0 + j;
                    const int v = (jj < fp_dim) ? fp[(size_t)g * fp_dim + jj] : 0;
                    if (v != 0) { lj[cnt] = j; lv[cnt] = (float)v; ++cnt; }
                }
                consumed = j;
                more = consumed < kEbJ;
            }
            s_cnt[tid] = cnt;
            for (int q = 0; q < cnt; ++q) { s_g[tid * 4 + q] = g; s_j[tid * 4 + q] = lj[q]; s_v[tid * 4 + q] = lv[q]; }
            const int any_more = __syncthreads_or((g < G && more) ? 1 : 0);
            // ordered replay: molecule order, then bit order
            for (int t = 0; t < 256; ++t) {
                const int c = s_cnt[t];
                for (int q = 0; q < c; ++q) {
                    const int gg = s_g[t * 4 + q], jj = s_j[t * 4 + q];
                    const float v = s_v[t * 4 + q];
                    const float* drow = d_out + (size_t)gg * out_dim + c0;
                    for (int cc = tid; cc < kEbC && c0 + cc < out_dim; cc += blockDim.x)
                        acc[jj][cc] = fmaf(v, drow[cc], acc[jj][cc]);
                }
            }
            __syncthreads();
            if (!any_more) break;
        }
    }
    for (int idx = tid; idx < kEbJ * kEbC; idx += blockDim.x) {
        const int j = idx / kEbC, cc = idx - j * kEbC;
        if (j0 + j < fp_dim && c0 + cc < out_dim) d_Wt[(size_t)(j0 + j) * out_dim + c0 + cc] += acc[j][cc];
    }
}

}  // namespace adkf

using namespace adkf;

extern "C" int adkf_readout_weighted_fwd(const float* scores, const float* values, const int32_t* graph_ptr,
                                         int32_t num_graphs, int32_t heads, int32_t head_dim, int32_t mode,
                                         float* weights, float* out, void* stream) {
    if (num_graphs <= 0) return 0;
    if (!scores || !values || !graph_ptr || !weights || !out || heads <= 0 || head_dim <= 0) return ADKF_ERR_BADARG;
    readout_weighted_fwd_kernel<<<num_graphs, 256, 0, (cudaStream_t)stream>>>(scores, values, graph_ptr, heads,
                                                                              head_dim, mode, weights, out);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_readout_weighted_bwd(const float* weights, const float* values, const int32_t* graph_ptr,
                                         int32_t num_graphs, int32_t heads, int32_t head_dim, int32_t mode,
                                         const float* d_out, float* d_scores, float* d_values, void* stream) {
    if (num_graphs <= 0) return 0;
    if (!weights || !values || !graph_ptr || !d_out || !d_scores || !d_values) return ADKF_ERR_BADARG;
    readout_weighted_bwd_kernel<<<num_graphs, 256, 0, (cudaStream_t)stream>>>(weights, values, graph_ptr, heads,
                                                                              head_dim, mode, d_out, d_scores, d_values);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_readout_max_fwd(const float* x, const int32_t* graph_ptr, int32_t num_graphs, int32_t dim,
                                    float* out, int32_t* argmax, void* stream) {
    if (num_graphs <= 0) return 0;
    if (!x || !graph_ptr || !out || !argmax || dim <= 0) return ADKF_ERR_BADARG;
    readout_max_fwd_kernel<<<num_graphs, 256, 0, (cudaStream_t)stream>>>(x, graph_ptr, dim, out, argmax);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_readout_max_bwd(const float* d_out, const int32_t* argmax, const int32_t* graph_ptr,
                                    int32_t num_graphs, int32_t dim, float* d_x, void* stream) {
    if (num_graphs <= 0) return 0;
    if (!d_out || !argmax || !graph_ptr || !d_x || dim <= 0) return ADKF_ERR_BADARG;
    readout_max_bwd_kernel<<<num_graphs, 256, 0, (cudaStream_t)stream>>>(d_out, argmax, graph_ptr, dim, d_x);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_ecfp_linear_fwd(const int32_t* fp, int32_t num_graphs, int32_t fp_dim, const float* Wt,
                                    int32_t out_dim, float* out, void* stream) {
    if (num_graphs <= 0) return 0;
    if (!fp || !Wt || !out || fp_dim <= 0 || out_dim <= 0 || (out_dim & 3)) return ADKF_ERR_BADARG;
    ecfp_linear_fwd_kernel<<<num_graphs, 256, 0, (cudaStream_t)stream>>>(fp, fp_dim, Wt, out_dim, out);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_ecfp_linear_bwd(const int32_t* fp, int32_t num_graphs, int32_t fp_dim, const float* d_out,
                                    int32_t out_dim, float* d_Wt, void* stream) {
    if (num_graphs <= 0) return 0;
    if (!fp || !d_out || !d_Wt || fp_dim <= 0 || out_dim <= 0) return ADKF_ERR_BADARG;
    ecfp_linear_bwd_kernel<<<dim3((fp_dim + kEbJ - 1) / kEbJ, (out_dim + kEbC - 1) / kEbC), 256, 0,
                             (cudaStream_t)stream>>>(fp, num_graphs, fp_dim, d_out, out_dim, d_Wt);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}
