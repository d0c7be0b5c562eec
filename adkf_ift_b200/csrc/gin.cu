// gin.cu -- MoleculeNet branch of ADKF-IFT (config 4): the GIN molecule encoder's graph kernels (sm_100a).
//
// Replaces, batched over all graphs of all tasks of a meta-batch:
//   GINConv.forward / message / propagate(aggr="add")     MoleculeNet/chem_lib/models/encoder.py:40-58
//        agg_i = sum_{j -> i} (x_j + E1[bond_type_ji] + E2[bond_dir_ji]) + (x_i + E1[4] + E2[0])      (self loop, :42-48)
//   atom / chirality embedding sum                        encoder.py:269
//   BatchNorm1d in train mode, ReLU, dropout              encoder.py:273-281 -- statistics PER encoder call of the reference,
//        i.e. per (task, support|query) segment of the merged batch, not over the whole merged batch
//   global_mean_pool                                      encoder.py:325,358
// The dense 300 -> 600 -> 300 MLP of every layer runs on the tcgen05 3xTF32 GEMM (gemm_tf32x3.cu); embedding-table gradients
// are one-hot^T x gradient products on the same GEMM (deterministic split-K) -- no float atomics anywhere.
#include "common.cuh"
#include "../../include/adkf_b200.h"

namespace adkf {

// ---- CSR of a PyG edge_index [2, E] (row 0 = source j, row 1 = target i), by target or by source --------------------------
__global__ void ei_count_kernel(const int64_t* __restrict__ key, int64_t E, int32_t* __restrict__ cnt) {
    for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < E; e += (int64_t)gridDim.x * blockDim.x)
        atomicAdd(&cnt[key[e]], 1);
}
__global__ void ei_scan_kernel(const int32_t* __restrict__ cnt, int V, int32_t* __restrict__ rowptr) {
    __shared__ int32_t part[1024];
    const int tid = threadIdx.x, per = (V + 1023) / 1024;
    const int beg = min(tid * per, V), end = min(beg + per, V);
    int s = 0;
    for (int i = beg; i < end; ++i) s += cnt[i];
    part[tid] = s;
    __syncthreads();
    if (tid == 0) {
        int run = 0;
        for (int i = 0; i < 1024; ++i) { const int c = part[i]; part[i] = run; run += c; }
        rowptr[V] = run;
    }
    __syncthreads();
    int run = part[tid];
    for (int i = beg; i < end; ++i) { rowptr[i] = run; run += cnt[i]; }
}
__global__ void ei_fill_kernel(const int64_t* __restrict__ key, int64_t E, const int32_t* __restrict__ rowptr,
                               int32_t* __restrict__ cursor, int32_t* __restrict__ col_eid) {
    for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < E; e += (int64_t)gridDim.x * blockDim.x) {
        const int v = (int)key[e];
        col_eid[rowptr[v] + atomicAdd(&cursor[v], 1)] = (int32_t)e;
    }
}
// sort every row by edge id (deterministic layout), then resolve the other endpoint
__global__ void ei_finalize_kernel(const int64_t* __restrict__ other, int V, const int32_t* __restrict__ rowptr,
                                   int32_t* __restrict__ col_eid, int32_t* __restrict__ col_other) {
    for (int v = blockIdx.x * blockDim.x + threadIdx.x; v < V; v += gridDim.x * blockDim.x) {
        const int beg = rowptr[v], end = rowptr[v + 1];
        for (int i = beg + 1; i < end; ++i) {
            const int32_t k = col_eid[i];
            int j = i - 1;
            while (j >= beg && col_eid[j] > k) { col_eid[j + 1] = col_eid[j]; --j; }
            col_eid[j + 1] = k;
        }
        for (int i = beg; i < end; ++i) col_other[i] = (int32_t)other[col_eid[i]];
    }
}

// ---- embedding sum: out[v] = T1[idx[v,0]] + T2[idx[v,1]] ---------------------------------------------------------------
__global__ void embed_sum2_kernel(const int64_t* __restrict__ idx, int64_t rows, int D, const float* __restrict__ T1,
                                  const float* __restrict__ T2, float* __restrict__ out) {
    const int64_t total = rows * D;
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < total; p += (int64_t)gridDim.x * blockDim.x) {
        const int64_t v = p / D;
        const int c = (int)(p - v * D);
        out[p] = T1[idx[2 * v] * D + c] + T2[idx[2 * v + 1] * D + c];
    }
}

// ---- GIN aggregation (forward, by-target CSR) and its transpose (backward, by-source CSR, no edge terms) -------------------
//   out[i] = x[i] + self_term + sum_{p in row i} ( x[col[p]] + E1[attr[eid[p],0]] + E2[attr[eid[p],1]] )
__global__ void gin_gather_kernel(const float* __restrict__ x, int V, int D, const int32_t* __restrict__ rowptr,
                                  const int32_t* __restrict__ col, const int32_t* __restrict__ col_eid,
                                  const int64_t* __restrict__ edge_attr, const float* __restrict__ E1,
                                  const float* __restrict__ E2, float* __restrict__ out) {
    const int64_t total = (int64_t)V * D;
    const bool edges = edge_attr != nullptr;
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < total; p += (int64_t)gridDim.x * blockDim.x) {
        const int i = (int)(p / D), c = (int)(p - (int64_t)i * D);
        float acc = 0.f;
        const int beg = rowptr[i], end = rowptr[i + 1];
        for (int q = beg; q < end; ++q) {
            float m = x[(int64_t)col[q] * D + c];
            if (edges) {
                const int64_t e = col_eid[q];
                m += E1[edge_attr[2 * e] * D + c] + E2[edge_attr[2 * e + 1] * D + c];
            }
            acc += m;
        }
        float self = x[p];
        if (edges) self += E1[4 * D + c] + E2[c];       // self-loop attr = (4, 0), appended after the real edges (encoder.py:44-47)
        out[p] = acc + self;
    }
}

// ---- per-segment BatchNorm1d (+ReLU, +dropout) ---------------------------------------------------------------------------
// mode 0: training statistics per segment (biased variance for the normalisation, like torch); mean / rstd [S, D] saved
// mode 1: evaluation with the running statistics
// act: 1 = ReLU after the affine.  keep < 1: inverted dropout with a counter-based hash of (seed, element index).
__device__ __forceinline__ uint32_t hash32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
    return x;
}
__device__ __forceinline__ float dropout_scale(uint64_t seed, int64_t p, float keep) {
    if (keep >= 1.f) return 1.f;
    const uint32_t h = hash32((uint32_t)p ^ hash32((uint32_t)(p >> 32) + (uint32_t)seed) ^ (uint32_t)(seed >> 32));
    return ((h >> 8) * (1.f / 16777216.f) < keep) ? 1.f / keep : 0.f;
}

constexpr int kBnCols = 32, kBnRows = 8;

__global__ void __launch_bounds__(kBnCols * kBnRows)
seg_bn_fwd_kernel(const float* __restrict__ x, int D, const int32_t* __restrict__ seg_ptr, const float* __restrict__ gamma,
                  const float* __restrict__ beta, const float* __restrict__ run_mean, const float* __restrict__ run_var,
                  float eps, int mode, int act, float keep, uint64_t seed, float* __restrict__ y,
                  float* __restrict__ save_mean, float* __restrict__ save_rstd, float* __restrict__ save_uvar) {
    const int s = blockIdx.x;
    const int c = blockIdx.y * kBnCols + (threadIdx.x & (kBnCols - 1));
    const int rl = threadIdx.x / kBnCols;
    const int r0 = seg_ptr[s], r1 = seg_ptr[s + 1], n = r1 - r0;
    __shared__ float red[kBnRows][kBnCols];
    const bool col_ok = c < D;
    float mean, rstd;
    if (mode == 0) {
        float acc = 0.f;
        if (col_ok) for (int r = r0 + rl; r < r1; r += kBnRows) acc += x[(int64_t)r * D + c];
        red[rl][threadIdx.x & (kBnCols - 1)] = acc;
        __syncthreads();
        float tot = 0.f;
#pragma unroll
        for (int k = 0; k < kBnRows; ++k) tot += red[k][threadIdx.x & (kBnCols - 1)];
        mean = tot / (float)max(n, 1);
        __syncthreads();
        acc = 0.f;
        if (col_ok) for (int r = r0 + rl; r < r1; r += kBnRows) { const float d = x[(int64_t)r * D + c] - mean; acc += d * d; }
        red[rl][threadIdx.x & (kBnCols - 1)] = acc;
        __syncthreads();
        tot = 0.f;
#pragma unroll
        for (int k = 0; k < kBnRows; ++k) tot += red[k][threadIdx.x & (kBnCols - 1)];
        const float var = tot / (float)max(n, 1);
        rstd = rsqrtf(var + eps);
        if (col_ok && rl == 0) {
            save_mean[(int64_t)s * D + c] = mean;
            save_rstd[(int64_t)s * D + c] = rstd;
            save_uvar[(int64_t)s * D + c] = tot / (float)max(n - 1, 1);   // unbiased, for the running estimate
        }
    } else {
        mean = col_ok ? run_mean[c] : 0.f;
        rstd = col_ok ? rsqrtf(run_var[c] + eps) : 0.f;
    }
    if (!col_ok) return;
    const float g = gamma[c] * rstd, b = beta[c] - mean * g;
    for (int r = r0 + rl; r < r1; r += kBnRows) {
        const int64_t p = (int64_t)r * D + c;
        float v = fmaf(x[p], g, b);
        if (act == 1) v = fmaxf(v, 0.f);
        y[p] = v * dropout_scale(seed, p, keep);
    }
}

// dy -> dx, per-segment dgamma / dbeta partials [S, D].  y is the forward OUTPUT (ReLU mask: y != 0 before dropout is
// recovered as "pre-dropout value > 0", i.e. xhat * gamma + beta > 0, recomputed here).
__global__ void __launch_bounds__(kBnCols * kBnRows)
seg_bn_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dy, int D, const int32_t* __restrict__ seg_ptr,
                  const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ save_mean,
                  const float* __restrict__ save_rstd, int act, float keep, uint64_t seed, float* __restrict__ dx,
                  float* __restrict__ dgamma_part, float* __restrict__ dbeta_part) {
    const int s = blockIdx.x;
    const int cl = threadIdx.x & (kBnCols - 1);
    const int c = blockIdx.y * kBnCols + cl;
    const int rl = threadIdx.x / kBnCols;
    const int r0 = seg_ptr[s], r1 = seg_ptr[s + 1], n = r1 - r0;
    __shared__ float red1[kBnRows][kBnCols], red2[kBnRows][kBnCols];
    const bool col_ok = c < D;
    const float mean = col_ok ? save_mean[(int64_t)s * D + c] : 0.f, rstd = col_ok ? save_rstd[(int64_t)s * D + c] : 0.f;
    const float g = col_ok ? gamma[c] : 0.f, b = col_ok ? beta[c] : 0.f;
    float s1 = 0.f, s2 = 0.f;
    if (col_ok)
        for (int r = r0 + rl; r < r1; r += kBnRows) {
            const int64_t p = (int64_t)r * D + c;
            const float xh = (x[p] - mean) * rstd;
            float d = dy[p] * dropout_scale(seed, p, keep);
            if (act == 1 && !(fmaf(xh, g, b) > 0.f)) d = 0.f;
            s1 += d;
            s2 += d * xh;
        }
    red1[rl][cl] = s1; red2[rl][cl] = s2;
    __syncthreads();
    float t1 = 0.f, t2 = 0.f;
#pragma unroll
    for (int k = 0; k < kBnRows; ++k) { t1 += red1[k][cl]; t2 += red2[k][cl]; }
    if (!col_ok) return;
    if (rl == 0) { dbeta_part[(int64_t)s * D + c] = t1; dgamma_part[(int64_t)s * D + c] = t2; }
    const float inv_n = 1.f / (float)max(n, 1);
    for (int r = r0 + rl; r < r1; r += kBnRows) {
        const int64_t p = (int64_t)r * D + c;
        const float xh = (x[p] - mean) * rstd;
        float d = dy[p] * dropout_scale(seed, p, keep);
        if (act == 1 && !(fmaf(xh, g, b) > 0.f)) d = 0.f;
        dx[p] = g * rstd * (d - inv_n * (t1 + xh * t2));
    }
}

// eval-mode backward is never needed (no gradients flow in evaluation).

// ---- mean pool over the nodes of each graph ------------------------------------------------------------------------------
__global__ void seg_mean_fwd_kernel(const float* __restrict__ x, int D, const int32_t* __restrict__ graph_ptr, int G,
                                    float* __restrict__ out) {
    const int64_t total = (int64_t)G * D;
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < total; p += (int64_t)gridDim.x * blockDim.x) {
        const int g = (int)(p / D), c = (int)(p - (int64_t)g * D);
        const int r0 = graph_ptr[g], r1 = graph_ptr[g + 1];
        float acc = 0.f;
        for (int r = r0; r < r1; ++r) acc += x[(int64_t)r * D + c];
        out[p] = acc / (float)max(r1 - r0, 1);
    }
}
__global__ void seg_mean_bwd_kernel(const float* __restrict__ dout, int D, const int32_t* __restrict__ graph_ptr, int G,
                                    float* __restrict__ dx) {
    const int g = blockIdx.x;
    const int r0 = graph_ptr[g], r1 = graph_ptr[g + 1];
    const float inv = 1.f / (float)max(r1 - r0, 1);
    for (int c = threadIdx.x; c < D; c += blockDim.x) {
        const float v = dout[(int64_t)g * D + c] * inv;
        for (int r = r0; r < r1; ++r) dx[(int64_t)r * D + c] = v;
    }
}

}  // namespace adkf

using namespace adkf;

static inline int grid_for(int64_t n, int block = 256, int cap = 148 * 16) {
    int64_t g = (n + block - 1) / block;
    if (g > cap) g = cap;
    return g < 1 ? 1 : (int)g;
}

extern "C" int adkf_csr_from_edge_index(const int64_t* edge_index, int64_t E, int32_t V, int32_t by_source, int32_t* rowptr,
                                        int32_t* col_other, int32_t* col_eid, int32_t* scratch /*[V]*/, void* stream) {
    if (V <= 0) return 0;
    if (!rowptr || !scratch || E < 0 || E > 0x7fffffff) return ADKF_ERR_BADARG;
    if (E > 0 && (!edge_index || !col_other || !col_eid)) return ADKF_ERR_BADARG;
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t* key = edge_index + (by_source ? 0 : E);
    const int64_t* other = edge_index + (by_source ? E : 0);
    cudaMemsetAsync(scratch, 0, sizeof(int32_t) * (size_t)V, s);
    if (E > 0) ei_count_kernel<<<grid_for(E), 256, 0, s>>>(key, E, scratch);
    ei_scan_kernel<<<1, 1024, 0, s>>>(scratch, V, rowptr);
    if (E > 0) {
        cudaMemsetAsync(scratch, 0, sizeof(int32_t) * (size_t)V, s);
        ei_fill_kernel<<<grid_for(E), 256, 0, s>>>(key, E, rowptr, scratch, col_eid);
        ei_finalize_kernel<<<grid_for(V), 256, 0, s>>>(other, V, rowptr, col_eid, col_other);
    }
    adkf_count_launches(4);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_embed_sum2(const int64_t* idx, int64_t rows, int32_t D, const float* T1, const float* T2, float* out,
                               void* stream) {
    if (rows <= 0) return 0;
    if (!idx || !T1 || !T2 || !out || D <= 0) return ADKF_ERR_BADARG;
    embed_sum2_kernel<<<grid_for(rows * D), 256, 0, (cudaStream_t)stream>>>(idx, rows, D, T1, T2, out);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_gin_gather(const float* x, int32_t V, int32_t D, const int32_t* rowptr, const int32_t* col,
                               const int32_t* col_eid, const int64_t* edge_attr, const float* E1, const float* E2, float* out,
                               void* stream) {
    if (V <= 0) return 0;
    if (!x || !rowptr || !out || D <= 0) return ADKF_ERR_BADARG;
    if (edge_attr && (!E1 || !E2 || !col_eid)) return ADKF_ERR_BADARG;
    gin_gather_kernel<<<grid_for((int64_t)V * D), 256, 0, (cudaStream_t)stream>>>(x, V, D, rowptr, col, col_eid, edge_attr, E1,
                                                                                 E2, out);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_seg_bn_fwd(const float* x, int32_t D, const int32_t* seg_ptr, int32_t S, const float* gamma,
                               const float* beta, const float* run_mean, const float* run_var, float eps, int32_t mode,
                               int32_t act, float keep, uint64_t seed, float* y, float* save_mean, float* save_rstd,
                               float* save_uvar, void* stream) {
    if (S <= 0) return 0;
    if (!x || !seg_ptr || !gamma || !beta || !y || D <= 0) return ADKF_ERR_BADARG;
    if (mode == 0 && (!save_mean || !save_rstd || !save_uvar)) return ADKF_ERR_BADARG;
    if (mode != 0 && (!run_mean || !run_var)) return ADKF_ERR_BADARG;
    seg_bn_fwd_kernel<<<dim3(S, (D + kBnCols - 1) / kBnCols), kBnCols * kBnRows, 0, (cudaStream_t)stream>>>(
        x, D, seg_ptr, gamma, beta, run_mean, run_var, eps, mode, act, keep, seed, y, save_mean, save_rstd, save_uvar);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_seg_bn_bwd(const float* x, const float* dy, int32_t D, const int32_t* seg_ptr, int32_t S,
                               const float* gamma, const float* beta, const float* save_mean, const float* save_rstd,
                               int32_t act, float keep, uint64_t seed, float* dx, float* dgamma_part, float* dbeta_part,
                               void* stream) {
    if (S <= 0) return 0;
    if (!x || !dy || !seg_ptr || !gamma || !beta || !save_mean || !save_rstd || !dx || !dgamma_part || !dbeta_part || D <= 0)
        return ADKF_ERR_BADARG;
    seg_bn_bwd_kernel<<<dim3(S, (D + kBnCols - 1) / kBnCols), kBnCols * kBnRows, 0, (cudaStream_t)stream>>>(
        x, dy, D, seg_ptr, gamma, beta, save_mean, save_rstd, act, keep, seed, dx, dgamma_part, dbeta_part);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_seg_mean_fwd(const float* x, int32_t D, const int32_t* graph_ptr, int32_t G, float* out, void* stream) {
    if (G <= 0) return 0;
    if (!x || !graph_ptr || !out || D <= 0) return ADKF_ERR_BADARG;
    seg_mean_fwd_kernel<<<grid_for((int64_t)G * D), 256, 0, (cudaStream_t)stream>>>(x, D, graph_ptr, G, out);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_seg_mean_bwd(const float* dout, int32_t D, const int32_t* graph_ptr, int32_t G, float* dx, void* stream) {
    if (G <= 0) return 0;
    if (!dout || !graph_ptr || !dx || D <= 0) return ADKF_ERR_BADARG;
    seg_mean_bwd_kernel<<<G, 128, 0, (cudaStream_t)stream>>>(dout, D, graph_ptr, G, dx);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}
