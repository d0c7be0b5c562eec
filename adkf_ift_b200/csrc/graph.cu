// graph.cu -- FS-Mol graph feature extractor kernels (sm_100a): CSR build (bit-exact integer path) and the
// fused PNA relational message passing of one GNN block (all 4 towers), forward and backward.
//
// Replaces  GNN.forward's bidirectional expansion (fs_mol/modules/gnn.py:540-544), RelationalMP.forward
// (gnn.py:129-150) and RelationalMultiAggrMP._aggregate_messages (gnn.py:197-265) -- in the reference a few
// thousand tiny launches (gather, cat, 12 Linear, relu, 5 torch_scatter atomics kernels per tower-layer).
// Here: one launch per block forward; messages never touch HBM; aggregation is CSR-ordered (no float
// atomics -> run-to-run bit-stable).
//
// Layouts: h [V,128] fp32 (tower k = columns 32k..32k+31); W [4][3][192][64], b [4][3][192] fp32 (the
// reference's mp_layers.k.message_fns.t._layers.0.{weight,bias} stacked); CSR over target nodes with the
// edges of a row sorted by the reference's concatenated message index (type-major) so that same-type edges
// are adjacent; out [V, 4*768] = per tower [sum|mean|std|max] x [1, amp, att].
#include "common.cuh"
#include "../../include/adkf_b200.h"

namespace adkf {

constexpr int kTin = 32;        // tower input width
constexpr int kMsg = 192;       // 3 * per_head_dim
constexpr int kPmd = 64;        // per_head_dim (partial message dim)
constexpr int kTypes = 3;
constexpr int kTowers = 4;
constexpr int kHid = 128;
constexpr int kWld = 65;        // padded row stride of W[o][.] in shared memory (bank-conflict free both ways)
constexpr int kOutTower = 768;
constexpr int kMpWarps = 8;
constexpr int kMbufDeg = 32;    // mean-part messages buffered per row for the std pass
constexpr float kSmall = 1e-7f;
constexpr float kDelta = 1.1515f;

// ---------------------------------------------------------------------------------------------
// CSR build
// ---------------------------------------------------------------------------------------------
struct AdjLists { const int64_t* adj[kTypes]; int64_t e[kTypes]; };

__device__ __forceinline__ void decode_edge(const AdjLists& A, int64_t eid, int& src, int& tgt, int& type) {
    int64_t off = 0;
    int t = 0;
    for (; t < kTypes - 1; ++t) {
        if (eid < off + 2 * A.e[t]) break;
        off += 2 * A.e[t];
    }
    const int64_t i = eid - off;
    type = t;
    if (i < A.e[t]) { src = (int)A.adj[t][2 * i]; tgt = (int)A.adj[t][2 * i + 1]; }
    else { const int64_t j = i - A.e[t]; src = (int)A.adj[t][2 * j + 1]; tgt = (int)A.adj[t][2 * j]; }
}

__global__ void csr_count_kernel(AdjLists A, int64_t etot, int32_t* degree) {
    for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < etot; e += (int64_t)gridDim.x * blockDim.x) {
        int s, t, ty;
        decode_edge(A, e, s, t, ty);
        atomicAdd(&degree[t], 1);   // integer atomics: result independent of order
    }
}

// single-block exclusive scan (V up to a few million: V/1024 sequential chunks)
__global__ void csr_scan_kernel(const int32_t* __restrict__ degree, int V, int32_t* __restrict__ rowptr) {
    __shared__ int32_t warp_tot[32];
    __shared__ int32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int base = 0; base < V; base += blockDim.x) {
        const int i = base + threadIdx.x;
        const int32_t v = (i < V) ? degree[i] : 0;
        int32_t x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane == 31) warp_tot[wid] = x;
        __syncthreads();
        if (wid == 0) {
            int32_t w = (lane < (blockDim.x >> 5)) ? warp_tot[lane] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int32_t y = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += y; }
            warp_tot[lane] = w;
        }
        __syncthreads();
        const int32_t prev = (wid > 0 ? warp_tot[wid - 1] : 0) + carry;
        if (i < V) rowptr[i] = prev + x - v;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry = prev + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) rowptr[V] = carry;
}

__global__ void csr_fill_kernel(AdjLists A, int64_t etot, const int32_t* __restrict__ rowptr,
                                int32_t* __restrict__ cursor, int32_t* __restrict__ col_eid) {
    for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < etot; e += (int64_t)gridDim.x * blockDim.x) {
        int s, t, ty;
        decode_edge(A, e, s, t, ty);
        const int pos = rowptr[t] + atomicAdd(&cursor[t], 1);
        col_eid[pos] = (int32_t)e;
    }
}

// canonical order: sort the (short) rows by message index, then materialise src/type/tgt and eid -> position
__global__ void csr_finalize_kernel(AdjLists A, int V, const int32_t* __restrict__ rowptr,
                                    int32_t* __restrict__ col_eid, int32_t* __restrict__ col_src,
                                    uint8_t* __restrict__ col_type, int32_t* __restrict__ col_tgt,
                                    int32_t* __restrict__ pos_of_eid) {
    for (int v = blockIdx.x * blockDim.x + threadIdx.x; v < V; v += gridDim.x * blockDim.x) {
        const int beg = rowptr[v], end = rowptr[v + 1];
        for (int i = beg + 1; i < end; ++i) {   // insertion sort (degree is tiny)
            const int32_t key = col_eid[i];
            int j = i - 1;
            while (j >= beg && col_eid[j] > key) { col_eid[j + 1] = col_eid[j]; --j; }
            col_eid[j + 1] = key;
        }
        for (int i = beg; i < end; ++i) {
            int s, t, ty;
            decode_edge(A, col_eid[i], s, t, ty);
            col_src[i] = s;
            col_type[i] = (uint8_t)ty;
            col_tgt[i] = t;
            pos_of_eid[col_eid[i]] = i;
        }
    }
}

__global__ void csr_reverse_kernel(AdjLists A, int64_t etot, const int32_t* __restrict__ col_eid,
                                   const int32_t* __restrict__ pos_of_eid, int32_t* __restrict__ rev_pos) {
    for (int64_t p = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; p < etot; p += (int64_t)gridDim.x * blockDim.x) {
        const int64_t eid = col_eid[p];
        int64_t off = 0;
        int t = 0;
        for (; t < kTypes - 1; ++t) {
            if (eid < off + 2 * A.e[t]) break;
            off += 2 * A.e[t];
        }
        const int64_t i = eid - off;
        const int64_t r = (i < A.e[t]) ? i + A.e[t] : i - A.e[t];
        rev_pos[p] = pos_of_eid[off + r];
    }
}

// ---------------------------------------------------------------------------------------------
// PNA message passing
// ---------------------------------------------------------------------------------------------
// W of one tower -> shared memory, layout [type][o][kWld]
__device__ __forceinline__ void load_tower_weights(const float* __restrict__ W, const float* __restrict__ b, int k,
                                                   float* Ws, float* bs) {
    const float* Wk = W + (size_t)k * kTypes * kMsg * 64;
    for (int idx = threadIdx.x; idx < kTypes * kMsg * 64; idx += blockDim.x) {
        const int to = idx >> 6, i = idx & 63;
        Ws[to * kWld + i] = Wk[idx];
    }
    for (int idx = threadIdx.x; idx < kTypes * kMsg; idx += blockDim.x) bs[idx] = b[(size_t)k * kTypes * kMsg + idx];
}

// q[c] = b[o] + sum_i W[o][32+i] x_t[i]   for o = lane + 32 c  (the target-node half of the message MLP)
__device__ __forceinline__ void target_half(const float* Wt, const float* bt, float xt, int lane, float (&q)[6]) {
#pragma unroll
    for (int c = 0; c < 6; ++c) q[c] = bt[lane + 32 * c];
#pragma unroll 8
    for (int i = 0; i < kTin; ++i) {
        const float xv = __shfl_sync(0xffffffffu, xt, i);
#pragma unroll
        for (int c = 0; c < 6; ++c) q[c] = fmaf(Wt[(lane + 32 * c) * kWld + kTin + i], xv, q[c]);
    }
}

// acc[g][c] = q[c] + sum_i W[o][i] x_s^g[i]   for CNT same-type edges sharing the weight loads
template <int CNT>
__device__ __forceinline__ void source_half(const float* Wt, const float (&xs)[4], const float (&q)[6], int lane,
                                            float (&acc)[4][6]) {
#pragma unroll
    for (int g = 0; g < CNT; ++g)
#pragma unroll
        for (int c = 0; c < 6; ++c) acc[g][c] = q[c];
#pragma unroll 8
    for (int i = 0; i < kTin; ++i) {
        float w[6];
#pragma unroll
        for (int c = 0; c < 6; ++c) w[c] = Wt[(lane + 32 * c) * kWld + i];
#pragma unroll
        for (int g = 0; g < CNT; ++g) {
            const float xv = __shfl_sync(0xffffffffu, xs[g], i);
#pragma unroll
            for (int c = 0; c < 6; ++c) acc[g][c] = fmaf(w[c], xv, acc[g][c]);
        }
    }
}

__device__ __forceinline__ void source_half_dyn(int cnt, const float* Wt, const float (&xs)[4], const float (&q)[6],
                                                int lane, float (&acc)[4][6]) {
    switch (cnt) {
        case 1: source_half<1>(Wt, xs, q, lane, acc); break;
        case 2: source_half<2>(Wt, xs, q, lane, acc); break;
        case 3: source_half<3>(Wt, xs, q, lane, acc); break;
        default: source_half<4>(Wt, xs, q, lane, acc); break;
    }
}

constexpr size_t kMpSmemFwd = (size_t)(kTypes * kMsg * kWld + kTypes * kMsg + kMpWarps * kMbufDeg * kPmd) * sizeof(float);
constexpr size_t kMpSmemBwd = (size_t)(kTypes * kMsg * kWld + kTypes * kMsg) * sizeof(float);

__global__ void __launch_bounds__(kMpWarps * 32, 1)
pna_fwd_kernel(const float* __restrict__ h, const float* __restrict__ W, const float* __restrict__ b,
               const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col_src,
               const uint8_t* __restrict__ col_type, int V, float* __restrict__ out,
               uint8_t* __restrict__ aux_cnt, uint8_t* __restrict__ aux_arg) {
    extern __shared__ __align__(16) float smem[];
    float* Ws = smem;
    float* bs = Ws + kTypes * kMsg * kWld;
    float* mbuf_all = bs + kTypes * kMsg;
    const int k = blockIdx.y;
    load_tower_weights(W, b, k, Ws, bs);
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    float* mbuf = mbuf_all + (size_t)wid * kMbufDeg * kPmd;

    for (int node = blockIdx.x * kMpWarps + wid; node < V; node += gridDim.x * kMpWarps) {
        const int beg = rowptr[node], end = rowptr[node + 1], deg = end - beg;
        const float xt = h[(size_t)node * kHid + k * kTin + lane];
        float q[6];
        int cur_type = -1;
        float sum[2] = {0.f, 0.f}, macc[2] = {0.f, 0.f}, mx[2] = {0.f, 0.f};
        int arg[2] = {0, 0};
        int e = beg;
        while (e < end) {
            const int tau = col_type[e];
            int cnt = 1;
            while (cnt < 4 && e + cnt < end && col_type[e + cnt] == tau) ++cnt;
            const float* Wt = Ws + tau * kMsg * kWld;
            if (tau != cur_type) { target_half(Wt, bs + tau * kMsg, xt, lane, q); cur_type = tau; }
            float xs[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int g = 0; g < 4; ++g)
                if (g < cnt) xs[g] = h[(size_t)col_src[e + g] * kHid + k * kTin + lane];
            float acc[4][6];
            source_half_dyn(cnt, Wt, xs, q, lane, acc);
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                if (g < cnt) {
                    const int idx = e + g - beg;
                    float m[6];
#pragma unroll
                    for (int c = 0; c < 6; ++c) m[c] = fmaxf(acc[g][c], 0.f);
                    sum[0] += m[0]; sum[1] += m[1];
                    macc[0] += m[2]; macc[1] += m[3];
                    if (idx < kMbufDeg) { mbuf[idx * kPmd + lane] = m[2]; mbuf[idx * kPmd + 32 + lane] = m[3]; }
                    if (idx == 0 || m[4] > mx[0]) { mx[0] = m[4]; arg[0] = idx; }
                    if (idx == 0 || m[5] > mx[1]) { mx[1] = m[5]; arg[1] = idx; }
                }
            }
            e += cnt;
        }
        const float fdeg = (float)max(deg, 1);
        const float mean[2] = {macc[0] / fdeg, macc[1] / fdeg};   // true division, like scatter_mean
        float ssum[2] = {0.f, 0.f};
        int cpos[2] = {0, 0};
        __syncwarp();
        for (int idx = 0; idx < deg; ++idx) {
            float mm[2];
            if (idx < kMbufDeg) {
                mm[0] = mbuf[idx * kPmd + lane];
                mm[1] = mbuf[idx * kPmd + 32 + lane];
            } else {
                // rare high-degree row: recompute the mean-part message of this edge
                const int ee = beg + idx;
                const int tau = col_type[ee];
                const float* Wt = Ws + tau * kMsg * kWld;
                float q2[6];
                target_half(Wt, bs + tau * kMsg, xt, lane, q2);
                float xs[4] = {h[(size_t)col_src[ee] * kHid + k * kTin + lane], 0.f, 0.f, 0.f};
                float acc[4][6];
                source_half<1>(Wt, xs, q2, lane, acc);
                mm[0] = fmaxf(acc[0][2], 0.f);
                mm[1] = fmaxf(acc[0][3], 0.f);
            }
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                // no FMA contraction: m^2 - mean^2 must be exactly 0 when m == mean (degree-1 rows), as in torch
                const float v = __fsub_rn(__fmul_rn(mm[j], mm[j]), __fmul_rn(mean[j], mean[j]));
                if (v > 0.f) { ssum[j] += v; ++cpos[j]; }
                ssum[j] += kSmall;
            }
        }
        __syncwarp();
        const float logd = logf((float)deg + 1.f);
        const float amp = logd / kDelta, att = kDelta / (logd + kSmall);
        float* o = out + (size_t)node * (kTowers * kOutTower) + (size_t)k * kOutTower;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const float vals[4] = {sum[j], mean[j], sqrtf(ssum[j]), (deg > 0) ? mx[j] : 0.f};
#pragma unroll
            for (int a = 0; a < 4; ++a) {
                const int col = a * kPmd + j * 32 + lane;
                o[col] = vals[a];
                o[256 + col] = amp * vals[a];
                o[512 + col] = att * vals[a];
            }
            if (aux_cnt) aux_cnt[((size_t)node * kTowers + k) * kPmd + j * 32 + lane] = (uint8_t)min(cpos[j], 255);
            if (aux_arg) aux_arg[((size_t)node * kTowers + k) * kPmd + j * 32 + lane] = (uint8_t)min(arg[j], 255);
        }
    }
}

// backward stage 1: per (target node, tower) recompute the pre-activations and emit
//   d_pre[e][k][192]  (gradient at the message pre-activation, after the relu mask)
//   d_xs[e][k][32]    (its pull-back to the SOURCE node slice, gathered by source in stage 2)
//   d_xt[v][k*32..]   (pull-back to the target node slice, summed over the row here)
__global__ void __launch_bounds__(kMpWarps * 32, 1)
pna_bwd_edges_kernel(const float* __restrict__ h, const float* __restrict__ W, const float* __restrict__ b,
                     const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col_src,
                     const uint8_t* __restrict__ col_type, int V, const float* __restrict__ fwd_out,
                     const uint8_t* __restrict__ aux_cnt, const uint8_t* __restrict__ aux_arg,
                     const float* __restrict__ d_out, float* __restrict__ d_pre, float* __restrict__ d_xs,
                     float* __restrict__ d_xt) {
    extern __shared__ __align__(16) float smem[];
    float* Ws = smem;
    float* bs = Ws + kTypes * kMsg * kWld;
    const int k = blockIdx.y;
    load_tower_weights(W, b, k, Ws, bs);
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int node = blockIdx.x * kMpWarps + wid; node < V; node += gridDim.x * kMpWarps) {
        const int beg = rowptr[node], end = rowptr[node + 1], deg = end - beg;
        float dxt_acc = 0.f;
        if (deg > 0) {
            const float xt = h[(size_t)node * kHid + k * kTin + lane];
            const float logd = logf((float)deg + 1.f);
            const float amp = logd / kDelta, att = kDelta / (logd + kSmall);
            const float* go = d_out + (size_t)node * (kTowers * kOutTower) + (size_t)k * kOutTower;
            const float* fo = fwd_out + (size_t)node * (kTowers * kOutTower) + (size_t)k * kOutTower;
            float g_sum[2], g_mean[2], g_std[2], g_max[2], mean[2], stdv[2], cpos[2];
            int arg[2];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int col = j * 32 + lane;
                g_sum[j] = go[col] + amp * go[256 + col] + att * go[512 + col];
                g_mean[j] = go[kPmd + col] + amp * go[256 + kPmd + col] + att * go[512 + kPmd + col];
                g_std[j] = go[2 * kPmd + col] + amp * go[256 + 2 * kPmd + col] + att * go[512 + 2 * kPmd + col];
                g_max[j] = go[3 * kPmd + col] + amp * go[256 + 3 * kPmd + col] + att * go[512 + 3 * kPmd + col];
                mean[j] = fo[kPmd + col];
                stdv[j] = fo[2 * kPmd + col];
                cpos[j] = (float)aux_cnt[((size_t)node * kTowers + k) * kPmd + col];
                arg[j] = aux_arg[((size_t)node * kTowers + k) * kPmd + col];
            }
            const float invdeg = 1.f / (float)deg;
            float q[6];
            int cur_type = -1;
            int e = beg;
            while (e < end) {
                const int tau = col_type[e];
                int cnt = 1;
                while (cnt < 4 && e + cnt < end && col_type[e + cnt] == tau) ++cnt;
                const float* Wt = Ws + tau * kMsg * kWld;
                if (tau != cur_type) { target_half(Wt, bs + tau * kMsg, xt, lane, q); cur_type = tau; }
                float xs[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int g = 0; g < 4; ++g)
                    if (g < cnt) xs[g] = h[(size_t)col_src[e + g] * kHid + k * kTin + lane];
                float acc[4][6];
                source_half_dyn(cnt, Wt, xs, q, lane, acc);
#pragma unroll
                for (int g = 0; g < 4; ++g) {
                    if (g < cnt) {
                        const int idx = e + g - beg;
                        float dp[6];
#pragma unroll
                        for (int j = 0; j < 2; ++j) {
                            const float m_mean = fmaxf(acc[g][2 + j], 0.f);
                            const float cc = 0.5f * g_std[j] / stdv[j];
                            const float ind = (__fsub_rn(__fmul_rn(m_mean, m_mean), __fmul_rn(mean[j], mean[j])) > 0.f) ? 1.f : 0.f;
                            const float dm_mean = g_mean[j] * invdeg + cc * (2.f * m_mean * ind - 2.f * mean[j] * cpos[j] * invdeg);
                            dp[j] = (acc[g][j] > 0.f) ? g_sum[j] : 0.f;
                            dp[2 + j] = (acc[g][2 + j] > 0.f) ? dm_mean : 0.f;
                            dp[4 + j] = (acc[g][4 + j] > 0.f && (idx == arg[j] || (idx >= 255 && arg[j] == 255))) ? g_max[j] : 0.f;
                        }
                        float* dpo = d_pre + ((size_t)(e + g) * kTowers + k) * kMsg;
#pragma unroll
                        for (int c = 0; c < 6; ++c) dpo[lane + 32 * c] = dp[c];
                        // pull back through W: lanes now index the INPUT feature i
                        float dxs = 0.f, dxt = 0.f;
#pragma unroll
                        for (int c = 0; c < 6; ++c) {
#pragma unroll 8
                            for (int l = 0; l < 32; ++l) {
                                const float v = __shfl_sync(0xffffffffu, dp[c], l);
                                const float* wr = Wt + (l + 32 * c) * kWld;
                                dxs = fmaf(wr[lane], v, dxs);
                                dxt = fmaf(wr[kTin + lane], v, dxt);
                            }
                        }
                        d_xs[((size_t)(e + g) * kTowers + k) * kTin + lane] = dxs;
                        dxt_acc += dxt;
                    }
                }
                e += cnt;
            }
        }
        d_xt[(size_t)node * kHid + k * kTin + lane] = dxt_acc;
    }
}

// backward stage 2: d_h[v] = d_xt[v] + sum over the outgoing edges of v (= reverses of its incoming edges)
__global__ void pna_bwd_nodes_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ rev_pos, int V,
                                     const float* __restrict__ d_xs, const float* __restrict__ d_xt,
                                     float* __restrict__ d_h) {
    const int c = threadIdx.x & (kHid - 1);
    const int sub = threadIdx.x / kHid;
    const int per = blockDim.x / kHid;
    for (int v = blockIdx.x * per + sub; v < V; v += gridDim.x * per) {
        float acc = d_xt[(size_t)v * kHid + c];
        const int beg = rowptr[v], end = rowptr[v + 1];
        for (int p = beg; p < end; ++p) acc += d_xs[(size_t)rev_pos[p] * kHid + c];
        d_h[(size_t)v * kHid + c] = acc;
    }
}

// backward stage 3: per (edge chunk, tower, type) partial  dW = sum_e d_pre_e (x) [x_s ; x_t],  db = sum_e d_pre_e
constexpr int kDwBatch = 8;
__global__ void __launch_bounds__(256)
pna_bwd_weights_kernel(const float* __restrict__ h, const int32_t* __restrict__ col_src,
                       const int32_t* __restrict__ col_tgt, const uint8_t* __restrict__ col_type, int E,
                       const float* __restrict__ d_pre, float* __restrict__ partials) {
    const int chunk = blockIdx.x, nchunk = gridDim.x;
    const int kt = blockIdx.y, k = kt / kTypes, tau = kt % kTypes;
    const int e_beg = (int)((int64_t)E * chunk / nchunk), e_end = (int)((int64_t)E * (chunk + 1) / nchunk);
    __shared__ float dp[kDwBatch][kMsg];
    __shared__ float xin[kDwBatch][64];
    __shared__ int list[256];
    __shared__ int nlist;
    const int tid = threadIdx.x, og = tid >> 3, ig = tid & 7;
    float acc[6][8], bacc[6];
#pragma unroll
    for (int a = 0; a < 6; ++a) {
        bacc[a] = 0.f;
#pragma unroll
        for (int q = 0; q < 8; ++q) acc[a][q] = 0.f;
    }
    for (int base = e_beg; base < e_end; base += 256) {
        // compact the edges of this type inside the 256-edge window (order preserved -> deterministic)
        if (tid == 0) nlist = 0;
        __syncthreads();
        const int e = base + tid;
        const bool match = e < e_end && col_type[e] == tau;
        const unsigned bal = __ballot_sync(0xffffffffu, match);
        __shared__ int warp_base[8];
        if ((tid & 31) == 0) warp_base[tid >> 5] = __popc(bal);
        __syncthreads();
        if (tid == 0) {
            int run = 0;
            for (int w = 0; w < 8; ++w) { const int c = warp_base[w]; warp_base[w] = run; run += c; }
            nlist = run;
        }
        __syncthreads();
        if (match) list[warp_base[tid >> 5] + __popc(bal & ((1u << (tid & 31)) - 1u))] = e;
        __syncthreads();
        const int n = nlist;
        for (int b0 = 0; b0 < n; b0 += kDwBatch) {
            const int nb = min(kDwBatch, n - b0);
            for (int idx = tid; idx < nb * kMsg; idx += 256) {
                const int g = idx / kMsg, o = idx - g * kMsg;
                dp[g][o] = d_pre[((size_t)list[b0 + g] * kTowers + k) * kMsg + o];
            }
            for (int idx = tid; idx < nb * 64; idx += 256) {
                const int g = idx >> 6, i = idx & 63;
                const int ee = list[b0 + g];
                const int nodeid = (i < kTin) ? col_src[ee] : col_tgt[ee];
                xin[g][i] = h[(size_t)nodeid * kHid + k * kTin + (i & 31)];
            }
            __syncthreads();
            for (int g = 0; g < nb; ++g) {
                float xv[8], dv[6];
#pragma unroll
                for (int q = 0; q < 8; ++q) xv[q] = xin[g][ig * 8 + q];
#pragma unroll
                for (int a = 0; a < 6; ++a) dv[a] = dp[g][og + 32 * a];
#pragma unroll
                for (int a = 0; a < 6; ++a) {
                    bacc[a] += dv[a];
#pragma unroll
                    for (int q = 0; q < 8; ++q) acc[a][q] = fmaf(dv[a], xv[q], acc[a][q]);
                }
            }
            __syncthreads();
        }
    }
    // partial layout: [chunk][kt][192*64 weights | 192 biases]
    float* P = partials + ((size_t)chunk * (kTowers * kTypes) + kt) * (kMsg * 64 + kMsg);
#pragma unroll
    for (int a = 0; a < 6; ++a) {
        const int o = og + 32 * a;
#pragma unroll
        for (int q = 0; q < 8; ++q) P[o * 64 + ig * 8 + q] = acc[a][q];
        if (ig == 0) P[kMsg * 64 + o] = bacc[a];
    }
}

__global__ void pna_bwd_reduce_kernel(const float* __restrict__ partials, int nchunk, float* __restrict__ d_W,
                                      float* __restrict__ d_b) {
    const int per = kMsg * 64 + kMsg;
    const int total = kTowers * kTypes * per;
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += gridDim.x * blockDim.x) {
        float acc = 0.f;
        for (int c = 0; c < nchunk; ++c) acc += partials[(size_t)c * total + idx];
        const int kt = idx / per, r = idx - kt * per;
        if (r < kMsg * 64) d_W[(size_t)kt * kMsg * 64 + r] = acc;
        else d_b[kt * kMsg + (r - kMsg * 64)] = acc;
    }
}

}  // namespace adkf

using namespace adkf;

extern "C" int adkf_graph_csr_build(const int64_t* adj0, int64_t e0, const int64_t* adj1, int64_t e1,
                                    const int64_t* adj2, int64_t e2, int32_t num_nodes, int32_t* rowptr,
                                    int32_t* col_src, uint8_t* col_type, int32_t* col_tgt, int32_t* col_eid,
                                    int32_t* rev_pos, int32_t* degree, int32_t* scratch /*[2E]*/, void* stream) {
    if (num_nodes < 0 || e0 < 0 || e1 < 0 || e2 < 0) return ADKF_ERR_BADARG;
    if (!rowptr || !degree) return ADKF_ERR_BADARG;
    const int64_t etot = 2 * (e0 + e1 + e2);
    if (etot > 0x7fffffff) return ADKF_ERR_TOOLARGE;
    if (etot > 0 && (!col_src || !col_type || !col_tgt || !col_eid || !rev_pos || !scratch)) return ADKF_ERR_BADARG;
    cudaStream_t s = (cudaStream_t)stream;
    AdjLists A;
    A.adj[0] = adj0; A.adj[1] = adj1; A.adj[2] = adj2;
    A.e[0] = e0; A.e[1] = e1; A.e[2] = e2;
    cudaMemsetAsync(degree, 0, sizeof(int32_t) * (size_t)num_nodes, s);
    const int eb = (int)((etot + 255) / 256) > 0 ? (int)((etot + 255) / 256) : 1;
    if (etot > 0) csr_count_kernel<<<min(eb, 4096), 256, 0, s>>>(A, etot, degree);
    csr_scan_kernel<<<1, 1024, 0, s>>>(degree, num_nodes, rowptr);
    if (etot > 0) {
        // `scratch` doubles as the fill cursor (first V entries are not enough when V > 2E, so use degree itself
        // as cursor after zeroing: it counts back up to the degree)
        cudaMemsetAsync(degree, 0, sizeof(int32_t) * (size_t)num_nodes, s);
        csr_fill_kernel<<<min(eb, 4096), 256, 0, s>>>(A, etot, rowptr, degree, col_eid);
        const int vb = (num_nodes + 255) / 256;
        csr_finalize_kernel<<<max(min(vb, 4096), 1), 256, 0, s>>>(A, num_nodes, rowptr, col_eid, col_src, col_type,
                                                                col_tgt, scratch);
        csr_reverse_kernel<<<min(eb, 4096), 256, 0, s>>>(A, etot, col_eid, scratch, rev_pos);
    }
    adkf_count_launches(5);
    ADKF_LAUNCH_CHECK();
    return 0;
}

static int mp_grid_x(int V) {
    int gx = (V + kMpWarps - 1) / kMpWarps;
    if (gx > 37) gx = 37;  // 37 CTAs x 4 towers = 148 SMs, one resident CTA per SM (150 KB of weights in smem)
    return gx < 1 ? 1 : gx;
}

extern "C" int adkf_mp_pna_fwd(const float* h, const float* W, const float* b, const int32_t* rowptr,
                               const int32_t* col_src, const uint8_t* col_type, int32_t num_nodes, float* out,
                               uint8_t* aux_cnt, uint8_t* aux_arg, void* stream) {
    if (num_nodes <= 0) return 0;
    if (!h || !W || !b || !rowptr || !out) return ADKF_ERR_BADARG;
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e = cudaFuncSetAttribute((const void*)pna_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)kMpSmemFwd);
    if (e != cudaSuccess) return (int)e;
    pna_fwd_kernel<<<dim3(mp_grid_x(num_nodes), kTowers), kMpWarps * 32, kMpSmemFwd, s>>>(
        h, W, b, rowptr, col_src, col_type, num_nodes, out, aux_cnt, aux_arg);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_mp_pna_bwd(const float* h, const float* W, const float* b, const int32_t* rowptr,
                               const int32_t* col_src, const int32_t* col_tgt, const uint8_t* col_type,
                               const int32_t* rev_pos, int32_t num_nodes, int32_t num_edges, const float* fwd_out,
                               const uint8_t* aux_cnt, const uint8_t* aux_arg, const float* d_out, float* d_pre,
                               float* d_xs, float* d_xt, float* d_h, float* d_W, float* d_b, float* partials,
                               int32_t num_partials, void* stream) {
    if (num_nodes <= 0) return 0;
    if (!h || !W || !b || !rowptr || !fwd_out || !aux_cnt || !aux_arg || !d_out || !d_xt || !d_h || !d_W || !d_b ||
        !partials || num_partials <= 0)
        return ADKF_ERR_BADARG;
    if (num_edges > 0 && (!col_src || !col_tgt || !col_type || !rev_pos || !d_pre || !d_xs)) return ADKF_ERR_BADARG;
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e = cudaFuncSetAttribute((const void*)pna_bwd_edges_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kMpSmemBwd);
    if (e != cudaSuccess) return (int)e;
    pna_bwd_edges_kernel<<<dim3(mp_grid_x(num_nodes), kTowers), kMpWarps * 32, kMpSmemBwd, s>>>(
        h, W, b, rowptr, col_src, col_type, num_nodes, fwd_out, aux_cnt, aux_arg, d_out, d_pre, d_xs, d_xt);
    const int per = 256 / kHid;
    int nb = (num_nodes + per - 1) / per;
    if (nb > 148 * 16) nb = 148 * 16;
    pna_bwd_nodes_kernel<<<nb, 256, 0, s>>>(rowptr, rev_pos, num_nodes, d_xs, d_xt, d_h);
    pna_bwd_weights_kernel<<<dim3(num_partials, kTowers * kTypes), 256, 0, s>>>(h, col_src, col_tgt, col_type,
                                                                               num_edges, d_pre, partials);
    pna_bwd_reduce_kernel<<<148 * 2, 256, 0, s>>>(partials, num_partials, d_W, d_b);
    adkf_count_launches(4);
    ADKF_LAUNCH_CHECK();
    return 0;
}
