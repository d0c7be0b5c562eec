// pna_combine.cu -- the three PNA degree scalers folded out of the aggregate (fs_mol/modules/gnn.py:242-263 followed by
// `msg_out_projection`, gnn.py:492-495).  The reference materialises, per tower, [agg | amp(v) * agg | att(v) * agg]
// ([V, 4 x 768] = 12 KB per node and block) and multiplies by W_out [128, 3072].  amp / att are PER-NODE scalars, so
//     out[v] = agg[v] W_1^T + amp(v) * agg[v] W_amp^T + att(v) * agg[v] W_att^T + b
// with agg [V, 4 x 256] the unscaled aggregates and W_* the matching column blocks of W_out: one GEMM
// P = agg @ [W_1; W_amp; W_att]^T ([V,1024] x [1024,384], same flops) and this combine.  The aggregate written, re-read by
// the GEMM and read again by the backward pass shrinks 3x (1.64 GB -> 0.55 GB per block at V = 133,728).
//   adkf_pna_degree_scalers : amp[v] = log(deg+1)/1.1515, att[v] = 1.1515/(log(deg+1)+1e-7), deg = incoming edges of all types
//   adkf_pna_combine_fwd    : out[v,c] = P[v,c] + amp[v] P[v,128+c] + att[v] P[v,256+c] + bias[c]      (HBM: 4*(384+128)*V)
//   adkf_pna_combine_bwd    : dP = [g | amp g | att g], dbias = colsum(g) (deterministic two-pass)       (HBM: 4*(128+384)*V)
// One warp per row, one float4 per lane (128 columns).
#include "common.cuh"
#include "../../include/adkf_b200.h"

namespace adkf {

constexpr int kPcThreads = 256;
constexpr int kPcWarps = kPcThreads / 32;
constexpr float kPcSmall = 1e-7f, kPcDelta = 1.1515f;

__global__ void __launch_bounds__(256)
pna_degree_scalers_kernel(const int32_t* __restrict__ tptr, int V, float* __restrict__ amp, float* __restrict__ att) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= V) return;
    int deg = 0;
#pragma unroll
    for (int t = 0; t < 3; ++t) deg += tptr[(size_t)t * V + v + 1] - tptr[(size_t)t * V + v];
    const float logd = logf((float)deg + 1.f);
    amp[v] = logd / kPcDelta;
    att[v] = kPcDelta / (logd + kPcSmall);
}

__global__ void __launch_bounds__(kPcThreads)
pna_combine_fwd_kernel(const float* __restrict__ P, const float* __restrict__ amp, const float* __restrict__ att,
                       const float* __restrict__ bias, int64_t V, float* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = (int64_t)blockIdx.x * kPcWarps + (threadIdx.x >> 5);
    const int64_t nwarps = (int64_t)gridDim.x * kPcWarps;
    const float4 b = bias ? reinterpret_cast<const float4*>(bias)[lane] : make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t v = warp; v < V; v += nwarps) {
        const float4* p = reinterpret_cast<const float4*>(P + v * 384);
        const float4 p0 = __ldg(p + lane), p1 = __ldg(p + 32 + lane), p2 = __ldg(p + 64 + lane);
        const float a = amp[v], t = att[v];
        float4 o;
        o.x = p0.x + a * p1.x + t * p2.x + b.x;
        o.y = p0.y + a * p1.y + t * p2.y + b.y;
        o.z = p0.z + a * p1.z + t * p2.z + b.z;
        o.w = p0.w + a * p1.w + t * p2.w + b.w;
        reinterpret_cast<float4*>(out + v * 128)[lane] = o;
    }
}

__global__ void __launch_bounds__(kPcThreads)
pna_combine_bwd_kernel(const float* __restrict__ g, const float* __restrict__ amp, const float* __restrict__ att, int64_t V,
                       int rows_per_block, float* __restrict__ dP, float* __restrict__ partial) {
    __shared__ float4 red[kPcWarps][32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
    const int64_t r1 = min(V, r0 + (int64_t)rows_per_block);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int64_t v = r0 + w; v < r1; v += kPcWarps) {
        const float4 x = __ldg(reinterpret_cast<const float4*>(g + v * 128) + lane);
        const float a = amp[v], t = att[v];
        float4* q = reinterpret_cast<float4*>(dP + v * 384);
        q[lane] = x;
        q[32 + lane] = make_float4(a * x.x, a * x.y, a * x.z, a * x.w);
        q[64 + lane] = make_float4(t * x.x, t * x.y, t * x.z, t * x.w);
        acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w;
    }
    if (partial == nullptr) return;
    red[w][lane] = acc;
    __syncthreads();
    if (w == 0) {
        float4 s = red[0][lane];
#pragma unroll
        for (int k = 1; k < kPcWarps; ++k) {
            const float4 u = red[k][lane];
            s.x += u.x; s.y += u.y; s.z += u.z; s.w += u.w;
        }
        reinterpret_cast<float4*>(partial + (size_t)blockIdx.x * 128)[lane] = s;
    }
}

// dbias[c] = sum_b partial[b][c], c < 128, fixed order
__global__ void __launch_bounds__(256)
pna_combine_finish_kernel(const float* __restrict__ partial, int nblk, float* __restrict__ dbias) {
    __shared__ float red[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + tx;
    float s = 0.f;
    for (int b = ty; b < nblk; b += 8) s += partial[(size_t)b * 128 + c];
    red[ty][tx] = s;
    __syncthreads();
    if (ty == 0) {
        float t = red[0][tx];
#pragma unroll
        for (int k = 1; k < 8; ++k) t += red[k][tx];
        dbias[c] = t;
    }
}

}  // namespace adkf

using namespace adkf;

static int pc_blocks(int64_t V) {
    const int64_t want = (V + 63) / 64;
    return (int)(want < 1 ? 1 : (want > 148 * 8 ? 148 * 8 : want));
}

static bool pc_misaligned(const void* a, const void* b = nullptr, const void* c = nullptr) {
    return ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(c)) & 15) != 0;
}

extern "C" int adkf_pna_degree_scalers(const int32_t* tptr, int32_t V, float* amp, float* att, void* stream) {
    if (V <= 0) return 0;
    if (!tptr || !amp || !att) return ADKF_ERR_BADARG;
    (void)cudaGetLastError();
    pna_degree_scalers_kernel<<<(V + 255) / 256, 256, 0, (cudaStream_t)stream>>>(tptr, V, amp, att);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int64_t adkf_pna_combine_bwd_ws_floats(int64_t V) { return (int64_t)pc_blocks(V) * 128; }

extern "C" int adkf_pna_combine_fwd(const float* P, const float* amp, const float* att, const float* bias, int64_t V, float* out,
                                    void* stream) {
    if (V <= 0) return 0;
    if (!P || !amp || !att || !out || pc_misaligned(P, out, bias)) return ADKF_ERR_BADARG;
    (void)cudaGetLastError();
    const int64_t want = (V + kPcWarps - 1) / kPcWarps;
    const int grid = (int)(want > 148 * 8 ? 148 * 8 : want);
    pna_combine_fwd_kernel<<<grid, kPcThreads, 0, (cudaStream_t)stream>>>(P, amp, att, bias, V, out);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_pna_combine_bwd(const float* g, const float* amp, const float* att, int64_t V, float* dP, float* dbias,
                                    float* ws, int64_t ws_floats, void* stream) {
    cudaStream_t s = (cudaStream_t)stream;
    if (V <= 0) {
        if (dbias) cudaMemsetAsync(dbias, 0, 128 * sizeof(float), s);
        return 0;
    }
    if (!g || !amp || !att || !dP || pc_misaligned(g, dP, ws)) return ADKF_ERR_BADARG;
    const int nblk = pc_blocks(V);
    if (dbias && (!ws || ws_floats < (int64_t)nblk * 128)) return ADKF_ERR_BADARG;
    (void)cudaGetLastError();
    const int rows_per_block = (int)((V + nblk - 1) / nblk);
    pna_combine_bwd_kernel<<<nblk, kPcThreads, 0, s>>>(g, amp, att, V, rows_per_block, dP, dbias ? ws : nullptr);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    if (dbias) {
        pna_combine_finish_kernel<<<4, 256, 0, s>>>(ws, nblk, dbias);
        adkf_count_launches(1);
        ADKF_LAUNCH_CHECK();
    }
    return 0;
}
