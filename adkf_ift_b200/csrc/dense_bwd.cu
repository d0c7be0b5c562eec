// dense_bwd.cu -- the elementwise half of a dense layer's backward (autograd of nn.Linear + ReLU / LeakyReLU in
// fs_mol/modules/gnn.py:79-95, mlp.py:6-22): ONE pass over dY that
//   * applies the activation derivative from the saved activation output Y (dZ = dY * act'(Y)), and
//   * accumulates the bias gradient db[c] = sum_r dZ[r, c]
// so that dY is read once (torch did where -> mul -> sum: three reads, two writes).  HBM-bound: algorithmic bytes
// 4*M*N (dY) + 4*M*N (Y, when there is an activation) + 4*M*N (dZ, when there is an activation).
// The column sums are deterministic: per-block partials in a workspace, then a fixed-order second pass.
#include "common.cuh"
#include "../../include/adkf_b200.h"

namespace adkf {

constexpr int kDbThreads = 256;

// block b handles rows [b*rows_per_block, ...); thread (ty, tx): column groups tx, tx + gx, ... (float4), rows ty, ty + gy, ...
__global__ void __launch_bounds__(kDbThreads)
act_bwd_colsum_kernel(const float* __restrict__ dY, int64_t ldy, const float* __restrict__ Y, int64_t ldyy, int64_t M, int N,
                      int act, float* __restrict__ dZ, int64_t ldz, int rows_per_block, int gx, float* __restrict__ partial) {
    extern __shared__ float4 red[];                        // [gy][gx] partials of the column group being reduced
    const int gy = kDbThreads / gx;
    const int tx = threadIdx.x % gx, ty = threadIdx.x / gx;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
    const int64_t r1 = min(M, r0 + rows_per_block);
    const int ncg = N >> 2;
    for (int cg0 = 0; cg0 < ncg; cg0 += gx) {
        const int cg = cg0 + tx;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        if (cg < ncg) {
#pragma unroll 4
            for (int64_t r = r0 + ty; r < r1; r += gy) {
                float4 g = *reinterpret_cast<const float4*>(dY + r * ldy + 4 * cg);
                if (act != 0) {
                    const float4 y = *reinterpret_cast<const float4*>(Y + r * ldyy + 4 * cg);
                    const float neg = (act == 1) ? 0.f : 0.01f;
                    g.x *= (y.x > 0.f) ? 1.f : neg;
                    g.y *= (y.y > 0.f) ? 1.f : neg;
                    g.z *= (y.z > 0.f) ? 1.f : neg;
                    g.w *= (y.w > 0.f) ? 1.f : neg;
                    *reinterpret_cast<float4*>(dZ + r * ldz + 4 * cg) = g;
                }
                acc.x += g.x; acc.y += g.y; acc.z += g.z; acc.w += g.w;
            }
        }
        if (partial != nullptr) {
            red[ty * gx + tx] = acc;
            __syncthreads();
            if (ty == 0 && cg < ncg) {
                float4 s = red[tx];
                for (int k = 1; k < gy; ++k) {
                    const float4 v = red[k * gx + tx];
                    s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
                }
                *reinterpret_cast<float4*>(partial + (size_t)blockIdx.x * N + 4 * cg) = s;
            }
            __syncthreads();
        }
    }
}

// db[c] = sum_b partial[b][c] in a fixed order: 32 columns x 8 row groups per block, row group g sums b = g, g + 8, ...
__global__ void __launch_bounds__(256)
colsum_finish_kernel(const float* __restrict__ partial, int nblk, int N, float* __restrict__ db) {
    __shared__ float red[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + tx;
    float s = 0.f;
    if (c < N)
        for (int b = ty; b < nblk; b += 8) s += partial[(size_t)b * N + c];
    red[ty][tx] = s;
    __syncthreads();
    if (ty == 0 && c < N) {
        float t = red[0][tx];
#pragma unroll
        for (int k = 1; k < 8; ++k) t += red[k][tx];
        db[c] = t;
    }
}

}  // namespace adkf

using namespace adkf;

static int db_blocks(int64_t M) {
    const int64_t want = (M + 63) / 64;                    // at least 64 rows per block
    return (int)(want < 1 ? 1 : (want > 148 * 8 ? 148 * 8 : want));
}

extern "C" int64_t adkf_act_bwd_colsum_ws_floats(int64_t M, int32_t N) { return (int64_t)db_blocks(M) * N; }

extern "C" int adkf_act_bwd_colsum(const float* dY, int64_t ldy, const float* Y, int64_t ldyy, int64_t M, int32_t N, int32_t act,
                                   float* dZ, int64_t ldz, float* db, float* ws, int64_t ws_floats, void* stream) {
    if (M <= 0 || N <= 0) return 0;
    if (!dY || N % 4 || ldy % 4 || (reinterpret_cast<uintptr_t>(dY) & 15) || act < 0 || act > 2) return ADKF_ERR_BADARG;
    if (act != 0 && (!Y || !dZ || ldyy % 4 || ldz % 4 || (reinterpret_cast<uintptr_t>(Y) & 15) || (reinterpret_cast<uintptr_t>(dZ) & 15)))
        return ADKF_ERR_BADARG;
    if (act == 0 && !db) return 0;
    const int nblk = db_blocks(M);
    if (db && (!ws || ws_floats < (int64_t)nblk * N)) return ADKF_ERR_BADARG;
    cudaStream_t s = (cudaStream_t)stream;
    (void)cudaGetLastError();
    const int ncg = N / 4;
    int gx = 1;
    while (gx < ncg && gx < kDbThreads) gx <<= 1;          // power of two <= 256 so that gy = 256 / gx is exact
    const int rows_per_block = (int)((M + nblk - 1) / nblk);
    act_bwd_colsum_kernel<<<nblk, kDbThreads, kDbThreads * sizeof(float4), s>>>(dY, ldy, Y, ldyy, M, N, act, dZ, ldz, rows_per_block,
                                                                             gx, db ? ws : nullptr);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    if (db) {
        colsum_finish_kernel<<<(N + 31) / 32, 256, 0, s>>>(ws, nblk, N, db);
        adkf_count_launches(1);
        ADKF_LAUNCH_CHECK();
    }
    return 0;
}

// ---- adkf_act_bwd_colsum_noalloc: the same two kernels with the column-sum partials in a file-scope device array --------------
// For callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): no workspace
// argument, no cudaMalloc / cudaFree / allocating library call anywhere on this path.  Every GPU that loads the library holds the
// array (kNoAllocMaxFloats * 4 bytes = 4.6 MiB) until exit.  All calls share it: two calls that OVERLAP on the GPU (different
// streams without an ordering between them) race on it; calls on one stream are ordered and safe.  Nothing in it has to start at
// zero: act_bwd_colsum_kernel writes every partial[b][c], b < nblk, c < N before colsum_finish_kernel reads exactly those.
namespace adkf {
constexpr int64_t kNoAllocMaxFloats = 1184ll * 1024;       // db_blocks(M) <= 148 * 8 = 1184 row blocks x N <= 1024 columns at full height
static __device__ __align__(256) float g_colsum_scratch[kNoAllocMaxFloats];
}  // namespace adkf

// device address of g_colsum_scratch on the CURRENT device (in host code the symbol's name is only a host-side stand-in, not a
// device pointer): looked up once per device with that device current, return code checked
static int noalloc_scratch(float** out) {
    constexpr int kMaxDev = 64;
    static float* ptrs[kMaxDev] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= kMaxDev) return ADKF_ERR_BADARG;
    if (ptrs[dev] == nullptr) {
        float* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, adkf::g_colsum_scratch);
        if (e != cudaSuccess) return (int)e;
        if (p == nullptr) return (int)cudaErrorInvalidSymbol;
        ptrs[dev] = p;
    }
    *out = ptrs[dev];
    return 0;
}

extern "C" int64_t adkf_act_bwd_colsum_noalloc_max_floats(void) { return adkf::kNoAllocMaxFloats; }

extern "C" int adkf_act_bwd_colsum_noalloc(const float* dY, int64_t ldy, const float* Y, int64_t ldyy, int64_t M, int32_t N,
                                           int32_t act, float* dZ, int64_t ldz, float* db, void* stream) {
    if (M <= 0 || N <= 0) return 0;
    // the partials of this problem must fit the static array: ADKF_ERR_TOOLARGE, nothing is launched
    if (db != nullptr && adkf_act_bwd_colsum_ws_floats(M, N) > adkf::kNoAllocMaxFloats) return ADKF_ERR_TOOLARGE;
    float* ws = nullptr;
    if (db != nullptr) {
        const int rc = noalloc_scratch(&ws);
        if (rc) return rc;
    }
    return adkf_act_bwd_colsum(dY, ldy, Y, ldyy, M, N, act, dZ, ldz, db, ws, ws ? adkf::kNoAllocMaxFloats : 0, stream);
}
