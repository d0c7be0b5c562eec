// rezero.cu -- the ReZero residual of a GNN block, v' = v + alpha * f(v) with a learned scalar alpha (fs_mol/modules/gnn.py:497-498,
// 512-513), forward and backward, as float4 streaming kernels:
//   adkf_rezero_fwd : out = h + alpha * x                                   (HBM: 3 x 4 n bytes)
//   adkf_rezero_bwd : dx = alpha * g  and  dalpha = <g, x>  in ONE pass over g (HBM: 3 x 4 n bytes; torch needed a mul kernel and a
//                     cuBLAS dot: g read twice), deterministic: per-block partials in a fixed-size workspace, fixed-order finish.
// alpha is read from the device (1 float): nothing synchronises.
#include <algorithm>
#include "common.cuh"
#include "../../include/adkf_b200.h"

namespace adkf {

constexpr int kRzThreads = 256, kRzBlocks = 148 * 8;

__global__ void __launch_bounds__(kRzThreads)
rezero_fwd_kernel(const float4* __restrict__ h, const float4* __restrict__ x, const float* __restrict__ alpha, int64_t n4,
                  float4* __restrict__ out) {
    const float a = __ldg(alpha);
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 hv = __ldg(h + i), xv = __ldg(x + i);
        out[i] = make_float4(fmaf(a, xv.x, hv.x), fmaf(a, xv.y, hv.y), fmaf(a, xv.z, hv.z), fmaf(a, xv.w, hv.w));
    }
}

__global__ void __launch_bounds__(kRzThreads)
rezero_bwd_kernel(const float4* __restrict__ g, const float4* __restrict__ x, const float* __restrict__ alpha, int64_t n4,
                  float4* __restrict__ dx, double* __restrict__ partial) {
    __shared__ double red[kRzThreads / 32];
    const float a = __ldg(alpha);
    double acc = 0.0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 gv = __ldg(g + i), xv = __ldg(x + i);
        if (dx != nullptr) dx[i] = make_float4(a * gv.x, a * gv.y, a * gv.z, a * gv.w);
        acc += (double)(gv.x * xv.x + gv.y * xv.y) + (double)(gv.z * xv.z + gv.w * xv.w);
    }
    if (partial == nullptr) return;
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < kRzThreads / 32; ++w) s += red[w];
        partial[blockIdx.x] = s;
    }
}

__global__ void rezero_finish_kernel(const double* __restrict__ partial, int nblk, float* __restrict__ dalpha) {
    __shared__ double red[32];
    double s = 0.0;
    for (int b = threadIdx.x; b < nblk; b += blockDim.x) s += partial[b];        // fixed assignment of blocks to threads
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += red[w];
        dalpha[0] = (float)t;
    }
}

}  // namespace adkf

using namespace adkf;

static int rz_blocks(int64_t n4) { return (int)std::max<int64_t>(1, std::min<int64_t>((n4 + kRzThreads - 1) / kRzThreads, kRzBlocks)); }

extern "C" int64_t adkf_rezero_bwd_ws_doubles(void) { return kRzBlocks; }

extern "C" int adkf_rezero_fwd(const float* h, const float* x, const float* alpha, int64_t n, float* out, void* stream) {
    if (n <= 0) return 0;
    if (!h || !x || !alpha || !out || n % 4 || ((reinterpret_cast<uintptr_t>(h) | reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15))
        return ADKF_ERR_BADARG;
    (void)cudaGetLastError();
    rezero_fwd_kernel<<<rz_blocks(n / 4), kRzThreads, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float4*>(h), reinterpret_cast<const float4*>(x),
                                                                             alpha, n / 4, reinterpret_cast<float4*>(out));
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

// dx may be NULL (x needs no gradient); dalpha may be NULL (then ws is unused).  ws: adkf_rezero_bwd_ws_doubles() doubles.
extern "C" int adkf_rezero_bwd(const float* g, const float* x, const float* alpha, int64_t n, float* dx, float* dalpha, double* ws,
                               void* stream) {
    if (n <= 0) return 0;
    if (!g || !x || !alpha || n % 4 || (dalpha != nullptr && ws == nullptr) ||
        ((reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dx)) & 15))
        return ADKF_ERR_BADARG;
    if (dx == nullptr && dalpha == nullptr) return 0;
    (void)cudaGetLastError();
    cudaStream_t s = (cudaStream_t)stream;
    const int nblk = rz_blocks(n / 4);
    rezero_bwd_kernel<<<nblk, kRzThreads, 0, s>>>(reinterpret_cast<const float4*>(g), reinterpret_cast<const float4*>(x), alpha, n / 4,
                                                  reinterpret_cast<float4*>(dx), dalpha ? ws : nullptr);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    if (dalpha != nullptr) {
        rezero_finish_kernel<<<1, 256, 0, s>>>(ws, nblk, dalpha);
        adkf_count_launches(1);
        ADKF_LAUNCH_CHECK();
    }
    return 0;
}
