// chol_blocked.cu -- large-pool GP path (N up to a few thousand rows): blocked Cholesky with the trailing updates on the
// tcgen05 3xTF32 GEMM, joint predictive NLL, and the analytic expected-improvement sweep over a candidate pool.
//
// Replaces, for the bayes_opt / predictive_performance experiments of the reference:
//   likelihood(model(x_test)).log_prob(y_test) on ~2.25k test molecules   bayes_opt/predictive_performance/adkt.py:93-96
//   (gpytorch switches to CG/Lanczos above 800 rows -- stochastic; here the factorisation is exact)
//   botorch ExpectedImprovement(model, best_f, maximize=False) called once per pool molecule   bayes_opt/bo_utils.py:368-378
//
// Layout: J is the joint (support rows first, then query rows) covariance K(u) + sig2 I in a padded [Np, Np] row-major
// buffer, Np = N rounded up to 128; the padding is the identity, so the factorisation runs on whole 128-blocks.
// Right-looking blocked Cholesky, panel width 128:
//   potrf128   one CTA: 128 x 128 diagonal block in shared memory -> L11 (written back) and X11 = L11^-1
//   trsm       L21 = A21 X11^T               -> adkf_gemm_tf32x3 (forward form, in place: one CTA owns a row tile)
//   syrk       A22 <- A22 - L21 L21^T        -> adkf_gemm_tf32x3 with the subtract-from-D epilogue
// Because the support rows come first, the leading n x n block of L is the Cholesky factor of the support covariance:
// ONE factorisation yields both the joint NLL and the support NLL, and F = -log p(y_q | y_s) is their difference.
#include "common.cuh"
#include "../../include/adkf_b200.h"

namespace adkf {

constexpr int kNB = 128;
constexpr int kNBld = kNB + 1;
constexpr float kHalfLog2PiC = 0.9189385332046727f;

__device__ __forceinline__ float kernel_value(float u, float s, float l, int kind) {
    if (kind == ADKF_KERNEL_MATERN52) {
        const float r = sqrtf(fmaxf(u / (l * l), 1e-30f));
        return s * (1.f + 2.2360679774997896f * r + (5.f / 3.f) * r * r) * expf(-2.2360679774997896f * r);
    }
    if (kind == ADKF_KERNEL_RBF) return s * expf(-0.5f * u / (l * l));
    if (kind == ADKF_KERNEL_LINEAR) return s * l * u;
    if (kind == ADKF_KERNEL_POLY1) return s * (u + l);
    if (kind == ADKF_KERNEL_POLY2) return s * (u + l) * (u + l);
    return s * u;
}

// J[i][j] = K(u_ij) + (i == j)(sig2 + jitter); identity on the padding
__global__ void build_cov_kernel(const float* __restrict__ U, int N, int Np, int kind, const float* __restrict__ phi,
                                 float jitter, float* __restrict__ J, const int32_t* __restrict__ gate) {
    if (gate != nullptr && *gate == 0) return;
    const float sig2 = softplusf(phi[0]) + 1e-4f, s = softplusf(phi[1]), l = softplusf(phi[2]);
    const size_t total = (size_t)Np * Np;
    for (size_t idx = blockIdx.x * (size_t)blockDim.x + threadIdx.x; idx < total; idx += (size_t)gridDim.x * blockDim.x) {
        const int i = (int)(idx / Np), j = (int)(idx - (size_t)i * Np);
        float v;
        if (i < N && j < N) {
            v = kernel_value(U[(size_t)i * N + j], s, l, kind);
            if (i == j) v += sig2 + jitter;
        } else {
            v = (i == j) ? 1.f : 0.f;
        }
        J[idx] = v;
    }
}

// diagonal block: Cholesky in shared memory, L11 written back (upper triangle zeroed), X = L11^-1 to Xout [128,128]
__global__ void __launch_bounds__(512, 1)
potrf128_kernel(float* __restrict__ A, int ld, float* __restrict__ Xout, int32_t* __restrict__ status,
                const int32_t* __restrict__ gate) {
    if (gate != nullptr && *gate == 0) return;
    extern __shared__ float sm[];
    float* L = sm;                    // [128][129]
    float* X = sm + kNB * kNBld;      // [128][129]
    const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, wid = tid >> 5, nw = nt >> 5;
    for (int idx = tid; idx < kNB * kNB; idx += nt) {
        const int i = idx >> 7, j = idx & 127;
        L[i * kNBld + j] = A[(size_t)i * ld + j];
    }
    __syncthreads();
    bool ok = true;
    for (int k = 0; k < kNB; ++k) {
        const float d = L[k * kNBld + k];
        if (!(d > 0.f) || !(d < 3.0e38f)) { ok = false; break; }   // uniform: same shared value in every thread
        const float piv = sqrtf(d), inv = 1.f / piv;
        for (int i = k + 1 + tid; i < kNB; i += nt) L[i * kNBld + k] *= inv;
        __syncthreads();
        if (tid == 0) L[k * kNBld + k] = piv;
        for (int i = k + 1 + wid; i < kNB; i += nw) {
            const float lik = L[i * kNBld + k];
            for (int j = k + 1 + lane; j <= i; j += 32) L[i * kNBld + j] -= lik * L[j * kNBld + k];
        }
        __syncthreads();
    }
    if (!ok) {
        if (tid == 0) atomicOr(status, ADKF_ST_NOT_PD);
        // leave a harmless identity factor so that the rest of the sweep stays finite
        for (int idx = tid; idx < kNB * kNB; idx += nt) {
            const int i = idx >> 7, j = idx & 127;
            A[(size_t)i * ld + j] = (i == j) ? 1.f : 0.f;
            Xout[idx] = (i == j) ? 1.f : 0.f;
        }
        return;
    }
    // X = L^-1, one thread per column (a thread only re-reads its own column)
    for (int c = tid; c < kNB; c += nt) {
        for (int i = 0; i < c; ++i) X[i * kNBld + c] = 0.f;
        X[c * kNBld + c] = 1.f / L[c * kNBld + c];
        for (int i = c + 1; i < kNB; ++i) {
            float acc = 0.f;
            for (int k = c; k < i; ++k) acc += L[i * kNBld + k] * X[k * kNBld + c];
            X[i * kNBld + c] = -acc / L[i * kNBld + i];
        }
    }
    __syncthreads();
    for (int idx = tid; idx < kNB * kNB; idx += nt) {
        const int i = idx >> 7, j = idx & 127;
        A[(size_t)i * ld + j] = (j <= i) ? L[i * kNBld + j] : 0.f;
        Xout[idx] = X[i * kNBld + j];
    }
}

// w = L^-1 y by blocks (one CTA): r_k = y_k - sum_{j<k} L_kj w_j ; w_k = X_kk r_k.  Then
//   out[0] = NLL of all N rows, out[1] = NLL of the first n_sup rows, out[2] = out[0] - out[1], out[3] = log det J.
__global__ void __launch_bounds__(512, 1)
blocked_solve_nll_kernel(const float* __restrict__ Lm, int ld, const float* __restrict__ Xblk, const float* __restrict__ y,
                         int N, int Np, int n_sup, float* __restrict__ w, float* __restrict__ out, const int32_t* __restrict__ gate) {
    if (gate != nullptr && *gate == 0) return;
    __shared__ float r[kNB];
    __shared__ double red[4][16];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int nblk = Np / kNB;
    for (int k = 0; k < nblk; ++k) {
        // 4 threads per row split the columns of the already solved part
        const int row = tid >> 2, part = tid & 3;
        const int gi = k * kNB + row;
        float acc = 0.f;
        const float* Lr = Lm + (size_t)gi * ld;
        for (int j = part; j < k * kNB; j += 4) acc += Lr[j] * w[j];
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        acc += __shfl_xor_sync(0xffffffffu, acc, 2);
        if (part == 0) r[row] = ((gi < N) ? y[gi] : 0.f) - acc;
        __syncthreads();
        if (tid < kNB) {
            const float* Xr = Xblk + (size_t)k * kNB * kNB + (size_t)tid * kNB;
            float a2 = 0.f;
            for (int j = 0; j <= tid; ++j) a2 += Xr[j] * r[j];
            w[k * kNB + tid] = a2;
        }
        __threadfence_block();
        __syncthreads();
    }
    double q_all = 0.0, q_sup = 0.0, ld_all = 0.0, ld_sup = 0.0;
    for (int i = tid; i < N; i += blockDim.x) {
        const double wi = w[i], li = log((double)Lm[(size_t)i * ld + i]);
        q_all += wi * wi; ld_all += li;
        if (i < n_sup) { q_sup += wi * wi; ld_sup += li; }
    }
    double v[4] = {q_all, q_sup, ld_all, ld_sup};
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        v[a] = warp_sum(v[a]);
        if (lane == 0) red[a][wid] = v[a];
    }
    __syncthreads();
    if (tid == 0) {
        double s[4] = {0, 0, 0, 0};
        for (int a = 0; a < 4; ++a)
            for (int q = 0; q < (int)(blockDim.x >> 5); ++q) s[a] += red[a][q];
        const double nll_all = 0.5 * s[0] + s[2] + N * (double)kHalfLog2PiC;
        const double nll_sup = 0.5 * s[1] + s[3] + n_sup * (double)kHalfLog2PiC;
        out[0] = (float)nll_all;
        out[1] = (float)nll_sup;
        out[2] = (float)(nll_all - nll_sup);
        out[3] = (float)(2.0 * s[2]);
    }
}

// analytic EI for minimisation on the latent posterior (botorch ExpectedImprovement(maximize=False), bo_utils.py:368):
//   sigma^2 = var - noise (posterior of f, not of y), z = (best - mu) / sigma, ei = sigma (z Phi(z) + phi(z))
__global__ void ei_pool_kernel(const float* __restrict__ mean, const float* __restrict__ var, int m, float noise, float best_f,
                               const uint8_t* __restrict__ exclude, float* __restrict__ ei) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < m; i += gridDim.x * blockDim.x) {
        if (exclude != nullptr && exclude[i]) { ei[i] = -INFINITY; continue; }
        const float sigma = sqrtf(fmaxf(var[i] - noise, 1e-9f));   // botorch clamps the variance at 1e-9
        const float z = (best_f - mean[i]) / sigma;
        const float pdf = 0.3989422804014327f * expf(-0.5f * z * z);
        const float cdf = 0.5f * erfcf(-z * 0.7071067811865476f);
        ei[i] = sigma * (pdf + z * cdf);
    }
}

}  // namespace adkf

using namespace adkf;

extern "C" int64_t adkf_chol_blocked_ws_floats(int32_t N) {
    const int64_t Np = ((int64_t)N + kNB - 1) / kNB * kNB;
    return Np * Np + (Np / kNB) * kNB * kNB + Np + 8;     // J | X blocks | w | out
}

namespace adkf {
int gemm_tf32x3_gated(const float* A, int32_t transA, int64_t lda, const float* B, int32_t transB, int64_t ldb, int64_t M, int32_t N,
                      int32_t K, int32_t act, float* D, int64_t ldd, const int32_t* gate, cudaStream_t stream);

// Between two attempts of the jitter ladder: the next attempt runs (gate = 1) only if the previous one flagged NOT_PD, and
// then the flag is replaced by ADKF_ST_JITTER (psd_safe_cholesky: success with jitter is a warning, not an error).
__global__ void ladder_gate_kernel(int32_t* __restrict__ status, int32_t* __restrict__ gate) {
    const bool failed = (*status & ADKF_ST_NOT_PD) != 0;
    *gate = failed ? 1 : 0;
    if (failed) *status = ADKF_ST_JITTER;
}
__global__ void ladder_finish_kernel(int32_t* __restrict__ status) {
    if (*status & ADKF_ST_NOT_PD) *status = ADKF_ST_NOT_PD;        // not positive definite even with the largest jitter
}
}  // namespace adkf

// one attempt; gate == nullptr: unconditional
static int joint_nll_attempt(const float* U, int32_t N, int32_t n_sup, const float* y, int32_t kernel, const float* phi, float jitter,
                             float* ws, float* out4, int32_t* status, const int32_t* gate, cudaStream_t s) {
    const int Np = (N + kNB - 1) / kNB * kNB, nblk = Np / kNB;
    float* J = ws;
    float* X = J + (size_t)Np * Np;
    float* w = X + (size_t)nblk * kNB * kNB;
    static bool attr_done = false;
    const int smem = 2 * kNB * kNBld * (int)sizeof(float);
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute((const void*)potrf128_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        if (e != cudaSuccess) return (int)e;
        attr_done = true;
    }
    build_cov_kernel<<<148 * 8, 256, 0, s>>>(U, N, Np, kernel & 0xff, phi, jitter, J, gate);
    int launches = 2;
    for (int k = 0; k < nblk; ++k) {
        float* Akk = J + (size_t)k * kNB * Np + (size_t)k * kNB;
        float* Xk = X + (size_t)k * kNB * kNB;
        potrf128_kernel<<<1, 512, smem, s>>>(Akk, Np, Xk, status, gate);
        ++launches;
        const int64_t M = (int64_t)Np - (int64_t)(k + 1) * kNB;
        if (M > 0) {
            float* A21 = Akk + (size_t)kNB * Np;
            int rc = gemm_tf32x3_gated(A21, 0, Np, Xk, 0, kNB, M, kNB, kNB, 0, A21, Np, gate, s);
            if (rc) return rc;
            float* A22 = A21 + kNB;
            rc = gemm_tf32x3_gated(A21, 0, Np, A21, 0, Np, M, (int32_t)M, kNB, ADKF_EPI_SUB_FROM_D, A22, Np, gate, s);
            if (rc) return rc;
        }
    }
    blocked_solve_nll_kernel<<<1, 512, 0, s>>>(J, Np, X, y, N, Np, n_sup, w, out4, gate);
    adkf_count_launches(launches);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_gp_joint_nll_blocked(const float* U, int32_t N, int32_t n_sup, const float* y, int32_t kernel,
                                         const float* phi, float jitter, float* ws, int64_t ws_floats, float* out4,
                                         int32_t* status, void* stream) {
    if (N <= 0) return 0;
    if (!U || !y || !phi || !ws || !out4 || !status || n_sup < 0 || n_sup > N) return ADKF_ERR_BADARG;
    if ((kernel & 0xff) < 0 || (kernel & 0xff) > ADKF_KERNEL_TANIMOTO) return ADKF_ERR_BADARG;
    if (ws_floats < adkf_chol_blocked_ws_floats(N)) return ADKF_ERR_BADARG;
    cudaStream_t s = (cudaStream_t)stream;
    cudaMemsetAsync(status, 0, sizeof(int32_t), s);
    return joint_nll_attempt(U, N, n_sup, y, kernel, phi, jitter, ws, out4, status, nullptr, s);
}

// gpytorch's psd_safe_cholesky ladder (plain attempt, then diagonal jitter 1e-6, 1e-5, 1e-4) entirely on the device: the three
// jittered attempts are enqueued behind the first one and every kernel of an attempt returns at once unless the attempt before
// it flagged ADKF_ST_NOT_PD -- no host synchronisation inside the op.  status: 0, ADKF_ST_JITTER, or
// ADKF_ST_NOT_PD if even 1e-4 did not help.
extern "C" int adkf_gp_joint_nll_blocked_ladder(const float* U, int32_t N, int32_t n_sup, const float* y, int32_t kernel,
                                                const float* phi, float* ws, int64_t ws_floats, float* out4, int32_t* status,
                                                void* stream) {
    if (N <= 0) return 0;
    if (!U || !y || !phi || !ws || !out4 || !status || n_sup < 0 || n_sup > N) return ADKF_ERR_BADARG;
    if ((kernel & 0xff) < 0 || (kernel & 0xff) > ADKF_KERNEL_TANIMOTO) return ADKF_ERR_BADARG;
    if (ws_floats < adkf_chol_blocked_ws_floats(N)) return ADKF_ERR_BADARG;
    cudaStream_t s = (cudaStream_t)stream;
    const int64_t Np = ((int64_t)N + kNB - 1) / kNB * kNB;
    int32_t* gates = reinterpret_cast<int32_t*>(ws + Np * Np + (Np / kNB) * kNB * kNB + Np);      // the 8 spare words of the workspace
    cudaMemsetAsync(status, 0, sizeof(int32_t), s);
    int rc = joint_nll_attempt(U, N, n_sup, y, kernel, phi, 0.f, ws, out4, status, nullptr, s);
    const float jit[3] = {1e-6f, 1e-5f, 1e-4f};
    for (int a = 0; a < 3 && rc == 0; ++a) {
        ladder_gate_kernel<<<1, 1, 0, s>>>(status, gates + a);
        adkf_count_launches(1);
        rc = joint_nll_attempt(U, N, n_sup, y, kernel, phi, jit[a], ws, out4, status, gates + a, s);
    }
    ladder_finish_kernel<<<1, 1, 0, s>>>(status);
    adkf_count_launches(1);
    return rc;
}

extern "C" int adkf_ei_pool(const float* mean, const float* var, int32_t m, float noise, float best_f,
                            const uint8_t* exclude, float* ei, void* stream) {
    if (m <= 0) return 0;
    if (!mean || !var || !ei) return ADKF_ERR_BADARG;
    ei_pool_kernel<<<(m + 255) / 256, 256, 0, (cudaStream_t)stream>>>(mean, var, m, noise, best_f, exclude, ei);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}
