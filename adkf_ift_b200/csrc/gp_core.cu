// gp_core.cu -- batched per-task Gaussian-process tail of ADKF-IFT, one CTA per task (sm_100a).
//
// Replaces, for T tasks per launch:
//   reinit_gp_params / compute_median_lengthscale_init   fs_mol/models/adaptive_dkt.py:88-131
//   ExactMarginalLogLikelihood + priors (gpytorch)        call sites adaptive_dkt.py:121-126,176
//   fit_gpytorch_scipy (botorch, SciPy L-BFGS-B)          fs_mol/utils/adaptive_dkt_utils.py:91
//   ExactGP eval-mode prediction + GaussianLikelihood     adaptive_dkt.py:198-203
//   cauchy_hypergradient's GP-side algebra                fs_mol/utils/cauchy_hypergradient.py:41-161
//
// Data layout: task t owns the N_t x N_t squared-distance matrix at D2 + d2_off[t] (row-major, ld = N_t,
// support rows/cols first).  The n x n working matrices (A -> L -> P = A^-1, X = L^-1 -> S = dK/dl) live in
// shared memory (row stride n|1 to keep column walks bank-conflict free) when they fit, else in an
// L2-resident global workspace; the joint (n+m) x (n+m) matrices of the IFT stage always live in the
// workspace.  All closed-form algebra follows SURVEY.md App. B and is mirrored in oracle/gp_closed_form.py.
#include <cstdio>
#include "common.cuh"
#include "../../include/adkf_b200.h"

namespace adkf {

constexpr float kSqrt5 = 2.2360679774997896f;
constexpr float kTau = 0.25f;                 // LogNormal prior scale
constexpr float kHalfLog2Pi = 0.9189385332046727f;
constexpr int kGpThreads = 512;
constexpr int kSmemBudgetFloats = 50 * 1024;  // 200 KB of matrices+vectors in shared memory

struct KParts { float K, Ku, Kl, Kll, Kul; };

// K = s k(u; l) and its partials w.r.t. the pair statistic u and the third hyperparameter l (see
// oracle/gp_closed_form.py::kernel_parts).  u is the squared distance for the stationary kernels, the plain dot
// product for linear / polynomial / cossim (gpytorch LinearKernel: v x.x'; PolynomialKernel: (x.x' + c)^p;
// fs_mol/utils/gp_utils.py:22-43) and the Tanimoto similarity itself for the Tanimoto kernel (gp_utils.py:106-134).
__device__ __forceinline__ KParts kernel_parts(float u, float s, float l, int kind) {
    KParts p;
    if (kind == ADKF_KERNEL_MATERN52) {
        const float il2 = 1.f / (l * l);
        const float r = sqrtf(fmaxf(u * il2, 1e-30f));
        const float e = expf(-kSqrt5 * r);
        const float h = 1.f + kSqrt5 * r;
        const float r2 = r * r;
        p.K = s * (h + (5.f / 3.f) * r2) * e;
        p.Ku = -s * (5.f / 6.f) * il2 * h * e;
        p.Kl = s * (5.f / 3.f) * (r2 / l) * h * e;
        p.Kll = -(5.f / 3.f) * s * il2 * e * r2 * (3.f + 3.f * kSqrt5 * r - 5.f * r2);
        p.Kul = (5.f / 6.f) * s * il2 / l * e * (2.f + 2.f * kSqrt5 * r - 5.f * r2);
    } else if (kind == ADKF_KERNEL_RBF) {
        const float il2 = 1.f / (l * l);
        const float k = expf(-0.5f * u * il2);
        p.K = s * k;
        p.Ku = -0.5f * p.K * il2;
        p.Kl = p.K * u * il2 / l;
        p.Kll = p.K * (u * u * il2 * il2 * il2 - 3.f * u * il2 * il2);
        p.Kul = (p.K * il2 / l) * (1.f - 0.5f * u * il2);
    } else if (kind == ADKF_KERNEL_LINEAR) {          // K = s v u, l = v
        p.K = s * l * u; p.Ku = s * l; p.Kl = s * u; p.Kll = 0.f; p.Kul = s;
    } else if (kind == ADKF_KERNEL_POLY1) {           // K = s (u + c), l = c
        p.K = s * (u + l); p.Ku = s; p.Kl = s; p.Kll = 0.f; p.Kul = 0.f;
    } else if (kind == ADKF_KERNEL_POLY2) {           // K = s (u + c)^2
        const float b = u + l;
        p.K = s * b * b; p.Ku = 2.f * s * b; p.Kl = p.Ku; p.Kll = 2.f * s; p.Kul = 2.f * s;
    } else {                                          // cossim (variance fixed to 1) and Tanimoto: K = s u, no third parameter
        p.K = s * u; p.Ku = s; p.Kl = 0.f; p.Kll = 0.f; p.Kul = 0.f;
    }
    return p;
}
__device__ __forceinline__ float kernel_only(float u, float s, float l, int kind) {
    if (kind == ADKF_KERNEL_MATERN52) {
        const float r = sqrtf(fmaxf(u / (l * l), 1e-30f));
        return s * (1.f + kSqrt5 * r + (5.f / 3.f) * r * r) * expf(-kSqrt5 * r);
    }
    if (kind == ADKF_KERNEL_RBF) return s * expf(-0.5f * u / (l * l));
    if (kind == ADKF_KERNEL_LINEAR) return s * l * u;
    if (kind == ADKF_KERNEL_POLY1) return s * (u + l);
    if (kind == ADKF_KERNEL_POLY2) return s * (u + l) * (u + l);
    return s * u;
}
// d2K/du2 for the stationary kernels (u = squared distance, l = lengthscale): what the directional derivative of dK/du along a
// change of the DISTANCES needs (ARD: the lengthscales act on u itself, adkf_gp_ift mode 4)
__device__ __forceinline__ float kernel_uu(float u, float s, float l, int kind) {
    const float il2 = 1.f / (l * l);
    if (kind == ADKF_KERNEL_MATERN52) {
        const float r = sqrtf(fmaxf(u * il2, 1e-30f));
        return s * (25.f / 12.f) * il2 * il2 * expf(-kSqrt5 * r);
    }
    if (kind == ADKF_KERNEL_RBF) return 0.25f * s * il2 * il2 * expf(-0.5f * u * il2);
    return 0.f;
}
__device__ __forceinline__ bool kind_is_stationary(int kind) { return kind == ADKF_KERNEL_MATERN52 || kind == ADKF_KERNEL_RBF; }
// parameters the inner problem optimises: bit i <-> raw parameter i.  cossim / Tanimoto have no third parameter; the
// caller can freeze more through the ADKF_KERNEL_FREEZE_* bits of the `kernel` argument (ExactTanimotoGP keeps the noise
// fixed for classification, gp_utils.py:143).
__device__ __forceinline__ int free_mask_of(int kernel_arg) {
    const int kind = kernel_arg & 0xff;
    int m = 7;
    if (kind == ADKF_KERNEL_COSSIM || kind == ADKF_KERNEL_TANIMOTO) m &= ~4;
    if (kernel_arg & ADKF_KERNEL_FREEZE_NOISE) m &= ~1;
    if (kernel_arg & ADKF_KERNEL_FREEZE_OUTPUTSCALE) m &= ~2;
    if (kernel_arg & ADKF_KERNEL_FREEZE_THIRD) m &= ~4;
    return m;
}

struct Hyper { float sig2, s, l, t[3], c[3]; };
__device__ __forceinline__ Hyper make_hyper(const float* phi) {
    Hyper h;
    h.sig2 = softplusf(phi[0]) + 1e-4f;   // GreaterThan(1e-4)
    h.s = softplusf(phi[1]);
    h.l = softplusf(phi[2]);
#pragma unroll
    for (int i = 0; i < 3; ++i) { h.t[i] = sigmoidf_(phi[i]); h.c[i] = h.t[i] * (1.f - h.t[i]); }
    return h;
}
__device__ __forceinline__ float inv_softplusf(float x) { return x + logf(-expm1f(-x)); }

// log LogNormal(x; loc, tau) and its first two derivatives in x
__device__ __forceinline__ void lognormal_terms(float x, float loc, float& lp, float& d1, float& d2) {
    const float lx = logf(x), it2 = 1.f / (kTau * kTau);
    lp = -lx - logf(kTau) - kHalfLog2Pi - 0.5f * (lx - loc) * (lx - loc) * it2;
    const float q = (lx - loc) * it2;
    d1 = -(1.f + q) / x;
    d2 = (1.f + q - it2) / (x * x);
}

// ---------------------------------------------------------------------------------------------
// block-cooperative dense building blocks (generic pointers: shared or global)
// ---------------------------------------------------------------------------------------------
// In-place lower Cholesky (right-looking).  Returns false (uniformly) on a non-positive pivot.
__device__ bool chol_lower(float* M, int n, int ld) {
    const int tid = threadIdx.x, nt = blockDim.x, lane = tid & 31, wid = tid >> 5, nw = nt >> 5;
    for (int k = 0; k < n; ++k) {
        const float d = M[(size_t)k * ld + k];
        if (!(d > 0.f) || !(d < 3.0e38f)) return false;  // same value in every thread -> uniform exit
        const float piv = sqrtf(d), inv = 1.f / piv;
        for (int i = k + 1 + tid; i < n; i += nt) M[(size_t)i * ld + k] *= inv;
        __syncthreads();
        if (tid == 0) M[(size_t)k * ld + k] = piv;
        for (int i = k + 1 + wid; i < n; i += nw) {
            const float lik = M[(size_t)i * ld + k];
            float* row = M + (size_t)i * ld;
            for (int j = k + 1 + lane; j <= i; j += 32) row[j] -= lik * M[(size_t)j * ld + k];
        }
        __syncthreads();
    }
    return true;
}

// w <- L^-1 w  (forward substitution by warp 0; w in shared memory).  Caller syncs afterwards.
__device__ void fwd_solve_warp0(const float* L, int n, int ld, float* w) {
    if ((threadIdx.x >> 5) != 0) return;
    const int lane = threadIdx.x & 31;
    for (int k = 0; k < n; ++k) {
        if (lane == 0) w[k] = w[k] / L[(size_t)k * ld + k];
        __syncwarp();
        const float wk = w[k];
        for (int i = k + 1 + lane; i < n; i += 32) w[i] -= L[(size_t)i * ld + k] * wk;
        __syncwarp();
    }
}

// X <- L^-1 (lower triangular, upper part zeroed); one thread per column, no synchronisation needed
// because a thread only ever re-reads its own column.
__device__ void tri_inverse(const float* L, float* X, int n, int ld) {
    for (int c = threadIdx.x; c < n; c += blockDim.x) {
        for (int i = 0; i < c; ++i) X[(size_t)i * ld + c] = 0.f;
        X[(size_t)c * ld + c] = 1.f / L[(size_t)c * ld + c];
        for (int i = c + 1; i < n; ++i) {
            const float* Li = L + (size_t)i * ld;
            float acc = 0.f;
            for (int k = i - 1; k >= c; --k) acc += Li[k] * X[(size_t)k * ld + c];
            X[(size_t)i * ld + c] = -acc / Li[i];
        }
    }
}

// X <- L^-1 with one WARP per column (right-looking substitution): lane l keeps the residual entries of rows l, l+32, ...
// in registers, every step broadcasts x_k by shuffle and updates all later rows at once, so the dependent chain per
// column is n short steps instead of the n^2/2 serial multiply-adds of the thread-per-column version (clock64 probe:
// 64 x 64 tile 100 k -> ~10 k cycles).  n <= 32 Q.
template <int Q>
__device__ void tri_inverse_cols(const float* L, float* X, int n, int ld) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int c = wid; c < n; c += nw) {
        float r[Q];
#pragma unroll
        for (int q = 0; q < Q; ++q) r[q] = (lane + 32 * q == c) ? 1.f : 0.f;
        for (int i = lane; i < c; i += 32) X[(size_t)i * ld + c] = 0.f;
        for (int k = c; k < n; ++k) {
            float rk = 0.f;
#pragma unroll
            for (int q = 0; q < Q; ++q) if (q == (k >> 5)) rk = r[q];
            rk = __shfl_sync(0xffffffffu, rk, k & 31);
            const float xk = rk / L[(size_t)k * ld + k];
            if (lane == (k & 31)) X[(size_t)k * ld + c] = xk;
#pragma unroll
            for (int q = 0; q < Q; ++q) {
                const int i = lane + 32 * q;
                if (i > k && i < n) r[q] = fmaf(-L[(size_t)i * ld + k], xk, r[q]);
            }
        }
    }
}
__device__ void tri_inverse_fast(const float* L, float* X, int n, int ld) {
    if (n <= 64) tri_inverse_cols<2>(L, X, n, ld);
    else if (n <= 128) tri_inverse_cols<4>(L, X, n, ld);
    else if (n <= 256) tri_inverse_cols<8>(L, X, n, ld);
    else tri_inverse(L, X, n, ld);
}

// P <- X^T X (full symmetric), X lower triangular.
__device__ void xtx(const float* X, float* P, int n, int ld) {
    for (int idx = threadIdx.x; idx < n * n; idx += blockDim.x) {
        const int i = idx / n, j = idx - i * n;
        float acc = 0.f;
        for (int k = max(i, j); k < n; ++k) acc += X[(size_t)k * ld + i] * X[(size_t)k * ld + j];
        P[(size_t)i * ld + j] = acc;
    }
}

// out_i = sum_j M[i][j] v[j]   (thread per row)
__device__ void matvec(const float* M, int n, int ld, const float* v, float* out) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float* r = M + (size_t)i * ld;
        float acc = 0.f;
        for (int j = 0; j < n; ++j) acc += r[j] * v[j];
        out[i] = acc;
    }
}

// ---------------------------------------------------------------------------------------------
// shared scalar state
// ---------------------------------------------------------------------------------------------
struct GpShared {
    double red[16 * 32];
    double sums[16];
    float phi[3], phi_try[3];
    float f, f_try, nll, logdet;
    float g_psi[3], g[3], H[9];
    float lam, jitter;
    float l0, ls_loc;
    int ok, hist[256];
    unsigned prefix;
    int krank, cnt, flag;
    float v[3], dF[3], pd[3];
};

struct GpProblem {
    const float* D2;   // n x n block (top-left of the task matrix)
    int ldd, n;
    const float* y;
    int kind;
    float noise_loc, ls_loc;
    int has_lsp;
    int free_mask;   // bit i set: raw parameter i is optimised (frozen ones get g = 0, H row/col = identity)
};

// A (+ jitter) -> Cholesky in M1 with the psd_safe_cholesky retry policy.  Returns false if not PD after
// the three jitter retries.  sh.jitter holds the jitter finally used.
__device__ bool build_and_factor(const GpProblem& pb, const Hyper& h, float* M1, int ld, GpShared& sh) {
    const int n = pb.n;
    const float jit[4] = {0.f, 1e-6f, 1e-5f, 1e-4f};
    for (int a = 0; a < 4; ++a) {
        for (int idx = threadIdx.x; idx < n * n; idx += blockDim.x) {
            const int i = idx / n, j = idx - i * n;
            float k = kernel_only(pb.D2[(size_t)i * pb.ldd + j], h.s, h.l, pb.kind);
            if (i == j) k += h.sig2 + jit[a];
            M1[(size_t)i * ld + j] = k;
        }
        __syncthreads();
        const bool ok = chol_lower(M1, n, ld);
        __syncthreads();
        if (ok) {
            if (threadIdx.x == 0) sh.jitter = jit[a];
            __syncthreads();
            return true;
        }
    }
    return false;
}

// f only (line-search trial).  Writes sh.f_try; returns false when A is not PD.
__device__ bool gp_eval_f(const GpProblem& pb, const float* phi, float* M1, int ld, float* vecw, GpShared& sh) {
    const int n = pb.n;
    const Hyper h = make_hyper(phi);
    if (!build_and_factor(pb, h, M1, ld, sh)) return false;
    for (int i = threadIdx.x; i < n; i += blockDim.x) vecw[i] = pb.y[i];
    __syncthreads();
    fwd_solve_warp0(M1, n, ld, vecw);
    __syncthreads();
    float part[2] = {0.f, 0.f};
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        part[0] += vecw[i] * vecw[i];
        part[1] += logf(M1[(size_t)i * ld + i]);
    }
    block_sum_multi<2>(part, sh.red, sh.sums);
    if (threadIdx.x == 0) {
        const float nll = 0.5f * (float)sh.sums[0] + (float)sh.sums[1] + n * kHalfLog2Pi;
        float lp = 0.f, d1, d2, lpl = 0.f;
        if (!isnan(pb.noise_loc)) lognormal_terms(h.sig2, pb.noise_loc, lp, d1, d2);   // NaN loc = no noise prior
        if (pb.has_lsp) { float a, b; lognormal_terms(h.l, pb.ls_loc, lpl, a, b); }
        sh.f_try = (nll - lp - lpl) / n;
    }
    __syncthreads();
    return true;
}

// f, gradient and Hessian w.r.t. the raw parameters.  On return M1 = P = A^-1, M2 = S = dK/dl,
// vec[0..n) = alpha, vec[n..2n) = u = S alpha, vec[2n..3n) = w = P alpha.  Results in sh.f/g/H/g_psi/nll.
__device__ bool gp_eval_full(const GpProblem& pb, const float* phi, float* M1, float* M2, int ld, float* vec,
                             GpShared& sh, bool want_hess) {
    const int n = pb.n;
    const Hyper h = make_hyper(phi);
    float* alpha = vec;
    float* u = vec + n;
    float* w = vec + 2 * n;
    float* q = vec + 3 * n;
    if (!build_and_factor(pb, h, M1, ld, sh)) return false;
    float part_ld[1] = {0.f};
    for (int i = threadIdx.x; i < n; i += blockDim.x) part_ld[0] += logf(M1[(size_t)i * ld + i]);
    block_sum_multi<1>(part_ld, sh.red, sh.sums);
    const float logdet = 2.f * (float)sh.sums[0];
    tri_inverse_fast(M1, M2, n, ld);
    __syncthreads();
    xtx(M2, M1, n, ld);            // M1 <- P
    __syncthreads();
    matvec(M1, n, ld, pb.y, alpha);
    // M2 <- S = dK/dl (diagonal is exactly 0: r = 0)
    for (int idx = threadIdx.x; idx < n * n; idx += blockDim.x) {
        const int i = idx / n, j = idx - i * n;
        const KParts kp = kernel_parts(pb.D2[(size_t)i * pb.ldd + j], h.s, h.l, pb.kind);
        M2[(size_t)i * ld + j] = kp.Kl;
    }
    __syncthreads();
    matvec(M2, n, ld, alpha, u);
    matvec(M1, n, ld, alpha, w);
    __syncthreads();
    matvec(M1, n, ld, u, q);
    // element-wise sums
    float part[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) part[k] = 0.f;
    // 0 trP 1 e1=sum P^2 2 aa 3 ya 4 aPa 5 e2=sum S*Gam 6 e3=sum S2*Gam 7 trPPS 8 trPSPS 9 wu 10 aSa 11 uPu
    for (int idx = threadIdx.x; idx < n * n; idx += blockDim.x) {
        const int i = idx / n, j = idx - i * n;
        const float p = M1[(size_t)i * ld + j];
        const float gam = 0.5f * (p - alpha[i] * alpha[j]);
        part[1] += p * p;
        part[5] += M2[(size_t)i * ld + j] * gam;
        if (want_hess) {
            const KParts kp = kernel_parts(pb.D2[(size_t)i * pb.ldd + j], h.s, h.l, pb.kind);
            part[6] += kp.Kll * gam;
        }
        if (i == j) part[0] += p;
    }
    __syncthreads();  // q complete
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        part[2] += alpha[i] * alpha[i];
        part[3] += pb.y[i] * alpha[i];
        part[4] += alpha[i] * w[i];
        part[9] += w[i] * u[i];
        part[10] += alpha[i] * u[i];
        part[11] += u[i] * q[i];
    }
    if (want_hess) {
        // n^3 part: (PS)_ij and (PS)_ji without materialising PS (P, S symmetric)
        for (int idx = threadIdx.x; idx < n * n; idx += blockDim.x) {
            const int i = idx / n, j = idx - i * n;
            const float* Pi = M1 + (size_t)i * ld;
            const float* Si = M2 + (size_t)i * ld;
            float ps = 0.f, sp = 0.f;
            for (int k = 0; k < n; ++k) {
                ps += Pi[k] * M2[(size_t)k * ld + j];
                sp += Si[k] * M1[(size_t)k * ld + j];
            }
            part[7] += ps * Pi[j];
            part[8] += ps * sp;
        }
    }
    block_sum_multi<12>(part, sh.red, sh.sums);
    if (threadIdx.x == 0) {
        const double* S = sh.sums;
        const double trP = S[0], e1 = S[1], aa = S[2], ya = S[3], aPa = S[4], e2 = S[5], e3 = S[6];
        const double trPPS = S[7], trPSPS = S[8], wu = S[9], aSa = S[10], uPu = S[11];
        const double sig2 = h.sig2, s = h.s;
        const float nll = (float)(0.5 * ya + 0.5 * logdet + n * (double)kHalfLog2Pi);
        float lp = 0.f, d1n = 0.f, d2n = 0.f, lpl = 0.f, d1l = 0.f, d2l = 0.f;
        if (!isnan(pb.noise_loc)) lognormal_terms(h.sig2, pb.noise_loc, lp, d1n, d2n);   // bo_utils.py:425-430: optional
        if (pb.has_lsp) lognormal_terms(h.l, pb.ls_loc, lpl, d1l, d2l);
        sh.nll = nll;
        sh.logdet = logdet;
        sh.f = (nll - lp - lpl) / n;
        double gp[3];
        gp[0] = (0.5 * trP - 0.5 * aa - d1n) / n;
        gp[1] = (0.5 * (n - sig2 * trP) / s - 0.5 * (ya - sig2 * aa) / s) / n;
        gp[2] = (e2 - d1l) / n;
        for (int i = 0; i < 3; ++i) { sh.g_psi[i] = (float)gp[i]; sh.g[i] = (float)(h.t[i] * gp[i]); }
        if (want_hess) {
            double Hp[9];
            Hp[0] = -0.5 * e1 + aPa - d2n;
            Hp[1] = Hp[3] = (-0.5 * (trP - sig2 * e1) + (aa - sig2 * aPa)) / s;
            Hp[2] = Hp[6] = -0.5 * trPPS + wu;
            Hp[4] = (-0.5 * (n - 2.0 * sig2 * trP + sig2 * sig2 * e1) + (ya - 2.0 * sig2 * aa + sig2 * sig2 * aPa)) / (s * s);
            Hp[5] = Hp[7] = (0.5 * sig2 * trPPS + 0.5 * aSa - sig2 * wu) / s;
            Hp[8] = -0.5 * trPSPS + e3 + uPu - d2l;
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) {
                    double v = Hp[i * 3 + j] / n * h.t[i] * h.t[j];
                    if (i == j) v += h.c[i] * gp[i];
                    sh.H[i * 3 + j] = (float)v;
                }
        }
        for (int i = 0; i < 3; ++i) {
            if (pb.free_mask & (1 << i)) continue;
            sh.g[i] = 0.f;
            sh.g_psi[i] = 0.f;
            if (want_hess) {
                for (int j = 0; j < 3; ++j) sh.H[i * 3 + j] = sh.H[j * 3 + i] = 0.f;
                sh.H[i * 3 + i] = 1.f;
            }
        }
    }
    __syncthreads();
    return true;
}

// 3x3 solve (H + lam I) p = -g by Cholesky in double; returns false if not PD.
__device__ bool solve3_spd(const float* H, float lam, const float* g, float* p) {
    double a00 = H[0] + lam, a10 = H[3], a11 = H[4] + lam, a20 = H[6], a21 = H[7], a22 = H[8] + lam;
    if (!(a00 > 0)) return false;
    const double l00 = sqrt(a00), l10 = a10 / l00, l20 = a20 / l00;
    const double d1 = a11 - l10 * l10;
    if (!(d1 > 0)) return false;
    const double l11 = sqrt(d1), l21 = (a21 - l20 * l10) / l11;
    const double d2 = a22 - l20 * l20 - l21 * l21;
    if (!(d2 > 0)) return false;
    const double l22 = sqrt(d2);
    // condition guard: reject near-singular systems
    const double mx = fmax(a00, fmax(a11, a22));
    if (d1 < 1e-9 * mx || d2 < 1e-9 * mx) return false;
    const double b0 = -g[0], b1 = -g[1], b2 = -g[2];
    const double y0 = b0 / l00, y1 = (b1 - l10 * y0) / l11, y2 = (b2 - l20 * y0 - l21 * y1) / l22;
    const double x2 = y2 / l22, x1 = (y1 - l21 * x2) / l11, x0 = (y0 - l10 * x1 - l20 * x2) / l00;
    p[0] = (float)x0; p[1] = (float)x1; p[2] = (float)x2;
    return true;
}
// general 3x3 solve H x = b (partial pivoting, double) -- torch.linalg.solve semantics for the IFT step
__device__ bool solve3_general(const float* H, const float* b, float* x) {
    double A[3][4] = {{H[0], H[1], H[2], b[0]}, {H[3], H[4], H[5], b[1]}, {H[6], H[7], H[8], b[2]}};
    for (int c = 0; c < 3; ++c) {
        int piv = c;
        for (int r = c + 1; r < 3; ++r) if (fabs(A[r][c]) > fabs(A[piv][c])) piv = r;
        if (fabs(A[piv][c]) < 1e-30) return false;
        if (piv != c) for (int k = 0; k < 4; ++k) { double tmp = A[c][k]; A[c][k] = A[piv][k]; A[piv][k] = tmp; }
        for (int r = c + 1; r < 3; ++r) {
            const double m = A[r][c] / A[c][c];
            for (int k = c; k < 4; ++k) A[r][k] -= m * A[c][k];
        }
    }
    double xs[3];
    for (int r = 2; r >= 0; --r) {
        double acc = A[r][3];
        for (int k = r + 1; k < 3; ++k) acc -= A[r][k] * xs[k];
        xs[r] = acc / A[r][r];
    }
    x[0] = (float)xs[0]; x[1] = (float)xs[1]; x[2] = (float)xs[2];
    return true;
}

// lower median of { D2_ij : i<j, D2_ij > 0 } by 4-pass radix select on the float bit pattern
// (torch.median semantics, adaptive_dkt.py:128-131).  Result -> sh.l0 = sqrt(0.5 * median).
__device__ void median_lengthscale(const GpProblem& pb, GpShared& sh) {
    const int n = pb.n;
    if (threadIdx.x == 0) { sh.cnt = 0; sh.prefix = 0u; }
    __syncthreads();
    int local = 0;
    for (int idx = threadIdx.x; idx < n * n; idx += blockDim.x) {
        const int i = idx / n, j = idx - i * n;
        if (i < j && pb.D2[(size_t)i * pb.ldd + j] > 0.f) ++local;
    }
    local = (int)warp_sum((float)local);  // counts < 2^24: exact in fp32
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(&sh.cnt, local);
    __syncthreads();
    if (sh.cnt == 0) {  // degenerate: all support points identical
        if (threadIdx.x == 0) sh.l0 = 1.f;
        __syncthreads();
        return;
    }
    if (threadIdx.x == 0) sh.krank = (sh.cnt - 1) / 2;
    __syncthreads();
    unsigned mask = 0u;
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 24 - 8 * pass;
        for (int b = threadIdx.x; b < 256; b += blockDim.x) sh.hist[b] = 0;
        __syncthreads();
        const unsigned prefix = sh.prefix;
        for (int idx = threadIdx.x; idx < n * n; idx += blockDim.x) {
            const int i = idx / n, j = idx - i * n;
            if (i < j) {
                const float v = pb.D2[(size_t)i * pb.ldd + j];
                if (v > 0.f) {
                    const unsigned bits = __float_as_uint(v);
                    if ((bits & mask) == prefix) atomicAdd(&sh.hist[(bits >> shift) & 255u], 1);
                }
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            int k = sh.krank, b = 0;
            for (; b < 256; ++b) {
                if (k < sh.hist[b]) break;
                k -= sh.hist[b];
            }
            sh.krank = k;
            sh.prefix = prefix | ((unsigned)b << shift);
        }
        mask |= (255u << shift);
        __syncthreads();
    }
    if (threadIdx.x == 0) sh.l0 = sqrtf(0.5f * __uint_as_float(sh.prefix));
    __syncthreads();
}

// carve the working matrices out of shared memory when they fit, else out of the global workspace
struct Carve { float *M1, *M2, *vec; int ld; };
__device__ __forceinline__ Carve carve_workspace(int n, float* smem_f, float* ws_task, bool& in_smem) {
    Carve c;
    c.ld = n | 1;
    const size_t need = (size_t)2 * n * c.ld + 4 * (size_t)n;
    in_smem = need <= (size_t)kSmemBudgetFloats;
    float* base = in_smem ? smem_f : ws_task;
    c.M1 = base;
    c.M2 = base + (size_t)n * c.ld;
    c.vec = base + (size_t)2 * n * c.ld;
    return c;
}

// =============================================================================================
// kernel 1: inner fit
// =============================================================================================
__global__ void __launch_bounds__(kGpThreads, 1)
gp_fit_kernel(const int32_t* __restrict__ z_off, const int32_t* __restrict__ n_sup,
              const int64_t* __restrict__ d2_off, const float* __restrict__ D2, const float* __restrict__ y,
              int kind, float noise_init, float noise_prior_loc, int ls_prior, int init_mode, int max_iter,
              float gtol, float* __restrict__ phi_io, float* __restrict__ aux, int32_t* __restrict__ status,
              float* __restrict__ ws, const int64_t* __restrict__ ws_off) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    GpShared& sh = *reinterpret_cast<GpShared*>(smem_raw);
    float* smem_f = reinterpret_cast<float*>(smem_raw + ((sizeof(GpShared) + 15) / 16) * 16);
    const int t = blockIdx.x;
    const int N = z_off[t + 1] - z_off[t];
    GpProblem pb;
    pb.n = n_sup[t];
    pb.ldd = N;
    pb.D2 = D2 + d2_off[t];
    pb.y = y + z_off[t];
    pb.kind = kind & 0xff;
    pb.free_mask = free_mask_of(kind);
    pb.noise_loc = noise_prior_loc;
    bool in_smem;
    if (ws_off[gridDim.x] < 0) { if (threadIdx.x == 0) status[t] = ADKF_ST_BADWS; return; }
    Carve cv = carve_workspace(pb.n, smem_f, ws + ws_off[t], in_smem);
    float* auxt = aux + (size_t)t * ADKF_AUX;
    int st = ADKF_ST_OK;
    const bool stationary = kind_is_stationary(pb.kind);
    if (!stationary) ls_prior = 0;   // the lengthscale prior / median initialisation only exist for matern and rbf

    if (stationary) median_lengthscale(pb, sh);   // (adaptive_dkt.py:92-101)
    else if (threadIdx.x == 0) sh.l0 = 0.6931471805599453f;   // softplus(0): gpytorch's default variance / offset
    __syncthreads();
    if (threadIdx.x == 0) {
        if (ls_prior == 2) sh.ls_loc = auxt[ADKF_AUX_LS_LOC];
        else sh.ls_loc = logf(sh.l0) + kTau * kTau;
        if (init_mode == 1) {
            sh.phi[0] = inv_softplusf(noise_init - 1e-4f);
            sh.phi[1] = 0.f;
            sh.phi[2] = stationary ? inv_softplusf(sh.l0) : 0.f;
            if (pb.kind == ADKF_KERNEL_COSSIM || pb.kind == ADKF_KERNEL_TANIMOTO) sh.phi[2] = 0.5413248546129181f;  // softplus^-1(1)
        } else {
            for (int i = 0; i < 3; ++i) sh.phi[i] = phi_io[t * 3 + i];
        }
        sh.lam = 0.f;
    }
    __syncthreads();
    pb.has_lsp = ls_prior != 0;
    pb.ls_loc = sh.ls_loc;

    int iters = 0, nfev = 1;
    bool ok = gp_eval_full(pb, sh.phi, cv.M1, cv.M2, cv.ld, cv.vec, sh, true);
    if (!ok) st |= ADKF_ST_NOT_PD;
    bool converged = false;
    while (ok && iters < max_iter) {
        float gmax = fmaxf(fabsf(sh.g[0]), fmaxf(fabsf(sh.g[1]), fabsf(sh.g[2])));
        if (gmax <= gtol) { converged = true; break; }
        bool accepted = false;
        for (int attempt = 0; attempt < 12 && !accepted; ++attempt) {
            // every thread computes the same 3x3 step (identical inputs -> identical result, no sync needed)
            float p[3];
            float lam = sh.lam;
            const float hmax = fmaxf(1.f, fmaxf(fabsf(sh.H[0]), fmaxf(fabsf(sh.H[4]), fabsf(sh.H[8]))));
            bool pd = solve3_spd(sh.H, lam, sh.g, p);
            if (pd) {
                float mx = fmaxf(fabsf(p[0]), fmaxf(fabsf(p[1]), fabsf(p[2])));
                if (mx > 2.f) { const float sc = 2.f / mx; p[0] *= sc; p[1] *= sc; p[2] *= sc; }
            }
            __syncthreads();
            if (!pd) {
                if (threadIdx.x == 0) sh.lam = fmaxf(4.f * lam, 1e-3f * hmax);
                __syncthreads();
                continue;
            }
            if (threadIdx.x == 0) for (int i = 0; i < 3; ++i) sh.phi_try[i] = sh.phi[i] + p[i];
            __syncthreads();
            const bool tok = gp_eval_f(pb, sh.phi_try, cv.M1, cv.ld, cv.vec, sh);
            ++nfev;
            const float gp = sh.g[0] * p[0] + sh.g[1] * p[1] + sh.g[2] * p[2];
            // Armijo with a little slack for fp32 noise in f near the optimum
            const bool good = tok && isfinite(sh.f_try) && (sh.f_try <= sh.f + 1e-4f * gp + 2e-7f * fabsf(sh.f));
            __syncthreads();
            if (good) {
                if (threadIdx.x == 0) {
                    for (int i = 0; i < 3; ++i) sh.phi[i] = sh.phi_try[i];
                    sh.lam = (lam > 1e-8f) ? 0.25f * lam : 0.f;
                }
                accepted = true;
            } else {
                if (threadIdx.x == 0) sh.lam = fmaxf(4.f * lam, 1e-3f * hmax);
            }
            __syncthreads();
        }
        if (!accepted) break;
        ++iters;
        ok = gp_eval_full(pb, sh.phi, cv.M1, cv.M2, cv.ld, cv.vec, sh, true);
        ++nfev;
        if (!ok) st |= ADKF_ST_NOT_PD;
    }
    if (ok && max_iter > 0 && !converged) {
        const float gmax = fmaxf(fabsf(sh.g[0]), fmaxf(fabsf(sh.g[1]), fabsf(sh.g[2])));
        // a line search that can no longer decrease f in fp32 at a tiny gradient is converged for our purposes
        if (gmax > fmaxf(gtol, 5e-5f)) st |= ADKF_ST_NOT_CONVERGED;
    }
    if (threadIdx.x == 0) {
        if (sh.jitter > 0.f) st |= ADKF_ST_JITTER;
        for (int i = 0; i < 3; ++i) phi_io[t * 3 + i] = sh.phi[i];
        auxt[ADKF_AUX_F] = sh.f;
        for (int i = 0; i < 3; ++i) auxt[ADKF_AUX_G + i] = sh.g[i];
        for (int i = 0; i < 9; ++i) auxt[ADKF_AUX_H + i] = sh.H[i];
        auxt[ADKF_AUX_L0] = sh.l0;
        auxt[ADKF_AUX_LS_LOC] = sh.ls_loc;
        auxt[ADKF_AUX_ITERS] = (float)iters;
        auxt[ADKF_AUX_JITTER] = sh.jitter;
        auxt[ADKF_AUX_NFEV] = (float)nfev;
        status[t] = st;
    }
}

// =============================================================================================
// kernel 2: posterior predictive mean / variance of the query rows
// =============================================================================================
__global__ void __launch_bounds__(kGpThreads, 1)
gp_predict_kernel(const int32_t* __restrict__ z_off, const int32_t* __restrict__ n_sup,
                  const int64_t* __restrict__ d2_off, const float* __restrict__ D2, const float* __restrict__ y,
                  int kind, const float* __restrict__ phi, float* __restrict__ mean, float* __restrict__ var,
                  int32_t* __restrict__ status, float* __restrict__ ws, const int64_t* __restrict__ ws_off) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    GpShared& sh = *reinterpret_cast<GpShared*>(smem_raw);
    float* smem_f = reinterpret_cast<float*>(smem_raw + ((sizeof(GpShared) + 15) / 16) * 16);
    const int t = blockIdx.x;
    const int N = z_off[t + 1] - z_off[t];
    GpProblem pb;
    pb.n = n_sup[t];
    pb.ldd = N;
    pb.D2 = D2 + d2_off[t];
    pb.y = y + z_off[t];
    kind &= 0xff;
    pb.kind = kind;
    pb.free_mask = 7;
    const int n = pb.n, m = N - n;
    bool in_smem;
    if (ws_off[gridDim.x] < 0) { if (threadIdx.x == 0) status[t] = ADKF_ST_BADWS; return; }
    Carve cv = carve_workspace(n, smem_f, ws + ws_off[t], in_smem);
    float ph[3] = {phi[t * 3], phi[t * 3 + 1], phi[t * 3 + 2]};
    const Hyper h = make_hyper(ph);
    if (!build_and_factor(pb, h, cv.M1, cv.ld, sh)) {
        for (int j = threadIdx.x; j < m; j += blockDim.x) {
            mean[z_off[t] + n + j] = nanf("");
            var[z_off[t] + n + j] = nanf("");
        }
        if (threadIdx.x == 0) status[t] = ADKF_ST_NOT_PD;
        return;
    }
    tri_inverse_fast(cv.M1, cv.M2, n, cv.ld);     // M2 = X = L^-1
    float* w = cv.vec;                       // w = X y
    float* alpha = cv.vec + n;               // alpha = X^T w
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        float acc = 0.f;
        for (int k = 0; k <= i; ++k) acc += cv.M2[(size_t)i * cv.ld + k] * pb.y[k];
        w[i] = acc;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        float acc = 0.f;
        for (int k = i; k < n; ++k) acc += cv.M2[(size_t)k * cv.ld + i] * w[k];
        alpha[i] = acc;
    }
    __syncthreads();
    // one warp per query point: k_j (n values, from D2 row n+j), mean = k.alpha, var = s + sig2 - |X k|^2
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    float* kbuf_all = cv.M1;                 // L no longer needed: reuse as per-warp k buffers (nw * n <= n*ld if nw <= n)
    const bool kbuf_ok = nw <= n;
    for (int j = wid; j < m; j += nw) {
        const float* d2row = pb.D2 + (size_t)(n + j) * pb.ldd;
        float mu = 0.f;
        float* kb = kbuf_all + (size_t)wid * n;
        for (int i = lane; i < n; i += 32) {
            const float k = kernel_only(d2row[i], h.s, h.l, kind);
            if (kbuf_ok) kb[i] = k;
            mu += k * alpha[i];
        }
        mu = warp_sum(mu);
        __syncwarp();
        float q = 0.f;
        for (int i = lane; i < n; i += 32) {   // (X k)_i = sum_{c<=i} X[i][c] k_c
            float acc = 0.f;
            const float* Xi = cv.M2 + (size_t)i * cv.ld;
            if (kbuf_ok) {
                for (int c = 0; c <= i; ++c) acc += Xi[c] * kb[c];
            } else {
                for (int c = 0; c <= i; ++c) acc += Xi[c] * kernel_only(d2row[c], h.s, h.l, kind);
            }
            q += acc * acc;
        }
        q = warp_sum(q);
        if (lane == 0) {
            mean[z_off[t] + n + j] = mu;
            // prior variance k(x*, x*) from the diagonal of the pair statistic (0 -> s for the stationary kernels)
            var[z_off[t] + n + j] = kernel_only(d2row[n + j], h.s, h.l, kind) + h.sig2 - q;
        }
        __syncwarp();
    }
    if (threadIdx.x == 0) status[t] = (sh.jitter > 0.f) ? ADKF_ST_JITTER : ADKF_ST_OK;
}

// ---------------------------------------------------------------------------------------------
// Blocked dense kernels for the joint (support + query) covariance of gp_ift_kernel: N = n + m ~ 320 does not fit
// in shared memory, so the matrix stays in the (L2-resident) global workspace and is processed in 64 x 64 tiles that
// are staged in three shared-memory buffers.  Everything is built from one CTA-wide tile product
// C = beta C + alpha A B^T (r1c ncu + in-kernel clock64 probe: the unblocked thread-per-column code spent 35 M cycles
// = 18 ms per step in these three phases at ~1 FMA per clock).
// ---------------------------------------------------------------------------------------------
constexpr int kJb = 64;            // tile edge
constexpr int kJld = kJb + 1;      // padded leading dimension of a tile in shared memory
constexpr int kJTileFloats = kJb * kJld;

// dst[r][c] (or dst[c][r] when TR) <- src[(r0 + r) * ld + c0 + c], zero outside the N x N matrix
template <bool TR>
__device__ __forceinline__ void jt_load(float* dst, const float* src, int ld, int r0, int c0, int N) {
    for (int idx = threadIdx.x; idx < kJb * kJb; idx += blockDim.x) {
        const int r = idx >> 6, c = idx & 63;
        const float v = (r0 + r < N && c0 + c < N) ? src[(size_t)(r0 + r) * ld + c0 + c] : 0.f;
        if (TR) dst[c * kJld + r] = v; else dst[r * kJld + c] = v;
    }
}
// dst[(r0 + r) * ld + c0 + c] <- tile[r][c] (or tile[c][r] when TR), inside the N x N matrix only
template <bool TR>
__device__ __forceinline__ void jt_store(float* dst, int ld, int r0, int c0, int N, const float* tile) {
    for (int idx = threadIdx.x; idx < kJb * kJb; idx += blockDim.x) {
        const int r = idx >> 6, c = idx & 63;
        if (r0 + r < N && c0 + c < N) dst[(size_t)(r0 + r) * ld + c0 + c] = TR ? tile[c * kJld + r] : tile[r * kJld + c];
    }
}
// C = beta C + alpha A B^T on 64 x 64 shared-memory tiles, 512 threads: thread (ti, tj) owns rows 2ti, 2ti+1 and columns
// tj, tj+16, tj+32, tj+48 (A reads are two broadcasts per warp, B reads hit 16 distinct banks).  Caller synchronises.
__device__ __forceinline__ void jt_mma(float* C, const float* A, const float* B, float alpha, float beta) {
    const int ti = threadIdx.x >> 4, tj = threadIdx.x & 15;
    if (ti >= 32) return;
    const float* a0 = A + (2 * ti) * kJld;
    const float* a1 = a0 + kJld;
    const float* b0 = B + tj * kJld;
    float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll 8
    for (int k = 0; k < kJb; ++k) {
        const float x0 = a0[k], x1 = a1[k];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const float bv = b0[c * 16 * kJld + k];
            acc[0][c] = fmaf(x0, bv, acc[0][c]);
            acc[1][c] = fmaf(x1, bv, acc[1][c]);
        }
    }
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            float* o = C + (2 * ti + a) * kJld + tj + 16 * c;
            *o = (beta == 0.f) ? alpha * acc[a][c] : fmaf(beta, *o, alpha * acc[a][c]);
        }
}

// In-place blocked Cholesky of the lower triangle of J1 (N x N, leading dimension ld); the inverses of the diagonal
// tiles, needed for the panel solves anyway, are kept as the diagonal tiles of X = L^-1 in J2.  Returns false
// (uniformly) on a non-positive pivot.
__device__ long long g_jprof[8];
#define JP(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) { long long c_ = clock64(); g_jprof[i] += c_ - jp0; jp0 = c_; } } while (0)
__device__ bool joint_chol_blocked(float* J1, float* J2, int N, int ld, float* T0, float* T1, float* T2) {
    const int nb = (N + kJb - 1) / kJb;
    long long jp0 = clock64();
    for (int p = 0; p < nb; ++p) {
        const int r0 = p * kJb, nv = min(kJb, N - r0);
        jt_load<false>(T0, J1, ld, r0, r0, N);
        for (int idx = threadIdx.x; idx < kJTileFloats; idx += blockDim.x) T1[idx] = 0.f;
        __syncthreads();
        JP(0);
        const bool ok = chol_lower(T0, nv, kJld);
        __syncthreads();
        JP(1);
        if (!ok) return false;
        tri_inverse_cols<2>(T0, T1, nv, kJld);
        __syncthreads();
        JP(2);
        jt_store<false>(J1, ld, r0, r0, N, T0);
        jt_store<false>(J2, ld, r0, r0, N, T1);
        __syncthreads();
        for (int i = p + 1; i < nb; ++i) {                 // panel: L[ip] = A[ip] L[pp]^-T
            jt_load<false>(T2, J1, ld, i * kJb, r0, N);
            __syncthreads();
            jt_mma(T0, T2, T1, 1.f, 0.f);
            __syncthreads();
            jt_store<false>(J1, ld, i * kJb, r0, N, T0);
            __syncthreads();
        }
        JP(3);
        for (int i = p + 1; i < nb; ++i) {                 // trailing update: A[ij] -= L[ip] L[jp]^T
            jt_load<false>(T0, J1, ld, i * kJb, r0, N);
            for (int j = p + 1; j <= i; ++j) {
                jt_load<false>(T1, J1, ld, j * kJb, r0, N);
                jt_load<false>(T2, J1, ld, i * kJb, j * kJb, N);
                __syncthreads();
                jt_mma(T2, T0, T1, -1.f, 1.f);
                __syncthreads();
                jt_store<false>(J1, ld, i * kJb, j * kJb, N, T2);
                __syncthreads();
            }
        }
        JP(4);
    }
    return true;
}

// Off-diagonal tiles of X = L^-1 (lower): X[ij] = -X[ii] sum_{k=j}^{i-1} L[ik] X[kj]; L in J1, X in J2.
__device__ void joint_trtri_blocked(const float* J1, float* J2, int N, int ld, float* T0, float* T1, float* T2) {
    const int nb = (N + kJb - 1) / kJb;
    for (int j = 0; j < nb; ++j) {
        for (int i = j + 1; i < nb; ++i) {
            for (int k = j; k < i; ++k) {                  // T2 = S^T = sum_k X[kj]^T L[ik]^T
                jt_load<true>(T0, J2, ld, k * kJb, j * kJb, N);
                jt_load<false>(T1, J1, ld, i * kJb, k * kJb, N);
                __syncthreads();
                jt_mma(T2, T0, T1, 1.f, k == j ? 0.f : 1.f);
                __syncthreads();
            }
            jt_load<false>(T0, J2, ld, i * kJb, i * kJb, N);
            __syncthreads();
            jt_mma(T1, T0, T2, -1.f, 0.f);
            __syncthreads();
            jt_store<false>(J2, ld, i * kJb, j * kJb, N, T1);
            __syncthreads();
        }
    }
}

// P = X^T X (full symmetric) into J1 from the lower-triangular X in J2: P[ij] = sum_{k >= i} X[ki]^T X[kj], i >= j.
__device__ void joint_lauum_blocked(float* J1, const float* J2, int N, int ld, float* T0, float* T1, float* T2) {
    const int nb = (N + kJb - 1) / kJb;
    for (int i = 0; i < nb; ++i) {
        for (int j = 0; j <= i; ++j) {
            for (int k = i; k < nb; ++k) {
                jt_load<true>(T0, J2, ld, k * kJb, i * kJb, N);
                jt_load<true>(T1, J2, ld, k * kJb, j * kJb, N);
                __syncthreads();
                jt_mma(T2, T0, T1, 1.f, k == i ? 0.f : 1.f);
                __syncthreads();
            }
            jt_store<false>(J1, ld, i * kJb, j * kJb, N, T2);
            if (i != j) jt_store<true>(J1, ld, j * kJb, i * kJb, N, T2);
            __syncthreads();
        }
    }
}

// build_and_factor for the joint matrix on the blocked path (same psd_safe_cholesky retry policy)
__device__ bool build_and_factor_blocked(const GpProblem& pb, const Hyper& h, float* J1, float* J2, int ld, GpShared& sh,
                                         float* T0, float* T1, float* T2) {
    const int n = pb.n;
    const float jit[4] = {0.f, 1e-6f, 1e-5f, 1e-4f};
    for (int a = 0; a < 4; ++a) {
        for (int idx = threadIdx.x; idx < n * n; idx += blockDim.x) {
            const int i = idx / n, j = idx - i * n;
            float k = kernel_only(pb.D2[(size_t)i * pb.ldd + j], h.s, h.l, pb.kind);
            if (i == j) k += h.sig2 + jit[a];
            J1[(size_t)i * ld + j] = k;
        }
        __syncthreads();
        const bool ok = joint_chol_blocked(J1, J2, n, ld, T0, T1, T2);
        __syncthreads();
        if (ok) {
            if (threadIdx.x == 0) sh.jitter = jit[a];
            __syncthreads();
            return true;
        }
    }
    return false;
}

// IFT correction on the support block:  Wm_ss += sign (1/n) [ Gamdot o Ku + Gam o Kudot ], the derivative of df/dD2 along the
// psi-space direction sh.pd (= t o raw direction).  sign -1: hypergradient (direction v = H^-1 dF/dphi, App. B);
// sign +1: the mixed second derivative d/deps df/dD2(phi + eps u) for a caller-supplied u (mode 3: double backward of f).
//   ddot (mode 4, else false): Wt holds on entry a direction dD2 of the pair statistics themselves (support block); then
//   Adot += Ku o dD2 and Kudot += Kuu o dD2, and sums[0..1] receive sum_ij gamdot_ij delta_ij and
//   sum_ij (gamdot_ij K_ij + gam_ij Kdot_ij) -- the pieces of d/deps of the (noise, outputscale) gradient components.
__device__ void ift_correction(const Carve& cv, int n, int N, const float* __restrict__ D2t, const Hyper& h, int kind,
                               float* __restrict__ Wt, const float* __restrict__ alpha, GpShared& sh, bool in_smem,
                               float* smem_f, float* small_ws, float sign, bool ddot = false) {
    const float pd0 = sh.pd[0], pd1 = sh.pd[1], pd2 = sh.pd[2];
    // M2 <- Adot = pd0 I + pd1 K/s + pd2 S (+ Ku o dD2)   (M2 currently holds S)
    for (int idx = threadIdx.x; idx < n * n; idx += blockDim.x) {
        const int i = idx / n, j = idx - i * n;
        const float u = D2t[(size_t)i * N + j];
        const float k = kernel_only(u, h.s, h.l, kind);
        float v = pd1 * k / h.s + pd2 * cv.M2[(size_t)i * cv.ld + j];
        if (ddot) v += kernel_parts(u, h.s, h.l, kind).Ku * Wt[(size_t)i * N + j];
        if (i == j) v += pd0;
        cv.M2[(size_t)i * cv.ld + j] = v;
    }
    __syncthreads();
    float dsum[2] = {0.f, 0.f};
    // one warp per row i: X_i = P_i: Adot ; Pdot_i: = -X_i P ; adot_i = -X_i . alpha
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    float* adot = cv.vec + n;           // u, w, q no longer needed
    float* xrow_all = cv.vec + 2 * n;   // needs nw * n floats: placed after vec in the workspace
    // first pass: adot for every row (needs full X rows) -- compute X row into per-warp buffer
    float* xbuf = (in_smem ? (smem_f + (size_t)2 * n * cv.ld + 4 * (size_t)n) : (small_ws + (size_t)2 * n * cv.ld + 4 * (size_t)n));
    (void)xrow_all;
    for (int i = wid; i < n; i += nw) {
        float* xr = xbuf + (size_t)wid * n;
        const float* Pi = cv.M1 + (size_t)i * cv.ld;
        float ad = 0.f;
        for (int j = lane; j < n; j += 32) {
            float acc = 0.f;
            for (int k = 0; k < n; ++k) acc += Pi[k] * cv.M2[(size_t)k * cv.ld + j];
            xr[j] = acc;
            ad += acc * alpha[j];
        }
        ad = warp_sum(ad);
        if (lane == 0) adot[i] = -ad;
        __syncwarp();
    }
    __syncthreads();
    for (int i = wid; i < n; i += nw) {
        float* xr = xbuf + (size_t)wid * n;
        const float* Pi = cv.M1 + (size_t)i * cv.ld;
        // recompute X row (cheaper than storing an n x n X)
        for (int j = lane; j < n; j += 32) {
            float acc = 0.f;
            for (int k = 0; k < n; ++k) acc += Pi[k] * cv.M2[(size_t)k * cv.ld + j];
            xr[j] = acc;
        }
        __syncwarp();
        for (int j = lane; j < n; j += 32) {
            float pdot = 0.f;
            for (int k = 0; k < n; ++k) pdot += xr[k] * cv.M1[(size_t)k * cv.ld + j];
            pdot = -pdot;
            const float gamdot = 0.5f * (pdot - adot[i] * alpha[j] - alpha[i] * adot[j]);
            const float gam = 0.5f * (Pi[j] - alpha[i] * alpha[j]);
            const float u = D2t[(size_t)i * N + j];
            const KParts kp = kernel_parts(u, h.s, h.l, kind);
            float kudot = pd1 * kp.Ku / h.s + pd2 * kp.Kul;
            if (ddot) {
                const float dd = Wt[(size_t)i * N + j];
                kudot += kernel_uu(u, h.s, h.l, kind) * dd;
                const float kdot = pd1 * kp.K / h.s + pd2 * kp.Kl + kp.Ku * dd;
                if (i == j) dsum[0] += gamdot;
                dsum[1] += gamdot * kp.K + gam * kdot;
                Wt[(size_t)i * N + j] = (gamdot * kp.Ku + gam * kudot) / n;
            } else {
                Wt[(size_t)i * N + j] += sign * (gamdot * kp.Ku + gam * kudot) / n;
            }
        }
        __syncwarp();
    }
    if (ddot) {
        __syncthreads();
        block_sum_multi<2>(dsum, sh.red, sh.sums);
    }
}

// =============================================================================================
// kernel 3: IFT hypergradient pieces (F, dF/dphi, H, v, Wm)
// =============================================================================================
__global__ void __launch_bounds__(kGpThreads, 1)
gp_ift_kernel(const int32_t* __restrict__ z_off, const int32_t* __restrict__ n_sup,
              const int64_t* __restrict__ d2_off, const float* __restrict__ D2, const float* __restrict__ y,
              int kind, float noise_prior_loc, int ls_prior, int ignore_corr, const float* __restrict__ phi,
              float* __restrict__ aux, float* __restrict__ Wm, int32_t* __restrict__ status,
              float* __restrict__ ws, const int64_t* __restrict__ ws_off) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    GpShared& sh = *reinterpret_cast<GpShared*>(smem_raw);
    float* smem_f = reinterpret_cast<float*>(smem_raw + ((sizeof(GpShared) + 15) / 16) * 16);
    const int t = blockIdx.x;
    const bool prof = (ignore_corr & 256) != 0 && blockIdx.x == 0 && threadIdx.x == 0;   // development probe
    ignore_corr &= 255;
    long long tk[8];
    if (prof) tk[0] = clock64();
    const int N = z_off[t + 1] - z_off[t];
    const float* D2t = D2 + d2_off[t];
    float* Wt = Wm + d2_off[t];
    const float* yt = y + z_off[t];
    GpProblem pb;
    pb.n = n_sup[t];
    pb.ldd = N;
    pb.D2 = D2t;
    pb.y = yt;
    pb.free_mask = free_mask_of(kind);
    kind &= 0xff;
    pb.kind = kind;
    pb.noise_loc = noise_prior_loc;
    float* auxt = aux + (size_t)t * ADKF_AUX;
    if (!kind_is_stationary(kind)) ls_prior = 0;
    pb.has_lsp = ls_prior != 0;
    pb.ls_loc = auxt[ADKF_AUX_LS_LOC];
    const int n = pb.n;
    int st = ADKF_ST_OK;
    // workspace: [J1 N x ldN][J2 N x ldN][beta N][small n-matrices if they do not fit in smem]
    const int ldN = N | 1;
    if (ws_off[gridDim.x] < 0) { if (threadIdx.x == 0) status[t] = ADKF_ST_BADWS; return; }
    float* wst = ws + ws_off[t];
    float* J1 = wst;
    float* J2 = wst + (size_t)N * ldN;
    float* beta = wst + (size_t)2 * N * ldN;
    float* small_ws = beta + N;
    bool in_smem;
    Carve cv = carve_workspace(n, smem_f, small_ws, in_smem);
    if (threadIdx.x < 3) sh.phi[threadIdx.x] = phi[t * 3 + threadIdx.x];
    __syncthreads();
    const Hyper h = make_hyper(sh.phi);

    // ---- inner quantities at phi: P (M1), S (M2), alpha, H ------------------------------------
    bool ok = gp_eval_full(pb, sh.phi, cv.M1, cv.M2, cv.ld, cv.vec, sh, true);
    const float jit_inner = sh.jitter;
    float* alpha = cv.vec;
    if (!ok) {
        // the reference raises here (gpytorch NotPSDError); the batched launch flags the task, gives it a zero gradient and a
        // NaN loss so that the caller can drop it from the meta-batch mean (adaptive_dkt_utils.train_step)
        if (threadIdx.x == 0) { status[t] = ADKF_ST_NOT_PD; auxt[ADKF_AUX_FOUTER] = nanf(""); auxt[ADKF_AUX_F] = nanf(""); }
        for (int idx = threadIdx.x; idx < N * N; idx += blockDim.x) Wt[idx] = 0.f;
        return;
    }
    const float nll_inner = sh.nll;
    if (prof) tk[1] = clock64();

    if (ignore_corr == 2) {
        // inner loss only: Wm_ss = df/dD2 = Gamma o Ku / n, zero elsewhere (first-order autograd of f = -MLL)
        for (int idx = threadIdx.x; idx < N * N; idx += blockDim.x) {
            const int a = idx / N, b = idx - a * N;
            float w = 0.f;
            if (a < n && b < n) {
                const KParts kp = kernel_parts(D2t[idx], h.s, h.l, kind);
                w = 0.5f * (cv.M1[(size_t)a * cv.ld + b] - alpha[a] * alpha[b]) * kp.Ku / n;
            }
            Wt[idx] = w;
        }
        if (threadIdx.x == 0) {
            auxt[ADKF_AUX_F] = sh.f;
            for (int i = 0; i < 3; ++i) auxt[ADKF_AUX_G + i] = sh.g[i];
            for (int i = 0; i < 9; ++i) auxt[ADKF_AUX_H + i] = sh.H[i];
            status[t] = (jit_inner > 0.f) ? ADKF_ST_JITTER : ADKF_ST_OK;
        }
        return;
    }

    if (ignore_corr == 3) {
        // mixed second derivative of the inner loss along the raw-space direction u = aux[V..V+2]:
        // Wm_ss = d/deps df/dD2(phi + eps u), zero elsewhere; aux F / G / H as in mode 2 (double backward of f = -MLL:
        // H u is formed by the caller from aux H)
        for (int idx = threadIdx.x; idx < N * N; idx += blockDim.x) Wt[idx] = 0.f;
        if (threadIdx.x == 0) {
            for (int i = 0; i < 3; ++i) sh.pd[i] = (pb.free_mask & (1 << i)) ? h.t[i] * auxt[ADKF_AUX_V + i] : 0.f;
            auxt[ADKF_AUX_F] = sh.f;
            for (int i = 0; i < 3; ++i) auxt[ADKF_AUX_G + i] = sh.g[i];
            for (int i = 0; i < 9; ++i) auxt[ADKF_AUX_H + i] = sh.H[i];
            status[t] = (jit_inner > 0.f) ? ADKF_ST_JITTER : ADKF_ST_OK;
        }
        __syncthreads();
        ift_correction(cv, n, N, D2t, h, kind, Wt, alpha, sh, in_smem, smem_f, small_ws, 1.f);
        return;
    }
    if (ignore_corr == 4) {
        // ARD building block: directional derivative of the inner gradient along (u_noise, u_outputscale) = aux[V..V+1] AND a
        // direction dD2 of the pair statistics (support block of Wm on entry).  Out: Wm_ss = d/deps df/dD2, aux[DF_DPHI..+1] =
        // d/deps of the raw (noise, outputscale) gradient components.  The caller contracts Wm_ss with the per-dimension pair
        // differences for the lengthscale components (adkf_ard_pairquad): no (d+2)^2 Hessian is ever formed.
        if (threadIdx.x == 0) {
            sh.pd[0] = (pb.free_mask & 1) ? h.t[0] * auxt[ADKF_AUX_V + 0] : 0.f;
            sh.pd[1] = (pb.free_mask & 2) ? h.t[1] * auxt[ADKF_AUX_V + 1] : 0.f;
            sh.pd[2] = 0.f;
        }
        __syncthreads();
        ift_correction(cv, n, N, D2t, h, kind, Wt, alpha, sh, in_smem, smem_f, small_ws, 1.f, true);
        if (threadIdx.x == 0) {
            // g_raw = t o g_psi:  d g_raw = t dg_psi + c u g_psi.   n g_psi_noise = tr(Gamma) - dlogp/dsig2,
            // n g_psi_s = sum(Gamma o K) / s
            float lp = 0.f, d1 = 0.f, d2 = 0.f;
            if (!isnan(pb.noise_loc)) lognormal_terms(h.sig2, pb.noise_loc, lp, d1, d2);
            const float u0 = auxt[ADKF_AUX_V + 0], u1 = auxt[ADKF_AUX_V + 1];
            const float gpsi0 = sh.g_psi[0], gpsi1 = sh.g_psi[1];
            const float dg0 = ((float)sh.sums[0] - d2 * sh.pd[0]) / n;
            const float dg1 = ((float)sh.sums[1] / h.s - (gpsi1 * n) * sh.pd[1] / h.s) / n;
            auxt[ADKF_AUX_DF_DPHI + 0] = (pb.free_mask & 1) ? h.t[0] * dg0 + h.c[0] * u0 * gpsi0 : 0.f;
            auxt[ADKF_AUX_DF_DPHI + 1] = (pb.free_mask & 2) ? h.t[1] * dg1 + h.c[1] * u1 * gpsi1 : 0.f;
            auxt[ADKF_AUX_F] = sh.f;
            for (int i = 0; i < 3; ++i) auxt[ADKF_AUX_G + i] = sh.g[i];
            status[t] = (jit_inner > 0.f) ? ADKF_ST_JITTER : ADKF_ST_OK;
        }
        return;
    }

    // ---- joint covariance J = K + sig2 I over support+query: Cholesky, inverse, beta --------------
    // blocked path when three 64 x 64 tiles fit behind the inner problem's shared-memory working set
    const size_t smem_used = in_smem ? ((size_t)2 * n * cv.ld + 4 * (size_t)n + (size_t)(blockDim.x >> 5) * n) : 0;
    const bool blocked = smem_used + 3 * (size_t)kJTileFloats <= (size_t)(kSmemBudgetFloats + kGpThreads / 32 * 256);
    float* T0 = smem_f + smem_used;
    float* T1 = T0 + kJTileFloats;
    float* T2 = T1 + kJTileFloats;
    {
        GpProblem pj = pb;
        pj.n = N;
        const bool okj = blocked ? build_and_factor_blocked(pj, h, J1, J2, ldN, sh, T0, T1, T2)
                                 : build_and_factor(pj, h, J1, ldN, sh);
        if (!okj) {
            if (threadIdx.x == 0) { status[t] = ADKF_ST_NOT_PD; auxt[ADKF_AUX_FOUTER] = nanf(""); }
            for (int idx = threadIdx.x; idx < N * N; idx += blockDim.x) Wt[idx] = 0.f;
            return;
        }
    }
    float part1[1] = {0.f};
    for (int i = threadIdx.x; i < N; i += blockDim.x) part1[0] += logf(J1[(size_t)i * ldN + i]);
    block_sum_multi<1>(part1, sh.red, sh.sums);
    const float logdetJ = 2.f * (float)sh.sums[0];
    if (prof) tk[2] = clock64();
    if (blocked) joint_trtri_blocked(J1, J2, N, ldN, T0, T1, T2);
    else tri_inverse(J1, J2, N, ldN);
    __syncthreads();
    if (prof) tk[3] = clock64();
    if (blocked) joint_lauum_blocked(J1, J2, N, ldN, T0, T1, T2);
    else xtx(J2, J1, N, ldN);   // J1 <- J^-1
    __syncthreads();
    if (prof) tk[4] = clock64();
    matvec(J1, N, ldN, yt, beta);
    __syncthreads();

    // ---- F and dF/dpsi = sum_ab G_ab dJ_ab/dpsi ;  G = 0.5 (J^-1 - beta beta^T) - pad(Gamma) ----------
    float part[4] = {0.f, 0.f, 0.f, 0.f};  // yb, trG, sum G*K/s, sum G*Kl
    for (int idx = threadIdx.x; idx < N * N; idx += blockDim.x) {
        const int a = idx / N, b = idx - a * N;
        float G = 0.5f * (J1[(size_t)a * ldN + b] - beta[a] * beta[b]);
        if (a < n && b < n) G -= 0.5f * (cv.M1[(size_t)a * cv.ld + b] - alpha[a] * alpha[b]);
        const KParts kp = kernel_parts(D2t[(size_t)a * N + b], h.s, h.l, kind);
        part[2] += G * kp.K;
        part[3] += G * kp.Kl;
        if (a == b) part[1] += G;
        Wt[idx] = G * kp.Ku;     // direct term dF/dD2
    }
    for (int i = threadIdx.x; i < N; i += blockDim.x) part[0] += yt[i] * beta[i];
    block_sum_multi<4>(part, sh.red, sh.sums);
    if (threadIdx.x == 0) {
        const float nllJ = (float)(0.5 * sh.sums[0] + 0.5 * logdetJ + N * (double)kHalfLog2Pi);
        auxt[ADKF_AUX_FOUTER] = nllJ - nll_inner;
        sh.dF[0] = h.t[0] * (float)sh.sums[1];
        sh.dF[1] = h.t[1] * (float)(sh.sums[2] / h.s);
        sh.dF[2] = h.t[2] * (float)sh.sums[3];
        for (int i = 0; i < 3; ++i) if (!(pb.free_mask & (1 << i))) sh.dF[i] = 0.f;   // frozen: not an inner variable
        bool solved = solve3_general(sh.H, sh.dF, sh.v);
        sh.flag = 0;
        if (!solved) { sh.v[0] = sh.v[1] = sh.v[2] = 0.f; sh.flag = 1; }
        for (int i = 0; i < 3; ++i) {
            sh.pd[i] = h.t[i] * sh.v[i];
            auxt[ADKF_AUX_DF_DPHI + i] = sh.dF[i];
            auxt[ADKF_AUX_V + i] = sh.v[i];
            auxt[ADKF_AUX_G + i] = sh.g[i];
        }
        for (int i = 0; i < 9; ++i) auxt[ADKF_AUX_H + i] = sh.H[i];
        auxt[ADKF_AUX_F] = sh.f;
    }
    __syncthreads();
    if (sh.flag && ignore_corr == 0) {
        // singular inner Hessian: the reference's torch.linalg.solve raises (cauchy_hypergradient.py:136).  Dropping the IFT
        // correction silently would hand back a different gradient, so the task is flagged and taken out like a non-PD one.
        if (threadIdx.x == 0) { status[t] = ADKF_ST_HESS_DAMPED; auxt[ADKF_AUX_FOUTER] = nanf(""); }
        for (int idx = threadIdx.x; idx < N * N; idx += blockDim.x) Wt[idx] = 0.f;
        return;
    }
    if (prof) tk[5] = clock64();

    // ---- IFT correction on the support block:  Wm_ss -= (1/n) [ Gamdot o Ku + Gam o Kudot ] ---------
    if (ignore_corr == 0) ift_correction(cv, n, N, D2t, h, kind, Wt, alpha, sh, in_smem, smem_f, small_ws, -1.f);
    if (prof) {
        tk[6] = clock64();
        printf("joint chol kcycles: load %lld chol_tile %lld inv_tile %lld store+panel %lld trailing %lld\n", g_jprof[0] / 1000,
               g_jprof[1] / 1000, g_jprof[2] / 1000, g_jprof[3] / 1000, g_jprof[4] / 1000);
        for (int i_ = 0; i_ < 8; ++i_) g_jprof[i_] = 0;
        printf("gp_ift n=%d N=%d kcycles: inner %lld  joint chol %lld  tri_inverse %lld  xtx %lld  G+sums %lld  ift corr %lld\n", n, N,
               (tk[1] - tk[0]) / 1000, (tk[2] - tk[1]) / 1000, (tk[3] - tk[2]) / 1000, (tk[4] - tk[3]) / 1000,
               (tk[5] - tk[4]) / 1000, (tk[6] - tk[5]) / 1000);
    }
    if (threadIdx.x == 0) {
        if (jit_inner > 0.f || sh.jitter > 0.f) st |= ADKF_ST_JITTER;
        status[t] = st;
    }
}

}  // namespace adkf

// =============================================================================================
// host side
// =============================================================================================
using namespace adkf;

static size_t gp_smem_bytes() { return ((sizeof(GpShared) + 15) / 16) * 16 + (size_t)(kSmemBudgetFloats + kGpThreads / 32 * 256) * sizeof(float); }

static int64_t small_ws_floats(int n) {
    const int ld = n | 1;
    return (int64_t)2 * n * ld + 4 * (int64_t)n + (int64_t)(kGpThreads / 32) * n;
}

extern "C" int64_t adkf_gp_ws_floats(int32_t T, const int32_t* z_off, const int32_t* n_sup, int32_t which) {
    // which: 0 fit, 1 predict, 2 ift.  Layout: [T+1 int64 offsets (as 2 floats each)] then per-task slices.
    int64_t total = 2 * (int64_t)(T + 1);
    for (int t = 0; t < T; ++t) {
        const int N = z_off[t + 1] - z_off[t];
        int64_t need = small_ws_floats(n_sup[t]);
        if (which == 2) need += (int64_t)2 * N * (N | 1) + N;
        total += (need + 3) & ~(int64_t)3;
    }
    return total;
}

// The per-task workspace offsets are computed on the device from z_off / n_sup (no host sync, no H2D).
__global__ void gp_ws_offsets_kernel(int T, const int32_t* z_off, const int32_t* n_sup, int which, int64_t* off, int64_t ws_floats) {
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        int64_t acc = 2 * (int64_t)(T + 1);
        for (int t = 0; t < T; ++t) {
            off[t] = acc;
            const int N = z_off[t + 1] - z_off[t];
            const int n = n_sup[t];
            int64_t need = (int64_t)2 * n * (n | 1) + 4 * (int64_t)n + (int64_t)(kGpThreads / 32) * n;
            if (which == 2) need += (int64_t)2 * N * (N | 1) + N;
            acc += (need + 3) & ~(int64_t)3;
        }
        // the caller's workspace is checked ON THE DEVICE (z_off / n_sup are device arrays): an undersized or stale buffer flags
        // every task with ADKF_ST_BADWS instead of writing out of bounds
        off[T] = (acc <= ws_floats) ? acc : -1;
    }
}

static int set_smem_attr(const void* fn) {
    static bool done[3] = {false, false, false};
    (void)done;
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gp_smem_bytes());
    return (int)e;
}

extern "C" int adkf_gp_fit(int32_t T, const int32_t* z_off, const int32_t* n_sup, const int64_t* d2_off,
                           const float* D2, const float* y, int32_t kernel, float noise_init,
                           float noise_prior_loc, int32_t ls_prior, int32_t init_mode, int32_t max_iter,
                           float gtol, float* phi_io, float* aux, int32_t* status, float* ws,
                           int64_t ws_floats, void* stream) {
    if (T <= 0) return 0;
    if (!z_off || !n_sup || !d2_off || !D2 || !y || !phi_io || !aux || !status || !ws) return ADKF_ERR_BADARG;
    if ((kernel & 0xff) < 0 || (kernel & 0xff) > ADKF_KERNEL_TANIMOTO) return ADKF_ERR_BADARG;
    cudaStream_t s = (cudaStream_t)stream;
    int e = set_smem_attr((const void*)gp_fit_kernel);
    if (e) return e;
    int64_t* off = reinterpret_cast<int64_t*>(ws);
    if (ws_floats < 2 * ((int64_t)T + 1)) return ADKF_ERR_BADARG;
    gp_ws_offsets_kernel<<<1, 1, 0, s>>>(T, z_off, n_sup, 0, off, ws_floats);
    gp_fit_kernel<<<T, kGpThreads, gp_smem_bytes(), s>>>(z_off, n_sup, d2_off, D2, y, kernel, noise_init,
                                                        noise_prior_loc, ls_prior, init_mode, max_iter, gtol,
                                                        phi_io, aux, status, ws, off);
    adkf_count_launches(2);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_gp_predict(int32_t T, const int32_t* z_off, const int32_t* n_sup, const int64_t* d2_off,
                               const float* D2, const float* y, int32_t kernel, const float* phi, float* mean,
                               float* var, int32_t* status, float* ws, int64_t ws_floats, void* stream) {
    if (T <= 0) return 0;
    if (!z_off || !n_sup || !d2_off || !D2 || !y || !phi || !mean || !var || !status || !ws) return ADKF_ERR_BADARG;
    cudaStream_t s = (cudaStream_t)stream;
    int e = set_smem_attr((const void*)gp_predict_kernel);
    if (e) return e;
    int64_t* off = reinterpret_cast<int64_t*>(ws);
    if (ws_floats < 2 * ((int64_t)T + 1)) return ADKF_ERR_BADARG;
    gp_ws_offsets_kernel<<<1, 1, 0, s>>>(T, z_off, n_sup, 1, off, ws_floats);
    gp_predict_kernel<<<T, kGpThreads, gp_smem_bytes(), s>>>(z_off, n_sup, d2_off, D2, y, kernel, phi, mean, var,
                                                            status, ws, off);
    adkf_count_launches(2);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_gp_ift(int32_t T, const int32_t* z_off, const int32_t* n_sup, const int64_t* d2_off,
                           const float* D2, const float* y, int32_t kernel, float noise_prior_loc,
                           int32_t ls_prior, int32_t mode, const float* phi, float* aux,
                           float* Wm, int32_t* status, float* ws, int64_t ws_floats, void* stream) {
    if (T <= 0) return 0;
    if (!z_off || !n_sup || !d2_off || !D2 || !y || !phi || !aux || !Wm || !status || !ws) return ADKF_ERR_BADARG;
    if ((mode & 255) < 0 || (mode & 255) > 4) return ADKF_ERR_BADARG;   // bit 8: development probe (phase timings)
    cudaStream_t s = (cudaStream_t)stream;
    int e = set_smem_attr((const void*)gp_ift_kernel);
    if (e) return e;
    int64_t* off = reinterpret_cast<int64_t*>(ws);
    if (ws_floats < 2 * ((int64_t)T + 1)) return ADKF_ERR_BADARG;
    gp_ws_offsets_kernel<<<1, 1, 0, s>>>(T, z_off, n_sup, 2, off, ws_floats);
    gp_ift_kernel<<<T, kGpThreads, gp_smem_bytes(), s>>>(z_off, n_sup, d2_off, D2, y, kernel, noise_prior_loc,
                                                        ls_prior, mode, phi, aux, Wm, status,
                                                        ws, off);
    adkf_count_launches(2);
    ADKF_LAUNCH_CHECK();
    return 0;
}
