// gemm_common.cuh -- pieces shared by the tcgen05 3xTF32 GEMM kernels (gemm_tf32x3.cu: one CTA per 128 x 128 tile, also the
// per-task Gram; gemm_pair.cu: cta_group::2 CTA pairs on 256 x 128 tiles): PTX wrappers, the epilogue descriptor, the fp32 ->
// (hi, lo) split and the coalesced tile store.
#pragma once
#include <cuda.h>
#include "common.cuh"
#include "../../include/adkf_b200.h"

namespace adkf {

constexpr int kBM = 128, kBN = 128, kBK = 32;        // tile; 32 floats = 128 B = one swizzle row
constexpr int kTileBytes = kBM * kBK * 4;            // 16 KB

// ---- PTX wrappers ---------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0, spins = 0;
    while (true) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done)
            : "r"(bar), "r"(parity)
            : "memory");
        if (done) break;
        if (++spins > (1u << 28)) __trap();   // never hang the GPU on a protocol error
    }
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
        : "memory");
}
// A operand in tensor memory (lane = row of the tile, one tf32 per 32-bit column, 8 columns per MMA), B through its
// shared-memory descriptor
__device__ __forceinline__ void umma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(acc)
        : "memory");
}
// tcgen05.mma / tcgen05.commit issued by ONE elected lane of a CONVERGED warp (elect.sync inside the statement; every lane executes
// the call).  Under `if (lane == 0)` ptxas cannot prove the issuing thread unique and wraps each tcgen05.mma in an ELECT + vote +
// branch loop: ~110 cycles of issue per MMA measured on the pair kernel (gemm_pair.cu), more than its tensor-pipe time.
__device__ __forceinline__ void umma_tf32_ts_e(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p, q;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(acc)
        : "memory");
}
__device__ __forceinline__ void umma_commit_e(uint32_t bar) {
    asm volatile(
        "{\n\t.reg .pred q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}"
        ::"r"(bar)
        : "memory");
}
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,"
        "%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
          "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]), "r"(r[19]),
          "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]), "r"(r[28]), "r"(r[29]),
          "r"(r[30]), "r"(r[31])
        : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// Register budget per warpgroup (setmaxnreg): the kernel starts with 96 registers per thread (640 threads); the control
// and B-splitter warpgroups give registers back, the two drain warpgroups (64 accumulators per thread) take them.
template <int N> __device__ __forceinline__ void reg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N> __device__ __forceinline__ void reg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// K-major, SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start address >> 4, leading byte
// offset 1 (unused for swizzled K-major), stride byte offset 1024 B (one 8-row swizzle atom), version 1, layout 2.
__device__ __forceinline__ uint64_t smem_desc_k128(uint32_t addr) {
    return (uint64_t)((addr >> 4) & 0x3FFF) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor: D = f32 (1<<4), A = B = tf32 (2<<7, 2<<10), both K-major, N>>3 at bit 17, M>>4 at bit 24
constexpr uint32_t kIdescTf32 = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(kBN >> 3) << 17) | ((uint32_t)(kBM >> 4) << 24);

__device__ __forceinline__ uint32_t rna_tf32(float x) {
    uint32_t u;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
    return u;
}

struct GemmEpilogue {
    // mode 0: Y = act(acc + bias) (or a split-K partial)       mode 1: per-task squared distances
    int mode;
    float* out;
    int64_t ldo;
    int64_t M;
    int N, K;
    const float* bias;
    int act;              // 0 none, 1 relu, 2 leaky_relu(0.01), 3 subtract the product from what `out` holds,
                          // 4 / 5: multiply by relu' / leaky_relu'(0.01) of mask[row, col] (the saved activation output of the
                          // layer below: the input-gradient GEMM of layer l+1 emits dZ_l directly, fs_mol/modules/gnn.py:79-95)
    const float* mask;
    int64_t ldm;
    // split-K (mode 0): CTA z handles k-blocks [z * kb_per_split, ...) and writes to out + z * split_stride
    int kb_per_split;
    int64_t split_stride;
    // batching (mode 0): blockIdx.z = batch * splits + split.  Batch b reads A at (mn + b*a_mn, k + b*a_k), B likewise,
    // writes D + b*d_off (+ bias + b*bias_off): column blocks / row blocks of the SAME 2-D arrays, one tensor map each.
    int splits;
    int gx, gy, ntiles;   // persistent tile walk: tile -> (bx = tile % gx, by = (tile / gx) % gy, bz = tile / (gx gy))
    const int32_t* gate;  // optional device word: the launch is a no-op when *gate == 0 (device-side jitter ladder of the blocked
                          // Cholesky: later attempts run only if the previous one failed, without a host round trip)
    int tma_store;        // pair kernel: the output tile leaves through TMA bulk stores (see pair_store_tile_tma); else gemm_store_tile
    int debug;            // development probe (ADKF_GEMM_DEBUG): 1 = splitter skips the split (128 only A, 256 only B), 2 = no MMAs
    int a_mn, a_k, b_mn, b_k;
    int64_t d_off, bias_off;
    // gram
    const int32_t* z_off;
    const int32_t* n_sup;
    const int64_t* d2_off;
    const float* mu;      // [T, d]
    const float* nrm;     // [rows]
    int qq_block;
    int stat;             // ADKF_STAT_*
};

// hi = v rounded to tf32 (round-to-nearest, ties away: what cvt.rna.tf32.f32 does, written as two integer ops so that it
// stays on the ALU pipe), lo = v - hi (exact in fp32)
__device__ __forceinline__ float tf32_hi(float v) { return __uint_as_float((__float_as_uint(v) + 0x1000u) & 0xffffe000u); }
__device__ __forceinline__ void split4(const float4 v, float4& hi, float4& lo) {
    hi.x = tf32_hi(v.x); lo.x = v.x - hi.x;
    hi.y = tf32_hi(v.y); lo.y = v.y - hi.y;
    hi.z = tf32_hi(v.z); lo.z = v.z - hi.z;
    hi.w = tf32_hi(v.w); lo.w = v.w - hi.w;
}

// Epilogue of one 128 x 128 output tile held by the 8 drain warps of a CTA (warp w: TMEM lane quarter q = w & 3, column half
// colbase = 0 | 64; acc[j] = row q*32+lane, column colbase+j).  mode 0 of GemmEpilogue.
// Through a warp-private swizzled smem tile (32 rows x 32 columns per pass) so that every global store
// instruction writes 4 rows x 128 contiguous bytes (the lane-per-row layout of tcgen05.ld would touch 32
// different lines with 16 bytes each).  Bias and activation are applied AFTER the transpose, where a lane
// owns 4 fixed columns: one float4 bias load per pass instead of 64 scalar loads per row.
__device__ __forceinline__ void gemm_store_tile(const GemmEpilogue& ep, const int rowA0, const int rowB0, const int64_t zslab,
                                                const int batch, const float (&acc)[64], const int q, const int lane,
                                                const int colbase, float4* __restrict__ stg) {
    struct { int rowA0, rowB0, batch; int64_t zslab; } c = {rowA0, rowB0, batch, zslab};
    float* out = ep.out + c.zslab * ep.split_stride + (int64_t)c.batch * ep.d_off;
    const float* bias = (ep.bias != nullptr && !(ep.debug & 8)) ? ep.bias + (int64_t)c.batch * ep.bias_off : nullptr;
    const bool vec_ok = (ep.ldo % 4 == 0) && ((reinterpret_cast<uintptr_t>(out) & 15) == 0);
    const int ch = lane & 7, rsub = lane >> 3;
    const int act = ep.act;
#pragma unroll
    for (int p = 0; p < 2; ++p) {
        // act 4 / 5: the 8 mask vectors this lane needs for the pass are requested BEFORE the staging traffic, so
        // their latency hides behind it (a load inside the store loop costs one round trip per store)
        float4 mk4[8];
        if (act >= 4) {
            const int mcol = c.rowB0 + colbase + p * 32 + ch * 4;
            const int64_t mrow = (int64_t)c.rowA0 + q * 32 + rsub;
            const bool mvec = (ep.ldm % 4 == 0) && ((reinterpret_cast<uintptr_t>(ep.mask) & 15) == 0) && mcol + 4 <= ep.N;
#pragma unroll
            for (int jj = 0; jj < 8; ++jj) {
                const int64_t r = mrow + jj * 4;
                mk4[jj] = make_float4(1.f, 1.f, 1.f, 1.f);
                if (r < ep.M && mcol < ep.N) {
                    const float* mk = ep.mask + r * ep.ldm + mcol;
                    if (mvec) {
                        mk4[jj] = __ldg(reinterpret_cast<const float4*>(mk));
                    } else {
                        mk4[jj].x = mk[0];
                        if (mcol + 1 < ep.N) mk4[jj].y = mk[1];
                        if (mcol + 2 < ep.N) mk4[jj].z = mk[2];
                        if (mcol + 3 < ep.N) mk4[jj].w = mk[3];
                    }
                }
            }
        }
#pragma unroll
        for (int j = 0; j < 8; ++j)
            stg[lane * 8 + (j ^ (lane & 7))] =
                make_float4(acc[p * 32 + 4 * j], acc[p * 32 + 4 * j + 1], acc[p * 32 + 4 * j + 2], acc[p * 32 + 4 * j + 3]);
        __syncwarp();
        const int col = c.rowB0 + colbase + p * 32 + ch * 4;
        const bool full = vec_ok && col + 4 <= ep.N;
        float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
        if (bias != nullptr) {
            if (col + 4 <= ep.N && ((reinterpret_cast<uintptr_t>(bias + col) & 15) == 0)) {
                b4 = *reinterpret_cast<const float4*>(bias + col);
            } else {
                if (col < ep.N) b4.x = bias[col];
                if (col + 1 < ep.N) b4.y = bias[col + 1];
                if (col + 2 < ep.N) b4.z = bias[col + 2];
                if (col + 3 < ep.N) b4.w = bias[col + 3];
            }
        }
        const int64_t rbase = (int64_t)c.rowA0 + q * 32 + rsub;
        float* o = out + rbase * ep.ldo + col;
        const int64_t ostep = 4 * ep.ldo;
#pragma unroll
        for (int jj = 0; jj < 8; ++jj, o += ostep) {
            const int rr = jj * 4 + rsub;
            float4 x = stg[rr * 8 + (ch ^ (rr & 7))];
            x.x += b4.x; x.y += b4.y; x.z += b4.z; x.w += b4.w;
            if (act == 1) {
                x.x = fmaxf(x.x, 0.f); x.y = fmaxf(x.y, 0.f); x.z = fmaxf(x.z, 0.f); x.w = fmaxf(x.w, 0.f);
            } else if (act == 2) {
                x.x = (x.x > 0.f) ? x.x : 0.01f * x.x; x.y = (x.y > 0.f) ? x.y : 0.01f * x.y;
                x.z = (x.z > 0.f) ? x.z : 0.01f * x.z; x.w = (x.w > 0.f) ? x.w : 0.01f * x.w;
            }
            if (rbase + jj * 4 < ep.M && col < ep.N && !(ep.debug & 4)) {
                if (act >= 4) {
                    const float neg = (act == 4) ? 0.f : 0.01f;
                    const float4 y = mk4[jj];
                    x.x *= (y.x > 0.f) ? 1.f : neg; x.y *= (y.y > 0.f) ? 1.f : neg;
                    x.z *= (y.z > 0.f) ? 1.f : neg; x.w *= (y.w > 0.f) ? 1.f : neg;
                }
                if (full) {
                    if (act == 3) {                                    // D <- D - A.B^T (trailing update)
                        const float4 old = *reinterpret_cast<const float4*>(o);
                        x = make_float4(old.x - x.x, old.y - x.y, old.z - x.z, old.w - x.w);
                    }
                    *reinterpret_cast<float4*>(o) = x;
                } else {
                    const float xs[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                        if (col + e < ep.N) o[e] = (act == 3) ? o[e] - xs[e] : xs[e];
                }
            }
        }
        __syncwarp();
    }
}

}  // namespace adkf

// ---- host: tensor maps through the driver entry point (no link-time dependency on libcuda) --------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static inline EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (fn == nullptr) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// An operand with `rows` indices on the output side and `kdim` contraction indices.
//   transposed = 0: stored [rows, kdim] (ld floats per row)  -> box 32 k x 128 rows, SWIZZLE_128B (already K-major)
//   transposed = 1: stored [kdim, rows]                      -> box 128 rows x 32 k, un-swizzled; the splitter transposes
//   box_rows: rows of the output side fetched per box (128; 64 for the B half tile of a CTA pair)
static inline int make_map_2d(CUtensorMap* map, const float* ptr, int64_t rows, int64_t kdim, int64_t ld, int transposed,
                              int box_rows = adkf::kBM) {
    EncodeTiledFn fn = encode_fn();
    if (fn == nullptr) return (int)cudaErrorNotSupported;
    if ((reinterpret_cast<uintptr_t>(ptr) & 15) || (ld * 4) % 16) return ADKF_ERR_BADARG;
    const cuuint64_t dims[2] = {(cuuint64_t)(transposed ? rows : kdim), (cuuint64_t)(transposed ? kdim : rows)};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    const cuuint32_t box[2] = {(cuuint32_t)(transposed ? box_rows : adkf::kBK), (cuuint32_t)(transposed ? adkf::kBK : box_rows)};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = CUDA_SUCCESS;
    for (int attempt = 0; attempt < 2; ++attempt) {
        r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(ptr), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, transposed ? CU_TENSOR_MAP_SWIZZLE_NONE : CU_TENSOR_MAP_SWIZZLE_128B,
               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_ERROR_INVALID_CONTEXT) break;
        cudaFree(0);   // a host thread whose first CUDA call is this driver entry point (autograd workers) has no
                       // current context yet: bind the primary context through the runtime and retry
    }
    return r == CUDA_SUCCESS ? 0 : -(1000 + (int)r);   // tensor-map encode failure: -(1000 + CUresult)
}

static inline int num_sms() {
    static int n = 0;
    if (n == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
            n = 148;
    }
    return n;
}

// Output map of the TMA-store epilogue: D [rows, cols] fp32 with row stride ld, box 32 columns (128 B) x 32 rows, SWIZZLE_128B (the
// layout of the warp-private staging tile).  Rows / columns past the extents are clipped by the hardware.
static inline int make_out_map_2d(CUtensorMap* map, float* ptr, int64_t rows, int64_t cols, int64_t ld) {
    EncodeTiledFn fn = encode_fn();
    if (fn == nullptr) return (int)cudaErrorNotSupported;
    if ((reinterpret_cast<uintptr_t>(ptr) & 15) || (ld * 4) % 16) return ADKF_ERR_BADARG;
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)ld * 4};
    const cuuint32_t box[2] = {32, 32};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = CUDA_SUCCESS;
    for (int attempt = 0; attempt < 2; ++attempt) {
        r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, ptr, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_ERROR_INVALID_CONTEXT) break;
        cudaFree(0);
    }
    return r == CUDA_SUCCESS ? 0 : -(1000 + (int)r);
}
