// pna_fused.cu -- one kernel per GNN block for the tower-style relational message passing of the PNA / MultiAggr layers
// (RelationalMP.forward + RelationalMultiAggrMP._aggregate_messages, fs_mol/modules/gnn.py:129-150, 197-265):
//
//   forward   gather [h_src | h_tgt] -> message MLP on the tensor cores (tcgen05, 3xTF32) -> bias + ReLU -> sum | mean | pseudo-std |
//             max over the edges arriving at each node, written as the UNSCALED aggregate [V, 4 x 256] (+ 2 aux bytes / channel)
//   backward  the same gather + MLP (the pre-activations are RECOMPUTED, bit for bit), epilogue = chain rule of the four
//             aggregators + ReLU mask -> d pre [2E, 768] in the (type, target) order of the weight / input gradient GEMMs
//
// What it replaces (pna_gemm.cu path): pna_gather_inputs writes XG [2E,256] (0.30 GB at C2), the batched GEMM reads it and writes
// pre [2E,768] (0.91 GB), pna_aggregate_fwd reads pre (0.91 GB) and writes [V,1024]; pna_dpre reads pre again.  Here neither XG nor
// pre reaches HBM in the forward pass: h [V,128] (68 MB) stays in L2, the gathered rows go from registers straight into TENSOR
// memory as the split A operand, the pre-activations live in TMEM / shared memory only.
//
// Work decomposition.  Edges in CSR order (target-major, type-major inside a node: the reference's message order).  A tile is
// 128 consecutive CSR rows starting at row 112 j; it OWNS the nodes whose first row lies in [112 j, 112 j + 112), so with an
// in-degree of at most 16 (host-checked; molecules have <= 6) every owned node has all its edges inside the tile.  The 192 message
// channels of a tower split into three 64-wide chunks (sum part | mean + std part | max part); a work item is (tower, chunk, tile)
// and a persistent CTA walks a contiguous range of items, so it keeps ONE (tower, chunk) weight slice in shared memory: for each
// of the 3 edge types a [64 x 64] block of W, pre-split once into tf32 hi / lo K-major SWIZZLE_128B tiles (96 KB).
// Rows of different edge types share a tile, so for every type present in the tile the full A tile is multiplied with that type's
// weights into its own accumulator pair and the epilogue picks, per row, the accumulators of the row's type.
//
// Numerics.  Identical to the unfused path bit for bit in the forward: same hi/lo split, same MMA order per accumulator
// (hi.hi -> big; lo.hi, hi.lo -> small), pre = (big + small) + bias with round-to-nearest adds, same reduction order per node.
//
// CTA = 384 threads:  warp 0 MMA issuer | warp 1 TMEM allocator | warps 4-7 gather + split (row = TMEM lane) |
//                     warps 8-11 drain (tcgen05.ld) + segment reduction / d-pre epilogue.
// TMEM (512 columns): type t: big 128 t .. +64, small 128 t + 64 .. +64;  A operand 384 .. 512 = [hi k0-31 | lo k0-31 | hi k32-63 | lo k32-63].
#include <cuda.h>
#include <algorithm>
#include "gemm_common.cuh"

namespace adkf {
namespace pf {

constexpr int kRows = 128, kStride = 112;
constexpr int kThreads = 384;
constexpr int kWTile = 64 * 32 * 4;                  // one [64 n x 32 k] K-major SWIZZLE_128B tile
constexpr int kWBytes = 12 * kWTile;                 // [type 3][hi|lo][k-block 2]
constexpr int kStgStride = 65;                       // floats per staged row (bank-conflict-free for lane = row and lane = column)
constexpr int kStgBytes = kRows * kStgStride * 4;    // 33,280
constexpr size_t kSmem = 1024 + kWBytes + 2 * kStgBytes + 2048;
constexpr int kTmemCols = 512, kACol = 384;
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
constexpr float kSmall = 1e-7f;

struct Args {
    const float* h;            // [V,128]
    const float* W;            // [4][3][192][64]
    const float* b;            // [4][3][192]
    const int32_t* rowptr;     // [V+1] CSR over target nodes
    const int32_t* col_src;    // [E]
    const int32_t* col_tgt;    // [E]
    const uint8_t* col_type;   // [E]
    const int32_t* tile_node0; // [ntiles+1] first node owned by tile j (first v with rowptr[v] >= 112 j); [ntiles] = V
    const int32_t* csr2ts;     // [E] (type, target)-sorted position of CSR row p (backward only)
    int V, E, ntiles;
    float* out;                // [V,1024]   forward: written; backward: read (mean, std)
    uint8_t* aux_cnt;          // [V,4,64]
    uint8_t* aux_arg;          // [V,4,64]
    const float* d_out;        // [V,1024]   backward
    float* d_pre;              // [E,768]    backward
};

__device__ __forceinline__ void tmem_ld16_nowait(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
// the loaded registers are defined only after tcgen05.wait::ld: this empty statement makes every later use depend on a point
// after the wait (the compiler may not move arithmetic on r[] above it)
__device__ __forceinline__ void touch16(uint32_t (&r)[16]) {
    asm volatile("" : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
                      "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void bar_sync_drain() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

// 32 gathered values of one row -> hi / lo tf32 in 64 consecutive TMEM columns
__device__ __forceinline__ void split32_to_tmem(const float4 (&x)[8], uint32_t taddr) {
    uint32_t hi[32], lo[32];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float v[4] = {x[j].x, x[j].y, x[j].z, x[j].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float hv = tf32_hi(v[e]);
            hi[4 * j + e] = __float_as_uint(hv);
            lo[4 * j + e] = __float_as_uint(v[e] - hv);
        }
    }
    tmem_st32(taddr, hi);
    tmem_st32(taddr + 32, lo);
}

template <bool BWD>
__global__ void __launch_bounds__(kThreads, 1) pna_fused_kernel(const Args a) {
    extern __shared__ unsigned char smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    unsigned char* base_ptr = smem_raw + (base - raw);
    unsigned char* w_ptr = base_ptr;                                         // 96 KB of weight tiles
    float* stg = reinterpret_cast<float*>(base_ptr + kWBytes);               // 2 x [128][65]
    unsigned char* misc = base_ptr + kWBytes + 2 * kStgBytes;
    float* bias_s = reinterpret_cast<float*>(misc);                          // [3][64]
    volatile uint8_t* row_type = reinterpret_cast<volatile uint8_t*>(misc + 768);        // [2][128]
    volatile uint32_t* present = reinterpret_cast<volatile uint32_t*>(misc + 768 + 256); // [2][4]
    const uint32_t bars = base + kWBytes + 2 * kStgBytes + 768 + 256 + 32;   // 8-byte aligned
    const uint32_t a_full = bars, a_free = bars + 8, acc_full = bars + 16, acc_free = bars + 24;
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(misc + 768 + 256 + 32 + 32);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        mbar_init(a_full, 4);
        mbar_init(a_free, 1);
        mbar_init(acc_full, 1);
        mbar_init(acc_free, 4);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32((const void*)tmem_slot)),
                     "n"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    const int64_t total = 12ll * a.ntiles;
    const int64_t per = (total + gridDim.x - 1) / gridDim.x;
    const int64_t i0 = (int64_t)blockIdx.x * per, i1 = (i0 + per < total) ? i0 + per : total;
    uint32_t it = 0;                                    // tiles processed by this CTA so far (barrier phases), same in every role

    for (int64_t seg0 = i0; seg0 < i1;) {
        const int combo = (int)(seg0 / a.ntiles);
        const int64_t seg_end = ((int64_t)(combo + 1) * a.ntiles < i1) ? (int64_t)(combo + 1) * a.ntiles : i1;
        const int tower = combo / 3, chunk = combo - 3 * tower;
        const int tile0 = (int)(seg0 - (int64_t)combo * a.ntiles), tile1 = (int)(seg_end - (int64_t)combo * a.ntiles);
        const int ntile_seg = tile1 - tile0;

        // ---- weights of (tower, chunk) for the 3 edge types: fp32 -> hi / lo tf32, K-major SWIZZLE_128B ----
        __syncthreads();                                // every MMA of the previous segment has retired (the drain warps waited for them)
        for (int f = threadIdx.x; f < 3 * 64 * 16; f += kThreads) {
            const int t = f >> 10, n = (f >> 4) & 63, j16 = f & 15, kb = j16 >> 3, j = j16 & 7;
            const float4 v = __ldg(reinterpret_cast<const float4*>(a.W + ((size_t)((tower * 3 + t) * 192 + chunk * 64 + n)) * 64) + j16);
            float4 hi, lo;
            split4(v, hi, lo);
            const int off = n * 128 + ((j ^ (n & 7)) << 4);
            *reinterpret_cast<float4*>(w_ptr + ((t * 2 + 0) * 2 + kb) * kWTile + off) = hi;
            *reinterpret_cast<float4*>(w_ptr + ((t * 2 + 1) * 2 + kb) * kWTile + off) = lo;
        }
        if (threadIdx.x < 192)
            bias_s[threadIdx.x] = a.b[(size_t)(tower * 3 + (threadIdx.x >> 6)) * 192 + chunk * 64 + (threadIdx.x & 63)];
        fence_async_smem();
        __syncthreads();

        if (warp == 0) {
            // ===== MMA issuer =====
            for (int i = 0; i < ntile_seg; ++i) {
                const uint32_t cur = it + (uint32_t)i, par = cur & 1;
                mbar_wait(a_full, par);
                mbar_wait(acc_free, par ^ 1);
                tc_fence_after();
                const uint32_t pres = present[par * 4 + 0] | present[par * 4 + 1] | present[par * 4 + 2] | present[par * 4 + 3];
                if (lane == 0) {
                    for (int t = 0; t < 3; ++t) {
                        if (!((pres >> t) & 1u)) continue;
                        const uint32_t big = tmem + (uint32_t)(t * 128), small = big + 64;
#pragma unroll
                        for (int kb = 0; kb < 2; ++kb) {
                            const uint64_t b_hi = smem_desc_k128(base + ((t * 2 + 0) * 2 + kb) * kWTile);
                            const uint64_t b_lo = smem_desc_k128(base + ((t * 2 + 1) * 2 + kb) * kWTile);
                            const uint32_t a_hi = tmem + (uint32_t)(kACol + kb * 64), a_lo = a_hi + 32;
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                const uint64_t adv = (uint64_t)(2 * k);
                                const uint32_t ak = (uint32_t)(8 * k);
                                umma_tf32_ts(big, a_hi + ak, b_hi + adv, kIdesc, (kb | k) ? 1u : 0u);
                                umma_tf32_ts(small, a_lo + ak, b_hi + adv, kIdesc, (kb | k) ? 1u : 0u);
                                umma_tf32_ts(small, a_hi + ak, b_lo + adv, kIdesc, 1u);
                            }
                        }
                    }
                    umma_commit(a_free);
                    umma_commit(acc_full);
                }
                __syncwarp();
            }
        } else if (warp >= 4 && warp < 8) {
            // ===== gather + split: row = TMEM lane =====
            const int row = (warp - 4) * 32 + lane;
            const uint32_t tmem_a = tmem + ((uint32_t)((warp - 4) * 32) << 16) + (uint32_t)kACol;
            for (int i = 0; i < ntile_seg; ++i) {
                const uint32_t cur = it + (uint32_t)i, par = cur & 1;
                const int64_t p = (int64_t)(tile0 + i) * kStride + row;
                const bool valid = p < a.E;
                int src = 0, tgt = 0, ty = 3;
                if (valid) { src = a.col_src[p]; tgt = a.col_tgt[p]; ty = a.col_type[p]; }
                float4 xs[8], xt[8];
                const float4* hs = reinterpret_cast<const float4*>(a.h + (size_t)src * 128 + tower * 32);
                const float4* ht = reinterpret_cast<const float4*>(a.h + (size_t)tgt * 128 + tower * 32);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    xs[j] = valid ? __ldg(hs + j) : make_float4(0.f, 0.f, 0.f, 0.f);
                    xt[j] = valid ? __ldg(ht + j) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
                const uint32_t m0 = __ballot_sync(0xffffffffu, ty == 0), m1 = __ballot_sync(0xffffffffu, ty == 1),
                               m2 = __ballot_sync(0xffffffffu, ty == 2);
                mbar_wait(a_free, par ^ 1);              // the MMAs of the previous tile have read the A operand
                tc_fence_after();
                row_type[par * 128 + row] = (uint8_t)ty;
                if (lane == 0) present[par * 4 + (warp - 4)] = (m0 ? 1u : 0u) | (m1 ? 2u : 0u) | (m2 ? 4u : 0u);
                split32_to_tmem(xs, tmem_a);
                split32_to_tmem(xt, tmem_a + 64);
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(a_full);
            }
        } else if (warp >= 8) {
            // ===== drain + segment epilogue =====
            const int q = warp - 8, row = q * 32 + lane;
            const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
            for (int i = 0; i < ntile_seg; ++i) {
                const uint32_t cur = it + (uint32_t)i, par = cur & 1;
                const int tile = tile0 + i;
                const int64_t row0 = (int64_t)tile * kStride;
                float* st = stg + (size_t)par * (kRows * kStgStride);
                mbar_wait(acc_full, par);
                tc_fence_after();
                const int rt = row_type[par * 128 + row];
                const uint32_t pres = present[par * 4 + 0] | present[par * 4 + 1] | present[par * 4 + 2] | present[par * 4 + 3];
#pragma unroll 1
                for (int piece = 0; piece < 4; ++piece) {
                    uint32_t vb[3][16], vs[3][16];
#pragma unroll
                    for (int t = 0; t < 3; ++t)
                        if ((pres >> t) & 1u) {
                            tmem_ld16_nowait(trow + (uint32_t)(t * 128 + piece * 16), vb[t]);
                            tmem_ld16_nowait(trow + (uint32_t)(t * 128 + 64 + piece * 16), vs[t]);
                        }
                    tmem_ld_wait();
                    float x[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) x[j] = 0.f;
#pragma unroll
                    for (int t = 0; t < 3; ++t)
                        if ((pres >> t) & 1u) {
                            touch16(vb[t]);
                            touch16(vs[t]);
                            if (rt == t) {
#pragma unroll
                                for (int j = 0; j < 16; ++j) x[j] = __uint_as_float(vb[t][j]) + __uint_as_float(vs[t][j]);
                            }
                        }
                    if (rt < 3) {
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const float pre = x[j] + bias_s[rt * 64 + piece * 16 + j];
                            st[row * kStgStride + piece * 16 + j] = BWD ? pre : fmaxf(pre, 0.f);
                        }
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(acc_free);    // the accumulators may be overwritten by the next tile
                bar_sync_drain();                         // the staged tile is complete

                const int n0 = a.tile_node0[tile], n1 = a.tile_node0[tile + 1];
                for (int item = q; item < 2 * (n1 - n0); item += 4) {
                    const int v = n0 + (item >> 1), col = (item & 1) * 32 + lane;
                    const int r0 = (int)(a.rowptr[v] - row0), deg = a.rowptr[v + 1] - a.rowptr[v];
                    const int r1 = min(r0 + deg, kRows);                     // in-degree <= 16 is checked by the host; never read past the tile
                    const size_t ob = (size_t)v * 1024 + (size_t)tower * 256;
                    const size_t xb = ((size_t)v * 4 + tower) * 64 + col;
                    const float* sc = st + col;
                    if (!BWD) {
                        float* o = a.out + ob;
                        if (chunk == 0) {
                            float s = 0.f;
                            for (int r = r0; r < r1; ++r) s += sc[r * kStgStride];
                            o[col] = s;
                        } else if (chunk == 1) {
                            float macc = 0.f;
                            for (int r = r0; r < r1; ++r) macc += sc[r * kStgStride];
                            const float mean = macc / (float)max(deg, 1);        // true division, like scatter_mean
                            float ssum = 0.f;
                            int cpos = 0;
                            for (int r = r0; r < r1; ++r) {
                                const float mm = sc[r * kStgStride];
                                // no FMA contraction: m^2 - mean^2 must be exactly 0 when m == mean (degree-1 rows), as in torch
                                const float d = __fsub_rn(__fmul_rn(mm, mm), __fmul_rn(mean, mean));
                                if (d > 0.f) { ssum += d; ++cpos; }
                                ssum += kSmall;
                            }
                            o[64 + col] = mean;
                            o[128 + col] = sqrtf(ssum);
                            a.aux_cnt[xb] = (uint8_t)min(cpos, 255);
                        } else {
                            float mx = 0.f;
                            int arg = 0, idx = 0;
                            for (int r = r0; r < r1; ++r, ++idx) {
                                const float m = sc[r * kStgStride];
                                if (idx == 0 || m > mx) { mx = m; arg = idx; }
                            }
                            o[192 + col] = (deg > 0) ? mx : 0.f;
                            a.aux_arg[xb] = (uint8_t)min(arg, 255);
                        }
                    } else if (deg > 0) {
                        const float* go = a.d_out + ob;
                        const float* fo = a.out + ob;
                        const int32_t* tsr = a.csr2ts + row0;
                        const size_t cb = (size_t)tower * 192 + (size_t)chunk * 64 + col;
                        if (chunk == 0) {
                            const float g = go[col];
                            for (int r = r0; r < r1; ++r)
                                a.d_pre[(size_t)tsr[r] * 768 + cb] = (sc[r * kStgStride] > 0.f) ? g : 0.f;
                        } else if (chunk == 1) {
                            const float g_mean = go[64 + col], g_std = go[128 + col], mean = fo[64 + col], stdv = fo[128 + col];
                            const float cpos = (float)a.aux_cnt[xb];
                            const float invdeg = 1.f / (float)deg;
                            const float cc = 0.5f * g_std / stdv;
                            for (int r = r0; r < r1; ++r) {
                                const float a_mean = sc[r * kStgStride];
                                const float m_mean = fmaxf(a_mean, 0.f);
                                const float ind = (__fsub_rn(__fmul_rn(m_mean, m_mean), __fmul_rn(mean, mean)) > 0.f) ? 1.f : 0.f;
                                const float dm = g_mean * invdeg + cc * (2.f * m_mean * ind - 2.f * mean * cpos * invdeg);
                                a.d_pre[(size_t)tsr[r] * 768 + cb] = (a_mean > 0.f) ? dm : 0.f;
                            }
                        } else {
                            const float g = go[192 + col];
                            const int arg = a.aux_arg[xb];
                            int idx = 0;
                            for (int r = r0; r < r1; ++r, ++idx) {
                                const bool hit = sc[r * kStgStride] > 0.f && (idx == arg || (idx >= 255 && arg == 255));
                                a.d_pre[(size_t)tsr[r] * 768 + cb] = hit ? g : 0.f;
                            }
                        }
                    }
                }
            }
            tc_fence_before();
        }
        it += (uint32_t)ntile_seg;
        seg0 = seg_end;
    }

    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kTmemCols) : "memory");
    }
}

template <bool BWD>
static int launch(const Args& a, cudaStream_t s) {
    static bool done = false;
    if (!done) {
        cudaError_t e = cudaFuncSetAttribute((const void*)pna_fused_kernel<BWD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmem);
        if (e != cudaSuccess) return (int)e;
        done = true;
    }
    const int64_t total = 12ll * a.ntiles;
    const unsigned grid = (unsigned)std::min<int64_t>(total, num_sms());
    pna_fused_kernel<BWD><<<grid, kThreads, kSmem, s>>>(a);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

}  // namespace pf
}  // namespace adkf

using namespace adkf;

extern "C" int32_t adkf_pna_fused_num_tiles(int64_t E) { return (int32_t)(E / pf::kStride + 1); }
extern "C" int32_t adkf_pna_fused_max_degree(void) { return pf::kRows - pf::kStride; }

extern "C" int adkf_pna_fused_fwd(const float* h, const float* W, const float* b, const int32_t* rowptr, const int32_t* col_src,
                                  const int32_t* col_tgt, const uint8_t* col_type, const int32_t* tile_node0, int32_t V, int64_t E,
                                  float* out, uint8_t* aux_cnt, uint8_t* aux_arg, void* stream) {
    if (V <= 0) return 0;
    if (!h || !W || !b || !rowptr || !tile_node0 || !out || !aux_cnt || !aux_arg || E < 0 || E > 0x7fffffff) return ADKF_ERR_BADARG;
    if (E > 0 && (!col_src || !col_tgt || !col_type)) return ADKF_ERR_BADARG;
    (void)cudaGetLastError();
    pf::Args a = {};
    a.h = h; a.W = W; a.b = b; a.rowptr = rowptr; a.col_src = col_src; a.col_tgt = col_tgt; a.col_type = col_type;
    a.tile_node0 = tile_node0; a.V = V; a.E = (int)E; a.ntiles = adkf_pna_fused_num_tiles(E);
    a.out = out; a.aux_cnt = aux_cnt; a.aux_arg = aux_arg;
    return pf::launch<false>(a, (cudaStream_t)stream);
}

extern "C" int adkf_pna_fused_dpre(const float* h, const float* W, const float* b, const int32_t* rowptr, const int32_t* col_src,
                                   const int32_t* col_tgt, const uint8_t* col_type, const int32_t* tile_node0, const int32_t* csr2ts,
                                   int32_t V, int64_t E, const float* fwd_out, const uint8_t* aux_cnt, const uint8_t* aux_arg,
                                   const float* d_out, float* d_pre, void* stream) {
    if (V <= 0 || E <= 0) return 0;
    if (!h || !W || !b || !rowptr || !col_src || !col_tgt || !col_type || !tile_node0 || !csr2ts || !fwd_out || !aux_cnt || !aux_arg ||
        !d_out || !d_pre || E > 0x7fffffff)
        return ADKF_ERR_BADARG;
    (void)cudaGetLastError();
    pf::Args a = {};
    a.h = h; a.W = W; a.b = b; a.rowptr = rowptr; a.col_src = col_src; a.col_tgt = col_tgt; a.col_type = col_type;
    a.tile_node0 = tile_node0; a.csr2ts = csr2ts; a.V = V; a.E = (int)E; a.ntiles = adkf_pna_fused_num_tiles(E);
    a.out = const_cast<float*>(fwd_out); a.aux_cnt = const_cast<uint8_t*>(aux_cnt); a.aux_arg = const_cast<uint8_t*>(aux_arg);
    a.d_out = d_out; a.d_pre = d_pre;
    return pf::launch<true>(a, (cudaStream_t)stream);
}
