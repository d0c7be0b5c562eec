// gemm_tf32x3.cu -- fp32-accurate GEMM on the 5th-generation tensor cores (sm_100a): tcgen05.mma kind::tf32 with the
// 3xTF32 error-compensated split, operands staged by TMA, accumulator in TMEM.
//
//   D[M,N] = A[M,K] . B[N,K]^T          (both operands K-major, fp32 in HBM)
//
// Used for
//   (1) the per-task Gram / squared-distance matrix of the GP inputs (gpytorch sq_dist inside MaternKernel / RBFKernel,
//       call sites fs_mol/utils/gp_utils.py:28,31): A = B = Z_task - mean(support rows), epilogue
//       D2 = |a|^2 + |b|^2 - 2 a.b, clamp_min(0), exact zero diagonal -- "the one dense contraction that goes on tensor
//       cores" of the north star;
//   (2) the dense layers of the extractor forward, y = act(x W^T + b) (nn.Linear in fs_mol/modules/gnn.py:79-95,
//       477-515, graph_readout.py:224-259, fs_mol/models/adaptive_dkt.py:61-65).
//
// Why 3xTF32: the parity target is 1e-4 relative on GP outputs at d = 2048 (SURVEY section 7 hard part (b)); a single TF32
// pass carries 2^-11 per operand.  Each fp32 value v is split as hi = rna_tf32(v), lo = v - hi (exact in fp32), and
// a.b ~ hi.hi + lo.hi + hi.lo (fp32 accumulation in TMEM), which leaves ~2^-21 per product -- fp32-class accuracy at a third
// of the TF32 rate.
//
// The tensor core adds into its fp32 accumulator with round-toward-zero (measured here: a 1408-long contraction kept in
// ONE TMEM accumulator comes out 1e-5 low, biased toward zero, growing linearly with K).  So the accumulation is split:
//   * the dominant hi.hi products go to a double-buffered TMEM accumulator that is DRAINED every kGroup k-blocks
//     (16 MMAs) into fp32 registers with ordinary round-to-nearest adds (bias <~ 6e-7);
//   * the two correction products lo.hi + hi.lo are 2^-11 smaller and stay in a third accumulator until the end.
//
// Operand layouts.  An operand is "K-major" when the contraction index is the contiguous one in HBM (x[M,K], W[N,K]) and
// "transposed" when the contraction index is the slow one (dY^T in dW = dY^T x, W^T in dX = dY W).  The tensor core only
// ever sees K-major SWIZZLE_128B tiles: a transposed operand is fetched by TMA as an un-swizzled [32 k][128 mn] box and the
// splitter warps, which touch every element anyway, write hi/lo TRANSPOSED into the K-major tiles.  One MMA path, three
// operand combinations (A,B) = (K,K) forward, (K,T) input gradient, (T,T) weight gradient (split-K over the node dimension,
// deterministic two-pass reduction).
//
// CTA = persistent over 128 x 128 output tiles, 16 warps, two smem rings (raw: TMA -> splitter, 4 x 32 KB; split B: splitter ->
// MMA, 2 x 32 KB) + 32 KB of epilogue staging; the split A operand lives in tensor memory (2 stages x 64 columns):
//   warp 0     TMA producer      cp.async.bulk.tensor.2d -> raw ring
//   warp 1     MMA issuer        one elected lane: 4 k-steps x 3 tcgen05.mma (A from TMEM, B from smem) per k-block,
//                                tcgen05.commit frees the split stage
//   warp 2     TMEM allocator    512 columns: big[2] x 128, small x 128, A operand 2 x (32 hi + 32 lo)
//   warps 4-7  splitter          raw A tile -> hi/lo rows in TMEM (tcgen05.st); raw B tile -> hi/lo K-major smem tiles
//                                (generic proxy -> fence.proxy.async)
//   warps 8-15 drain + epilogue  tcgen05.ld of the accumulator rows (warp w owns TMEM lanes 32(w%4).., 64 columns each),
//              RN accumulation in registers, fused epilogue, stores.
// Every mbarrier wait is bounded (trap after ~2^28 polls) so that a protocol bug cannot hang the device.
#include <cuda.h>
#include <algorithm>
#include <cstdlib>
#include <string>
#include "gemm_common.cuh"
#include "../../include/adkf_b200.h"

namespace adkf {

constexpr int kRaw = 4;                              // raw ring depth (TMA destination: A_raw, B_raw)
constexpr int kSplit = 2;                            // split ring depth (MMA source: B_hi, B_lo in smem; A_hi, A_lo in TMEM)
constexpr int kRawBytes = 2 * kTileBytes;
constexpr int kSplitBytes = 2 * kTileBytes;
constexpr int kRingBytes = kRaw * kRawBytes + kSplit * kSplitBytes;   // 192 KB
constexpr int kStageBytes = 8 * 4096;                // epilogue staging: 8 drain warps x (32 rows x 32 columns) fp32
constexpr int kGemmThreads = 640;                    // 5 warpgroups: control, A splitter, B splitter, 2 x drain
constexpr int kTmemCols = 512;                       // big[0], big[1], small (128 fp32 columns each), A operand 2 x 64
constexpr int kSmallCol = 2 * kBN;                   // first TMEM column of the correction accumulator
constexpr int kAopCol = 3 * kBN;                     // first TMEM column of the A operand stages (hi 32 + lo 32 columns each)
constexpr int kGroup = 4;                            // k-blocks accumulated in TMEM between two drains
constexpr size_t kGemmSmem = (size_t)kRingBytes + kStageBytes + 1024 /*align*/ + 256 /*barriers*/;

// One operand tile, raw -> (hi, lo) in the K-major SWIZZLE_128B layout the MMA descriptors describe (row r = 128 B, its
// logical 16 B chunk j stored at chunk j ^ (r & 7)).  t = 0..127 (the four splitter warps).
//   TR = false: the raw tile already has that layout (TMA SWIZZLE_128B box of 128 rows x 32 floats); float4 p of the tile
//               is row p / 8, physical chunk p % 8.  `mu4` (gram centring) is the slice of the task mean for this thread's
//               logical chunk, which is the same for all its rows: (t & 7) ^ ((t >> 3) & 7).
//   TR = true : the raw tile is [32 k][128 mn] floats, un-swizzled; thread t owns output row mn = t and gathers 4
//               consecutive k per store (reads are conflict-free: a warp reads 32 consecutive floats).
// All loads of a tile are issued before the first store: the splitter runs one warp per scheduler, so its throughput is
// set by how many shared-memory loads it keeps in flight (ncu r1b: a load -> convert -> store chain per float4 left the
// splitter busy 76% of the time and the tensor pipe at 31%).
template <bool TR>
__device__ __forceinline__ void split_tile(const unsigned char* __restrict__ raw, unsigned char* __restrict__ hi_t,
                                           unsigned char* __restrict__ lo_t, int t, const float4 mu4) {
    float4* __restrict__ hi4 = reinterpret_cast<float4*>(hi_t);
    float4* __restrict__ lo4 = reinterpret_cast<float4*>(lo_t);
    if (!TR) {
        const float4* __restrict__ r4 = reinterpret_cast<const float4*>(raw);
        constexpr int kPer = (kTileBytes / 16) / 128;                 // 1024 float4 / 128 threads = 8
        float4 v[kPer];
#pragma unroll
        for (int i = 0; i < kPer; ++i) v[i] = r4[i * 128 + t];
#pragma unroll
        for (int i = 0; i < kPer; ++i) {
            float4 x = v[i];
            x.x -= mu4.x; x.y -= mu4.y; x.z -= mu4.z; x.w -= mu4.w;
            float4 hi, lo;
            split4(x, hi, lo);
            hi4[i * 128 + t] = hi;
            lo4[i * 128 + t] = lo;
        }
    } else {
        const float* __restrict__ rf = reinterpret_cast<const float*>(raw);
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            float4 v[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int c = half * 4 + q;
                v[q].x = rf[(4 * c + 0) * kBM + t];
                v[q].y = rf[(4 * c + 1) * kBM + t];
                v[q].z = rf[(4 * c + 2) * kBM + t];
                v[q].w = rf[(4 * c + 3) * kBM + t];
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int c = half * 4 + q;
                float4 hi, lo;
                split4(v[q], hi, lo);
                const int p = t * 8 + (c ^ (t & 7));
                hi4[p] = hi;
                lo4[p] = lo;
            }
        }
    }
}

// The A operand goes to TENSOR memory instead of shared memory: splitter thread t (warp 4 + t/32, which may touch TMEM lanes
// 32 (t/32)..) owns row t of the tile, splits its 32 k values and writes hi to 32 TMEM columns and lo to the next 32
// (tcgen05.st 32x32b: lane = row, register i = column i).  That removes the A_hi / A_lo stores and the three A reads per
// k-step of the MMAs from the shared-memory pipe, which is what bounded this kernel (ncu r1e: smem pipe 95%+; 224 KB of
// shared-memory traffic per k-block against 1107 tensor-pipe cycles -> 144 KB).
//   TR = false: the raw tile is the SWIZZLE_128B box (row t = 128 B, logical chunk j at physical chunk j ^ (t & 7)); the 8
//               lanes of a quarter warp read 8 different chunks -> conflict-free.
//   TR = true : the raw tile is [32 k][128 mn]; lane reads of one k are 32 consecutive floats.
// mu_k (gram centring, else nullptr) points at the task mean for this k-block; kleft = K - k0.
template <bool TR>
__device__ __forceinline__ void split_rows_to_tmem(const unsigned char* __restrict__ raw, uint32_t taddr, int t,
                                                   const float* __restrict__ mu_k, int kleft) {
    float x[kBK];
    if (!TR) {
        const float4* __restrict__ r4 = reinterpret_cast<const float4*>(raw) + t * 8;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float4 v = r4[j ^ (t & 7)];
            x[4 * j] = v.x; x[4 * j + 1] = v.y; x[4 * j + 2] = v.z; x[4 * j + 3] = v.w;
        }
    } else {
        const float* __restrict__ rf = reinterpret_cast<const float*>(raw) + t;
#pragma unroll
        for (int k = 0; k < kBK; ++k) x[k] = rf[k * kBM];
    }
    if (mu_k != nullptr) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
            if (4 * j + 3 < kleft) {
                const float4 m = *reinterpret_cast<const float4*>(mu_k + 4 * j);
                x[4 * j] -= m.x; x[4 * j + 1] -= m.y; x[4 * j + 2] -= m.z; x[4 * j + 3] -= m.w;
            }
    }
    uint32_t hi[kBK], lo[kBK];
#pragma unroll
    for (int k = 0; k < kBK; ++k) {
        const float h = tf32_hi(x[k]);
        hi[k] = __float_as_uint(h);
        lo[k] = __float_as_uint(x[k] - h);
    }
    tmem_st32(taddr, hi);
    tmem_st32(taddr + kBK, lo);
    tmem_st_wait();
}

// One output tile of the persistent loop: everything a warp role needs, a pure function of the tile index (so every
// role takes the same skip decisions without talking to the others).
struct TileCoord {
    int rowA0, rowB0;      // first output row / column (mode 1: absolute row of Z)
    int task, Nt, nsup;    // mode 1
    int ti, tj;            // mode 1 tile coordinates inside the task
    int kb0, nk;           // first k-block and number of k-blocks
    int batch;
    int64_t zslab;         // blockIdx.z of the non-persistent formulation (split-K slab index)
    bool skip;
};

__device__ __forceinline__ TileCoord tile_coord(const GemmEpilogue& ep, int tile) {
    TileCoord c;
    const int bx = tile % ep.gx, by = (tile / ep.gx) % ep.gy, bz = tile / (ep.gx * ep.gy);
    c.task = 0; c.Nt = 0; c.nsup = 0; c.ti = by; c.tj = bx; c.kb0 = 0; c.batch = 0; c.zslab = bz; c.skip = false;
    c.nk = (ep.K + kBK - 1) / kBK;
    if (ep.mode == 1) {
        c.task = bz;
        const int r0 = ep.z_off[bz];
        c.Nt = ep.z_off[bz + 1] - r0;
        c.nsup = ep.n_sup[bz];
        // symmetric: lower tiles only; the query x query block is not needed for meta-test
        c.skip = (bx > by) || (by * kBM >= c.Nt) || (!ep.qq_block && bx * kBN >= c.nsup);
        c.rowA0 = r0 + by * kBM;
        c.rowB0 = r0 + bx * kBN;
    } else {
        c.rowA0 = by * kBM;
        c.rowB0 = bx * kBN;
        const int split = bz % ep.splits;
        c.batch = bz / ep.splits;
        c.kb0 = split * ep.kb_per_split;
        c.nk = min(c.nk, c.kb0 + ep.kb_per_split) - c.kb0;              // host guarantees >= 1
    }
    return c;
}

// Persistent: gridDim.x = min(tiles, SMs); CTA b walks tiles b, b + gridDim.x, ...  The smem rings, the double-buffered
// TMEM accumulators and all mbarrier phases run on across tiles, so the loads / splits / MMAs of tile t+1 overlap the
// drain and the global stores of tile t (what makes the short-K GEMMs -- K = 64 message MLP, K = 128 BOOM / input
// gradient -- store-bound instead of launch- and latency-bound: ncu r1c had them at 6-10% tensor-pipe utilisation with one
// CTA per tile).
template <bool TA, bool TB>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tf32x3_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, GemmEpilogue ep) {
    if (ep.gate != nullptr && *ep.gate == 0) return;     // gated launch (uniform over the grid: nothing has been set up yet)
    extern __shared__ unsigned char smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;                  // SWIZZLE_128B tiles need 1024 B alignment
    unsigned char* base_ptr = smem_raw + (base - raw);
    const uint32_t split_base = base + kRaw * kRawBytes;
    unsigned char* split_ptr = base_ptr + kRaw * kRawBytes;
    const uint32_t bars = base + kRingBytes + kStageBytes;
    unsigned char* stage_ptr = base_ptr + kRingBytes;
    auto rawfull_bar = [&](int r) { return bars + 8u * r; };
    auto rawempty_bar = [&](int r) { return bars + 8u * (kRaw + r); };
    auto conv_bar = [&](int s) { return bars + 8u * (2 * kRaw + s); };              // split stage filled
    auto empty_bar = [&](int s) { return bars + 8u * (2 * kRaw + kSplit + s); };    // split stage consumed by the MMAs
    auto accfull_bar = [&](int b) { return bars + 8u * (2 * kRaw + 2 * kSplit + b); };
    auto accempty_bar = [&](int b) { return bars + 8u * (2 * kRaw + 2 * kSplit + 2 + b); };
    auto tileempty_bar = [&](int b) { return bars + 8u * (2 * kRaw + 2 * kSplit + 4 + b); };   // small[b] drained
    volatile uint32_t* tmem_slot =
        reinterpret_cast<volatile uint32_t*>(base_ptr + kRingBytes + kStageBytes + 8 * (2 * kRaw + 2 * kSplit + 6));

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ntiles = ep.ntiles;

    // ---- one-time setup -----------------------------------------------------------------------------------
    if (threadIdx.x == 0) {
        for (int r = 0; r < kRaw; ++r) {
            mbar_init(rawfull_bar(r), 1);
            mbar_init(rawempty_bar(r), 8);  // one arrive per splitter warp (4 A + 4 B)
        }
        for (int s = 0; s < kSplit; ++s) {
            mbar_init(conv_bar(s), 8);      // one arrive per splitter warp
            mbar_init(empty_bar(s), 1);     // tcgen05.commit
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(accfull_bar(b), 1);
            mbar_init(accempty_bar(b), 8);  // one arrive per drain warp
            mbar_init(tileempty_bar(b), 8);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32((const void*)tmem_slot)),
                     "n"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_acc = *tmem_slot;     // columns: big[0] 0..127, big[1] 128..255, small 256..383, A operand 384..511

    // (each setmaxnreg sits at the top of its role's branch so that the code it governs is dominated by it)
    if (warp == 0) {
        // ===== TMA producer =====
        reg_dec<40>();
        if (lane == 0) {
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
                const TileCoord c = tile_coord(ep, tile);
                if (c.skip) continue;
                const int a_mn0 = c.rowA0 + c.batch * ep.a_mn, b_mn0 = c.rowB0 + c.batch * ep.b_mn;
                const int a_k0 = c.batch * ep.a_k, b_k0 = c.batch * ep.b_k;
                for (int i = 0; i < c.nk; ++i, ++it) {
                    const int r = it % kRaw;
                    const uint32_t ph = (it / kRaw) & 1;
                    const int k = (c.kb0 + i) * kBK;
                    mbar_wait(rawempty_bar(r), ph ^ 1);
                    if (ep.debug & 16) { mbar_arrive(rawfull_bar(r)); continue; }
                    mbar_expect_tx(rawfull_bar(r), kRawBytes);
                    const uint32_t dst = base + r * kRawBytes;
                    if (TA) tma_load_2d(dst, &tmA, rawfull_bar(r), a_mn0, k + a_k0);
                    else    tma_load_2d(dst, &tmA, rawfull_bar(r), k + a_k0, a_mn0);
                    if (TB) tma_load_2d(dst + kTileBytes, &tmB, rawfull_bar(r), b_mn0, k + b_k0);
                    else    tma_load_2d(dst + kTileBytes, &tmB, rawfull_bar(r), k + b_k0, b_mn0);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        reg_dec<40>();
        uint32_t it = 0, gc = 0, tl = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            const TileCoord c = tile_coord(ep, tile);
            if (c.skip) continue;
            mbar_wait(tileempty_bar(0), (tl & 1) ^ 1);                     // small read out by the drain warps of the last tile
            for (int i = 0; i < c.nk; ++i, ++it) {
                const int s = it % kSplit;
                const uint32_t ph = (it / kSplit) & 1;
                const uint32_t g = gc + i / kGroup;
                const int b = g & 1;
                const bool group_first = (i % kGroup) == 0;
                const bool group_last = (i % kGroup) == kGroup - 1 || i == c.nk - 1;
                if (group_first) mbar_wait(accempty_bar(b), ((g >> 1) & 1) ^ 1);   // drained two groups ago
                mbar_wait(conv_bar(s), ph);
                tc_fence_after();
                {                                                // every lane runs this; one elected lane issues (umma_*_e)
                    const uint32_t st = split_base + s * kSplitBytes;
                    const uint64_t b_hi = smem_desc_k128(st), b_lo = smem_desc_k128(st + kTileBytes);
                    const uint32_t a_hi = tmem_acc + (uint32_t)(kAopCol + s * 2 * kBK), a_lo = a_hi + kBK;
                    const uint32_t acc_big = tmem_acc + (uint32_t)(b * kBN), acc_small = tmem_acc + (uint32_t)kSmallCol;
#pragma unroll
                    for (int k = 0; k < ((ep.debug & 2) ? 0 : kBK / 8); ++k) {   // UMMA_K = 8 tf32: 8 TMEM columns of A, 32 B of B
                        const uint64_t adv = (uint64_t)(2 * k);                  // (+2 in the descriptor's address field)
                        const uint32_t ak = (uint32_t)(8 * k);
                        umma_tf32_ts_e(acc_small, a_lo + ak, b_hi + adv, kIdescTf32, (i | k) ? 1u : 0u);
                        umma_tf32_ts_e(acc_small, a_hi + ak, b_lo + adv, kIdescTf32, 1u);
                        umma_tf32_ts_e(acc_big, a_hi + ak, b_hi + adv, kIdescTf32, (group_first && k == 0) ? 0u : 1u);
                    }
                    umma_commit_e(empty_bar(s));                             // frees the split stage when these MMAs retire
                    if (group_last) umma_commit_e(accfull_bar(b));
                }
                __syncwarp();
            }
            gc += (uint32_t)((c.nk + kGroup - 1) / kGroup);
            ++tl;
        }
    } else if (warp < 4) {
        reg_dec<40>();                                                 // TMEM allocator / spare warp: the whole warpgroup must execute it
    } else if (warp < 8) {
        // ===== A splitter: row t of the raw A tile -> hi / lo in tensor memory (warp w may touch TMEM lanes 32 (w % 4)..) =====
        const int t = threadIdx.x - 128;                               // 0..127
        const uint32_t tmem_a = tmem_acc + ((uint32_t)((warp - 4) * 32) << 16) + (uint32_t)kAopCol;
        uint32_t it = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            const TileCoord c = tile_coord(ep, tile);
            if (c.skip) continue;
            for (int i = 0; i < c.nk; ++i, ++it) {
                const int r = it % kRaw, s = it % kSplit;
                const int k0 = (c.kb0 + i) * kBK;
                mbar_wait(rawfull_bar(r), (it / kRaw) & 1);
                mbar_wait(empty_bar(s), ((it / kSplit) & 1) ^ 1);      // the MMAs that read this TMEM stage have retired
                tc_fence_after();
                if (!(ep.debug & (1 | 128)))
                    split_rows_to_tmem<TA>(base_ptr + (size_t)r * kRawBytes, tmem_a + (uint32_t)(s * 2 * kBK), t,
                                           ep.mode == 1 ? ep.mu + (size_t)c.task * ep.K + k0 : nullptr, ep.K - k0);
                tc_fence_before();                                     // the TMEM stores are ordered before the arrive below
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(conv_bar(s));
                    mbar_arrive(rawempty_bar(r));
                }
            }
        }
    } else if (warp < 12) {
        // ===== B splitter: raw B tile -> hi / lo K-major smem tiles (transposing where the operand is not K-major) =====
        reg_dec<72>();
        const int t = threadIdx.x - 256;                               // 0..127
        const int chunk = (t & 7) ^ ((t >> 3) & 7);                    // logical 16 B chunk of this thread's float4s
        uint32_t it = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            const TileCoord c = tile_coord(ep, tile);
            if (c.skip) continue;
            for (int i = 0; i < c.nk; ++i, ++it) {
                const int r = it % kRaw, s = it % kSplit;
                float4 mu4 = make_float4(0.f, 0.f, 0.f, 0.f);
                if (ep.mode == 1) {
                    const int k = (c.kb0 + i) * kBK + chunk * 4;
                    if (k + 3 < ep.K) mu4 = *reinterpret_cast<const float4*>(ep.mu + (size_t)c.task * ep.K + k);
                }
                mbar_wait(rawfull_bar(r), (it / kRaw) & 1);
                mbar_wait(empty_bar(s), ((it / kSplit) & 1) ^ 1);
                unsigned char* st = split_ptr + (size_t)s * kSplitBytes;
                if (!(ep.debug & (1 | 256)))
                    split_tile<TB>(base_ptr + (size_t)r * kRawBytes + kTileBytes, st, st + kTileBytes, t, mu4);
                fence_async_smem();                                    // generic-proxy writes -> visible to the MMA (async proxy)
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(conv_bar(s));
                    mbar_arrive(rawempty_bar(r));
                }
            }
        }
    } else {
        // ===== drain (RN accumulation of the TMEM partial sums in registers), then epilogue =====
        reg_inc<136>();
        const int q = warp & 3;                                        // TMEM lane quarter this warp may access
        const int half = (warp - 12) >> 2;                             // which 64 accumulator columns
        const int r_local = q * 32 + lane;
        const uint32_t trow = tmem_acc + ((uint32_t)(q * 32) << 16) + (uint32_t)(half * 64);
        const int colbase = half * 64;
        uint32_t gc = 0, tl = 0;
        for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
            const TileCoord c = tile_coord(ep, tile);
            if (c.skip) continue;
            float acc[64];
#pragma unroll
            for (int j = 0; j < 64; ++j) acc[j] = 0.f;
            const int ngroups = (c.nk + kGroup - 1) / kGroup;
            for (int gl = 0; gl < ngroups; ++gl) {
                const uint32_t g = gc + gl;
                const int b = g & 1;
                mbar_wait(accfull_bar(b), (g >> 1) & 1);
                tc_fence_after();
#pragma unroll
                for (int c0 = 0; c0 < ((ep.debug & 32) ? 0 : 64); c0 += 16) {
                    float v[16];
                    tmem_ld16(trow + (uint32_t)(b * kBN) + c0, v);
#pragma unroll
                    for (int j = 0; j < 16; ++j) acc[c0 + j] += v[j];
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(accempty_bar(b));
            }
#pragma unroll
            for (int c0 = 0; c0 < ((ep.debug & 32) ? 0 : 64); c0 += 16) {   // the small correction accumulator, once per tile
                float v[16];
                tmem_ld16(trow + (uint32_t)kSmallCol + c0, v);
#pragma unroll
                for (int j = 0; j < 16; ++j) acc[c0 + j] += v[j];
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tileempty_bar(0));
            gc += (uint32_t)ngroups;
            ++tl;

            if (ep.debug & 64) continue;
            if (ep.mode == 0) {
                gemm_store_tile(ep, c.rowA0, c.rowB0, c.zslab, c.batch, acc, q, lane, colbase,
                                reinterpret_cast<float4*>(stage_ptr + (size_t)(warp - 12) * 4096));
            } else {
                const int r0 = ep.z_off[c.task];
                const int r = c.ti * kBM + r_local;                    // row inside the task
                float* Dt = ep.out + ep.d2_off[c.task];
                const float ni = (r < c.Nt) ? ep.nrm[r0 + r] : 0.f;
#pragma unroll
                for (int c0 = 0; c0 < 64; c0 += 16) {
                    float v[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[j] = acc[c0 + j];
                    if (r < c.Nt) {
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const int cc = c.tj * kBN + colbase + c0 + j;
                            if (cc < c.Nt && (c.ti != c.tj || cc <= r)) {
                                const float d = pairstat_value(ep.stat, ni, ep.nrm[r0 + cc], v[j], r == cc);
                                Dt[(size_t)r * c.Nt + cc] = d;
                                Dt[(size_t)cc * c.Nt + r] = d;
                            }
                        }
                    }
                }
            }
        }
        tc_fence_before();
    }

    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_acc), "n"(kTmemCols) : "memory");
    }
}

}  // namespace adkf

using namespace adkf;

template <bool TA, bool TB>
static int gemm_attr() {
    static bool done = false;
    if (!done) {
        cudaError_t e = cudaFuncSetAttribute((const void*)gemm_tf32x3_kernel<TA, TB>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kGemmSmem);
        if (e != cudaSuccess) return (int)e;
        done = true;
    }
    return 0;
}

namespace adkf {
__global__ void splitk_reduce_kernel(const float* __restrict__ ws, int splits, int64_t n, float* __restrict__ out,
                                     int64_t d_off) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int b = blockIdx.y;                                  // batch: its `splits` slabs are consecutive in ws
    const float* w = ws + (int64_t)b * splits * n;
    float acc = 0.f;
    for (int s = 0; s < splits; ++s) acc += w[(int64_t)s * n + i];
    out[(int64_t)b * d_off + i] = acc;
}
}  // namespace adkf

// split-K plan.  The persistent kernel walks (tile, split) work items 148 at a time, so what matters is how full the LAST
// wave is: e.g. 66 tiles x 5 splits = 330 items = 2.23 waves runs as 3 waves at 74%.  Cost model in k-block units
// (one 128x128x32 k-block ~ 0.7 us per CTA): waves x (k-blocks per split + fixed tile cost) + the slab write/read of the
// reduce pass; every split keeps >= 8 k-blocks.  Deterministic: the plan depends on the shape only.
static int plan_splits_tiles(int64_t tiles, int nk, int workers = 148) {
    if (nk < 16 || tiles >= 4 * workers) return 1;
    const int smax = std::min(nk / 8, 64);
    double best = 1e300;
    int best_s = 1;
    for (int s = 1; s <= smax; ++s) {
        const int per = (nk + s - 1) / s;
        if ((nk + per - 1) / per != s) continue;                 // s would leave an empty split
        const int64_t items = tiles * s;
        const int64_t waves = (items + workers - 1) / workers;
        double t = (double)waves * (per + 6.0);
        if (s > 1) t += 4.0 + 0.036 * (double)items;
        if (t < best * 0.98) { best = t; best_s = s; }           // prefer fewer splits unless >2% better
    }
    return best_s;
}
// CTA pairs (gemm_pair.cu, 256 x 128 tiles, 74 workers) take every mode-0 product with more than one 128-row tile;
// ADKF_GEMM_IMPL=1 forces the one-CTA kernel (A/B measurements).
namespace adkf {
int gemm_pair_presplit(const float* W, int64_t rows, int K, int64_t ld, int corr, float* hi, float* co, cudaStream_t s);
int gemm_pair_launch_pre(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mc, const CUtensorMap& mo, const CUtensorMap& mm,
                         const GemmEpilogue& ep, int corr, cudaStream_t s);
int gemm_pair_launch(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mo, const CUtensorMap& mm, const GemmEpilogue& ep,
                     int transA, int transB, cudaStream_t s);
}
namespace adkf {
int g_gemm_impl = getenv("ADKF_GEMM_IMPL") ? atoi(getenv("ADKF_GEMM_IMPL")) : 2;          // 1: one CTA per tile, 2: CTA pairs
// pair kernel, correction products (ADKF_GEMM_CORR = tf32 | bf16 | mixed | fp16 | fp16all): 0 tf32 everywhere, 1 bf16 everywhere,
// 2 tf32 forward / bf16 backward (round-2 default until the fp16 scheme), 3 (default) scaled fp16 forward / bf16 backward, 4 fp16 everywhere
static int corr_from_env() {
    const char* e = getenv("ADKF_GEMM_CORR");
    if (!e) return 3;
    const std::string v(e);
    if (v == "tf32" || v == "t") return 0;
    if (v == "bf16" || v == "b") return 1;
    if (v == "mixed" || v == "m") return 2;
    if (v == "fp16all") return 4;
    return 3;
}
int g_gemm_corr16 = corr_from_env();
}
static bool use_pair(int64_t M) { return adkf::g_gemm_impl == 2 && M > kBM; }

extern "C" int adkf_gemm_configure(int32_t impl, int32_t corr16) {
    const int prev = adkf::g_gemm_impl | (adkf::g_gemm_corr16 << 8);
    if (impl == 1 || impl == 2) adkf::g_gemm_impl = impl;
    if (corr16 >= 0 && corr16 <= 4) adkf::g_gemm_corr16 = corr16;
    return prev;
}
static int64_t out_tiles(int64_t M, int32_t N) {
    const int64_t bm = use_pair(M) ? 2 * kBM : kBM;
    return ((M + bm - 1) / bm) * ((N + kBN - 1) / kBN);
}
static int plan_splits_batched(int64_t M, int32_t N, int32_t K, int batch) {
    return plan_splits_tiles(out_tiles(M, N) * batch, (K + kBK - 1) / kBK, use_pair(M) ? 74 : 148);
}
static int plan_splits(int64_t M, int32_t N, int32_t K) { return plan_splits_batched(M, N, K, 1); }

extern "C" int64_t adkf_gemm_splitk_ws_floats(int64_t M, int32_t N, int32_t K) {
    const int s = plan_splits(M, N, K);
    return s > 1 ? (int64_t)s * M * N : 0;
}

// batch > 1: batch b reads opA at (mn + b*a_mn, k + b*a_k) and opB at (mn + b*b_mn, k + b*b_k) of the SAME 2-D arrays
// (a_rows / b_rows = number of rows of the underlying arrays along the slow dimension, for the tensor-map bounds) and
// writes D + b*d_off.  Rows / columns past M / N of a batch are computed from whatever lies there and masked in the epilogue.
static int gemm_launch(const float* A, int transA, int64_t lda, int64_t a_slow, int64_t a_fast, const float* B, int transB,
                       int64_t ldb, int64_t b_slow, int64_t b_fast, int64_t M, int32_t N, int32_t K, const float* bias,
                       int32_t act, float* D, int64_t ldd, float* splitk_ws, int64_t splitk_ws_floats, int batch, int a_mn,
                       int a_k, int b_mn, int b_k, int64_t d_off, int64_t bias_off, cudaStream_t s, const float* mask = nullptr,
                       int64_t ldm = 0, const int32_t* gate = nullptr) {
    (void)cudaGetLastError();   // a stale error left in this host thread by someone else is not this launch's failure
    CUtensorMap ma, mb;
    // make_map_2d(ptr, rows (mn extent), kdim (k extent), ld, transposed): extents of the WHOLE array
    int rc = make_map_2d(&ma, A, transA ? a_fast : a_slow, transA ? a_slow : a_fast, lda, transA ? 1 : 0);
    if (rc) return rc;
    const bool pair = use_pair(M);
    rc = make_map_2d(&mb, B, transB ? b_fast : b_slow, transB ? b_slow : b_fast, ldb, transB ? 1 : 0, pair ? kBN / 2 : kBN);
    if (rc) return rc;
    int splits = (batch > 1) ? 1 : plan_splits(M, N, K);
    if (batch > 1 && splitk_ws != nullptr)          // batched split-K: same plan for every batch, ws = [batch][splits][M][N]
        splits = plan_splits_batched(M, N, K, batch);
    if (splits > 1 && (splitk_ws == nullptr || splitk_ws_floats < (int64_t)splits * batch * M * N || bias != nullptr ||
                       act != 0 || ldd != N))
        splits = 1;                                                  // no workspace (or fused epilogue): single pass
    const int nk = (K + kBK - 1) / kBK;
    GemmEpilogue ep = {};
    ep.mode = 0; ep.M = M; ep.N = N; ep.K = K; ep.bias = bias; ep.act = act; ep.mask = mask; ep.ldm = ldm; ep.gate = gate;
    ep.kb_per_split = (nk + splits - 1) / splits;
    static const int dbg = getenv("ADKF_GEMM_DEBUG") ? atoi(getenv("ADKF_GEMM_DEBUG")) : 0;
    ep.debug = dbg;
    ep.splits = splits; ep.a_mn = a_mn; ep.a_k = a_k; ep.b_mn = b_mn; ep.b_k = b_k; ep.bias_off = bias_off;
    if (splits > 1) {
        // blockIdx.z = b * splits + split  ->  out = ws + z * M*N ; the reduce pass sums `splits` consecutive slabs
        ep.out = splitk_ws; ep.ldo = N; ep.split_stride = M * (int64_t)N; ep.d_off = 0;
    } else {
        ep.out = D; ep.ldo = ldd; ep.split_stride = 0; ep.d_off = d_off;
    }
    ep.gx = (int)((N + kBN - 1) / kBN); ep.gy = (int)((M + (pair ? 2 : 1) * kBM - 1) / ((pair ? 2 : 1) * kBM));
    const int64_t ntiles64 = (int64_t)ep.gx * ep.gy * splits * batch;
    if (ntiles64 > 0x7fffffff) return ADKF_ERR_BADARG;
    ep.ntiles = (int)ntiles64;
    dim3 grid((unsigned)std::min<int64_t>(ntiles64, num_sms()));
    if (pair) {
        // TMA-store epilogue (gemm_pair.cu): plain / bias / activation outputs without split-K whose 32-column passes are all-in or
        // all-out (N % 32 == 0), column-batched at most; everything else (D -= .., split-K slabs, odd widths) keeps gemm_store_tile
        static const int tma_env = getenv("ADKF_GEMM_TMA_STORE") ? atoi(getenv("ADKF_GEMM_TMA_STORE")) : 1;
        CUtensorMap mo = ma, mm = ma;
        const bool bias_ok = bias == nullptr || ((reinterpret_cast<uintptr_t>(bias) & 15) == 0 && bias_off % 4 == 0);
        const bool mask_ok = act < 4 || (!transA && transB && mask != nullptr && ldm % 4 == 0 && (reinterpret_cast<uintptr_t>(mask) & 15) == 0);
        if (tma_env && act >= 0 && act != 3 && act <= 5 && splits == 1 && N % 32 == 0 && bias_ok && mask_ok &&
            (int64_t)(batch - 1) * d_off + N <= ldd && M < 0x7fffffff) {
            if (make_out_map_2d(&mo, D, M, ldd, ldd) == 0 &&
                (act < 4 || make_out_map_2d(&mm, const_cast<float*>(mask), M, N, ldm) == 0))
                ep.tma_store = 1;
        }
        if ((rc = gemm_pair_launch(ma, mb, mo, mm, ep, transA, transB, s))) return rc;
    } else if (!transA && !transB) {
        if ((rc = gemm_attr<false, false>())) return rc;
        gemm_tf32x3_kernel<false, false><<<grid, kGemmThreads, kGemmSmem, s>>>(ma, mb, ep);
    } else if (!transA && transB) {
        if ((rc = gemm_attr<false, true>())) return rc;
        gemm_tf32x3_kernel<false, true><<<grid, kGemmThreads, kGemmSmem, s>>>(ma, mb, ep);
    } else {
        if ((rc = gemm_attr<true, true>())) return rc;
        gemm_tf32x3_kernel<true, true><<<grid, kGemmThreads, kGemmSmem, s>>>(ma, mb, ep);
    }
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    if (splits > 1) {
        const int64_t n = M * (int64_t)N;
        splitk_reduce_kernel<<<dim3((unsigned)((n + 255) / 256), batch), 256, 0, s>>>(splitk_ws, splits, n, D, d_off);
        adkf_count_launches(1);
        ADKF_LAUNCH_CHECK();
    }
    return 0;
}

extern "C" int adkf_gemm_tf32x3(const float* A, int32_t transA, int64_t lda, const float* B, int32_t transB, int64_t ldb,
                                int64_t M, int32_t N, int32_t K, const float* bias, int32_t act, float* D, int64_t ldd,
                                float* splitk_ws, int64_t splitk_ws_floats, void* stream) {
    if (M <= 0 || N <= 0) return 0;
    if (!A || !B || !D || K <= 0 || act < 0 || act > 3) return ADKF_ERR_BADARG;
    if (transA && !transB) return ADKF_ERR_BADARG;                  // (T,K) is not on the path
    return gemm_launch(A, transA, lda, transA ? K : M, transA ? M : K, B, transB, ldb, transB ? K : N, transB ? N : K, M, N, K,
                       bias, act, D, ldd, splitk_ws, splitk_ws_floats, 1, 0, 0, 0, 0, 0, 0, (cudaStream_t)stream);
}

// adkf_gemm_tf32x3 without split-K whose launch is skipped on the device when *gate == 0 (chol_blocked.cu)
namespace adkf {
int gemm_tf32x3_gated(const float* A, int32_t transA, int64_t lda, const float* B, int32_t transB, int64_t ldb, int64_t M, int32_t N,
                      int32_t K, int32_t act, float* D, int64_t ldd, const int32_t* gate, cudaStream_t stream) {
    if (M <= 0 || N <= 0) return 0;
    return gemm_launch(A, transA, lda, transA ? K : M, transA ? M : K, B, transB, ldb, transB ? K : N, transB ? N : K, M, N, K,
                       nullptr, act, D, ldd, nullptr, 0, 1, 0, 0, 0, 0, 0, 0, stream, nullptr, 0, gate);
}
}  // namespace adkf

// D = (A.B^T) * act'(mask): the input-gradient GEMM of a layer fused with the activation derivative of the layer below
// (mask = that layer's saved activation OUTPUT [M,N], mask_act 1 relu / 2 leaky_relu(0.01)).  No split-K, no bias.
extern "C" int adkf_gemm_tf32x3_masked(const float* A, int32_t transA, int64_t lda, const float* B, int32_t transB, int64_t ldb,
                                       int64_t M, int32_t N, int32_t K, const float* mask, int64_t ldm, int32_t mask_act,
                                       float* D, int64_t ldd, void* stream) {
    if (M <= 0 || N <= 0) return 0;
    if (!A || !B || !D || !mask || K <= 0 || (mask_act != 1 && mask_act != 2) || ldm < N) return ADKF_ERR_BADARG;
    if (transA && !transB) return ADKF_ERR_BADARG;
    return gemm_launch(A, transA, lda, transA ? K : M, transA ? M : K, B, transB, ldb, transB ? K : N, transB ? N : K, M, N, K,
                       nullptr, 3 + mask_act, D, ldd, nullptr, 0, 1, 0, 0, 0, 0, 0, 0, (cudaStream_t)stream, mask, ldm);
}

// ---- pre-split weight images (gemm_pair.cu, PRE variant) ----------------------------------------------------------------------
extern "C" int adkf_gemm_presplit(const float* W, int64_t rows, int32_t K, int64_t ldw, int32_t corr, float* hi, float* co, void* stream) {
    if (rows <= 0 || K <= 0) return 0;
    if (!W || !hi || !co || K % 32 || ldw % 2 || (reinterpret_cast<uintptr_t>(W) & 7) || (reinterpret_cast<uintptr_t>(hi) & 15) ||
        (reinterpret_cast<uintptr_t>(co) & 15) || (corr != 1 && corr != 2))
        return ADKF_ERR_BADARG;
    (void)cudaGetLastError();
    const int rc = adkf::gemm_pair_presplit(W, rows, K, ldw, corr, hi, co, (cudaStream_t)stream);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return rc;
}

extern "C" int adkf_gemm_tf32x3_pre(const float* A, int64_t lda, const float* B_hi, const float* B_co, int32_t corr, int64_t M, int32_t N,
                                    int32_t K, const float* bias, int32_t act, const float* mask, int64_t ldm, int32_t mask_act, float* D,
                                    int64_t ldd, void* stream) {
    if (M <= 0 || N <= 0) return 0;
    if (!A || !B_hi || !B_co || !D || K <= 0 || K % 32 || M <= kBM || act < 0 || act > 2 || (corr != 1 && corr != 2)) return ADKF_ERR_BADARG;
    if (mask != nullptr && (act != 0 || bias != nullptr || (mask_act != 1 && mask_act != 2) || ldm < N)) return ADKF_ERR_BADARG;
    cudaStream_t s = (cudaStream_t)stream;
    (void)cudaGetLastError();
    CUtensorMap ma, mb, mc;
    int rc = make_map_2d(&ma, A, M, K, lda, 0);
    if (rc) return rc;
    if ((rc = make_map_2d(&mb, B_hi, N, K, K, 0, kBN / 2))) return rc;
    if ((rc = make_map_2d(&mc, B_co, N, K, K, 0, kBN / 2))) return rc;
    const int eact = mask != nullptr ? 3 + mask_act : act;
    GemmEpilogue ep = {};
    ep.mode = 0; ep.M = M; ep.N = N; ep.K = K; ep.bias = bias; ep.act = eact; ep.mask = mask; ep.ldm = ldm;
    ep.kb_per_split = K / kBK; ep.splits = 1; ep.out = D; ep.ldo = ldd;
    ep.gx = (N + kBN - 1) / kBN; ep.gy = (int)((M + 2 * kBM - 1) / (2 * kBM));
    const int64_t ntiles64 = (int64_t)ep.gx * ep.gy;
    if (ntiles64 > 0x7fffffff) return ADKF_ERR_BADARG;
    ep.ntiles = (int)ntiles64;
    CUtensorMap mo = ma, mm = ma;
    const bool bias_ok = bias == nullptr || (reinterpret_cast<uintptr_t>(bias) & 15) == 0;
    const bool mask_ok = mask == nullptr || (ldm % 4 == 0 && (reinterpret_cast<uintptr_t>(mask) & 15) == 0);
    if (N % 32 == 0 && bias_ok && mask_ok && N <= ldd && make_out_map_2d(&mo, D, M, ldd, ldd) == 0 &&
        (mask == nullptr || make_out_map_2d(&mm, const_cast<float*>(mask), M, N, ldm) == 0))
        ep.tma_store = 1;
    if ((rc = adkf::gemm_pair_launch_pre(ma, mb, mc, mo, mm, ep, corr, s))) return rc;
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int64_t adkf_gemm_batched_splitk_ws_floats(int64_t M, int32_t N, int32_t K, int32_t batch) {
    const int splits = plan_splits_batched(M, N, K, batch);
    return splits > 1 ? (int64_t)splits * batch * M * N : 0;
}

extern "C" int adkf_gemm_tf32x3_batched(const float* A, int32_t transA, int64_t lda, int64_t a_slow, int64_t a_fast,
                                        int32_t a_mn_stride, int32_t a_k_stride, const float* B, int32_t transB, int64_t ldb,
                                        int64_t b_slow, int64_t b_fast, int32_t b_mn_stride, int32_t b_k_stride, int64_t M,
                                        int32_t N, int32_t K, int32_t batch, const float* bias, int64_t bias_stride,
                                        int32_t act, float* D, int64_t ldd, int64_t d_stride, float* splitk_ws,
                                        int64_t splitk_ws_floats, void* stream) {
    if (M <= 0 || N <= 0 || batch <= 0) return 0;
    if (!A || !B || !D || K <= 0 || act < 0 || act > 3) return ADKF_ERR_BADARG;
    if (transA && !transB) return ADKF_ERR_BADARG;
    if (batch > 1 && (a_k_stride || b_k_stride) && K % kBK) return ADKF_ERR_BADARG;   // a partial k-block would read the next batch
    return gemm_launch(A, transA, lda, a_slow, a_fast, B, transB, ldb, b_slow, b_fast, M, N, K, bias, act, D, ldd, splitk_ws,
                       splitk_ws_floats, batch, a_mn_stride, a_k_stride, b_mn_stride, b_k_stride, d_stride, bias_stride,
                       (cudaStream_t)stream);
}

extern "C" int adkf_linear_tf32x3(const float* X, int64_t M, int32_t K, int64_t ldx, const float* W, int32_t N,
                                  int64_t ldw, const float* bias, int32_t act, float* Y, int64_t ldy, void* stream) {
    return adkf_gemm_tf32x3(X, 0, ldx, W, 0, ldw, M, N, K, bias, act, Y, ldy, nullptr, 0, stream);
}

extern "C" int adkf_gram_pairstat_tc(int32_t stat, const float* Z, int32_t d, int32_t T, int64_t total_rows,
                                     const int32_t* z_off, const int32_t* n_sup, const int64_t* d2_off, int32_t max_n,
                                     int32_t qq_block, float* mu_ws, float* nrm_ws, float* D2, void* stream) {
    if (T <= 0 || max_n <= 0) return 0;
    if (!Z || !z_off || !n_sup || !d2_off || !mu_ws || !nrm_ws || !D2 || d <= 0 || d % 4) return ADKF_ERR_BADARG;
    if (stat < ADKF_STAT_SQDIST || stat > ADKF_STAT_TANIMOTO) return ADKF_ERR_BADARG;
    cudaStream_t s = (cudaStream_t)stream;
    (void)cudaGetLastError();
    CUtensorMap mz;
    int rc = make_map_2d(&mz, Z, total_rows, d, d, 0);
    if (rc) return rc;
    rc = gemm_attr<false, false>();
    if (rc) return rc;
    launch_center_stats(Z, d, T, z_off, n_sup, max_n, mu_ws, nrm_ws, s, stat == ADKF_STAT_SQDIST);
    GemmEpilogue ep = {};
    ep.mode = 1; ep.out = D2; ep.K = d; ep.z_off = z_off; ep.n_sup = n_sup; ep.d2_off = d2_off; ep.mu = mu_ws;
    ep.nrm = nrm_ws; ep.qq_block = qq_block; ep.kb_per_split = (d + kBK - 1) / kBK; ep.stat = stat;
    const int nt = (max_n + kBM - 1) / kBM;
    ep.gx = nt; ep.gy = nt; ep.ntiles = nt * nt * T;
    gemm_tf32x3_kernel<false, false><<<dim3((unsigned)std::min(ep.ntiles, num_sms())), kGemmThreads, kGemmSmem, s>>>(mz, mz, ep);
    if (stat != ADKF_STAT_SQDIST && !qq_block) launch_pairstat_diag(stat, T, z_off, d2_off, max_n, nrm_ws, D2, s);
    adkf_count_launches(3);
    ADKF_LAUNCH_CHECK();
    return 0;
}

extern "C" int adkf_gram_sqdist_tc(const float* Z, int32_t d, int32_t T, int64_t total_rows, const int32_t* z_off,
                                   const int32_t* n_sup, const int64_t* d2_off, int32_t max_n, int32_t qq_block,
                                   float* mu_ws, float* nrm_ws, float* D2, void* stream) {
    return adkf_gram_pairstat_tc(ADKF_STAT_SQDIST, Z, d, T, total_rows, z_off, n_sup, d2_off, max_n, qq_block, mu_ws,
                                 nrm_ws, D2, stream);
}
