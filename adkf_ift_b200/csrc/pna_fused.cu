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
// Rows of different edge types share a tile, so the contraction is K-concatenated over the types: the A operand has one 64-wide slot
// per type, a row fills the slot of its own type and holds zeros in the others, and slot t is multiplied with W[type t] -- every row
// gets exactly its own products (plus exact zeros) in ONE accumulator pair; slots of types absent from the tile are skipped.
//
// Numerics.  Identical to the unfused path bit for bit in the forward: same hi/lo split, same MMA order per accumulator
// (hi.hi -> big; lo.hi, hi.lo -> small), pre = (big + small) + bias with round-to-nearest adds, same reduction order per node.
//
// CTA = 640 threads:  warp 0 MMA issuer | warp 1 TMEM allocator | warps 4-7 gather + split (row = TMEM lane) |
//                     warps 8-19 three epilogue warpgroups taking tiles round-robin: drain (tcgen05.ld) to shared memory, then the
//                     per-node reduction / d-pre chain rule.  The per-node part is latency-bound arithmetic (short dependent chains per
//                     node), so it is what the warps are spent on; the MMA warp gets its accumulators back right after the drain.
// TMEM (512 columns): accumulators big 0..63, small 64..127; A operand slot of type t at 128 + 128 t = [hi k0-31 | lo k0-31 | hi k32-63 | lo k32-63].
#include <cuda.h>
#include <algorithm>
#include "gemm_common.cuh"

namespace adkf {
namespace pf {

constexpr int kRows = 128, kStride = 112;
constexpr int kThreads = 640;                        // 5 warpgroups: control | gather + split | 3 x drain + segment epilogue
constexpr int kEpi = 3;                              // epilogue warpgroups (tiles round-robin)
constexpr int kWTile = 64 * 32 * 4;                  // one [64 n x 32 k] K-major SWIZZLE_128B tile
constexpr int kWBytes = 12 * kWTile;                 // [type 3][hi|lo][k-block 2]
constexpr int kStgStride = 65;                       // floats per staged row (bank-conflict-free for lane = row and lane = column)
constexpr int kStgBytes = kRows * kStgStride * 4;    // 33,280 per epilogue warpgroup
constexpr int kMisc = 5120;
constexpr size_t kSmem = 1024 + kWBytes + kEpi * kStgBytes + kMisc;
constexpr int kTmemCols = 512, kACol = 128;   // accumulators: big 0..63, small 64..127; A operand: 3 types x 128 columns
constexpr uint32_t kIdesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(64 >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
constexpr float kSmall = 1e-7f;
// misc block (byte offsets)
constexpr int oBias = 0;          // [2][192] float   (double-buffered by segment)
constexpr int oRowType = 1536;    // [2][128] u8      (by tile parity)
constexpr int oPresent = 1792;    // [2][4] u32
constexpr int oBars = 1824;       // a_full, a_free, acc_free, acc_full[3]
constexpr int oTmemSlot = 1872;
constexpr int oRp = 1888;         // [3][132] int     (per epilogue warpgroup)
constexpr int oTs = 3472;         // [3][128] int

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
    unsigned long long* dbg;   // development probe: per-CTA cycle counters [gridDim.x][16] (NULL = off)
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
// MMA / commit issued by ONE elected lane of a CONVERGED warp (elect.sync): under `if (lane == 0)` ptxas has to assume divergence
// and wraps every tcgen05.mma in an ELECT + branch loop (~8 instructions, ~60 cycles per MMA: with N = 64 the issue, not the
// tensor pipe, paced this kernel).
__device__ __forceinline__ void umma_tf32_ts_elect(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p, q;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
        ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(acc)
        : "memory");
}
__device__ __forceinline__ void umma_commit_elect(uint32_t bar) {
    asm volatile(
        "{\n\t.reg .pred q;\n\t"
        "elect.sync _|q, 0xffffffff;\n\t"
        "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}"
        ::"r"(bar)
        : "memory");
}
__device__ __forceinline__ long long clk() { return clock64(); }
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void bar_sync_wg(int id) { asm volatile("bar.sync %0, 128;" ::"r"(id) : "memory"); }
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
        ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
          "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}

// 32 gathered values of one row -> hi / lo tf32 in 64 consecutive TMEM columns (hi 0..31, lo 32..63), 16 values at a time; rows that
// do not own the slot (`mine` false) store zeros.  tcgen05.st is warp-collective, so the per-lane choice is a select, not a branch.
__device__ __forceinline__ void masked32_to_tmem(const float4 (&x)[8], const bool mine, uint32_t taddr) {
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        uint32_t hi[16], lo[16];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float4 q = x[half * 4 + j];
            const float v[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float hv = tf32_hi(v[e]);
                hi[4 * j + e] = mine ? __float_as_uint(hv) : 0u;
                lo[4 * j + e] = mine ? __float_as_uint(v[e] - hv) : 0u;
            }
        }
        tmem_st16(taddr + half * 16, hi);
        tmem_st16(taddr + 32 + half * 16, lo);
    }
}

// CTA b: tower b & 3, tile range `slot` = b >> 2 of gridDim.x / 4; three segments (one per 64-channel chunk) over the same tiles,
// so every CTA does the same mix of epilogue kinds (the mean + std chunk costs about twice the sum chunk).
template <bool BWD>
__global__ void __launch_bounds__(kThreads, 1) pna_fused_kernel(const Args a) {
    extern __shared__ unsigned char smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    unsigned char* base_ptr = smem_raw + (base - raw);
    unsigned char* w_ptr = base_ptr;                                         // 96 KB of weight tiles
    float* stg = reinterpret_cast<float*>(base_ptr + kWBytes);               // kEpi x [128][65]
    unsigned char* misc = base_ptr + kWBytes + kEpi * kStgBytes;
    float* bias_s = reinterpret_cast<float*>(misc + oBias);
    volatile uint8_t* row_type = reinterpret_cast<volatile uint8_t*>(misc + oRowType);
    volatile uint32_t* present = reinterpret_cast<volatile uint32_t*>(misc + oPresent);
    const uint32_t bars = base + kWBytes + kEpi * kStgBytes + oBars;
    const uint32_t a_full = bars, a_free = bars + 8, acc_free = bars + 16;
    auto acc_full = [&](uint32_t g) { return bars + 24 + 8 * g; };
    volatile uint32_t* tmem_slot = reinterpret_cast<volatile uint32_t*>(misc + oTmemSlot);
    int* rp_s = reinterpret_cast<int*>(misc + oRp);
    int* ts_s = reinterpret_cast<int*>(misc + oTs);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        mbar_init(a_full, 4);
        mbar_init(a_free, 1);
        mbar_init(acc_free, 4);
        for (uint32_t e = 0; e < (uint32_t)kEpi; ++e) mbar_init(acc_full(e), 1);
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

    const int tower = blockIdx.x & 3, slot = blockIdx.x >> 2, nslots = gridDim.x >> 2;
    const int per = (a.ntiles + nslots - 1) / nslots;
    const int tile0 = min(slot * per, a.ntiles), tile1 = min(tile0 + per, a.ntiles);
    const int nts = tile1 - tile0;                       // tiles per segment; 3 segments (chunks) per CTA

    if (warp < 4) {
        reg_dec<64>();                                        // pool: 640 threads x 96 = 128 x 480 = 64 + 144 + 3 x 88 (+ 8 spare)
        if (warp == 0) {
            // ===== MMA issuer =====
            long long tmr[3] = {0, 0, 0};
            const long long cstart = clk();
            for (int seg = 0; seg < 3; ++seg)
                for (int i = 0; i < nts; ++i) {
                    const uint32_t cur = (uint32_t)(seg * nts + i), par = cur & 1;
                    const long long c0 = clk();
                    mbar_wait(a_full, par);
                    const long long c1 = clk();
                    mbar_wait(acc_free, par ^ 1);
                    const long long c2 = clk();
                    tmr[0] += c1 - c0; tmr[1] += c2 - c1;
                    tc_fence_after();
                    const uint32_t pres = present[par * 4 + 0] | present[par * 4 + 1] | present[par * 4 + 2] | present[par * 4 + 3];
                    {
                        // K = 3 x 64: k-block pair t multiplies the rows of edge type t (the other rows hold zeros there) with W[t];
                        // every row's products land in ONE accumulator pair in the order the unfused GEMM adds them (+0 elsewhere)
                        const uint32_t big = tmem, small = tmem + 64;
                        uint32_t first = 0u;                                  // 0 until the first MMA of the tile has been issued
#pragma unroll
                        for (int t = 0; t < 3; ++t) {
                            if (!((pres >> t) & 1u)) continue;
#pragma unroll
                            for (int kb = 0; kb < 2; ++kb) {
                                const uint64_t b_hi = smem_desc_k128(base + ((t * 2 + 0) * 2 + kb) * kWTile);
                                const uint64_t b_lo = smem_desc_k128(base + ((t * 2 + 1) * 2 + kb) * kWTile);
                                const uint32_t a_hi = tmem + (uint32_t)(kACol + t * 128 + kb * 64), a_lo = a_hi + 32;
#pragma unroll
                                for (int k = 0; k < 4; ++k) {
                                    const uint64_t adv = (uint64_t)(2 * k);
                                    const uint32_t ak = (uint32_t)(8 * k);
                                    umma_tf32_ts_elect(big, a_hi + ak, b_hi + adv, kIdesc, first);
                                    umma_tf32_ts_elect(small, a_lo + ak, b_hi + adv, kIdesc, first);
                                    umma_tf32_ts_elect(small, a_hi + ak, b_lo + adv, kIdesc, 1u);
                                    first = 1u;
                                }
                            }
                        }
                        umma_commit_elect(a_free);
                        umma_commit_elect(acc_full(cur % kEpi));
                    }
                    __syncwarp();
                }
            if (a.dbg != nullptr && lane == 0) {
                a.dbg[blockIdx.x * 16 + 0] = (unsigned long long)tmr[0];            // MMA warp waiting for the A operand
                a.dbg[blockIdx.x * 16 + 1] = (unsigned long long)tmr[1];            // ... for the accumulators
                a.dbg[blockIdx.x * 16 + 2] = (unsigned long long)(clk() - cstart);  // role total
            }
        }
    } else if (warp < 8) {
        // ===== gather + split: row = TMEM lane; also loads the weight slice of each segment =====
        // latency chain per tile: the edge indices are fetched one tile ahead, the 16 row loads of a tile are in flight while
        // the MMAs of the previous tile run (the wait for the A operand comes after the loads were issued)
        reg_inc<144>();
        const int row = (warp - 4) * 32 + lane;
        const uint32_t tmem_a = tmem + ((uint32_t)((warp - 4) * 32) << 16) + (uint32_t)kACol;
        int ty_prev = -1;                                     // -1: tensor memory not initialised yet, write every slot
        long long tg[3] = {0, 0, 0};
        const long long gstart = clk();
        for (int seg = 0; seg < 3; ++seg) {
            const int chunk = seg;
            int src = 0, tgt = 0, ty = 3;
            if (nts > 0) {
                const int64_t p = (int64_t)tile0 * kStride + row;
                if (p < a.E) { src = a.col_src[p]; tgt = a.col_tgt[p]; ty = a.col_type[p]; }
            }
            for (int i = 0; i < nts; ++i) {
                const uint32_t cur = (uint32_t)(seg * nts + i), par = cur & 1;
                const bool valid = ty < 3;
                float4 xs[8], xt[8];
                const float4* hs = reinterpret_cast<const float4*>(a.h + (size_t)src * 128 + tower * 32);
                const float4* ht = reinterpret_cast<const float4*>(a.h + (size_t)tgt * 128 + tower * 32);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    xs[j] = valid ? __ldg(hs + j) : make_float4(0.f, 0.f, 0.f, 0.f);
                    xt[j] = valid ? __ldg(ht + j) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
                const int ty_cur = ty;
                src = 0; tgt = 0; ty = 3;
                if (i + 1 < nts) {
                    const int64_t p = (int64_t)(tile0 + i + 1) * kStride + row;
                    if (p < a.E) { src = a.col_src[p]; tgt = a.col_tgt[p]; ty = a.col_type[p]; }
                }
                const uint32_t m0 = __ballot_sync(0xffffffffu, ty_cur == 0), m1 = __ballot_sync(0xffffffffu, ty_cur == 1),
                               m2 = __ballot_sync(0xffffffffu, ty_cur == 2);
                const long long g0 = clk();
                mbar_wait(a_free, par ^ 1);              // the MMAs of the previous tile have read the A operand (and the weights)
                const long long g1 = clk();
                tg[0] += g1 - g0;
                tc_fence_after();
                if (i == 0) {
                    // weights of (tower, chunk) for the 3 edge types: fp32 -> hi / lo tf32, K-major SWIZZLE_128B
                    const int t128 = threadIdx.x - 128;
                    for (int f = t128; f < 3 * 64 * 16; f += 128) {
                        const int t = f >> 10, n = (f >> 4) & 63, j16 = f & 15, kb = j16 >> 3, j = j16 & 7;
                        const float4 v = __ldg(reinterpret_cast<const float4*>(a.W + ((size_t)((tower * 3 + t) * 192 + chunk * 64 + n)) * 64) + j16);
                        float4 hi, lo;
                        split4(v, hi, lo);
                        const int off = n * 128 + ((j ^ (n & 7)) << 4);
                        *reinterpret_cast<float4*>(w_ptr + ((t * 2 + 0) * 2 + kb) * kWTile + off) = hi;
                        *reinterpret_cast<float4*>(w_ptr + ((t * 2 + 1) * 2 + kb) * kWTile + off) = lo;
                    }
                    for (int f = t128; f < 192; f += 128)
                        bias_s[(seg & 1) * 192 + f] = a.b[(size_t)(tower * 3 + (f >> 6)) * 192 + chunk * 64 + (f & 63)];
                    fence_async_smem();
                }
                row_type[par * 128 + row] = (uint8_t)ty_cur;
                if (lane == 0) present[par * 4 + (warp - 4)] = (m0 ? 1u : 0u) | (m1 ? 2u : 0u) | (m2 ? 4u : 0u);
                // slot t of the A operand holds this row's (hi, lo) when the row has type t and zeros otherwise.  A slot is rewritten
                // only when some row of this warp needs it: its type is t now, or was t in the previous tile (to clear it).
#pragma unroll
                for (int t = 0; t < 3; ++t) {
                    const bool need = __ballot_sync(0xffffffffu, ty_cur == t || ty_prev == t || ty_prev < 0) != 0u;
                    if (!need) continue;
                    masked32_to_tmem(xs, ty_cur == t, tmem_a + t * 128);
                    masked32_to_tmem(xt, ty_cur == t, tmem_a + t * 128 + 64);
                }
                ty_prev = ty_cur;
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(a_full);
                tg[1] += clk() - g1;                     // (includes the wait for the row loads)
            }
        }
        if (a.dbg != nullptr && threadIdx.x == 128) {
            a.dbg[blockIdx.x * 16 + 3] = (unsigned long long)tg[0];                 // gather waiting for the MMAs
            a.dbg[blockIdx.x * 16 + 4] = (unsigned long long)tg[1];                 // load wait + split + tcgen05.st
            a.dbg[blockIdx.x * 16 + 5] = (unsigned long long)(clk() - gstart);
        }
    } else {
        // ===== drain + segment epilogue: two warpgroups, alternating tiles =====
        // While one warpgroup walks the nodes of tile i (the long part), the other one drains tile i+1 from tensor memory, so the
        // accumulators are handed back to the MMA warp as soon as they have been copied to shared memory.
        reg_dec<88>();
        const uint32_t g = (uint32_t)(warp - 8) >> 2;        // epilogue warpgroup
        const int q = warp & 3, row = q * 32 + lane;
        const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
        float* st = stg + (size_t)g * (kRows * kStgStride);
        int* rp = rp_s + g * 132;
        int* tsr = ts_s + g * 128;
        const float* sc = st + lane;                          // this lane's two columns: lane and lane + 32
        long long td[4] = {0, 0, 0, 0};
        const long long dstart = clk();
        for (int seg = 0; seg < 3; ++seg) {
            const int chunk = seg;
            const float* bias_c = bias_s + (seg & 1) * 192;
            for (int i = 0; i < nts; ++i) {
                const uint32_t cur = (uint32_t)(seg * nts + i), par = cur & 1;
                if (cur % kEpi != g) continue;
                const int tile = tile0 + i;
                const int64_t row0 = (int64_t)tile * kStride;
                const int n0 = a.tile_node0[tile], n1 = a.tile_node0[tile + 1], nn = n1 - n0;
                const bool rp_smem = nn <= 127;              // (more owned nodes than that = a run of isolated nodes: read rowptr directly)
                int rp_mine = 0, ts_mine = 0;
                if (rp_smem && row <= nn) rp_mine = min(max((int)(a.rowptr[n0 + row] - row0), 0), kRows);
                if (BWD) {
                    const int64_t p = row0 + row;
                    ts_mine = (p < a.E) ? a.csr2ts[p] : 0;
                    // node-level inputs of the chain rule -> L1, so that the per-node loop below does not wait on DRAM
                    if (row < nn) {
                        const size_t ob = (size_t)(n0 + row) * 1024 + (size_t)tower * 256;
                        if (chunk == 0) { prefetch_l1(a.d_out + ob); prefetch_l1(a.d_out + ob + 32); }
                        else if (chunk == 1) {
#pragma unroll
                            for (int c = 64; c < 192; c += 32) { prefetch_l1(a.d_out + ob + c); prefetch_l1(a.out + ob + c); }
                        } else { prefetch_l1(a.d_out + ob + 192); prefetch_l1(a.d_out + ob + 224); }
                    }
                }
                const long long d0 = clk();
                bar_sync_wg(1 + (int)g);                      // every warp of this group is done with the previous staged tile
                const long long d1 = clk();
                mbar_wait(acc_full(g), (cur / kEpi) & 1);
                const long long d2 = clk();
                td[0] += d1 - d0; td[1] += d2 - d1;
                tc_fence_after();
                const int rt = row_type[par * 128 + row];
#pragma unroll
                for (int pp = 0; pp < 2; ++pp) {
                    uint32_t vb[2][16], vs[2][16];
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        tmem_ld16_nowait(trow + (uint32_t)((2 * pp + h) * 16), vb[h]);
                        tmem_ld16_nowait(trow + (uint32_t)(64 + (2 * pp + h) * 16), vs[h]);
                    }
                    tmem_ld_wait();
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        touch16(vb[h]);
                        touch16(vs[h]);
                        if (rt < 3) {
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                const int c = (2 * pp + h) * 16 + j;
                                const float pre = (__uint_as_float(vb[h][j]) + __uint_as_float(vs[h][j])) + bias_c[rt * 64 + c];
                                st[row * kStgStride + c] = BWD ? pre : fmaxf(pre, 0.f);
                            }
                        }
                    }
                }
                rp[row] = rp_mine;
                if (BWD) tsr[row] = ts_mine;
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(acc_free);        // the accumulators may be overwritten by the next tile
                bar_sync_wg(1 + (int)g);                      // the staged tile and its metadata are complete
                const long long d3 = clk();
                td[2] += d3 - d2;

                // ---- per-node epilogue: warp q takes nodes q, q + 4, ...; lane owns channels lane and lane + 32 of the chunk ----
                // A node has 1-4 rows almost always: its first four rows are fetched with predicated, mutually independent loads
                // (so the short dependent chains of a node overlap instead of waiting on each other), the rest in a loop.  Sums run in
                // row order (adding +0.0f for an absent row changes nothing), so the result is the one of pna_aggregate_fwd_kernel.
                auto node_rows = [&](int vi, int& r0, int& r1) {
                    if (rp_smem) { r0 = rp[vi]; r1 = rp[vi + 1]; }
                    else {
                        const int b0 = a.rowptr[n0 + vi], b1 = a.rowptr[n0 + vi + 1];
                        r0 = min(max((int)(b0 - row0), 0), kRows); r1 = min(max((int)(b1 - row0), 0), kRows);
                    }
                };
                if (!BWD) {
                    float* o = a.out + (size_t)(n0 + q) * 1024 + (size_t)tower * 256 + lane;
                    uint8_t* xc = a.aux_cnt + (size_t)(n0 + q) * 256 + tower * 64 + lane;
                    uint8_t* xa = a.aux_arg + (size_t)(n0 + q) * 256 + tower * 64 + lane;
                    for (int vi = q; vi < nn; vi += 4, o += 4 * 1024, xc += 4 * 256, xa += 4 * 256) {
                        int r0, r1;
                        node_rows(vi, r0, r1);
                        const float* p = sc + r0 * kStgStride;
                        float v0[4], v1[4];
                        bool on[4];
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            on[k] = r0 + k < r1;
                            v0[k] = on[k] ? p[k * kStgStride] : 0.f;
                            v1[k] = on[k] ? p[k * kStgStride + 32] : 0.f;
                        }
                        if (chunk == 0) {
                            float s0 = 0.f, s1 = 0.f;
#pragma unroll
                            for (int k = 0; k < 4; ++k) { s0 += v0[k]; s1 += v1[k]; }
                            for (int r = r0 + 4; r < r1; ++r) { s0 += sc[r * kStgStride]; s1 += sc[r * kStgStride + 32]; }
                            o[0] = s0; o[32] = s1;
                        } else if (chunk == 1) {
                            float m0 = 0.f, m1 = 0.f;
#pragma unroll
                            for (int k = 0; k < 4; ++k) { m0 += v0[k]; m1 += v1[k]; }
                            for (int r = r0 + 4; r < r1; ++r) { m0 += sc[r * kStgStride]; m1 += sc[r * kStgStride + 32]; }
                            const float fd = (float)max(r1 - r0, 1);
                            const float mean0 = m0 / fd, mean1 = m1 / fd;                 // true division, like scatter_mean
                            // no FMA contraction: m^2 - mean^2 must be exactly 0 when m == mean (degree-1 rows), as in torch
                            const float sq0 = __fmul_rn(mean0, mean0), sq1 = __fmul_rn(mean1, mean1);
                            float s0 = 0.f, s1 = 0.f;
                            int c0 = 0, c1 = 0;
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                if (on[k]) {
                                    const float d0 = __fsub_rn(__fmul_rn(v0[k], v0[k]), sq0), d1 = __fsub_rn(__fmul_rn(v1[k], v1[k]), sq1);
                                    if (d0 > 0.f) { s0 += d0; ++c0; }
                                    if (d1 > 0.f) { s1 += d1; ++c1; }
                                    s0 += kSmall; s1 += kSmall;
                                }
                            for (int r = r0 + 4; r < r1; ++r) {
                                const float a0 = sc[r * kStgStride], a1 = sc[r * kStgStride + 32];
                                const float d0 = __fsub_rn(__fmul_rn(a0, a0), sq0), d1 = __fsub_rn(__fmul_rn(a1, a1), sq1);
                                if (d0 > 0.f) { s0 += d0; ++c0; }
                                if (d1 > 0.f) { s1 += d1; ++c1; }
                                s0 += kSmall; s1 += kSmall;
                            }
                            o[64] = mean0; o[96] = mean1;
                            o[128] = sqrtf(s0); o[160] = sqrtf(s1);
                            xc[0] = (uint8_t)min(c0, 255); xc[32] = (uint8_t)min(c1, 255);
                        } else {
                            float x0 = 0.f, x1 = 0.f;
                            int g0 = 0, g1 = 0;
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                if (on[k]) {
                                    if (k == 0 || v0[k] > x0) { x0 = v0[k]; g0 = k; }
                                    if (k == 0 || v1[k] > x1) { x1 = v1[k]; g1 = k; }
                                }
                            for (int r = r0 + 4; r < r1; ++r) {
                                const float a0 = sc[r * kStgStride], a1 = sc[r * kStgStride + 32];
                                if (a0 > x0) { x0 = a0; g0 = r - r0; }
                                if (a1 > x1) { x1 = a1; g1 = r - r0; }
                            }
                            o[192] = x0; o[224] = x1;                                    // (no rows: 0)
                            xa[0] = (uint8_t)min(g0, 255); xa[32] = (uint8_t)min(g1, 255);
                        }
                    }
                } else {
                    const float* go = a.d_out + (size_t)(n0 + q) * 1024 + (size_t)tower * 256 + lane;
                    const float* fo = a.out + (size_t)(n0 + q) * 1024 + (size_t)tower * 256 + lane;
                    const uint8_t* xc = a.aux_cnt + (size_t)(n0 + q) * 256 + tower * 64 + lane;
                    const uint8_t* xa = a.aux_arg + (size_t)(n0 + q) * 256 + tower * 64 + lane;
                    float* dp = a.d_pre + (size_t)tower * 192 + (size_t)chunk * 64 + lane;
                    for (int vi = q; vi < nn; vi += 4, go += 4 * 1024, fo += 4 * 1024, xc += 4 * 256, xa += 4 * 256) {
                        int r0, r1;
                        node_rows(vi, r0, r1);
                        if (r1 <= r0) continue;
                        // node-level inputs first (L1 hits thanks to the prefetch above), then the staged rows
                        float ga0, ga1, gb0 = 0.f, gb1 = 0.f, mean0 = 0.f, mean1 = 0.f, sd0 = 1.f, sd1 = 1.f;
                        int ax0 = 0, ax1 = 0;
                        if (chunk == 0) { ga0 = go[0]; ga1 = go[32]; }
                        else if (chunk == 1) {
                            ga0 = go[64]; ga1 = go[96]; gb0 = go[128]; gb1 = go[160];
                            mean0 = fo[64]; mean1 = fo[96]; sd0 = fo[128]; sd1 = fo[160];
                            ax0 = xc[0]; ax1 = xc[32];
                        } else { ga0 = go[192]; ga1 = go[224]; ax0 = xa[0]; ax1 = xa[32]; }
                        const float* p = sc + r0 * kStgStride;
                        float v0[4], v1[4];
                        int tr[4];
                        bool on[4];
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            on[k] = r0 + k < r1;
                            v0[k] = on[k] ? p[k * kStgStride] : 0.f;
                            v1[k] = on[k] ? p[k * kStgStride + 32] : 0.f;
                            tr[k] = on[k] ? tsr[r0 + k] : 0;
                        }
                        if (chunk == 0) {
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                if (on[k]) {
                                    float* d = dp + (size_t)tr[k] * 768;
                                    d[0] = (v0[k] > 0.f) ? ga0 : 0.f;
                                    d[32] = (v1[k] > 0.f) ? ga1 : 0.f;
                                }
                            for (int r = r0 + 4; r < r1; ++r) {
                                float* d = dp + (size_t)tsr[r] * 768;
                                d[0] = (sc[r * kStgStride] > 0.f) ? ga0 : 0.f;
                                d[32] = (sc[r * kStgStride + 32] > 0.f) ? ga1 : 0.f;
                            }
                        } else if (chunk == 1) {
                            const float invdeg = 1.f / (float)(r1 - r0);
                            const float cc0 = 0.5f * gb0 / sd0, cc1 = 0.5f * gb1 / sd1;
                            const float sq0 = __fmul_rn(mean0, mean0), sq1 = __fmul_rn(mean1, mean1);
                            const float cp0 = (float)ax0, cp1 = (float)ax1;
                            auto one = [&](float a0, float a1, int t) {
                                const float p0 = fmaxf(a0, 0.f), p1 = fmaxf(a1, 0.f);
                                const float i0 = (__fsub_rn(__fmul_rn(p0, p0), sq0) > 0.f) ? 1.f : 0.f;
                                const float i1 = (__fsub_rn(__fmul_rn(p1, p1), sq1) > 0.f) ? 1.f : 0.f;
                                const float dm0 = ga0 * invdeg + cc0 * (2.f * p0 * i0 - 2.f * mean0 * cp0 * invdeg);
                                const float dm1 = ga1 * invdeg + cc1 * (2.f * p1 * i1 - 2.f * mean1 * cp1 * invdeg);
                                float* d = dp + (size_t)t * 768;
                                d[0] = (a0 > 0.f) ? dm0 : 0.f;
                                d[32] = (a1 > 0.f) ? dm1 : 0.f;
                            };
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                if (on[k]) one(v0[k], v1[k], tr[k]);
                            for (int r = r0 + 4; r < r1; ++r) one(sc[r * kStgStride], sc[r * kStgStride + 32], tsr[r]);
                        } else {
                            auto one = [&](float a0, float a1, int t, int idx) {
                                const bool h0 = a0 > 0.f && (idx == ax0 || (idx >= 255 && ax0 == 255));
                                const bool h1 = a1 > 0.f && (idx == ax1 || (idx >= 255 && ax1 == 255));
                                float* d = dp + (size_t)t * 768;
                                d[0] = h0 ? ga0 : 0.f;
                                d[32] = h1 ? ga1 : 0.f;
                            };
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                if (on[k]) one(v0[k], v1[k], tr[k], k);
                            for (int r = r0 + 4; r < r1; ++r) one(sc[r * kStgStride], sc[r * kStgStride + 32], tsr[r], r - r0);
                        }
                    }
                }
                td[3] += clk() - d3;
            }
        }
        tc_fence_before();
        if (a.dbg != nullptr && (threadIdx.x == 256 || threadIdx.x == 384)) {      // (the first two groups)
            unsigned long long* o = a.dbg + blockIdx.x * 16 + 6 + 5 * g;
            o[0] = (unsigned long long)td[0];                                       // waiting for the other warps of the group
            o[1] = (unsigned long long)td[1];                                       // waiting for the accumulators
            o[2] = (unsigned long long)td[2];                                       // TMEM -> shared memory
            o[3] = (unsigned long long)td[3];                                       // per-node loop
            o[4] = (unsigned long long)(clk() - dstart);
        }
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
    // 4 towers x tile ranges: a multiple of 4 CTAs, at most one per SM
    const int slots = std::max(1, std::min(a.ntiles, num_sms() / 4));
    const unsigned grid = (unsigned)(4 * slots);
    pna_fused_kernel<BWD><<<grid, kThreads, kSmem, s>>>(a);
    adkf_count_launches(1);
    ADKF_LAUNCH_CHECK();
    return 0;
}

}  // namespace pf
}  // namespace adkf

using namespace adkf;

static unsigned long long* g_pna_fused_dbg = nullptr;
// development probe: device buffer of 148 x 16 cycle counters filled by the next launches (NULL switches it off)
extern "C" void adkf_pna_fused_set_debug(void* buf) { g_pna_fused_dbg = (unsigned long long*)buf; }

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
    a.out = out; a.aux_cnt = aux_cnt; a.aux_arg = aux_arg; a.dbg = g_pna_fused_dbg;
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
    a.d_out = d_out; a.d_pre = d_pre; a.dbg = g_pna_fused_dbg;
    return pf::launch<true>(a, (cudaStream_t)stream);
}
