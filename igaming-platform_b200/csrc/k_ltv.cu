// k_ltv.cu — LTV head (SURVEY K8 + K9), sm_100a: 256 -> 1024 -> 1024 -> 1024 -> 1 MLP in bf16 with fp32
// accumulation on the 5th-generation tensor cores, then the post-model steps of
// LTVPredictor.Predict (prediction/ltv.go:124-139) in fp64.
//
// k_mlp_gemm — one layer, out[M,N] = relu(A[M,K] * W[N,K]^T + bias) rounded to bf16.
//   Persistent, one CTA per SM, 128 x 256 output tiles, warp-specialised:
//     warp 0 (one lane)  TMA producer: cp.async.bulk.tensor 2D loads of the A (128 x 64) and W
//                        (256 x 64) k-blocks into a 3-stage 128B-swizzled smem ring (SASS UTMALDG)
//     warp 1 (one lane)  MMA issuer: tcgen05.mma.cta_group::1.kind::f16, M=128 N=256 K=16, smem
//                        descriptors advanced 32 B per K step, accumulators in TMEM (2 x 256 columns,
//                        double buffered so the epilogue of tile i overlaps the MMAs of tile i+1);
//                        tcgen05.commit releases smem stages / publishes the accumulator (SASS UTCHMMA)
//     warp 2             TMEM allocation (512 columns) and release
//     warps 4-7 (4-11)   epilogue: tcgen05.ld 32x32b.x32 (SASS LDTM) -> +bias, ReLU, bf16 -> swizzled smem
//                        staging -> cp.async.bulk.tensor store of 64-column x 32-row boxes (SASS UTMASTG)
//   Synchronisation is mbarrier-only inside the tile loop (full/empty per smem stage, full/empty per
//   TMEM accumulator).
//   Layer 3 runs the kHead variant: a CTA owns whole 128-row tiles and walks their 4 column tiles in order;
//   the epilogue thread that owns a row accumulates layer 4 (1024 -> 1) on the bf16-rounded activations in a
//   fixed order and finishes with the fp64 epilogue (churn, adjusted LTV, segment, survival,
//   next-best-action, confidence; prediction/ltv.go:228-382) — the layer-3 activations never reach HBM.
#include <cuda.h>
#include <cuda_bf16.h>

#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "igs_internal.cuh"
#include "ltv_math.cuh"

namespace {

// ------------------------------------------------------------------------------------------------
// PTX helpers
// ------------------------------------------------------------------------------------------------
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
// bounded: a hand-over that never comes is a bug (a lost arrive between the CTAs of a pair, say), and a trap beats a hung GPU
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok, spins = 0;
    do {
        asm volatile(
            "{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            " selp.u32 %0, 1, 0, p;\n}\n"
            : "=r"(ok)
            : "r"(bar), "r"(parity)
            : "memory");
        if (!ok && ++spins > (1u << 26)) __trap();
    } while (!ok);
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* tm, int c0, int c1, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
// The load of one CTA of a pair (`cta_group::2`): data into this CTA's shared memory, the bytes credited to a barrier
// that may live in the pair's other CTA (`bar` is a shared::cluster address, see mapa_rank0).
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* tm, int c0, int c1, uint32_t bar) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}
// shared::cluster address of the same shared-memory offset in CTA 0 of the cluster
__device__ __forceinline__ uint32_t mapa_rank0(uint32_t addr) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, 0;" : "=r"(r) : "r"(addr));
    return r;
}
// No memory is published with this arrive (it says "my tcgen05.ld of the accumulator have completed", which
// tcgen05.wait::ld has already established), so it carries no cluster-scope release: that would cost a fence per arrive.
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ bool elect_one() {        // one lane of a converged warp
    uint32_t pred;
    asm volatile("{\n .reg .pred p;\n elect.sync _|p, 0xffffffff;\n selp.u32 %0, 1, 0, p;\n}\n" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* tm, uint32_t src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(tm), "r"(src), "r"(c0), "r"(c1)
                 : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// pair form: arrives on the barrier at this offset in both CTAs once the pair's MMAs issued so far have retired
__device__ __forceinline__ void tc_commit_pair(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)0x3)
                 : "memory");
}
// one MMA across the pair: D rows 0-127 in CTA 0's TMEM, 128-255 in CTA 1's; A from each CTA's own shared memory,
// B's N rows split between them (descriptors are offsets, the same in both CTAs); issued by CTA 0 only
__device__ __forceinline__ void tc_mma_f16_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                                uint32_t accumulate) {
    asm volatile(
        "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
        " tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n}\n"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tc_mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                           uint32_t accumulate) {
    asm volatile(
        "{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n"
        " tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void tc_ld_32x32b_x32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ float bf16_bits_to_f32(uint32_t b) { return __uint_as_float(b << 16); }

// ------------------------------------------------------------------------------------------------
// GEMM layer
// ------------------------------------------------------------------------------------------------
constexpr int BM = 128, BN = 256, BK = 64, UMMA_K = 16, STAGES = 3;
constexpr uint32_t kSwizzleBytes = BK * 2;          // one k-block row = one swizzle atom row (64 or 128 bytes)
constexpr int kNonEpiWarps = 4;   // TMA producer, MMA issuer, TMEM allocator, spare
constexpr uint32_t kABytes = BM * BK * 2;     // 16 KB
constexpr uint32_t kBBytes = BN * BK * 2;     // 32 KB
constexpr uint32_t kTmemCols = 512;           // 2 accumulators x 256 fp32 columns
// output staging for the TMA store: the bf16 tile as 64-column x 32-row boxes (4 KB, 128B-swizzled), one set per
// epilogue warp — 128 x 256 x 2 B = 64 KB.  (Three smem stages feed the MMAs as well as four did: the fourth
// stage's 48 KB went here.)
constexpr uint32_t kOutBoxCols = 64, kOutBoxRows = 32, kOutBoxBytes = kOutBoxCols * kOutBoxRows * 2;
constexpr uint32_t kOutStageBytes = BM * BN * 2;
// a CTA of a pair (kCl = 2) keeps half of the W k-block, so a fourth stage fits
__host__ __device__ constexpr int gemm_stages(int cl) { return cl == 2 ? 4 : STAGES; }
__host__ __device__ constexpr uint32_t gemm_stage_bytes(int cl) { return kABytes + kBBytes / (uint32_t)cl; }
__host__ __device__ constexpr uint32_t gemm_smem(int cl) {
    return (uint32_t)gemm_stages(cl) * gemm_stage_bytes(cl) + kOutStageBytes + 1024 /* alignment slack */ + 256 /* barriers */ +
           4096 /* w4 as fp32 */ + 4096 /* bias */;
}
// instruction descriptor, kind::f16: D=F32 (bit 4), A=BF16 (bit 7), B=BF16 (bit 10), both K-major,
// N>>3 at bit 17, M>>4 at bit 24
__host__ __device__ constexpr uint32_t make_idesc(int m) { return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(m >> 4) << 24); }
// smem matrix descriptor (K-major, SWIZZLE_128B): LBO = 1 (unused for swizzled K-major), SBO = 1024 B
// between 8-row groups, descriptor version 1 (sm_100), layout type 2
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr) {
    const uint32_t lo = ((smem_addr >> 4) & 0x3FFFu) | (1u << 16);
    const uint32_t hi = ((8u * kSwizzleBytes) >> 4) | (1u << 14) | ((kSwizzleBytes == 128 ? 2u : 4u) << 29);
    return ((uint64_t)hi << 32) | lo;
}

struct GemmParams {
    CUtensorMap tm_a;     // A  [M, K] bf16, K contiguous; box 64 x 128
    CUtensorMap tm_b;     // W  [N, K] bf16, K contiguous; box 64 x 256
    CUtensorMap tm_out;   // out [M, N] bf16; box 64 x 32 (store)
    CUtensorMap tm_bh;    // W again, box 64 x 128: the half of a W k-block one CTA of a pair loads (kCl == 2)
    const float* bias;    // [N]
    uint16_t* out;        // [M, N] bf16 (null when the head is fused and nobody wants the activations)
    uint32_t M, N, K;
    // fused head (layer 3 only): layer 4 dot product + fp64 epilogue in the GEMM epilogue
    const uint16_t* w4;   // [N] bf16, null = no head
    float b4;
    const igs_ltv_row* rows;
    igs_ltv_result* res;
    float* raw;           // optional fp32 model output
};

// kHead = false: work items are 128 x 256 output tiles, round-robin over the CTAs.
// kHead = true : work items are 128-row tiles; a CTA walks the N/256 column tiles of its row tile in order, so
//               the thread that owns a row accumulates that row's layer-4 dot product in a fixed order and
//               finishes with the fp64 epilogue — the layer-3 activations never go to HBM.
// kEpiWarps = 4: one epilogue warp per TMEM lane quarter, all 256 columns each.
// kEpiWarps = 8: two per quarter, 128 columns each — for the K=256 layer, whose MMAs per tile are too short
//               to hide a 4-warp epilogue.
// kCl = 2: CTA pairs (a cluster of two = the two SMs of a TPC) on 256 x 256 tiles with `cta_group::2` MMAs.  CTA r of
// the pair owns rows r*128.. of the tile: it loads those rows of A and rows r*128.. of the W k-block (HALF of it), the
// accumulator rows sit in its own TMEM, and its epilogue warps drain them.  CTA 0 issues the MMAs for both; the TMA
// loads of both CTAs credit CTA 0's `bar_full`; the commits that free a stage / publish an accumulator arrive in both
// CTAs; both CTAs' epilogue warps arrive on CTA 0's `bar_tempty`.  Per MMA each SM takes 32 KB from L2 instead of 48.
template <bool kHead, int kEpiWarps, int kCl>
__global__ void __launch_bounds__((kNonEpiWarps + kEpiWarps) * 32, 1)
k_mlp_gemm(const __grid_constant__ GemmParams p) {
    constexpr int kGemmThreads = (kNonEpiWarps + kEpiWarps) * 32;
    static_assert(!kHead || kEpiWarps == 4, "the fused head keeps one thread per row");
    extern __shared__ uint8_t smem_raw[];
    // 1024-byte alignment for the 128B swizzle atoms
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t* smem = smem_raw + (base - raw);
    constexpr bool kPair = kCl == 2;
    constexpr int kStages = gemm_stages(kCl);
    constexpr uint32_t kBStage = kBBytes / kCl, kStage = gemm_stage_bytes(kCl);
    constexpr uint32_t kIdesc = make_idesc(BM * kCl);
    const uint32_t a_base = base;
    const uint32_t b_base = base + kStages * kABytes;
    const uint32_t out_base = base + kStages * kStage;              // 1024-aligned: stages are multiples of 1 KB
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * kStage + kOutStageBytes);
    const uint32_t bar_full = smem_u32(bars);             // [kStages]
    const uint32_t bar_empty = bar_full + 8 * kStages;    // [kStages]
    const uint32_t bar_tfull = bar_empty + 8 * kStages;   // [2]
    const uint32_t bar_tempty = bar_tfull + 16;           // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);
    float* w4s = reinterpret_cast<float*>(smem + kStages * kStage + kOutStageBytes + 256);
    float* bias_s = w4s + IGS_MLP_HIDDEN;
    if (kHead)
        for (int i = threadIdx.x; i < IGS_MLP_HIDDEN; i += kGemmThreads) w4s[i] = bf16_bits_to_f32(__ldg(p.w4 + i));
    for (uint32_t i = threadIdx.x; i < p.N; i += kGemmThreads) bias_s[i] = __ldg(p.bias + i);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(bar_tfull + 8 * a, 1); mbar_init(bar_tempty + 8 * a, kEpiWarps * kCl); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {                              // with pairs: one warp of EACH CTA, collectively
        if (kPair) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                         "r"(kTmemCols)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                         "r"(kTmemCols)
                         : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    tc_fence_before();
    __syncthreads();
    if (kPair) cluster_sync();                   // the peer's barriers and TMEM exist before anything touches them
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    // programmatic dependent launch: everything above (barriers, TMEM, the weights' bias / w4 rows) is independent of
    // the previous launch; its activations are not.  The next launch may begin its own prologue from here on.
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    // With pairs, the unit of scheduling is the pair: `unit` / `n_units` replace blockIdx.x / gridDim.x, a row tile
    // of the plain kernel becomes a pair of row tiles (this CTA takes the one of its rank), and row tiles past the
    // end (odd count) are computed on zero-filled loads and never stored.
    const uint32_t rank = kCl > 1 ? cluster_ctarank() : 0u;
    const uint32_t unit = blockIdx.x / kCl, n_units = gridDim.x / kCl;
    const uint32_t m_blks = (p.M + BM - 1) / BM, n_blks = p.N / BN, k_blks = p.K / BK;
    const uint32_t m_grps = (m_blks + kCl - 1) / kCl;
    const uint32_t n_tiles = m_grps * n_blks;
    // i-th work tile of this CTA -> tile index in units of (row-tile group, column tile), m-major
    const uint32_t my_tiles = kHead ? (m_grps > unit ? (m_grps - unit + n_units - 1) / n_units : 0) * n_blks
                                    : (n_tiles > unit ? (n_tiles - unit + n_units - 1) / n_units : 0);
    auto tile_of = [&](uint32_t i) -> uint32_t {     // returns (row tile of THIS CTA) * n_blks + column tile
        const uint32_t g = kHead ? (unit + (i / n_blks) * n_units) * n_blks + i % n_blks : unit + i * n_units;
        return ((g / n_blks) * kCl + rank) * n_blks + g % n_blks;
    };

    if (warp == 0) {
        if (lane == 0) {
            // pairs: every load of either CTA credits CTA 0's bar_full, which expects both CTAs' stages
            const uint32_t full0 = kPair ? mapa_rank0(bar_full) : bar_full;
            uint32_t s = 0, ph = 0;
            for (uint32_t i = 0; i < my_tiles; ++i) {
                const uint32_t t = tile_of(i);
                const int m0 = (int)(t / n_blks) * BM, n0 = (int)(t % n_blks) * BN;
                for (uint32_t kb = 0; kb < k_blks; ++kb) {
                    mbar_wait(bar_empty + 8 * s, ph ^ 1);
                    if (!kPair) {
                        mbar_expect_tx(bar_full + 8 * s, kStage);
                        tma_load_2d(a_base + s * kABytes, &p.tm_a, (int)(kb * BK), m0, bar_full + 8 * s);
                        tma_load_2d(b_base + s * kBStage, &p.tm_b, (int)(kb * BK), n0, bar_full + 8 * s);
                    } else {
                        if (rank == 0) mbar_expect_tx(bar_full + 8 * s, 2 * kStage);
                        tma_load_2d_pair(a_base + s * kABytes, &p.tm_a, (int)(kb * BK), m0, full0 + 8 * s);
                        tma_load_2d_pair(b_base + s * kBStage, &p.tm_bh, (int)(kb * BK), n0 + (int)rank * (BN / 2),
                                         full0 + 8 * s);
                    }
                    if (++s == kStages) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1 && kPair) {
        // The issuing thread is the critical path once the data keeps up: 4 MMAs of 128 cycles each per k-block leave
        // ~500 cycles for everything else it does per k-block.  So: the whole warp runs the loop (no divergent region
        // around instructions that want uniform operands), the stage index is a compile-time constant (K / 64 is a
        // multiple of the 4 stages), and a descriptor is one add on a base computed once.
        if (rank == 0) {
            const uint64_t adesc0 = make_desc(a_base), bdesc0 = make_desc(b_base);
            uint32_t ph = 0;
            for (uint32_t it = 0; it < my_tiles; ++it) {
                const uint32_t acc = it & 1;
                mbar_wait(bar_tempty + 8 * acc, ((it >> 1) & 1) ^ 1);     // epilogue (of both CTAs) drained this accumulator
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + acc * BN;
                for (uint32_t kb = 0; kb < k_blks; kb += kStages) {
#pragma unroll
                    for (int s = 0; s < kStages; ++s) {
                        mbar_wait(bar_full + 8 * s, ph);
                        tc_fence_after();
                        if (elect_one()) {
#pragma unroll
                            for (int k = 0; k < BK / UMMA_K; ++k)
                                tc_mma_f16_pair(tmem_d, adesc0 + (uint64_t)(s * (kABytes >> 4) + k * 2),
                                                bdesc0 + (uint64_t)(s * (kBStage >> 4) + k * 2), kIdesc,
                                                (kb | (uint32_t)(s | k)) != 0 ? 1u : 0u);
                            tc_commit_pair(bar_empty + 8 * s);             // frees the smem stage (in both CTAs) when the MMAs retire
                        }
                        __syncwarp();
                    }
                    ph ^= 1;
                }
                if (elect_one()) tc_commit_pair(bar_tfull + 8 * acc);      // accumulator complete
                __syncwarp();
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && rank == 0) {
            uint32_t s = 0, ph = 0, it = 0;
            for (; it < my_tiles; ++it) {
                const uint32_t acc = it & 1;
                mbar_wait(bar_tempty + 8 * acc, ((it >> 1) & 1) ^ 1);     // epilogue (of both CTAs) drained this accumulator
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + acc * BN;
                for (uint32_t kb = 0; kb < k_blks; ++kb) {
                    mbar_wait(bar_full + 8 * s, ph);
                    tc_fence_after();
                    const uint64_t adesc = make_desc(a_base + s * kABytes);
                    const uint64_t bdesc = make_desc(b_base + s * kBStage);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) {
                        // +32 bytes (2 x 16 B units) per K step inside the 128B swizzle atom
                        if (kPair)
                            tc_mma_f16_pair(tmem_d, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), kIdesc,
                                            (kb | (uint32_t)k) != 0 ? 1u : 0u);
                        else
                            tc_mma_f16(tmem_d, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), kIdesc,
                                       (kb | (uint32_t)k) != 0 ? 1u : 0u);
                    }
                    if (kPair) tc_commit_pair(bar_empty + 8 * s);          // frees the smem stage (in both CTAs) when the MMAs retire
                    else tc_commit(bar_empty + 8 * s);
                    if (++s == kStages) { s = 0; ph ^= 1; }
                }
                if (kPair) tc_commit_pair(bar_tfull + 8 * acc);            // accumulator complete
                else tc_commit(bar_tfull + 8 * acc);
            }
        }
    } else if (warp >= 4) {
        const int q = (warp - kNonEpiWarps) & 3;                           // TMEM lane quarter of this warp (= warp % 4)
        constexpr int kCols = BN / (kEpiWarps / 4);                        // columns handled by this warp
        const int c0 = ((warp - kNonEpiWarps) >> 2) * kCols;
        float dot = 0.f;                                                   // layer-4 partial of this thread's row
        for (uint32_t it = 0; it < my_tiles; ++it) {
            const uint32_t t = tile_of(it);
            const uint32_t acc = it & 1;
            const uint32_t m0 = (t / n_blks) * BM, n0 = (t % n_blks) * BN;
            mbar_wait(bar_tfull + 8 * acc, (it >> 1) & 1);
            tc_fence_after();
            const uint32_t row = m0 + q * 32 + lane;
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN;
            // this warp's staging: kCols / 64 boxes of 32 rows x 64 columns
            const uint32_t my_out = out_base + (uint32_t)(warp - kNonEpiWarps) * (kCols / kOutBoxCols) * kOutBoxBytes;
            const bool store = p.out != nullptr;
            if (store) {                                                   // the previous tile's store has read the staging
                if (lane == 0) bulk_wait_read0();
                __syncwarp();
            }
#pragma unroll 1
            for (int c = c0; c < c0 + kCols; c += 32) {
                uint32_t r[32];
                tc_ld_32x32b_x32(taddr + c, r);
                tc_wait_ld();
                uint32_t packed[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const float2 bb = *reinterpret_cast<const float2*>(bias_s + n0 + c + 2 * j);   // smem broadcast
                    float v0 = __fadd_rn(__uint_as_float(r[2 * j]), bb.x);
                    float v1 = __fadd_rn(__uint_as_float(r[2 * j + 1]), bb.y);
                    v0 = v0 > 0.f ? v0 : 0.f;
                    v1 = v1 > 0.f ? v1 : 0.f;
                    const __nv_bfloat162 b2 = __floats2bfloat162_rn(v0, v1);   // .x (low half) = v0
                    packed[j] = *reinterpret_cast<const uint32_t*>(&b2);
                    if (kHead) {                                               // layer 4 on the bf16-rounded activations
                        dot = __fmaf_rn(bf16_bits_to_f32(packed[j] & 0xFFFFu), w4s[n0 + c + 2 * j], dot);
                        dot = __fmaf_rn(bf16_bits_to_f32(packed[j] >> 16), w4s[n0 + c + 2 * j + 1], dot);
                    }
                }
                if (store) {
                    // row `lane` of box (c - c0) / 64, 16-byte chunks (c % 64) / 8 .. +3, 128B swizzle: chunk ^ (row & 7)
                    const uint32_t box = my_out + (uint32_t)((c - c0) / (int)kOutBoxCols) * kOutBoxBytes + (uint32_t)lane * 128u;
                    const uint32_t j0 = (uint32_t)(c % (int)kOutBoxCols) / 8u;
#pragma unroll
                    for (uint32_t j = 0; j < 4; ++j) {
                        const uint32_t addr = box + (((j0 + j) ^ ((uint32_t)lane & 7u)) << 4);
                        asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(packed[4 * j]),
                                     "r"(packed[4 * j + 1]), "r"(packed[4 * j + 2]), "r"(packed[4 * j + 3])
                                     : "memory");
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (kPair && rank != 0) mbar_arrive_cluster(mapa_rank0(bar_tempty + 8 * acc));
                else mbar_arrive(bar_tempty + 8 * acc);
            }
            if (store) {                                                   // whole lines to global, off the LSU path
                fence_async_smem();
                __syncwarp();
                if (lane == 0) {
#pragma unroll
                    for (int b = 0; b < kCols / (int)kOutBoxCols; ++b)
                        tma_store_2d(&p.tm_out, my_out + (uint32_t)b * kOutBoxBytes, (int)(n0 + c0 + b * (int)kOutBoxCols),
                                     (int)(m0 + q * 32));
                    bulk_commit();
                }
            }
            if (kHead && (t % n_blks) == n_blks - 1) {                        // row complete: model output + epilogue
                if (row < p.M) {
                    const float y = __fadd_rn(dot, p.b4);
                    if (p.raw) p.raw[row] = y;
                    const igs_ltv_row f = p.rows[row];
                    igs_ltv_result res;
                    ltv_epilogue(f, (double)y, &res);
                    p.res[row] = res;
                }
                dot = 0.f;
            }
        }
    }
    if (warp >= kNonEpiWarps && lane == 0) bulk_wait0();                   // outstanding output stores
    tc_fence_before();
    __syncthreads();
    if (kPair) cluster_sync();                   // the peer may still be signalling this CTA's barriers / using its TMEM
    if (warp == 2) {
        tc_fence_after();
        if (kPair)
            asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
        else
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
struct MlpDevice {
    uint16_t *w1 = nullptr, *w2 = nullptr, *w3 = nullptr, *w4 = nullptr;
    float *b1 = nullptr, *b2 = nullptr, *b3 = nullptr;
    float b4 = 0.f;
    igs_buf act[2];   // ping-pong activations of one row chunk
};

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (fn) return fn;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
        return nullptr;
    fn = (EncodeTiledFn)p;
    return fn;
}

// 2D bf16 row-major [rows, cols] tensor map with a (64 x box_rows) box and 128B swizzle
int make_tmap(igs_engine* e, CUtensorMap* tm, const void* ptr, uint64_t rows, uint64_t cols, uint32_t box_rows,
              uint32_t box_cols = BK, CUtensorMapSwizzle swz = kSwizzleBytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                                                                     : CU_TENSOR_MAP_SWIZZLE_64B) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) {
        igs_set_error(e, "cuTensorMapEncodeTiled is not available from the driver");
        return IGS_ERR_CUDA;
    }
    const cuuint64_t gdim[2] = {cols, rows};
    const cuuint64_t gstride[1] = {cols * 2};
    const cuuint32_t box[2] = {(cuuint32_t)box_cols, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        igs_set_error(e, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
        return IGS_ERR_CUDA;
    }
    return IGS_OK;
}

struct HeadArgs {
    const uint16_t* w4 = nullptr;
    float b4 = 0.f;
    const igs_ltv_row* rows = nullptr;
    igs_ltv_result* res = nullptr;
    float* raw = nullptr;
};

int launch_gemm(igs_engine* e, const uint16_t* a, const uint16_t* w, const float* bias, uint16_t* out, uint32_t M,
                uint32_t N, uint32_t K, cudaStream_t st, const HeadArgs* head = nullptr) {
    GemmParams p;
    memset(&p, 0, sizeof(p));
    int rc = make_tmap(e, &p.tm_a, a, M, K, BM);
    if (rc != IGS_OK) return rc;
    rc = make_tmap(e, &p.tm_b, w, N, K, BN);
    if (rc != IGS_OK) return rc;
    if (out) {
        rc = make_tmap(e, &p.tm_out, out, M, N, kOutBoxRows, kOutBoxCols, CU_TENSOR_MAP_SWIZZLE_128B);
        if (rc != IGS_OK) return rc;
    }
    p.bias = bias; p.out = out; p.M = M; p.N = N; p.K = K;
    if (head) { p.w4 = head->w4; p.b4 = head->b4; p.rows = head->rows; p.res = head->res; p.raw = head->raw; }
    // the short-K layer gets 8 epilogue warps; long-K layers hide a 4-warp epilogue behind their MMAs
    const bool wide_epi = !head && K <= 256;
    static const int pairs_knob = getenv("IGS_MLP_PAIRS") ? atoi(getenv("IGS_MLP_PAIRS")) : 1;   // tuning knob: 0 = no CTA pairs
    const uint32_t m_blks = (M + BM - 1) / BM;
    const bool pairs = pairs_knob != 0 && m_blks >= 2 * (uint32_t)e->sm_count;     // large M only: small ones gain nothing
    const int threads = (kNonEpiWarps + (wide_epi ? 8 : 4)) * 32;
    if (!pairs) {
        auto kern = head ? k_mlp_gemm<true, 4, 1> : wide_epi ? k_mlp_gemm<false, 8, 1> : k_mlp_gemm<false, 4, 1>;
        IGS_CUDA(e, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, gemm_smem(1)));
        const uint32_t items = head ? m_blks : m_blks * (N / BN);
        const uint32_t grid = items < (uint32_t)e->sm_count ? items : (uint32_t)e->sm_count;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(grid, 1, 1);
        cfg.blockDim = dim3((unsigned)threads, 1, 1);
        cfg.dynamicSmemBytes = gemm_smem(1);
        cfg.stream = st;
        cudaLaunchAttribute attr;
        attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr.val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = &attr;
        cfg.numAttrs = 1;
        IGS_CUDA(e, cudaLaunchKernelEx(&cfg, kern, p));
        IGS_LAUNCHED(e);
        return IGS_OK;
    }
    rc = make_tmap(e, &p.tm_bh, w, N, K, BN / 2);
    if (rc != IGS_OK) return rc;
    auto kern = head ? k_mlp_gemm<true, 4, 2> : wide_epi ? k_mlp_gemm<false, 8, 2> : k_mlp_gemm<false, 4, 2>;
    IGS_CUDA(e, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, gemm_smem(2)));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(e->sm_count & ~1), 1, 1);          // whole pairs; every pair has work (m_blks is large)
    cfg.blockDim = dim3((unsigned)threads, 1, 1);
    cfg.dynamicSmemBytes = gemm_smem(2);
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 2;
    // the schedule is static: a pair that does not fit in the first wave would double the run time, so launch exactly
    // the pairs that are co-resident (a GPC with an odd number of free SMs leaves one out)
    int& max_pairs = e->mlp_max_pairs[head ? 2 : wide_epi ? 1 : 0];
    if (max_pairs == 0) {
        cfg.numAttrs = 1;
        IGS_CUDA(e, cudaOccupancyMaxActiveClusters(&max_pairs, kern, &cfg));
        cfg.numAttrs = 2;
    }
    IGS_REQUIRE(e, max_pairs > 0, "no CTA pair of the MLP kernel fits on this device");
    if ((unsigned)max_pairs * 2u < cfg.gridDim.x) cfg.gridDim.x = (unsigned)max_pairs * 2u;
    IGS_CUDA(e, cudaLaunchKernelEx(&cfg, kern, p));
    IGS_LAUNCHED(e);
    return IGS_OK;
}

constexpr uint64_t kLtvChunkRows = 256 * 1024;   // 512 MB of bf16 activations per ping-pong buffer

}  // namespace

void igs_mlp_free(igs_engine* e) {
    MlpDevice* m = (MlpDevice*)e->d_mlp;
    if (!m) return;
    for (void* p : {(void*)m->w1, (void*)m->w2, (void*)m->w3, (void*)m->w4, (void*)m->b1, (void*)m->b2, (void*)m->b3})
        if (p) cudaFree(p);
    for (igs_buf& b : m->act)
        if (b.p) cudaFree(b.p);
    delete m;
    e->d_mlp = nullptr;
    e->has_mlp = false;
}

int igs_mlp_upload(igs_engine* e, const igs_mlp_desc* d) {
    IGS_REQUIRE(e, d->w1 && d->w2 && d->w3 && d->w4 && d->b1 && d->b2 && d->b3 && d->b4, "a weight / bias pointer is null");
    IGS_CUDA(e, cudaDeviceSynchronize());
    igs_mlp_free(e);
    MlpDevice* m = new MlpDevice();
    e->d_mlp = m;
    const size_t H = IGS_MLP_HIDDEN, I = IGS_MLP_IN;
    struct { uint16_t** dst; const uint16_t* src; size_t n; } ws[] = {
        {&m->w1, d->w1, H * I}, {&m->w2, d->w2, H * H}, {&m->w3, d->w3, H * H}, {&m->w4, d->w4, H}};
    for (auto& w : ws) {
        IGS_CUDA(e, cudaMalloc(w.dst, w.n * 2));
        IGS_CUDA(e, cudaMemcpy(*w.dst, w.src, w.n * 2, cudaMemcpyHostToDevice));
    }
    struct { float** dst; const float* src; } bs[] = {{&m->b1, d->b1}, {&m->b2, d->b2}, {&m->b3, d->b3}};
    for (auto& b : bs) {
        IGS_CUDA(e, cudaMalloc(b.dst, H * 4));
        IGS_CUDA(e, cudaMemcpy(*b.dst, b.src, H * 4, cudaMemcpyHostToDevice));
    }
    m->b4 = d->b4[0];
    e->has_mlp = true;
    return IGS_OK;
}

// Runs layers 1..n_layers (1..3) on rows [0, n) of d_x and leaves the bf16 activations of the last one
// in *h_out (a chunk buffer owned by the engine); n <= kLtvChunkRows.
static int run_hidden(igs_engine* e, MlpDevice* m, const uint16_t* d_x, uint32_t n, int n_layers, uint16_t** h_out,
                      cudaStream_t st) {
    const uint32_t H = IGS_MLP_HIDDEN;
    for (igs_buf& b : m->act) {
        int rc = igs_reserve(e, &b, (size_t)kLtvChunkRows * H * 2);
        if (rc != IGS_OK) return rc;
    }
    uint16_t* a0 = (uint16_t*)m->act[0].p;
    uint16_t* a1 = (uint16_t*)m->act[1].p;
    int rc = launch_gemm(e, d_x, m->w1, m->b1, a0, n, H, IGS_MLP_IN, st);
    if (rc != IGS_OK) return rc;
    *h_out = a0;
    if (n_layers >= 2) {
        rc = launch_gemm(e, a0, m->w2, m->b2, a1, n, H, H, st);
        if (rc != IGS_OK) return rc;
        *h_out = a1;
    }
    if (n_layers >= 3) {
        rc = launch_gemm(e, a1, m->w3, m->b3, a0, n, H, H, st);
        if (rc != IGS_OK) return rc;
        *h_out = a0;
    }
    return IGS_OK;
}

int igs_launch_ltv(igs_engine* e, const uint16_t* d_x, const igs_ltv_row* d_rows, uint64_t n, igs_ltv_result* d_out,
                   float* d_ltv_raw, cudaStream_t st) {
    if (!e->has_mlp) {
        igs_set_error(e, "no LTV model loaded (igs_load_mlp)");
        return IGS_ERR_STATE;
    }
    MlpDevice* m = (MlpDevice*)e->d_mlp;
    const uint32_t H = IGS_MLP_HIDDEN;
    for (igs_buf& b : m->act) {
        int rc = igs_reserve(e, &b, (size_t)kLtvChunkRows * H * 2);
        if (rc != IGS_OK) return rc;
    }
    uint16_t* a0 = (uint16_t*)m->act[0].p;
    uint16_t* a1 = (uint16_t*)m->act[1].p;
    for (uint64_t off = 0; off < n; off += kLtvChunkRows) {
        const uint32_t cnt = (uint32_t)(n - off < kLtvChunkRows ? n - off : kLtvChunkRows);
        int rc = launch_gemm(e, d_x + off * IGS_MLP_IN, m->w1, m->b1, a0, cnt, H, IGS_MLP_IN, st);
        if (rc != IGS_OK) return rc;
        rc = launch_gemm(e, a0, m->w2, m->b2, a1, cnt, H, H, st);
        if (rc != IGS_OK) return rc;
        HeadArgs head;                       // layer 3 + layer 4 + fp64 epilogue in one kernel
        head.w4 = m->w4; head.b4 = m->b4; head.rows = d_rows + off; head.res = d_out + off;
        head.raw = d_ltv_raw ? d_ltv_raw + off : nullptr;
        rc = launch_gemm(e, a1, m->w3, m->b3, nullptr, cnt, H, H, st, &head);
        if (rc != IGS_OK) return rc;
    }
    return IGS_OK;
}

int igs_launch_ltv_hidden(igs_engine* e, const uint16_t* d_x, uint64_t n, int layer, uint16_t* d_hidden,
                          cudaStream_t st) {
    if (!e->has_mlp) {
        igs_set_error(e, "no LTV model loaded (igs_load_mlp)");
        return IGS_ERR_STATE;
    }
    IGS_REQUIRE(e, layer >= 1 && layer <= 3, "layer must be 1..3");
    IGS_REQUIRE(e, n <= kLtvChunkRows, "at most 262144 rows per diagnostic call");
    if (n == 0) return IGS_OK;
    MlpDevice* m = (MlpDevice*)e->d_mlp;
    uint16_t* h = nullptr;
    int rc = run_hidden(e, m, d_x, (uint32_t)n, layer, &h, st);
    if (rc != IGS_OK) return rc;
    IGS_CUDA(e, cudaMemcpyAsync(d_hidden, h, n * IGS_MLP_HIDDEN * 2, cudaMemcpyDeviceToDevice, st));
    return IGS_OK;
}
