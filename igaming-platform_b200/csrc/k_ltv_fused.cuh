// k_ltv_fused.cuh — the three GEMM layers of the LTV MLP in one persistent kernel (included by k_ltv.cu,
// inside its anonymous namespace, after the PTX helpers).
//
// A CTA carries a PAIR of 128-row tiles (A, B) through all three layers in the order
//     L1(A) L1(B) L2(A) L2(B) L3(A) L3(B)          (each phase = the 4 column tiles of that layer)
// so that the activations a layer writes are read back by the same CTA one phase later, while they are still
// in L2: they go through three private 128 x 1024 bf16 scratch tiles per CTA (X, Y, Z; 114 MB for 148 CTAs)
// and never need to reach HBM.  Between the TMA stores of a phase and the TMA loads of the phase that
// consumes them sits one other phase of work and an mbarrier per scratch tile (arrived by the epilogue warps
// once their stores have completed, awaited by the producer before its first load), so neither the MMA
// stream nor the loads wait.  Warp roles, smem ring, TMEM double buffering and the epilogue arithmetic are
// those of k_mlp_gemm; layer 3 ends in the same fused head (layer 4 dot product + fp64 epilogue).
//      phase      0      1      2      3      4      5
//      layer      1      1      2      2      3      3
//      tile       A      B      A      B      A      B
//      reads      x      x      X      Y      Z      X
//      writes     X      Y      Z      X      -      -
constexpr int kFusedEpiWarps = 8;
constexpr int kFusedThreads = (kNonEpiWarps + kFusedEpiWarps) * 32;
constexpr int kFusedPhases = 6, kFusedColTiles = IGS_MLP_HIDDEN / BN;      // 4 column tiles per phase
constexpr uint32_t kFusedSmem = STAGES * kStageBytes + kOutStageBytes + 1024 + 256 + 4 * 4096;   // + 3 biases, w4

struct FusedParams {
    CUtensorMap tm_x;        // input  [M, 256]        box 64 x 128 (load)
    CUtensorMap tm_w[3];     // W1 [1024, 256], W2, W3 [1024, 1024]   box 64 x 256 (load)
    CUtensorMap tm_s_ld;     // scratch [grid * 3 * 128, 1024]        box 64 x 128 (load)
    CUtensorMap tm_s_st;     // scratch, box 64 x 32 (store)
    const float* bias[3];
    const uint16_t* w4;
    float b4;
    uint32_t M;
    uint32_t single;         // 1: groups of one row tile (consumer phase follows its producer directly)
    const igs_ltv_row* rows;
    igs_ltv_result* res;
    float* raw;
};

__device__ __forceinline__ int fused_layer(int ph) { return ph >> 1; }            // 0, 1, 2
__device__ __forceinline__ int fused_src(int ph) { return ph == 2 ? 0 : ph == 3 ? 1 : ph == 4 ? 2 : 0; }   // X Y Z X
__device__ __forceinline__ int fused_dst(int ph) { return ph == 0 ? 0 : ph == 1 ? 1 : ph == 2 ? 2 : 0; }   // X Y Z X

__global__ void __launch_bounds__(kFusedThreads, 1)
k_mlp_fused(const __grid_constant__ FusedParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t raw = smem_u32(smem_raw);
    const uint32_t base = (raw + 1023u) & ~1023u;
    uint8_t* smem = smem_raw + (base - raw);
    const uint32_t a_base = base;
    const uint32_t b_base = base + STAGES * kABytes;
    const uint32_t out_base = base + STAGES * kStageBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * kStageBytes + kOutStageBytes);
    const uint32_t bar_full = smem_u32(bars);             // [STAGES]
    const uint32_t bar_empty = bar_full + 8 * STAGES;     // [STAGES]
    const uint32_t bar_tfull = bar_empty + 8 * STAGES;    // [2]
    const uint32_t bar_tempty = bar_tfull + 16;           // [2]
    const uint32_t bar_ready = bar_tempty + 16;           // [3]  scratch tile X / Y / Z written and visible
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4 + 3);
    float* w4s = reinterpret_cast<float*>(smem + STAGES * kStageBytes + kOutStageBytes + 256);
    float* bias_s = w4s + IGS_MLP_HIDDEN;                 // [3][1024]
    for (int i = threadIdx.x; i < IGS_MLP_HIDDEN; i += kFusedThreads) {
        w4s[i] = bf16_bits_to_f32(__ldg(p.w4 + i));
#pragma unroll
        for (int l = 0; l < 3; ++l) bias_s[l * IGS_MLP_HIDDEN + i] = __ldg(p.bias[l] + i);
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(bar_tfull + 8 * a, 1); mbar_init(bar_tempty + 8 * a, kFusedEpiWarps); }
        for (int b = 0; b < 3; ++b) mbar_init(bar_ready + 8 * b, kFusedEpiWarps);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                     "r"(kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    const uint32_t m_blks = (p.M + BM - 1) / BM;
    const uint32_t n_pairs = p.single ? m_blks : (m_blks + 1) / 2;
    const uint32_t my_pairs = n_pairs > blockIdx.x ? (n_pairs - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    // row tile of (pair i of this CTA, phase ph); >= m_blks: the pair's second tile does not exist
    auto tile_of = [&](uint32_t i, int ph) -> uint32_t {
        const uint32_t g = blockIdx.x + i * gridDim.x;
        if (p.single) return (ph & 1) ? m_blks : g;          // one row tile per group: the B phases do not exist
        return g * 2 + (uint32_t)(ph & 1);
    };
    const int scratch_row0 = (int)(blockIdx.x * 3 * BM);   // this CTA's X, Y, Z in the scratch tensor

    if (warp == 0) {
        if (lane == 0) {
            uint32_t s = 0, ph_bit = 0;
            uint32_t ready_par[3] = {0, 0, 0};
            for (uint32_t i = 0; i < my_pairs; ++i) {
                for (int ph = 0; ph < kFusedPhases; ++ph) {
                    const uint32_t t = tile_of(i, ph);
                    if (t >= m_blks) continue;
                    const int layer = fused_layer(ph);
                    const uint32_t k_blks = (layer == 0 ? IGS_MLP_IN : IGS_MLP_HIDDEN) / BK;
                    const CUtensorMap* tm_a = layer == 0 ? &p.tm_x : &p.tm_s_ld;
                    const int a_row = layer == 0 ? (int)(t * BM) : scratch_row0 + fused_src(ph) * BM;
                    if (layer != 0) {        // the phase that wrote this scratch tile has stored it
                        const int b = fused_src(ph);
                        mbar_wait(bar_ready + 8 * b, ready_par[b]);
                        ready_par[b] ^= 1;
                        asm volatile("fence.proxy.async;" ::: "memory");
                    }
                    for (int n = 0; n < kFusedColTiles; ++n) {
                        for (uint32_t kb = 0; kb < k_blks; ++kb) {
                            mbar_wait(bar_empty + 8 * s, ph_bit ^ 1);
                            mbar_expect_tx(bar_full + 8 * s, kStageBytes);
                            tma_load_2d(a_base + s * kABytes, tm_a, (int)(kb * BK), a_row, bar_full + 8 * s);
                            tma_load_2d(b_base + s * kBBytes, &p.tm_w[layer], (int)(kb * BK), n * BN, bar_full + 8 * s);
                            if (++s == STAGES) { s = 0; ph_bit ^= 1; }
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            uint32_t s = 0, ph_bit = 0, it = 0;
            for (uint32_t i = 0; i < my_pairs; ++i) {
                for (int ph = 0; ph < kFusedPhases; ++ph) {
                    if (tile_of(i, ph) >= m_blks) continue;
                    const uint32_t k_blks = (fused_layer(ph) == 0 ? IGS_MLP_IN : IGS_MLP_HIDDEN) / BK;
                    for (int n = 0; n < kFusedColTiles; ++n, ++it) {
                        const uint32_t acc = it & 1;
                        mbar_wait(bar_tempty + 8 * acc, ((it >> 1) & 1) ^ 1);
                        tc_fence_after();
                        const uint32_t tmem_d = tmem_base + acc * BN;
                        for (uint32_t kb = 0; kb < k_blks; ++kb) {
                            mbar_wait(bar_full + 8 * s, ph_bit);
                            tc_fence_after();
                            const uint64_t adesc = make_desc(a_base + s * kABytes);
                            const uint64_t bdesc = make_desc(b_base + s * kBBytes);
#pragma unroll
                            for (int k = 0; k < BK / UMMA_K; ++k)
                                tc_mma_f16(tmem_d, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), kIdesc,
                                           (kb | (uint32_t)k) != 0 ? 1u : 0u);
                            tc_commit(bar_empty + 8 * s);
                            if (++s == STAGES) { s = 0; ph_bit ^= 1; }
                        }
                        tc_commit(bar_tfull + 8 * acc);
                    }
                }
            }
        }
    }
    else if (warp >= kNonEpiWarps) {
        const int ew = warp - kNonEpiWarps;                    // 0..7
        const int q = ew & 3;                                  // TMEM lane quarter
        const int half = ew >> 2;                              // which 128 columns in the storing layers
        const uint32_t my_out = out_base + (uint32_t)ew * 2u * kOutBoxBytes;
        uint32_t it = 0;
        for (uint32_t i = 0; i < my_pairs; ++i) {
            for (int ph = 0; ph < kFusedPhases; ++ph) {
                const uint32_t t = tile_of(i, ph);
                if (t >= m_blks) continue;
                const int layer = fused_layer(ph);
                const bool head = layer == 2;
                const float* bias_l = bias_s + layer * IGS_MLP_HIDDEN;
                const uint32_t row = t * BM + q * 32 + lane;
                // layers 1, 2: this warp converts and stores 128 columns; layer 3: warps 0-3 own a whole row each
                // (256 columns, the order of the layer-4 sum is the one k_mlp_gemm uses), warps 4-7 only hand the
                // accumulator back
                const int kCols = head ? BN : BN / 2;
                const int c0 = head ? 0 : half * (BN / 2);
                const bool active = !head || half == 0;
                float dot = 0.f;
                for (int n = 0; n < kFusedColTiles; ++n, ++it) {
                    const uint32_t acc = it & 1;
                    const uint32_t n0 = (uint32_t)n * BN;
                    mbar_wait(bar_tfull + 8 * acc, (it >> 1) & 1);
                    tc_fence_after();
                    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + acc * BN;
                    if (!head) {
                        if (lane == 0) bulk_wait_read0();          // the previous store has read the staging
                        __syncwarp();
                    }
                    if (active) {
#pragma unroll 1
                        for (int c = c0; c < c0 + kCols; c += 32) {
                            uint32_t r[32];
                            tc_ld_32x32b_x32(taddr + c, r);
                            tc_wait_ld();
                            uint32_t packed[16];
#pragma unroll
                            for (int j = 0; j < 16; ++j) {
                                const float2 bb = *reinterpret_cast<const float2*>(bias_l + n0 + c + 2 * j);
                                float v0 = __fadd_rn(__uint_as_float(r[2 * j]), bb.x);
                                float v1 = __fadd_rn(__uint_as_float(r[2 * j + 1]), bb.y);
                                v0 = v0 > 0.f ? v0 : 0.f;
                                v1 = v1 > 0.f ? v1 : 0.f;
                                const __nv_bfloat162 b2 = __floats2bfloat162_rn(v0, v1);
                                packed[j] = *reinterpret_cast<const uint32_t*>(&b2);
                                if (head) {
                                    dot = __fmaf_rn(bf16_bits_to_f32(packed[j] & 0xFFFFu), w4s[n0 + c + 2 * j], dot);
                                    dot = __fmaf_rn(bf16_bits_to_f32(packed[j] >> 16), w4s[n0 + c + 2 * j + 1], dot);
                                }
                            }
                            if (!head) {
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
                    }
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_tempty + 8 * acc);
                    if (!head) {
                        fence_async_smem();
                        __syncwarp();
                        if (lane == 0) {
                            const int dst_row = scratch_row0 + fused_dst(ph) * BM + q * 32;
#pragma unroll
                            for (int b = 0; b < 2; ++b)
                                tma_store_2d(&p.tm_s_st, my_out + (uint32_t)b * kOutBoxBytes,
                                             (int)(n0 + c0 + b * (int)kOutBoxCols), dst_row);
                            bulk_commit();
                            if (n == kFusedColTiles - 1) {     // the scratch tile is complete: publish it to the producer
                                bulk_wait0();
                                asm volatile("fence.proxy.async;" ::: "memory");
                                mbar_arrive(bar_ready + 8 * fused_dst(ph));
                            }
                        }
                    }
                }
                if (head && active && row < p.M) {                  // row complete: model output + fp64 epilogue
                    const float y = __fadd_rn(dot, p.b4);
                    if (p.raw) p.raw[row] = y;
                    const igs_ltv_row f = p.rows[row];
                    igs_ltv_result res;
                    ltv_epilogue(f, (double)y, &res);
                    p.res[row] = res;
                }
            }
        }
        if (lane == 0) bulk_wait0();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(kTmemCols) : "memory");
    }
}
