// conv_halo.cu — 3x3 stride-1 convolution forward / data-gradient with activation-halo reuse.
//
// Same GEMM view, operands and epilogue as conv_gemm.cu (reference sites: models/resnet.py:23-27 and the
// autograd dgrad of main.py:104). Two measured limiters of that kernel shape this one:
//   * bytes staged from L2 into shared memory per MMA (profiles/r1_ncu_tap_gemm_full.txt: ~45 B/clk/SM, the
//     L2->SM ceiling, tensor pipe 45 % active). Here the activation tile of a 64-channel chunk is loaded ONCE
//     with its halo and the 9 taps are issued as row-shifted shared-memory descriptors over it; tcgen05
//     applies the 128-byte swizzle to the absolute smem address, so a descriptor may start at any 128 B row
//     and step between 8-row groups by any SBO (verified on hardware: profiles/r1_umma_probe.log, P1).
//   * the single warp that issues tcgen05.mma is bound by its own instruction latency (~5 cycles per dependent
//     instruction): every instruction in the per-tap loop costs tensor time. The issue loops are therefore
//     compile-time unrolled (mode / residency are template parameters), use one elected region per
//     activation stage (36 or 12 MMAs), and ring positions come from uniform counters and power-of-two rings.
//
//   MODE 0 (W >= 16): tile = 16 rows x 8 cols of one image; halo box {64 ch, 10, 18, 1}; tap (a, b) of the
//                     3x3 window starts at row a*10 + b of the tile; 8-row groups are 10 rows apart (SBO 1280).
//   MODE 1 (W == 8) : tile = two whole 8x8 images; three boxes {64 ch, 10, 8, 2}, one per window row a,
//                     loaded at image row a-1; tap b starts at row b.
//   RES             : cin = cout = 64 — all nine 8 KB weight slabs are loaded once per CTA and stay resident.
//
// Warp roles (320 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer, warps 2..9 = epilogue.
#include "b2c_kernels.h"
#include "b2c_launch.h"
#include "b2c_ptx.cuh"

namespace b2c {

namespace {

constexpr int kAStageBytes = 23552;  // 180 rows x 128 B rounded up to 1024

template <int BN>
struct HaloCfg {
    static constexpr int kBBytes = BN * 128;
    static constexpr int kSA = BN == 256 ? 2 : 4;      // activation stages (power of two)
    static constexpr int kBRingBytes = (BN == 256 ? 128 : 112) * 1024;
    static constexpr int kSB = kBRingBytes / kBBytes;  // 14 / 7 / 4 weight-tile slots
    static constexpr int kTmemCols = 2 * BN;
    static constexpr int kBarBytes = 512;
    static constexpr int kStatBytes = 2 * 4 * BN * 4 + 2 * BN * 4;   // running sums + fused-BN-backward scale|shift
    static constexpr int kStoreBytes = 8 * 2048;       // per-epilogue-warp store staging (store_chunk_staged)
    static constexpr int kSmemBytes = 1024 + kSA * kAStageBytes + kBRingBytes + kBarBytes + kStatBytes + kStoreBytes;
};

template <int BN, bool kStats, int MODE, bool RES>
__global__ void __launch_bounds__(320, 1)
conv_halo_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                 const __grid_constant__ HaloParams p) {
    using Cfg = HaloCfg<BN>;
    constexpr int kSA = Cfg::kSA;
    constexpr int LOADS = MODE == 0 ? 1 : 3;   // activation boxes per 64-channel chunk
    constexpr int TPL = 9 / LOADS;             // taps served by one box
    constexpr uint32_t kABytes = MODE == 0 ? 180u * 128u : 160u * 128u;
    constexpr uint32_t kSB = Cfg::kSB;
    static_assert((kSA & (kSA - 1)) == 0, "the activation ring is a power of two");
    static_assert((2 * kSA + 2 * Cfg::kSB + 4) * 8 + 4 <= Cfg::kBarBytes, "barrier block");
    static_assert(!RES || Cfg::kSB >= 9, "resident weights need nine slots");

    pdl_trigger_big();
    extern __shared__ uint8_t smem_raw[];
    // (offset arithmetic on the shared pointer keeps the address space visible: LDS / STS, not generic accesses)
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sA = smem;
    uint8_t* sB = smem + kSA * kAStageBytes;
    uint8_t* aux = sB + Cfg::kBRingBytes;
    uint64_t* full_a = reinterpret_cast<uint64_t*>(aux);
    uint64_t* empty_a = full_a + kSA;
    uint64_t* full_b = empty_a + kSA;
    uint64_t* empty_b = full_b + Cfg::kSB;
    uint64_t* tfull = empty_b + Cfg::kSB;
    uint64_t* tempty = tfull + 2;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty + 2);
    float* stat = reinterpret_cast<float*>(aux + Cfg::kBarBytes);
    float* bwd_sc = stat + 2 * 4 * BN;          // [scale | shift] of this CTA's channels (fused BN backward)
    uint8_t* store_stage = aux + Cfg::kBarBytes + Cfg::kStatBytes;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&mapA);
        tma_prefetch_desc(&mapB);
    }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < kSA; ++s) {
                mbar_init(&full_a[s], 1);
                mbar_init(&empty_a[s], 1);
            }
            for (int s = 0; s < Cfg::kSB; ++s) {
                mbar_init(&full_b[s], 1);
                mbar_init(&empty_b[s], 1);
            }
            mbar_init(&tfull[0], 1);
            mbar_init(&tfull[1], 1);
            mbar_init(&tempty[0], 8);
            mbar_init(&tempty[1], 8);
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc<Cfg::kTmemCols>(tmem_ptr);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    pdl_wait();      // everything above (barriers, TMEM, descriptor prefetch) overlaps the previous kernel's tail

    const int total_tiles = p.num_m_tiles * p.num_n_tiles;
    const int per_img = p.tiles_h * p.tiles_w;

    if (warp == 0) {
        // ------------------------------ TMA producer (whole warp loops, one elected lane issues) ----------
        uint32_t a_cnt = 0, b_cnt = 0;   // stages / weight tiles issued so far (uniform across the warp)
        bool first = true;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            const int m_tile = tile / p.num_n_tiles;
            const int n_tile = tile - m_tile * p.num_n_tiles;
            int n0, h0 = 0, w0 = 0;
            if (MODE == 0) {
                n0 = m_tile / per_img;
                const int rem = m_tile - n0 * per_img;
                const int th = rem / p.tiles_w;
                h0 = th * 16;
                w0 = (rem - th * p.tiles_w) * 8;
            } else {
                n0 = m_tile * 2;
            }
            const int brow0 = n_tile * BN;
            for (int kc = 0; kc < p.k_chunks; ++kc) {
#pragma unroll
                for (int l = 0; l < LOADS; ++l) {
                    const uint32_t sa = a_cnt & (kSA - 1);
                    mbar_wait(&empty_a[sa], ((a_cnt / kSA) & 1u) ^ 1u, 21);
                    if (elect_one_sync()) {
                        mbar_arrive_expect_tx(&full_a[sa], kABytes);
                        tma_load_4d(sA + sa * kAStageBytes, &mapA, &full_a[sa], kc * 64, w0 - 1,
                                    MODE == 0 ? h0 - 1 : l - 1, n0);
                        if (!RES || first) {
#pragma unroll
                            for (int j = 0; j < TPL; ++j) {
                                const int a = MODE == 0 ? j / 3 : l;
                                const int b = MODE == 0 ? j % 3 : j;
                                const int r = p.flip ? 2 - a : a;
                                const int s = p.flip ? 2 - b : b;
                                const uint32_t bi = RES ? static_cast<uint32_t>(l * TPL + j) : ((b_cnt + j) % kSB);
                                if (!RES) mbar_wait(&empty_b[bi], (((b_cnt + j) / kSB) & 1u) ^ 1u, 22);
                                mbar_arrive_expect_tx(&full_b[bi], Cfg::kBBytes);
                                tma_load_2d(sB + bi * Cfg::kBBytes, &mapB, &full_b[bi], kc * 64,
                                            (r * 3 + s) * p.b_rows_per_tap + brow0);
                            }
                        }
                    }
                    __syncwarp();
                    ++a_cnt;
                    b_cnt += TPL;
                }
            }
            first = false;
        }
    } else if (warp == 1) {
        // ------------------------------ MMA issuer (whole warp loops, one elected lane issues) ------------
        constexpr uint32_t idesc = make_idesc_bf16(128, BN, 0, 0);
        uint32_t a_cnt = 0, b_cnt = 0;
        int acc = 0;
        uint32_t acc_phase = 0;
        bool first = true;
        const uint32_t sB_u32 = smem_u32(sB);
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            mbar_wait(&tempty[acc], acc_phase ^ 1u, 23);
            tc_fence_after();
            const uint32_t d = tmem_base + static_cast<uint32_t>(acc * BN);
            for (int kc = 0; kc < p.k_chunks; ++kc) {
#pragma unroll
                for (int l = 0; l < LOADS; ++l) {
                    const uint32_t sa = a_cnt & (kSA - 1);
                    mbar_wait(&full_a[sa], (a_cnt / kSA) & 1u, 24);
                    tc_fence_after();
                    if (elect_one_sync()) {
                        const uint64_t ad0 = make_smem_desc_sw128(smem_u32(sA + sa * kAStageBytes), 16, 1280);
#pragma unroll
                        for (int j = 0; j < TPL; ++j) {
                            const int off_rows = MODE == 0 ? (j / 3) * 10 + (j % 3) : j;
                            const uint32_t bi = RES ? static_cast<uint32_t>(l * TPL + j) : ((b_cnt + j) % kSB);
                            if (!RES) {
                                mbar_wait(&full_b[bi], ((b_cnt + j) / kSB) & 1u, 25);
                                tc_fence_after();
                            } else if (first) {
                                mbar_wait(&full_b[bi], 0, 25);
                                tc_fence_after();
                            }
                            const uint64_t ad = ad0 + static_cast<uint64_t>(off_rows * 8);   // 128 B rows, >> 4
                            const uint64_t bd = make_smem_desc_sw128(sB_u32 + bi * Cfg::kBBytes, 16, 1024);
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                umma_bf16(d, ad + 2u * k, bd + 2u * k, idesc, (kc | l | j | k) != 0 ? 1u : 0u);
                            }
                            if (!RES) umma_commit(&empty_b[bi]);
                        }
                        umma_commit(&empty_a[sa]);
                    }
                    __syncwarp();
                    ++a_cnt;
                    b_cnt += TPL;
                }
            }
            if (elect_one_sync()) umma_commit(&tfull[acc]);
            __syncwarp();
            if (++acc == 2) {
                acc = 0;
                acc_phase ^= 1u;
            }
            first = false;
        }
    } else {
        // ---------------------------------- epilogue --------------------------------------
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int g = row >> 3, j = row & 7;
        const int half = (warp - 2) >> 2;   // two warps per TMEM lane quarter, alternating 32-column chunks
        const int et = threadIdx.x - 64;
        // BN == 64: a warp owns the same 32 channels in every tile, so the per-channel sums run in registers (one
        // add + one fma per element) and the cross-lane reduction happens once, after the last tile.
        constexpr bool kRegStats = kStats && BN == 64;
        float rs1[32], rs2[32];
        if constexpr (kRegStats) {
#pragma unroll
            for (int e = 0; e < 32; ++e) { rs1[e] = 0.f; rs2[e] = 0.f; }
        }
        if (kStats) {
            for (int c = lane + 32 * half; c < BN; c += 64) {   // the chunks this warp owns
                stat[(0 * 4 + q) * BN + c] = 0.f;
                stat[(1 * 4 + q) * BN + c] = 0.f;
            }
        }
        const bool bwd_fused = kStats && p.bwd_y != nullptr;
        if (bwd_fused) {
            const int ch0 = (blockIdx.x % p.num_n_tiles) * BN;     // constant per CTA (see the launcher)
            for (int i = et; i < BN; i += 256) {
                bwd_sc[i] = __ldg(p.bwd_scale + ch0 + i);
                bwd_sc[BN + i] = __ldg(p.bwd_shift + ch0 + i);
            }
            named_bar_sync(1, 256);
        }
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
            const int m_tile = tile / p.num_n_tiles;
            const int n_tile = tile - m_tile * p.num_n_tiles;
            int n, h, w;
            if (MODE == 0) {
                n = m_tile / per_img;
                const int rem = m_tile - n * per_img;
                const int th = rem / p.tiles_w;
                h = th * 16 + g;
                w = (rem - th * p.tiles_w) * 8 + j;
            } else {
                n = m_tile * 2 + (g >> 3);
                h = g & 7;
                w = j;
            }
            const bool valid = n < p.N;
            const size_t off = (static_cast<size_t>(n * p.H + h) * p.W + w) * p.out_ld;
            const int col0 = n_tile * BN;
            unsigned long long rowoff[4];
            chunk_rowoffs(off, valid, lane, rowoff);
            uint8_t* stg_w = store_stage + (warp - 2) * 2048;
            mbar_wait(&tfull[acc], acc_phase, 26);
            tc_fence_after();
#pragma unroll 1
            for (int c = half; c < BN / 32; c += 2) {
                float v[32];
                tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                                  static_cast<uint32_t>(acc * BN + c * 32),
                              v);
                if (p.add_src != nullptr) {
                    uint4 au[4];
                    load_chunk_staged(stg_w, p.add_src, rowoff, col0 + c * 32, au, lane);
#pragma unroll
                    for (int g4 = 0; g4 < 4; ++g4) {
                        v[g4 * 8 + 0] += bf16lo(au[g4].x);
                        v[g4 * 8 + 1] += bf16hi(au[g4].x);
                        v[g4 * 8 + 2] += bf16lo(au[g4].y);
                        v[g4 * 8 + 3] += bf16hi(au[g4].y);
                        v[g4 * 8 + 4] += bf16lo(au[g4].z);
                        v[g4 * 8 + 5] += bf16hi(au[g4].z);
                        v[g4 * 8 + 6] += bf16lo(au[g4].w);
                        v[g4 * 8 + 7] += bf16hi(au[g4].w);
                    }
                }
                float sq[32];      // second statistic operand (BN > 64 only; 64-wide tiles accumulate in registers)
                if (kStats) {
                    if (bwd_fused) {
                        const float* scs = bwd_sc + c * 32;
                        const float* shs = bwd_sc + BN + c * 32;
                        uint4 yall[4];
                        load_chunk_staged(stg_w, p.bwd_y, rowoff, col0 + c * 32, yall, lane);
#pragma unroll
                        for (int g4 = 0; g4 < 4; ++g4) {
                            const uint4 yu = yall[g4];
                            const float yy[8] = {bf16lo(yu.x), bf16hi(yu.x), bf16lo(yu.y), bf16hi(yu.y),
                                                 bf16lo(yu.z), bf16hi(yu.z), bf16lo(yu.w), bf16hi(yu.w)};
#pragma unroll
                            for (int k8 = 0; k8 < 8; ++k8) {
                                const int e = g4 * 8 + k8;
                                const float dz = fmaf(yy[k8], scs[e], shs[e]) > 0.f ? v[e] : 0.f;
                                v[e] = dz;
                                if constexpr (kRegStats) {
                                    rs1[e] += dz;
                                    rs2[e] = fmaf(dz, yy[k8], rs2[e]);
                                } else {
                                    sq[e] = dz * yy[k8];
                                }
                            }
                        }
                    } else {
                        // (rows past the end were zero-filled by TMA, so they add nothing)
#pragma unroll
                        for (int e = 0; e < 32; ++e) {
                            if constexpr (kRegStats) {
                                rs1[e] += v[e];
                                rs2[e] = fmaf(v[e], v[e], rs2[e]);
                            } else {
                                sq[e] = v[e] * v[e];
                            }
                        }
                    }
                }
                {
                    uint4 u[4];
#pragma unroll
                    for (int gg = 0; gg < 4; ++gg) {
                        u[gg].x = pack_bf16x2(v[gg * 8 + 0], v[gg * 8 + 1]);
                        u[gg].y = pack_bf16x2(v[gg * 8 + 2], v[gg * 8 + 3]);
                        u[gg].z = pack_bf16x2(v[gg * 8 + 4], v[gg * 8 + 5]);
                        u[gg].w = pack_bf16x2(v[gg * 8 + 6], v[gg * 8 + 7]);
                    }
                    store_chunk_staged(stg_w, p.out, rowoff, col0 + c * 32, u, lane);
                }
                if constexpr (kStats && !kRegStats) {
                    const float s1 = warp_colsum32(v, lane);
                    const float s2 = warp_colsum32(sq, lane);
                    stat[(0 * 4 + q) * BN + c * 32 + lane] += s1;
                    stat[(1 * 4 + q) * BN + c * 32 + lane] += s2;
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[acc]);
            if (++acc == 2) {
                acc = 0;
                acc_phase ^= 1u;
            }
        }
        if constexpr (kRegStats) {
            const float s1 = warp_colsum32(rs1, lane);
            const float s2 = warp_colsum32(rs2, lane);
            stat[(0 * 4 + q) * BN + half * 32 + lane] = s1;
            stat[(1 * 4 + q) * BN + half * 32 + lane] = s2;
        }
        if (kStats) {
            named_bar_sync(1, 256);
            const int n_tile = blockIdx.x % p.num_n_tiles;
            const int prow = blockIdx.x / p.num_n_tiles;
            for (int col = et; col < BN; col += 256) {
                const float s1 = stat[(0 * 4 + 0) * BN + col] + stat[(0 * 4 + 1) * BN + col] +
                                 stat[(0 * 4 + 2) * BN + col] + stat[(0 * 4 + 3) * BN + col];
                const float s2 = stat[(1 * 4 + 0) * BN + col] + stat[(1 * 4 + 1) * BN + col] +
                                 stat[(1 * 4 + 2) * BN + col] + stat[(1 * 4 + 3) * BN + col];
                const size_t o = static_cast<size_t>(prow) * p.stat_ld + n_tile * BN + col;
                if (p.stat_ld == 0) {      // one accumulated row (flags bit 9): the caller zero-filled it
                    atomicAdd(p.part_sum + o, s1);
                    atomicAdd(p.part_sq + o, s2);
                } else {
                    p.part_sum[o] = s1;
                    p.part_sq[o] = s2;
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    if (warp == 1) tmem_dealloc<Cfg::kTmemCols>(tmem_base);
}

template <int BN, bool kStats, int MODE, bool RES>
cudaError_t launch_inst(const CUtensorMap& a, const CUtensorMap& b, const HaloParams& p, int grid,
                        cudaStream_t stream) {
    using Cfg = HaloCfg<BN>;
    auto kern = conv_halo_kernel<BN, kStats, MODE, RES>;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    B2C_LAUNCH(kern, grid, 320, Cfg::kSmemBytes, stream, a, b, p);
    return cudaGetLastError();
}

template <int BN, bool kStats>
cudaError_t launch_one(const CUtensorMap& a, const CUtensorMap& b, HaloParams p, int max_ctas, int* stat_rows,
                       cudaStream_t stream) {
    const int total = p.num_m_tiles * p.num_n_tiles;
    int grid = total < max_ctas ? total : max_ctas;
    grid -= grid % p.num_n_tiles;
    if (grid < p.num_n_tiles) grid = p.num_n_tiles;
    if (stat_rows != nullptr) *stat_rows = grid / p.num_n_tiles;
    p.b_resident = 0;
    if (p.mode == 0) {
        if constexpr (BN == 64) {
            if (p.k_chunks == 1 && p.num_n_tiles == 1) {
                p.b_resident = 1;
                return launch_inst<64, kStats, 0, true>(a, b, p, grid, stream);
            }
        }
        return launch_inst<BN, kStats, 0, false>(a, b, p, grid, stream);
    }
    return launch_inst<BN, kStats, 1, false>(a, b, p, grid, stream);
}

}  // namespace

cudaError_t launch_conv_halo(int bn, const CUtensorMap& a, const CUtensorMap& b, const HaloParams& p, int max_ctas,
                             int* stat_rows, cudaStream_t stream) {
    const bool stats = p.part_sum != nullptr;
    switch (bn) {
        case 64:
            return stats ? launch_one<64, true>(a, b, p, max_ctas, stat_rows, stream)
                         : launch_one<64, false>(a, b, p, max_ctas, stat_rows, stream);
        case 128:
            return stats ? launch_one<128, true>(a, b, p, max_ctas, stat_rows, stream)
                         : launch_one<128, false>(a, b, p, max_ctas, stat_rows, stream);
        case 256:
            return stats ? launch_one<256, true>(a, b, p, max_ctas, stat_rows, stream)
                         : launch_one<256, false>(a, b, p, max_ctas, stat_rows, stream);
        default:
            return cudaErrorInvalidValue;
    }
}

}  // namespace b2c
