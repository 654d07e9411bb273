// conv_gemm.cu — implicit-GEMM convolution forward / data-gradient on tcgen05 tensor cores.
//
// Replaces what `nn.Conv2d.forward` and its autograd dgrad resolve to in the reference
// (models/resnet.py:23-27,33-34,102-103; models/vgg.py:33; models/preact_resnet.py:19,21,70 and
// `loss.backward()` main.py:104), i.e. cuDNN / oneDNN library kernels. Here: one persistent,
// warp-specialised sm_100a kernel.
//
//   GEMM view      D[pixel, cout] = sum_{tap, cin} A[pixel shifted by tap, cin] * W[tap][cout, cin]
//   A operand      NHWC bf16 activations; one TMA 4-D box {64 ch, gW, Hb, Nb} = 128 pixels per k-step,
//                  shifted by the tap offset; out-of-image taps are zero-filled by the TMA unit
//                  (this is the conv's zero padding), stride-2 convs use the TMA traversal stride.
//   B operand      weights as [tap][cout][cin] bf16 (fprop) or [tap][cin][cout] (dgrad): 2-D TMA box.
//   accumulators   fp32 in TMEM, double buffered (2 x BN columns) so the epilogue of tile i overlaps
//                  the mainloop of tile i+1.
//   epilogue       tcgen05.ld -> optional residual add -> bf16 NHWC store, plus per-tile per-channel
//                  sum / sum-of-squares (fp32, taken before rounding) for the following BatchNorm.
//
// Warp roles (320 threads): warp 0 = TMA producer, warp 1 = TMEM owner + MMA issuer,
// warps 2..9 = epilogue (one TMEM lane quarter each).
#include "b2c_kernels.h"
#include "b2c_launch.h"
#include "b2c_ptx.cuh"

namespace b2c {

namespace {

constexpr int kBM = 128;            // GEMM rows (pixels) per tile
constexpr int kABytes = kBM * 128;  // 128 rows x 64 bf16

template <int BN>
struct TapCfg {
    static constexpr int kBBytes = BN * 128;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kStages = BN == 64 ? 8 : (BN == 128 ? 6 : 4);
    static constexpr int kTmemCols = 2 * BN;
    static constexpr int kBarBytes = 256;                       // (2*stages + 4) mbarriers + tmem ptr
    static constexpr int kStatBytes = 2 * 4 * BN * 4 + 2 * BN * 4;   // running sums + fused-BN-backward scale|shift           // [sum|sq][quarter][BN] fp32 running sums
    static constexpr int kStoreBytes = 8 * 2048;                // per-epilogue-warp store staging (32 rows x 64 B)
    static constexpr int kSmemBytes = 1024 + kStages * kStageBytes + kBarBytes + kStatBytes + kStoreBytes;
};

template <int BN, bool kStats>
__global__ void __launch_bounds__(320, 1)
tap_gemm_kernel(const __grid_constant__ CUtensorMap mapA0, const __grid_constant__ CUtensorMap mapA1,
                const __grid_constant__ CUtensorMap mapA2, const __grid_constant__ CUtensorMap mapA3,
                const __grid_constant__ CUtensorMap mapB, const __grid_constant__ TapGemmParams p) {
    using Cfg = TapCfg<BN>;
    pdl_trigger_big();
    extern __shared__ uint8_t smem_raw[];
    // (offset arithmetic on the shared pointer itself keeps the address space visible to the compiler: plain
    // loads / stores below become LDS / STS instead of generic accesses)
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* aux = smem + Cfg::kStages * Cfg::kStageBytes;
    uint64_t* full = reinterpret_cast<uint64_t*>(aux);
    uint64_t* empty = full + Cfg::kStages;
    uint64_t* tfull = empty + Cfg::kStages;
    uint64_t* tempty = tfull + 2;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty + 2);
    float* stat = reinterpret_cast<float*>(aux + Cfg::kBarBytes);
    float* bwd_sc = stat + 2 * 4 * BN;          // [scale | shift] of this CTA's channels (fused BN backward)
    uint8_t* store_stage = aux + Cfg::kBarBytes + Cfg::kStatBytes;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&mapA0);
        tma_prefetch_desc(&mapA1);
        tma_prefetch_desc(&mapA2);
        tma_prefetch_desc(&mapA3);
        tma_prefetch_desc(&mapB);
    }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < Cfg::kStages; ++s) {
                mbar_init(&full[s], 1);
                mbar_init(&empty[s], 1);
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

    const int total_tiles = p.num_m_tiles * p.num_n_tiles * p.num_phases;
    const int img = p.gH * p.gW;
    // tile -> (m_tile, phase, n_tile), n_tile fastest. The phase is rotated by m_tile so that a CTA whose tile
    // stride is a multiple of num_phases still sees every phase (they differ in tap count).
    auto decode = [&](int tile, int& m_tile, int& n_tile, int& ph) {
        const int rest = tile / p.num_n_tiles;
        n_tile = tile - rest * p.num_n_tiles;
        if (p.num_phases == 1) {
            m_tile = rest;
            ph = 0;
        } else {
            m_tile = rest / p.num_phases;
            ph = (rest - m_tile * p.num_phases + m_tile) % p.num_phases;
        }
    };

    if (warp == 0) {
        {
            // ------------------------------ TMA producer (whole warp loops, one elected lane issues) ----
            int stage = 0;
            uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                int m_tile, n_tile, ph;
                decode(tile, m_tile, n_tile, ph);
                const int p0 = m_tile * kBM;
                const int n0 = p0 / img;
                const int h0 = (p0 - n0 * img) / p.gW;
                const int t_end = p.phase_begin[ph + 1];
                for (int t = p.phase_begin[ph]; t < t_end; ++t) {
                    const int am = p.tap_a[t];
                    const CUtensorMap* ma = am == 0 ? &mapA0 : (am == 1 ? &mapA1 : (am == 2 ? &mapA2 : &mapA3));
                    const int cw = p.tap_dw[t];
                    const int ch = h0 * p.in_stride + p.tap_dh[t];
                    const int brow = p.tap_brow[t] + n_tile * BN;
                    const int nk = p.tap_kc[t] != 0 ? p.tap_kc[t] : p.k_chunks;
                    for (int kc = 0; kc < nk; ++kc) {
                        mbar_wait(&empty[stage], phase ^ 1u, 1);
                        if (elect_one_sync()) {
                            mbar_arrive_expect_tx(&full[stage], Cfg::kStageBytes);
                            uint8_t* sa = smem + stage * Cfg::kStageBytes;
                            if (p.a_split_w)     // w = 2*w2 + parity: {c, parity, w2, h, n}; -1 -> (1, -1)
                                tma_load_5d(sa, ma, &full[stage], kc * 64, cw & 1, cw >> 1, ch, n0);
                            else
                                tma_load_4d(sa, ma, &full[stage], kc * 64, cw, ch, n0);
                            tma_load_2d(sa + kABytes, &mapB, &full[stage], kc * 64, brow);
                        }
                        __syncwarp();
                        if (++stage == Cfg::kStages) {
                            stage = 0;
                            phase ^= 1u;
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        {
            // ------------------------------ MMA issuer (whole warp loops, one elected lane issues) ------
            constexpr uint32_t idesc = make_idesc_bf16(kBM, BN, 0, 0);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
                int m_tile, n_tile, ph;
                decode(tile, m_tile, n_tile, ph);
                int k_iters = 0;
                for (int t = p.phase_begin[ph]; t < p.phase_begin[ph + 1]; ++t)
                    k_iters += p.tap_kc[t] != 0 ? p.tap_kc[t] : p.k_chunks;
                mbar_wait(&tempty[acc], acc_phase ^ 1u, 2);
                tc_fence_after();
                const uint32_t d = tmem_base + static_cast<uint32_t>(acc * BN);
                for (int it = 0; it < k_iters; ++it) {
                    mbar_wait(&full[stage], phase, 3);
                    tc_fence_after();
                    if (elect_one_sync()) {
                        const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
                        const uint64_t ad = make_smem_desc_sw128(sa, 16, 1024);
                        const uint64_t bd = make_smem_desc_sw128(sa + kABytes, 16, 1024);
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            // advance 16 bf16 = 32 B along K inside the 128 B swizzle span
                            umma_bf16(d, ad + 2u * k, bd + 2u * k, idesc, (it | k) != 0 ? 1u : 0u);
                        }
                        umma_commit(&empty[stage]);
                    }
                    __syncwarp();
                    if (++stage == Cfg::kStages) {
                        stage = 0;
                        phase ^= 1u;
                    }
                }
                if (elect_one_sync()) umma_commit(&tfull[acc]);
                __syncwarp();
                if (++acc == 2) {
                    acc = 0;
                    acc_phase ^= 1u;
                }
            }
        }
    } else {
        // ---------------------------------- epilogue --------------------------------------
        const int q = warp & 3;  // TMEM lane quarter this warp may read
        const int row = q * 32 + lane;
        const int half = (warp - 2) >> 2;   // two warps per TMEM lane quarter, alternating 32-column chunks
        const int et = threadIdx.x - 64;  // 0..127
        // BN == 64: a warp owns the same 32 channels in every tile, so the per-channel sums run in registers (one
        // add + one fma per element) and the cross-lane reduction happens once, after the last tile.
        constexpr bool kRegStats = kStats && BN == 64;
        float rs1[32], rs2[32];
        if constexpr (kRegStats) {
#pragma unroll
            for (int e = 0; e < 32; ++e) { rs1[e] = 0.f; rs2[e] = 0.f; }
        }
        if (kStats) {
            // each (quarter, column) slot is owned by exactly one thread: no sync needed to clear it
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
            int m_tile, n_tile, ph;
            decode(tile, m_tile, n_tile, ph);
            mbar_wait(&tfull[acc], acc_phase, 4);
            tc_fence_after();
            const int pix = m_tile * kBM + row;
            const bool valid = pix < p.m_total;
            const int n = pix / img;
            const int rem = pix - n * img;
            const int i = rem / p.gW;
            const int j = rem - i * p.gW;
            const size_t off =
                (static_cast<size_t>(n * p.out_H + i * p.sy + p.phase_py[ph]) * p.out_W + (j * p.sx + p.phase_px[ph])) *
                p.out_ld;
            const int col0 = n_tile * BN;
            unsigned long long rowoff[4];
            chunk_rowoffs(off, valid, lane, rowoff);
            uint8_t* stg_w = store_stage + (warp - 2) * 2048;
            float* st = stat;
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
                    for (int g = 0; g < 4; ++g) {
                        u[g].x = pack_bf16x2(v[g * 8 + 0], v[g * 8 + 1]);
                        u[g].y = pack_bf16x2(v[g * 8 + 2], v[g * 8 + 3]);
                        u[g].z = pack_bf16x2(v[g * 8 + 4], v[g * 8 + 5]);
                        u[g].w = pack_bf16x2(v[g * 8 + 6], v[g * 8 + 7]);
                    }
                    store_chunk_staged(stg_w, p.out, rowoff, col0 + c * 32, u, lane);
                }
                if constexpr (kStats && !kRegStats) {
                    const float s1 = warp_colsum32(v, lane);
                    const float s2 = warp_colsum32(sq, lane);
                    st[(0 * 4 + q) * BN + c * 32 + lane] += s1;
                    st[(1 * 4 + q) * BN + c * 32 + lane] += s2;
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
            // gridDim.x is a multiple of num_n_tiles (see launch_one), so every tile of this CTA had
            // the same n_tile: one partial row per CTA, P = gridDim.x / num_n_tiles rows in all.
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

template <int BN, bool kStats>
cudaError_t launch_one(const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& a2, const CUtensorMap& a3,
                       const CUtensorMap& b, const TapGemmParams& p, int max_ctas, int* stat_rows, cudaStream_t stream) {
    using Cfg = TapCfg<BN>;
    auto kern = tap_gemm_kernel<BN, kStats>;
    static bool attr_done = false;  // per instantiation
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    const int total = p.num_m_tiles * p.num_n_tiles;
    int grid = total < max_ctas ? total : max_ctas;
    grid -= grid % p.num_n_tiles;  // keeps n_tile constant per CTA (per-CTA BatchNorm partial rows)
    if (grid < p.num_n_tiles) grid = p.num_n_tiles;
    if (stat_rows != nullptr) *stat_rows = grid / p.num_n_tiles;
    B2C_LAUNCH(kern, grid, 320, Cfg::kSmemBytes, stream, a0, a1, a2, a3, b, p);
    return cudaGetLastError();
}

}  // namespace

cudaError_t launch_tap_gemm4(int bn, const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& a2,
                             const CUtensorMap& a3, const CUtensorMap& b, const TapGemmParams& p, int max_ctas,
                             int* stat_rows, cudaStream_t stream) {
    const bool stats = p.part_sum != nullptr;
    switch (bn) {
        case 64:
            return stats ? launch_one<64, true>(a0, a1, a2, a3, b, p, max_ctas, stat_rows, stream)
                         : launch_one<64, false>(a0, a1, a2, a3, b, p, max_ctas, stat_rows, stream);
        case 128:
            return stats ? launch_one<128, true>(a0, a1, a2, a3, b, p, max_ctas, stat_rows, stream)
                         : launch_one<128, false>(a0, a1, a2, a3, b, p, max_ctas, stat_rows, stream);
        case 256:
            return stats ? launch_one<256, true>(a0, a1, a2, a3, b, p, max_ctas, stat_rows, stream)
                         : launch_one<256, false>(a0, a1, a2, a3, b, p, max_ctas, stat_rows, stream);
        default:
            return cudaErrorInvalidValue;
    }
}

cudaError_t launch_tap_gemm(int bn, const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& b,
                            const TapGemmParams& p, int max_ctas, int* stat_rows, cudaStream_t stream) {
    return launch_tap_gemm4(bn, a0, a1, a1, a1, b, p, max_ctas, stat_rows, stream);
}

}  // namespace b2c
