// conv_gemm2.cu — the per-tap implicit-GEMM convolution of conv_gemm.cu on a CTA PAIR (tcgen05.mma.cta_group::2).
//
// Same GEMM, operands, epilogue and reference sites as conv_gemm.cu (nn.Conv2d forward and its dgrad,
// models/resnet.py:23-27,33-34; main.py:104). What changes is who feeds the tensor core: two CTAs of a cluster work on
// two adjacent 128-pixel tiles (M = 256) and ONE weight tile; each CTA loads its own activation box and HALF of the
// weight tile (BN/2 rows), the leader CTA issues `tcgen05.mma.cta_group::2` for both. A single CTA is bound by its
// shared-memory operand port (33 / 50 / 67 % of the tensor peak for N = 64 / 128 / 256); the pair runs at the peak from
// N = 128 (profiles/r1g_umma2_probe.log).
//
// Cross-CTA synchronisation (all mbarriers live at the same offsets in both CTAs):
//   full[s]     in the LEADER: both CTAs' TMA loads (cp.async.bulk.tensor.cta_group::2) signal their bytes there
//   empty[s]    both CTAs, arrived by the leader's tcgen05.commit multicast: stage s has been consumed
//   tfull[a]    both CTAs, commit multicast: accumulator a is complete (each CTA reads its own 128 rows from TMEM)
//   tempty[a]   leader: its own epilogue warps drained accumulator a;  ptempty[a]: the peer's (remote arrive)
// Output phases (the four parity phases of a stride-2 dgrad, written through a strided output) and the second
// activation map (the fused 1x1 shortcut dgrad) work as in conv_gemm.cu: the two CTAs of a pair always run the SAME
// phase and weight tile on adjacent pixel tiles, so they share every B tile and tap sequence.
#include "b2c_kernels.h"
#include "b2c_launch.h"
#include "b2c_ptx.cuh"

namespace b2c {

namespace {

constexpr int kBM = 128;            // GEMM rows (pixels) per tile and CTA
constexpr int kABytes = kBM * 128;  // 128 rows x 64 bf16

template <int BN>
struct Tap2Cfg {
    static constexpr int kBBytes = (BN / 2) * 128;               // this CTA's half of the weight tile
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kStages = BN == 64 ? 8 : (BN == 128 ? 7 : 5);
    static constexpr int kTmemCols = 2 * BN;
    static constexpr int kBarBytes = 320;                        // (3*stages + 6) mbarriers + tmem ptr
    static constexpr int kStatBytes = 2 * 4 * BN * 4 + 2 * BN * 4;
    static constexpr int kStoreBytes = 8 * 2048;
    static constexpr int kSmemBytes = 1024 + kStages * kStageBytes + kBarBytes + kStatBytes + kStoreBytes;
    static_assert((3 * kStages + 6) * 8 + 4 <= kBarBytes, "barrier block");
};

__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the mbarrier at the same shared-memory offset in CTA `target` of the cluster
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* bar, uint32_t target) {
    uint32_t raddr;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(raddr) : "r"(smem_u32(bar)), "r"(target));
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(raddr) : "memory");
}
// TMA loads of a CTA pair: the data lands in THIS CTA's shared memory, the transaction bytes are signalled on the
// barrier at the same offset in the LEADER CTA (rank 0), which the MMA issuer waits on — no relay through the peer.
__device__ __forceinline__ uint32_t cluster_addr(const void* p, uint32_t rank) {
    uint32_t raddr;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(raddr) : "r"(smem_u32(p)), "r"(rank));
    return raddr;
}
__device__ __forceinline__ void tma2_load_4d(uint32_t dst_cluster, const CUtensorMap* m, uint32_t bar_cluster, int c0,
                                             int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst_cluster),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void tma2_load_5d(uint32_t dst_cluster, const CUtensorMap* m, uint32_t bar_cluster, int c0,
                                             int c1, int c2, int c3, int c4) {
    asm volatile(
        "cp.async.bulk.tensor.5d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];" ::"r"(dst_cluster),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4)
        : "memory");
}
__device__ __forceinline__ void tma2_load_2d(uint32_t dst_cluster, const CUtensorMap* m, uint32_t bar_cluster, int c0,
                                             int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%3, %4}], [%2];" ::"r"(dst_cluster),
        "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_cluster), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ void umma2_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                           uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive on `bar` in BOTH CTAs once every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void umma2_commit_both(uint64_t* bar) {
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
            smem_u32(bar)),
        "h"(static_cast<uint16_t>(3))
        : "memory");
}

template <int BN, bool kStats>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(320, 1)
tap_gemm2_kernel(const __grid_constant__ CUtensorMap mapA0, const __grid_constant__ CUtensorMap mapA1,
                 const __grid_constant__ CUtensorMap mapB, const __grid_constant__ TapGemmParams p) {
    using Cfg = Tap2Cfg<BN>;
    pdl_trigger_big();
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* aux = smem + Cfg::kStages * Cfg::kStageBytes;
    uint64_t* full = reinterpret_cast<uint64_t*>(aux);
    uint64_t* empty = full + Cfg::kStages;
    uint64_t* pfull = empty + Cfg::kStages;
    uint64_t* tfull = pfull + Cfg::kStages;
    uint64_t* tempty = tfull + 2;
    uint64_t* ptempty = tempty + 2;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(ptempty + 2);
    float* stat = reinterpret_cast<float*>(aux + Cfg::kBarBytes);
    float* bwd_sc = stat + 2 * 4 * BN;
    uint8_t* store_stage = aux + Cfg::kBarBytes + Cfg::kStatBytes;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_rank();          // 0 = leader (issues the MMAs), 1 = peer
    const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&mapA0);
        tma_prefetch_desc(&mapA1);
        tma_prefetch_desc(&mapB);
    }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < Cfg::kStages; ++s) {
                mbar_init(&full[s], 1);
                mbar_init(&empty[s], 1);
                mbar_init(&pfull[s], 1);
            }
            for (int a = 0; a < 2; ++a) {
                mbar_init(&tfull[a], 1);
                mbar_init(&tempty[a], 8);
                mbar_init(&ptempty[a], 8);
            }
            fence_mbar_init();
        }
        __syncwarp();
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)),
                     "n"(Cfg::kTmemCols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();      // both CTAs' barriers are initialised and their TMEM allocated before any remote arrive
    tc_fence_after();
    const uint32_t tmem_base = *tmem_ptr;
    pdl_wait();

    // pair tiles: (m_pair, phase, n_tile), n_tile fastest; this CTA's pixel tile is 2*m_pair + rank. The phase is
    // rotated by m_pair so that a pair whose tile stride is a multiple of num_phases still sees every phase (they
    // differ in tap count).
    const int m_pairs = (p.num_m_tiles + 1) >> 1;
    const int total_tiles = m_pairs * p.num_n_tiles * p.num_phases;
    const int img = p.gH * p.gW;
    auto decode = [&](int tile, int& m_tile, int& n_tile, int& ph) {
        const int rest = tile / p.num_n_tiles;
        n_tile = tile - rest * p.num_n_tiles;
        int mp;
        if (p.num_phases == 1) {
            mp = rest;
            ph = 0;
        } else {
            mp = rest / p.num_phases;
            ph = (rest - mp * p.num_phases + mp) % p.num_phases;
        }
        m_tile = 2 * mp + static_cast<int>(rank);
    };

    if (warp == 0) {
        // ------------------------------ TMA producer (both CTAs) ------------------------------
        int stage = 0;
        uint32_t phase = 0;
        for (int tile = pair; tile < total_tiles; tile += num_pairs) {
            int m_tile, n_tile, ph;
            decode(tile, m_tile, n_tile, ph);
            const int p0 = m_tile * kBM;             // past the last tile for an odd tile count: TMA zero-fills
            const int n0 = p0 / img;
            const int h0 = (p0 - n0 * img) / p.gW;
            const int t_end = p.phase_begin[ph + 1];
            for (int t = p.phase_begin[ph]; t < t_end; ++t) {
                const CUtensorMap* ma = p.tap_a[t] ? &mapA1 : &mapA0;
                const int cw = p.tap_dw[t];
                const int ch = h0 * p.in_stride + p.tap_dh[t];
                const int brow = p.tap_brow[t] + n_tile * BN + static_cast<int>(rank) * (BN / 2);
                for (int kc = 0; kc < p.k_chunks; ++kc) {
                    mbar_wait(&empty[stage], phase ^ 1u, 51);
                    if (elect_one_sync()) {
                        // the leader's barrier collects both CTAs' bytes
                        if (rank == 0) mbar_arrive_expect_tx(&full[stage], 2 * Cfg::kStageBytes);
                        uint8_t* sa = smem + stage * Cfg::kStageBytes;
                        const uint32_t bar = cluster_addr(&full[stage], 0);
                        if (p.a_split_w)
                            tma2_load_5d(cluster_addr(sa, rank), ma, bar, kc * 64, cw & 1, cw >> 1, ch, n0);
                        else
                            tma2_load_4d(cluster_addr(sa, rank), ma, bar, kc * 64, cw, ch, n0);
                        tma2_load_2d(cluster_addr(sa + kABytes, rank), &mapB, bar, kc * 64, brow);
                    }
                    __syncwarp();
                    if (++stage == Cfg::kStages) {
                        stage = 0;
                        phase ^= 1u;
                    }
                }
            }
        }
    } else if (warp == 1 && rank == 1) {
        // the peer's MMA warp has nothing to do: its TMA loads signal the leader's barriers directly
    } else if (warp == 1) {
        // ------------------------------ leader: MMA issuer for the pair (M = 256) ----------------
        constexpr uint32_t idesc = make_idesc_bf16(2 * kBM, BN, 0, 0);
        int stage = 0;
        uint32_t phase = 0;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int tile = pair; tile < total_tiles; tile += num_pairs) {
            int m_tile, n_tile, ph;
            decode(tile, m_tile, n_tile, ph);
            const int k_iters = (p.phase_begin[ph + 1] - p.phase_begin[ph]) * p.k_chunks;
            mbar_wait(&tempty[acc], acc_phase ^ 1u, 53);
            mbar_wait(&ptempty[acc], acc_phase ^ 1u, 54);
            tc_fence_after();
            const uint32_t d = tmem_base + static_cast<uint32_t>(acc * BN);
            for (int it = 0; it < k_iters; ++it) {
                mbar_wait(&full[stage], phase, 55);
                tc_fence_after();
                if (elect_one_sync()) {
                    const uint32_t sa = smem_u32(smem + stage * Cfg::kStageBytes);
                    const uint64_t ad = make_smem_desc_sw128(sa, 16, 1024);
                    const uint64_t bd = make_smem_desc_sw128(sa + kABytes, 16, 1024);
#pragma unroll
                    for (int k = 0; k < 4; ++k) umma2_bf16(d, ad + 2u * k, bd + 2u * k, idesc, (it | k) != 0 ? 1u : 0u);
                    umma2_commit_both(&empty[stage]);
                }
                __syncwarp();
                if (++stage == Cfg::kStages) {
                    stage = 0;
                    phase ^= 1u;
                }
            }
            if (elect_one_sync()) umma2_commit_both(&tfull[acc]);
            __syncwarp();
            if (++acc == 2) {
                acc = 0;
                acc_phase ^= 1u;
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
            const int ch0 = (pair % p.num_n_tiles) * BN;     // constant per CTA (see the launcher)
            for (int i = et; i < BN; i += 256) {
                bwd_sc[i] = __ldg(p.bwd_scale + ch0 + i);
                bwd_sc[BN + i] = __ldg(p.bwd_shift + ch0 + i);
            }
            named_bar_sync(1, 256);
        }
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int tile = pair; tile < total_tiles; tile += num_pairs) {
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
            if (lane == 0) {            // the leader's MMA warp waits for both CTAs' epilogues
                if (rank == 0) mbar_arrive(&tempty[acc]);
                else mbar_arrive_remote(&ptempty[acc], 0);
            }
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
            const int n_tile = pair % p.num_n_tiles;
            const int prow = (pair / p.num_n_tiles) * 2 + static_cast<int>(rank);
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
    cluster_sync_all();      // the peer's shared memory and barriers stay valid until the leader is done with them
    tc_fence_after();
    if (warp == 1)
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(Cfg::kTmemCols) : "memory");
}

template <int BN, bool kStats>
cudaError_t launch_one2(const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& b, const TapGemmParams& p,
                        int max_ctas, int* stat_rows, cudaStream_t stream) {
    using Cfg = Tap2Cfg<BN>;
    auto kern = tap_gemm2_kernel<BN, kStats>;
    static bool attr_done = false;  // per instantiation
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes);
        if (e != cudaSuccess) return e;
        attr_done = true;
    }
    const int total = ((p.num_m_tiles + 1) / 2) * p.num_n_tiles * p.num_phases;
    int pairs = total < max_ctas / 2 ? total : max_ctas / 2;
    pairs -= pairs % p.num_n_tiles;   // keeps n_tile constant per pair (per-CTA BatchNorm partial rows)
    if (pairs < p.num_n_tiles) pairs = p.num_n_tiles;
    if (stat_rows != nullptr) *stat_rows = 2 * (pairs / p.num_n_tiles);
    B2C_LAUNCH(kern, 2 * pairs, 320, Cfg::kSmemBytes, stream, a0, a1, b, p);
    return cudaGetLastError();
}

}  // namespace

// b: 2-D map of the weight matrix with a box of BN/2 rows (each CTA of a pair loads half a tile)
cudaError_t launch_tap_gemm2(int bn, const CUtensorMap& a0, const CUtensorMap& a1, const CUtensorMap& b,
                             const TapGemmParams& p, int max_ctas, int* stat_rows, cudaStream_t stream) {
    const bool stats = p.part_sum != nullptr;
    switch (bn) {
        case 64:
            return stats ? launch_one2<64, true>(a0, a1, b, p, max_ctas, stat_rows, stream)
                         : launch_one2<64, false>(a0, a1, b, p, max_ctas, stat_rows, stream);
        case 128:
            return stats ? launch_one2<128, true>(a0, a1, b, p, max_ctas, stat_rows, stream)
                         : launch_one2<128, false>(a0, a1, b, p, max_ctas, stat_rows, stream);
        case 256:
            return stats ? launch_one2<256, true>(a0, a1, b, p, max_ctas, stat_rows, stream)
                         : launch_one2<256, false>(a0, a1, b, p, max_ctas, stat_rows, stream);
        default:
            return cudaErrorInvalidValue;
    }
}

}  // namespace b2c
