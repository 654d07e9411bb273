// conv_halo2.cu — conv_halo.cu on a CTA PAIR: 3x3 stride-1 convolution forward / data-gradient with activation-halo
// reuse, issued as tcgen05.mma.cta_group::2 (M = 256: two adjacent pixel tiles, one per CTA of a 2-CTA cluster).
//
// Same GEMM view, halo boxes, tap descriptors and epilogue as conv_halo.cu (reference sites: models/resnet.py:23-27
// and the autograd dgrad of main.py:104). A single CTA is bound by its shared-memory operand port (33 / 50 % of the
// tensor peak for N = 64 / 128); the pair reaches 74 % / 100 % (profiles/r1g_umma2_probe.log), and the halo reuse
// keeps the operand traffic from L2 low enough to feed it (the per-tap pair kernel, conv_gemm2.cu, is L2-bound).
// Each CTA loads its own activation boxes and HALF of every weight tile (BN/2 rows); the leader CTA issues the MMAs.
//
// Cross-CTA synchronisation (mbarriers at the same offsets in both CTAs):
//   full_a/full_b   in the LEADER: both CTAs' TMA loads (cp.async.bulk.tensor.cta_group::2) signal their bytes there
//                   (a first version relayed the peer's local barriers with remote arrives: one relay per weight
//                   tile is slower than the 256 cycles its MMAs take)
//   empty_a/empty_b, tfull   arrived in BOTH CTAs by the leader's tcgen05.commit multicast
//   tempty (leader's own epilogue) / ptempty (the peer's epilogue, remote arrive)
//
// Warp roles (320 threads per CTA): warp 0 = TMA producer, warp 1 = MMA issuer (leader) / relay (peer),
// warps 2..9 = epilogue.
#include "b2c_kernels.h"
#include "b2c_launch.h"
#include "b2c_ptx.cuh"

namespace b2c {

namespace {

constexpr int kAStageBytes = 23552;  // 180 rows x 128 B rounded up to 1024

template <int BN>
struct Halo2Cfg {
    static constexpr int kBBytes = (BN / 2) * 128;     // this CTA's half of a weight tile
    static constexpr int kSA = BN == 256 ? 2 : 4;      // activation stages (power of two)
    static constexpr int kSB = BN == 64 ? 14 : (BN == 128 ? 7 : 4);   // weight half-tile slots
    static constexpr int kBRingBytes = kSB * kBBytes;
    static constexpr int kTmemCols = 2 * BN;
    static constexpr int kBarBytes = 640;
    static constexpr int kStatBytes = 2 * 4 * BN * 4 + 2 * BN * 4;   // running sums + fused-BN-backward scale|shift
    static constexpr int kStoreBytes = 8 * 2048;       // per-epilogue-warp store staging (store_chunk_staged)
    static constexpr int kSmemBytes = 1024 + kSA * kAStageBytes + kBRingBytes + kBarBytes + kStatBytes + kStoreBytes;
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
// arrive on the mbarrier at the same shared-memory offset in CTA `target` of the cluster (default .release.cta
// semantics: a cluster-scope release costs ~1 us per arrive and serialises the relay)
__device__ __forceinline__ void mbar_arrive_remote(uint64_t* bar, uint32_t target) {
    uint32_t raddr;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(raddr) : "r"(smem_u32(bar)), "r"(target));
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(raddr) : "memory");
}
// TMA loads of a CTA pair: the data lands in THIS CTA's shared memory, the transaction bytes are signalled on the
// barrier at the same offset in the LEADER CTA (rank 0), which the MMA issuer waits on — no relay through the peer.
__device__ __forceinline__ uint32_t leader_bar(uint64_t* bar) {
    uint32_t raddr;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(raddr) : "r"(smem_u32(bar)), "r"(0));
    return raddr;
}
__device__ __forceinline__ uint32_t own_cluster_addr(const void* p, uint32_t rank) {
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

// PRE: the epilogue operand (residual, or y of the fused BatchNorm backward) of the next tile is requested one tile
// ahead — its own instantiation, so launches without such an operand keep the plain epilogue.
template <int BN, bool kStats, int MODE, bool RES, bool PRE = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(320, 1)
conv_halo2_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                 const __grid_constant__ HaloParams p) {
    using Cfg = Halo2Cfg<BN>;
    constexpr int kSA = Cfg::kSA;
    constexpr int LOADS = MODE == 0 ? 1 : 3;   // activation boxes per 64-channel chunk
    constexpr int TPL = 9 / LOADS;             // taps served by one box
    constexpr uint32_t kABytes = MODE == 0 ? 180u * 128u : 160u * 128u;
    constexpr uint32_t kSB = Cfg::kSB;
    static_assert((kSA & (kSA - 1)) == 0, "the activation ring is a power of two");
    static_assert((3 * kSA + 3 * Cfg::kSB + 6) * 8 + 4 <= Cfg::kBarBytes, "barrier block");
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
    uint64_t* pfull_a = empty_a + kSA;
    uint64_t* full_b = pfull_a + kSA;
    uint64_t* empty_b = full_b + Cfg::kSB;
    uint64_t* pfull_b = empty_b + Cfg::kSB;
    uint64_t* tfull = pfull_b + Cfg::kSB;
    uint64_t* tempty = tfull + 2;
    uint64_t* ptempty = tempty + 2;
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(ptempty + 2);
    float* stat = reinterpret_cast<float*>(aux + Cfg::kBarBytes);
    float* bwd_sc = stat + 2 * 4 * BN;          // [scale | shift] of this CTA's channels (fused BN backward)
    uint8_t* store_stage = aux + Cfg::kBarBytes + Cfg::kStatBytes;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = cluster_rank();          // 0 = leader (issues the MMAs), 1 = peer
    const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&mapA);
        tma_prefetch_desc(&mapB);
    }
    if (warp == 1) {
        if (lane == 0) {
            for (int s = 0; s < kSA; ++s) {
                mbar_init(&full_a[s], 1);
                mbar_init(&empty_a[s], 1);
                mbar_init(&pfull_a[s], 1);
            }
            for (int s = 0; s < Cfg::kSB; ++s) {
                mbar_init(&full_b[s], 1);
                mbar_init(&empty_b[s], 1);
                mbar_init(&pfull_b[s], 1);
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
    pdl_wait();      // everything above (barriers, TMEM, descriptor prefetch) overlaps the previous kernel's tail

    // pair tiles: (m_pair, n_tile), n_tile fastest; this CTA's pixel tile is 2*m_pair + rank (one past the end for an
    // odd tile count: its boxes are out of range, TMA zero-fills, the epilogue stores nothing)
    const int total_tiles = ((p.num_m_tiles + 1) >> 1) * p.num_n_tiles;
    const int per_img = p.tiles_h * p.tiles_w;

    if (warp == 0) {
        // ------------------------------ TMA producer (whole warp loops, one elected lane issues) ----------
        uint32_t a_cnt = 0, b_cnt = 0;   // stages / weight tiles issued so far (uniform across the warp)
        bool first = true;
        for (int tile = pair; tile < total_tiles; tile += num_pairs) {
            const int m_pair = tile / p.num_n_tiles;
            const int n_tile = tile - m_pair * p.num_n_tiles;
            const int m_tile = 2 * m_pair + static_cast<int>(rank);
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
            const int brow0 = n_tile * BN + static_cast<int>(rank) * (BN / 2);
            for (int kc = 0; kc < p.k_chunks; ++kc) {
#pragma unroll
                for (int l = 0; l < LOADS; ++l) {
                    const uint32_t sa = a_cnt & (kSA - 1);
                    mbar_wait(&empty_a[sa], ((a_cnt / kSA) & 1u) ^ 1u, 21);
                    if (elect_one_sync()) {
                        // the leader's barrier collects both CTAs' bytes (a complete_tx that overtakes the expect_tx
                        // only drives the transaction count negative for a moment)
                        if (rank == 0) mbar_arrive_expect_tx(&full_a[sa], 2 * kABytes);
                        tma2_load_4d(own_cluster_addr(sA + sa * kAStageBytes, rank), &mapA, leader_bar(&full_a[sa]),
                                     kc * 64, w0 - 1, MODE == 0 ? h0 - 1 : l - 1, n0);
                        if (!RES || first) {
#pragma unroll
                            for (int j = 0; j < TPL; ++j) {
                                const int a = MODE == 0 ? j / 3 : l;
                                const int b = MODE == 0 ? j % 3 : j;
                                const int r = p.flip ? 2 - a : a;
                                const int s = p.flip ? 2 - b : b;
                                const uint32_t bi = RES ? static_cast<uint32_t>(l * TPL + j) : ((b_cnt + j) % kSB);
                                if (!RES) mbar_wait(&empty_b[bi], (((b_cnt + j) / kSB) & 1u) ^ 1u, 22);
                                if (rank == 0) mbar_arrive_expect_tx(&full_b[bi], 2 * Cfg::kBBytes);
                                tma2_load_2d(own_cluster_addr(sB + bi * Cfg::kBBytes, rank), &mapB,
                                             leader_bar(&full_b[bi]), kc * 64,
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
    } else if (warp == 1 && rank == 1) {
        // the peer's MMA warp has nothing to do: its TMA loads signal the leader's barriers directly
    } else if (warp == 1) {
        // ------------------------------ leader: MMA issuer for the pair (M = 256) ------------------------
        constexpr uint32_t idesc = make_idesc_bf16(256, BN, 0, 0);
        uint32_t a_cnt = 0, b_cnt = 0;
        int acc = 0;
        uint32_t acc_phase = 0;
        bool first = true;
        const uint32_t sB_u32 = smem_u32(sB);
        for (int tile = pair; tile < total_tiles; tile += num_pairs) {
            mbar_wait(&tempty[acc], acc_phase ^ 1u, 23);
            mbar_wait(&ptempty[acc], acc_phase ^ 1u, 29);
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
                                umma2_bf16(d, ad + 2u * k, bd + 2u * k, idesc, (kc | l | j | k) != 0 ? 1u : 0u);
                            }
                            if (!RES) umma2_commit_both(&empty_b[bi]);
                        }
                        umma2_commit_both(&empty_a[sa]);
                    }
                    __syncwarp();
                    ++a_cnt;
                    b_cnt += TPL;
                }
            }
            if (elect_one_sync()) umma2_commit_both(&tfull[acc]);
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
            const int ch0 = (pair % p.num_n_tiles) * BN;            // constant per pair (see the launcher)
            for (int i = et; i < BN; i += 256) {
                bwd_sc[i] = __ldg(p.bwd_scale + ch0 + i);
                bwd_sc[BN + i] = __ldg(p.bwd_shift + ch0 + i);
            }
            named_bar_sync(1, 256);
        }
        int acc = 0;
        uint32_t acc_phase = 0;
        // Row offsets of this lane's output rows in tile `tile` (rows past the end: ~0) and the tile's first column.
        auto tile_rows = [&](int tile, unsigned long long (&ro)[4], int& col0) {
            const int m_pair = tile / p.num_n_tiles;
            const int n_tile = tile - m_pair * p.num_n_tiles;
            const int m_tile = 2 * m_pair + static_cast<int>(rank);
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
            col0 = n_tile * BN;
            chunk_rowoffs(off, valid, lane, ro);
        };
        uint8_t* stg_w = store_stage + (warp - 2) * 2048;
        // 64-wide tiles: a warp reads ONE 32x32 chunk of the epilogue operand (the residual / add_src, or y of the fused
        // BatchNorm backward) per tile, and the tile's MMAs take ~0.8 us — less than a global load's latency, so a load
        // issued after the accumulator wait makes the epilogue, not the tensor pipe, set the pace (layer-1 dgrad: 102 -
        // 115 us against 68 us for the same GEMM as fprop). The operand of the NEXT tile is therefore requested one
        // tile ahead: its raw 16-byte pieces travel in registers across the accumulator wait and are parked in the
        // warp's shared-memory stage (the transpose the synchronous path does anyway) at the top of their own tile.
        const __nv_bfloat16* pre_src = nullptr;
        // (one operand at a time: with both a residual and a fused BatchNorm backward the synchronous loads stay)
        if constexpr (PRE) pre_src = bwd_fused ? (p.add_src != nullptr ? nullptr : p.bwd_y) : p.add_src;
        uint4 raw[4];
        unsigned long long rowoff[4];
        int col0 = 0;
        auto request = [&](const unsigned long long (&ro)[4], int col) {
#pragma unroll
            for (int it = 0; it < 4; ++it) {
                raw[it] = make_uint4(0u, 0u, 0u, 0u);
                if (ro[it] != ~0ull)
                    raw[it] = __ldg(reinterpret_cast<const uint4*>(pre_src + ro[it] + col + (lane & 3) * 8));
            }
        };
        if (pre_src != nullptr && pair < total_tiles) {
            tile_rows(pair, rowoff, col0);
            request(rowoff, col0 + half * 32);
        }
        for (int tile = pair; tile < total_tiles; tile += num_pairs) {
            if (pre_src != nullptr) {
                __syncwarp();                               // earlier reads of the stage are done
#pragma unroll
                for (int it = 0; it < 4; ++it) {
                    const int r = it * 8 + (lane >> 2);
                    *reinterpret_cast<uint4*>(stg_w + r * 64 + (((lane & 3) ^ ((r >> 1) & 3)) << 4)) = raw[it];
                }
                if (tile + num_pairs < total_tiles) {       // the next tile's operand: only `raw` outlives the wait
                    tile_rows(tile + num_pairs, rowoff, col0);
                    request(rowoff, col0 + half * 32);
                }
            }
            tile_rows(tile, rowoff, col0);
            mbar_wait(&tfull[acc], acc_phase, 26);
            tc_fence_after();
#pragma unroll 1
            for (int c = half; c < BN / 32; c += 2) {
                float v[32];
                tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                                  static_cast<uint32_t>(acc * BN + c * 32),
                              v);
                auto unpark = [&](uint4 (&u)[4]) {          // the operand parked at the top of this tile
                    __syncwarp();
                    const uint4* prow = reinterpret_cast<const uint4*>(stg_w + lane * 64);
                    const int sw = (lane >> 1) & 3;
#pragma unroll
                    for (int g4 = 0; g4 < 4; ++g4) u[g4] = prow[g4 ^ sw];
                };
                if (p.add_src != nullptr) {
                    uint4 au[4];
                    if (pre_src == p.add_src) {
                        unpark(au);
                    } else {
                        load_chunk_staged(stg_w, p.add_src, rowoff, col0 + c * 32, au, lane);
                    }
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
                        if (pre_src == p.bwd_y) {
                            unpark(yall);
                        } else {
                            load_chunk_staged(stg_w, p.bwd_y, rowoff, col0 + c * 32, yall, lane);
                        }
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

template <int BN, bool kStats, int MODE, bool RES, bool PRE = false>
cudaError_t launch_inst2(const CUtensorMap& a, const CUtensorMap& b, const HaloParams& p, int grid,
                        cudaStream_t stream) {
    using Cfg = Halo2Cfg<BN>;
    auto kern = conv_halo2_kernel<BN, kStats, MODE, RES, PRE>;
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
cudaError_t launch_one2(const CUtensorMap& a, const CUtensorMap& b, HaloParams p, int max_ctas, int* stat_rows,
                       cudaStream_t stream) {
    const int total = ((p.num_m_tiles + 1) / 2) * p.num_n_tiles;
    int pairs = total < max_ctas / 2 ? total : max_ctas / 2;
    pairs -= pairs % p.num_n_tiles;   // keeps n_tile constant per pair (per-CTA BatchNorm partial rows)
    if (pairs < p.num_n_tiles) pairs = p.num_n_tiles;
    if (stat_rows != nullptr) *stat_rows = 2 * (pairs / p.num_n_tiles);
    const int grid = 2 * pairs;
    p.b_resident = 0;
    if (p.mode == 0) {
        if constexpr (BN == 64) {
            if (p.k_chunks == 1 && p.num_n_tiles == 1) {
                p.b_resident = 1;
                // exactly one epilogue operand: request it a tile ahead (the 0.8 us of MMAs per 64-wide tile are
                // shorter than a global load's latency)
                if ((p.add_src != nullptr) != (p.bwd_y != nullptr))
                    return launch_inst2<64, kStats, 0, true, true>(a, b, p, grid, stream);
                return launch_inst2<64, kStats, 0, true>(a, b, p, grid, stream);
            }
        }
        return launch_inst2<BN, kStats, 0, false>(a, b, p, grid, stream);
    }
    return launch_inst2<BN, kStats, 1, false>(a, b, p, grid, stream);
}

}  // namespace

// b: 2-D map of the weight matrix with a box of bn/2 rows (each CTA of a pair loads half a tile)
cudaError_t launch_conv_halo2(int bn, const CUtensorMap& a, const CUtensorMap& b, const HaloParams& p, int max_ctas,
                             int* stat_rows, cudaStream_t stream) {
    const bool stats = p.part_sum != nullptr;
    switch (bn) {
        case 64:
            return stats ? launch_one2<64, true>(a, b, p, max_ctas, stat_rows, stream)
                         : launch_one2<64, false>(a, b, p, max_ctas, stat_rows, stream);
        case 128:
            return stats ? launch_one2<128, true>(a, b, p, max_ctas, stat_rows, stream)
                         : launch_one2<128, false>(a, b, p, max_ctas, stat_rows, stream);
        case 256:
            return stats ? launch_one2<256, true>(a, b, p, max_ctas, stat_rows, stream)
                         : launch_one2<256, false>(a, b, p, max_ctas, stat_rows, stream);
        default:
            return cudaErrorInvalidValue;
    }
}

}  // namespace b2c
