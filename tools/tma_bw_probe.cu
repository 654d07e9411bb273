// tma_bw_probe.cu — standalone measurement: sustained TMA global->shared fill rate per SM as a function of
// box shape and ring depth, all 148 SMs loading at once (one issuing thread per CTA, no consumer work).
// Answers: is a conv mainloop bound by L2->SM bandwidth, by per-load latency, or by the box shape?
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../pytorch-cifar_b200/csrc/b2c_ptx.cuh"

using namespace b2c;

#define CK(x)                                                                              \
    do {                                                                                   \
        cudaError_t e_ = (x);                                                              \
        if (e_ != cudaSuccess) {                                                           \
            printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); \
            exit(2);                                                                       \
        }                                                                                  \
    } while (0)

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn g_enc;

struct Args {
    int stages, box_bytes, stage_stride, iters, mode;  // mode 0: 4-D activation walk, 1: 2-D weight walk
    int bw, bh, bn;                                     // activation box pixels
    int W, H, N;
    int halo;                                           // start at (w0-1, h0-1)
    int b_rows;                                         // 2-D: rows per box; matrix rows total
    int b_total_rows;
};

__global__ void __launch_bounds__(32, 1)
tma_probe_kernel(const __grid_constant__ CUtensorMap map, Args a, long long* cycles) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + a.stages * a.stage_stride);
    if (threadIdx.x == 0) {
        for (int s = 0; s < a.stages; ++s) mbar_init(&bars[s], 1);
        fence_mbar_init();
    }
    __syncwarp();
    const long long t0 = clock64();
    const int tiles_w = a.W / a.bw, tiles_h = a.H / a.bh;
    const int per_img = tiles_w * tiles_h;
    for (int i = 0; i < a.iters; ++i) {
        const int st = i % a.stages;
        // slot reuse: the previous load into this slot was completion number (i / stages - 1) of its barrier
        if (i >= a.stages) mbar_wait(&bars[st], static_cast<uint32_t>((i / a.stages - 1) & 1), 200);
        const int t = blockIdx.x + i * gridDim.x;
        const int n = (t / per_img) % a.N;
        const int rem = t % per_img;
        const int h0 = (rem / tiles_w) * a.bh;
        const int w0 = (rem % tiles_w) * a.bw;
        if (elect_one_sync()) {
            mbar_arrive_expect_tx(&bars[st], a.box_bytes);
            if (a.mode == 0) {
                tma_load_4d(smem + st * a.stage_stride, &map, &bars[st], 0, w0 - a.halo, h0 - a.halo, n);
            } else {
                tma_load_2d(smem + st * a.stage_stride, &map, &bars[st], 0, (t * a.b_rows) % a.b_total_rows);
            }
        }
        __syncwarp();
    }
    for (int j = (a.iters > a.stages ? a.iters - a.stages : 0); j < a.iters; ++j)
        mbar_wait(&bars[j % a.stages], static_cast<uint32_t>((j / a.stages) & 1), 201);
    if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
}

__global__ void __launch_bounds__(32, 1)
tma_tight_kernel(const __grid_constant__ CUtensorMap map, int log2_stages, int box_bytes, int iters, int rows,
                 int total_rows, int prefetch, int loads_per_bar, long long* cycles) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int stages = 1 << log2_stages;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + stages * box_bytes * loads_per_bar);
    if (threadIdx.x == 0) {
        for (int s = 0; s < stages; ++s) mbar_init(&bars[s], 1);
        fence_mbar_init();
        if (prefetch) tma_prefetch_desc(&map);
    }
    __syncwarp();
    const long long t0 = clock64();
    int row = (blockIdx.x * rows) % total_rows;
    for (int i = 0; i < iters; ++i) {
        const int st = i & (stages - 1);
        if (i >= stages) mbar_wait(&bars[st], static_cast<uint32_t>(((i >> log2_stages) - 1) & 1), 200);
        if (elect_one_sync()) {
            mbar_arrive_expect_tx(&bars[st], box_bytes * loads_per_bar);
            int r2 = row;
            for (int l = 0; l < loads_per_bar; ++l) {
                tma_load_2d(smem + (st * loads_per_bar + l) * box_bytes, &map, &bars[st], 0, r2);
                r2 += rows;
                if (r2 >= total_rows) r2 -= total_rows;
            }
        }
        __syncwarp();
        row += rows * loads_per_bar;
        while (row >= total_rows) row -= total_rows;
    }
    for (int j = (iters > stages ? iters - stages : 0); j < iters; ++j)
        mbar_wait(&bars[j & (stages - 1)], static_cast<uint32_t>((j >> log2_stages) & 1), 201);
    if (threadIdx.x == 0) cycles[blockIdx.x] = clock64() - t0;
}

int main() {
    CK(cudaSetDevice(0));
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q));
    g_enc = (EncodeTiledFn)p;
    const int N = 1024, H = 32, W = 32, C = 64;
    __nv_bfloat16* act;
    CK(cudaMalloc(&act, (size_t)N * H * W * C * 2));
    CK(cudaMemset(act, 0, (size_t)N * H * W * C * 2));
    const int WR = 9 * 512;  // weight matrix rows
    __nv_bfloat16* wt;
    CK(cudaMalloc(&wt, (size_t)WR * 64 * 2));
    CK(cudaMemset(wt, 0, (size_t)WR * 64 * 2));
    long long* dcy;
    CK(cudaMalloc(&dcy, 148 * 8));
    CK(cudaFuncSetAttribute(tma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    int clk_khz = 0;
    CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0));

    struct Shape { const char* name; int bw, bh, bn, halo; };
    Shape shapes[] = {{"act 32x4 (16 KB, per-tap tile)", 32, 4, 1, 0}, {"act halo 10x18 (22.5 KB)", 8, 16, 1, 1},
                      {"act 8x16 no halo (16 KB)", 8, 16, 1, 0}, {"act 32x8 (32 KB)", 32, 8, 1, 0}};
    for (auto& s : shapes) {
        for (int stages : {2, 4, 8}) {
            const int bxw = s.bw + 2 * s.halo, bxh = s.bh + 2 * s.halo;
            const int bytes = bxw * bxh * s.bn * 128;
            const int stride = (bytes + 1023) / 1024 * 1024;
            if (stages * stride > 200 * 1024) continue;
            CUtensorMap m;
            cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
            cuuint64_t str[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
            cuuint32_t box[4] = {64, (cuuint32_t)bxw, (cuuint32_t)bxh, (cuuint32_t)s.bn};
            cuuint32_t es[4] = {1, 1, 1, 1};
            if (g_enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, act, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) { printf("encode failed\n"); continue; }
            Args a{stages, bytes, stride, 200, 0, s.bw, s.bh, s.bn, W, H, N, s.halo, 0, 0};
            tma_probe_kernel<<<148, 32, stages * stride + 1024 + 256>>>(m, a, dcy);
            CK(cudaDeviceSynchronize());
            tma_probe_kernel<<<148, 32, stages * stride + 1024 + 256>>>(m, a, dcy);
            CK(cudaDeviceSynchronize());
            std::vector<long long> cy(148);
            CK(cudaMemcpy(cy.data(), dcy, 148 * 8, cudaMemcpyDeviceToHost));
            double avg = 0;
            for (auto c : cy) avg += c;
            avg /= 148;
            printf("%-34s stages=%d : %.0f cyc/load, %.1f B/clk/SM, implied latency %.0f cyc\n", s.name, stages,
                   avg / a.iters, (double)bytes * a.iters / avg, avg / a.iters * stages);
        }
    }
    for (int rows : {64, 128, 256}) {
        for (int stages : {2, 4, 8}) {
            const int bytes = rows * 128;
            if (stages * bytes > 200 * 1024) continue;
            CUtensorMap m;
            cuuint64_t dims[2] = {64, (cuuint64_t)WR};
            cuuint64_t str[1] = {128};
            cuuint32_t box[2] = {64, (cuuint32_t)rows};
            cuuint32_t es[2] = {1, 1};
            if (g_enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, wt, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                      CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) { printf("encode failed\n"); continue; }
            Args a{stages, bytes, bytes, 400, 1, 0, 0, 0, 0, 0, 0, 0, rows, WR};
            tma_probe_kernel<<<148, 32, stages * bytes + 1024 + 256>>>(m, a, dcy);
            CK(cudaDeviceSynchronize());
            tma_probe_kernel<<<148, 32, stages * bytes + 1024 + 256>>>(m, a, dcy);
            CK(cudaDeviceSynchronize());
            std::vector<long long> cy(148);
            CK(cudaMemcpy(cy.data(), dcy, 148 * 8, cudaMemcpyDeviceToHost));
            double avg = 0;
            for (auto c : cy) avg += c;
            avg /= 148;
            printf("weights 2-D box %3d rows (%2d KB, L2-resident) stages=%d : %.0f cyc/load, %.1f B/clk/SM, implied latency %.0f cyc\n",
                   rows, bytes / 1024, stages, avg / a.iters, (double)bytes * a.iters / avg, avg / a.iters * stages);
        }
    }
    CK(cudaFuncSetAttribute(tma_tight_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    for (int rows : {64, 128, 256}) {
        for (int lpb : {1, 2}) {
            for (int pf : {0, 1}) {
                for (int ls : {2, 3}) {
                    const int bytes = rows * 128, stages = 1 << ls;
                    if (stages * bytes * lpb > 200 * 1024) continue;
                    CUtensorMap m;
                    cuuint64_t dims[2] = {64, (cuuint64_t)WR};
                    cuuint64_t str[1] = {128};
                    cuuint32_t box[2] = {64, (cuuint32_t)rows};
                    cuuint32_t es[2] = {1, 1};
                    g_enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, wt, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
                    const int iters = 400;
                    for (int rep = 0; rep < 2; ++rep) {
                        tma_tight_kernel<<<148, 32, stages * bytes * lpb + 1024 + 256>>>(m, ls, bytes, iters, rows, WR, pf, lpb, dcy);
                        CK(cudaDeviceSynchronize());
                    }
                    std::vector<long long> cy(148);
                    CK(cudaMemcpy(cy.data(), dcy, 148 * 8, cudaMemcpyDeviceToHost));
                    double avg = 0;
                    for (auto c : cy) avg += c;
                    avg /= 148;
                    printf("tight 2-D %3d rows x%d loads/barrier prefetch=%d stages=%d : %.0f cyc/iter, %.1f B/clk/SM\n", rows, lpb, pf,
                           stages, avg / iters, (double)bytes * lpb * iters / avg);
                }
            }
        }
    }
    // single CTA: is the limit per SM or chip-wide?
    {
        CUtensorMap m;
        cuuint64_t dims[2] = {64, (cuuint64_t)WR};
        cuuint64_t str[1] = {128};
        cuuint32_t box[2] = {64, 128};
        cuuint32_t es[2] = {1, 1};
        g_enc(&m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, wt, dims, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        for (int grid : {1, 16, 74, 148}) {
            for (int rep = 0; rep < 2; ++rep) {
                tma_tight_kernel<<<grid, 32, 8 * 16384 + 1024 + 256>>>(m, 3, 16384, 400, 128, WR, 1, 1, dcy);
                CK(cudaDeviceSynchronize());
            }
            std::vector<long long> cy(grid);
            CK(cudaMemcpy(cy.data(), dcy, grid * 8, cudaMemcpyDeviceToHost));
            double avg = 0;
            for (auto c : cy) avg += c;
            avg /= grid;
            printf("tight 2-D 128 rows grid=%3d : %.0f cyc/iter, %.1f B/clk/SM\n", grid, avg / 400, 16384.0 * 400 / avg);
        }
    }
    printf("clock %d kHz\ntma probe done\n", clk_khz);
    return 0;
}
