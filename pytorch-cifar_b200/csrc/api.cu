// api.cu — the extern "C" boundary of libb2c.so (declared in include/b2c.h).
// Host-side only: validates arguments, derives the GEMM view of each convolution, encodes the TMA
// tensor maps and launches the kernels on the caller's stream. No allocation, no synchronisation.
#include "../../include/b2c.h"

#include <cuda.h>
#include <cuda_runtime.h>

#include <mutex>

#include "b2c_elementwise.h"
#include "b2c_kernels.h"

namespace {

using b2c::TapGemmParams;
using b2c::WgradParams;

inline int cuda_rc(cudaError_t e) { return e == cudaSuccess ? 0 : -static_cast<int>(e); }
inline cudaStream_t as_stream(void* s) { return static_cast<cudaStream_t>(s); }

// cuTensorMapEncodeTiled is a driver entry point; resolve it through the runtime so that the
// library links (and loads on a machine without a GPU driver) with cudart alone.
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess) {
            fn = reinterpret_cast<EncodeTiledFn>(p);
        }
    });
    return fn;
}

// NHWC bf16 tensor [n, h, w, c] -> 4-D map, box {64 ch, bw, bh, bn} pixels, traversal stride `es` on w/h.
// ld: channel pitch of the tensor in memory (0 = c): a channel slice of a wider tensor is the same map with the
// slice's base pointer and the wide tensor's pitch.
int encode_nhwc(CUtensorMap* m, const void* base, int n, int h, int w, int c, int bw, int bh, int bn, int es,
                int ld = 0) {
    EncodeTiledFn enc = get_encode();
    if (enc == nullptr) return B2C_ERR_DRIVER;
    if (ld == 0) ld = c;
    cuuint64_t dims[4] = {static_cast<cuuint64_t>(c), static_cast<cuuint64_t>(w), static_cast<cuuint64_t>(h),
                          static_cast<cuuint64_t>(n)};
    cuuint64_t strides[3] = {static_cast<cuuint64_t>(ld) * 2, static_cast<cuuint64_t>(w) * ld * 2,
                             static_cast<cuuint64_t>(h) * w * ld * 2};
    cuuint32_t box[4] = {64u, static_cast<cuuint32_t>(bw * es), static_cast<cuuint32_t>(bh * es),
                         static_cast<cuuint32_t>(bn)};
    cuuint32_t estr[4] = {1u, static_cast<cuuint32_t>(es), static_cast<cuuint32_t>(es), 1u};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : B2C_ERR_DRIVER;
}

// Stride-2 view of an NHWC bf16 tensor for TMA: the W axis is split into (w/2, parity) so that a box takes ONE
// parity of bw consecutive w/2 positions — plain tiled addressing. A traversal stride on W instead makes the
// TMA unit fetch the whole 2*bw bounding row from L2 and drop every other pixel (ncu: 604 MB of L2->SM reads
// for 302 MB of operand on the 64->128 stride-2 layer). The H axis keeps the traversal stride (rows are
// separate lines anyway). Dims {c, parity, w/2, h, n}; coordinates of pixel column w: (w & 1, w >> 1).
int encode_nhwc_s2(CUtensorMap* m, const void* base, int n, int h, int w, int c, int bw, int bh, int bn, int ld = 0) {
    EncodeTiledFn enc = get_encode();
    if (enc == nullptr) return B2C_ERR_DRIVER;
    if (ld == 0) ld = c;
    cuuint64_t dims[5] = {static_cast<cuuint64_t>(c), 2u, static_cast<cuuint64_t>(w / 2), static_cast<cuuint64_t>(h),
                          static_cast<cuuint64_t>(n)};
    cuuint64_t strides[4] = {static_cast<cuuint64_t>(ld) * 2, static_cast<cuuint64_t>(ld) * 4,
                             static_cast<cuuint64_t>(w) * ld * 2, static_cast<cuuint64_t>(h) * w * ld * 2};
    cuuint32_t box[5] = {64u, 1u, static_cast<cuuint32_t>(bw), static_cast<cuuint32_t>(bh * 2),
                         static_cast<cuuint32_t>(bn)};
    cuuint32_t estr[5] = {1u, 1u, 1u, 2u, 1u};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : B2C_ERR_DRIVER;
}

void single_phase(b2c::TapGemmParams& p) {
    p.num_phases = 1;
    p.phase_begin[0] = 0;
    p.phase_begin[1] = static_cast<uint8_t>(p.num_taps);
    p.phase_py[0] = p.phase_px[0] = 0;
}

// Row-major bf16 matrix [rows, cols] -> 2-D map, box {64 cols, brows}.
int encode_2d(CUtensorMap* m, const void* base, long long rows, int cols, int brows) {
    EncodeTiledFn enc = get_encode();
    if (enc == nullptr) return B2C_ERR_DRIVER;
    cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
    cuuint64_t strides[1] = {static_cast<cuuint64_t>(cols) * 2};
    cuuint32_t box[2] = {64u, static_cast<cuuint32_t>(brows)};
    cuuint32_t estr[2] = {1u, 1u};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : B2C_ERR_DRIVER;
}

bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// pixel grid (gH x gW per image) must tile into boxes of `px` pixels: full rows, whole images.
bool grid_tiles(int gH, int gW, int px) {
    if (gW < 1 || gH < 1 || gW > px || (px % gW) != 0) return false;
    const int img = gH * gW;
    return (img % px) == 0 || (px % img) == 0;
}

// Operands of the fused BatchNorm backward of b2c_conv_dgrad_bn (null pointer = plain dgrad).
struct BnFuse {
    const void* y;         // the BatchNorm's input (the conv output of the forward pass), dx's shape
    const float* scale;    // forward scale / shift of that BatchNorm (relu mask = fma(y, scale, shift) > 0)
    const float* shift;
    float* partial;        // [rows][2][cin] fp32 partial sums for b2c_bn_bwd_finalize
    int* stat_rows;
};

// 3x3 stride-1 convs on square 32 / 16 / 8 pixel images go to the halo-reuse kernel (conv_halo.cu).
bool use_halo(const B2cConvDesc* d) {
    if (d->flags & 1) return false;
    if (d->ksize != 3 || d->stride != 1 || d->h != d->w) return false;
    return d->w == 32 || d->w == 16 || d->w == 8;
}

int launch_halo(const B2cConvDesc* d, const void* act, int act_c, int act_ld, const void* w, int gemm_n, int out_ld,
                int bn, bool flip, void* out, const void* add_src, float* part_sum, float* part_sq, int* stat_rows,
                cudaStream_t st, const BnFuse* bn_fuse = nullptr) {
    b2c::HaloParams p = {};
    p.N = d->n;
    p.H = d->h;
    p.W = d->w;
    p.mode = d->w == 8 ? 1 : 0;
    p.tiles_h = d->h / 16;
    p.tiles_w = d->w / 8;
    p.num_m_tiles = p.mode == 0 ? d->n * p.tiles_h * p.tiles_w : (d->n + 1) / 2;
    p.num_n_tiles = gemm_n / bn;
    p.k_chunks = act_c / 64;
    p.flip = flip ? 1 : 0;
    p.b_rows_per_tap = gemm_n;
    p.out = static_cast<__nv_bfloat16*>(out);
    p.add_src = static_cast<const __nv_bfloat16*>(add_src);
    p.out_ld = out_ld;
    p.part_sum = part_sum;
    p.part_sq = part_sq;
    p.cout = gemm_n;
    p.stat_ld = gemm_n;
    if (bn_fuse != nullptr) {
        p.bwd_y = static_cast<const __nv_bfloat16*>(bn_fuse->y);
        p.bwd_scale = bn_fuse->scale;
        p.bwd_shift = bn_fuse->shift;
        p.part_sum = bn_fuse->partial;
        p.part_sq = bn_fuse->partial + gemm_n;
        p.stat_ld = 2 * gemm_n;
        stat_rows = bn_fuse->stat_rows;
    }
    const bool acc = (d->flags & 512) != 0;       // statistics accumulated into one zero-filled row (fp32 atomics)
    if (acc) p.stat_ld = 0;
    p.debug = d->flags & 0x3e;
    CUtensorMap ma, mb;
    int rc = p.mode == 0 ? encode_nhwc(&ma, act, d->n, d->h, d->w, act_c, 10, 18, 1, 1, act_ld)
                         : encode_nhwc(&ma, act, d->n, d->h, d->w, act_c, 10, 8, 2, 1, act_ld);
    if (rc) return rc;
    const bool pair = (d->flags & 128) != 0;      // CTA-pair kernel (conv_halo2.cu): each CTA loads half a weight tile
    rc = encode_2d(&mb, w, 9LL * gemm_n, act_c, pair ? bn / 2 : bn);
    if (rc) return rc;
    const int max_ctas = d->max_ctas > 0 ? d->max_ctas : 148;
    rc = pair ? cuda_rc(b2c::launch_conv_halo2(bn, ma, mb, p, max_ctas, stat_rows, st))
              : cuda_rc(b2c::launch_conv_halo(bn, ma, mb, p, max_ctas, stat_rows, st));
    if (acc && stat_rows != nullptr) *stat_rows = 1;
    return rc;
}

// GEMM N tile: the widest tile that still gives every SM work. tcgen05.mma reads both operands from shared
// memory, so N = 256 halves the operand bytes per FLOP of N = 128 (measured +20-25 % on the 512-channel layers).
// `cap` is the widest tile the kernel family takes for this layer (the halo-reuse kernels keep N <= 128 up to 256
// channels). When even N = 128 leaves most SMs without a tile — the 4x4 / 8x8 layers at 128 images per GPU, the
// reference's own per-GPU batch at 8 GPUs (main_dist.py:111): 16 pixel tiles x 4 = 64 tiles on 148 SMs — N = 64
// doubles the tile count; a CTA pair still runs N = 64 MMAs at 74 % of the tensor peak (profiles/r1g_umma2_probe.log),
// which beats idle SMs.
int pick_bn(int requested, int n_dim, int cap, long long m_tiles) {
    if (requested != 0) return requested;
    if (cap >= 256 && n_dim % 256 == 0 && m_tiles * (n_dim / 256) >= 120) return 256;
    if (n_dim % 128 == 0 && (m_tiles * (n_dim / 128) >= 100 || m_tiles * (n_dim / 64) <= 16)) return 128;
    return 64;
}

int check_conv(const B2cConvDesc* d) {
    if (d == nullptr) return B2C_ERR_ARG;
    if (d->n < 1 || d->h < 1 || d->w < 1) return B2C_ERR_ARG;
    if (d->cin < 64 || d->cout < 64 || (d->cin % 64) != 0 || (d->cout % 64) != 0) return B2C_ERR_UNSUPPORTED;
    if (d->ksize != 1 && d->ksize != 3) return B2C_ERR_UNSUPPORTED;
    if (d->stride != 1 && d->stride != 2) return B2C_ERR_UNSUPPORTED;
    if (d->stride == 2 && ((d->h & 1) || (d->w & 1))) return B2C_ERR_UNSUPPORTED;
    if (d->block_n != 0 && d->block_n != 64 && d->block_n != 128 && d->block_n != 256) return B2C_ERR_ARG;
    if (d->x_ld != 0 && (d->x_ld < d->cin || (d->x_ld & 7))) return B2C_ERR_ARG;
    if (d->y_ld != 0 && (d->y_ld < d->cout || (d->y_ld & 7))) return B2C_ERR_ARG;
    return 0;
}

// channel pitch of the input-side tensor (x, dx) and of the output-side tensor (y, dy, add_src of the fprop)
inline int x_pitch(const B2cConvDesc* d) { return d->x_ld != 0 ? d->x_ld : d->cin; }
inline int y_pitch(const B2cConvDesc* d) { return d->y_ld != 0 ? d->y_ld : d->cout; }

}  // namespace

extern "C" {

int b2c_version(void) { return 100; }

const char* b2c_error_string(int code) {
    if (code == 0) return "success";
    if (code == B2C_ERR_UNSUPPORTED) return "b2c: shape or option not supported by the sm_100a kernels";
    if (code == B2C_ERR_DRIVER) return "b2c: cuTensorMapEncodeTiled unavailable or failed (no CUDA driver?)";
    if (code == B2C_ERR_ARG) return "b2c: invalid argument (null, misaligned or inconsistent)";
    if (code == B2C_ERR_TOO_LARGE)
        return "b2c: size exceeds the fixed in-library scratch of this no-allocation entry point "
               "(channels <= B2C_NOALLOC_MAX_CHANNELS = 512); use the two-call form with a caller buffer";
    if (code < 0) return cudaGetErrorString(static_cast<cudaError_t>(-code));
    return "b2c: unknown code";
}

int b2c_conv_fprop(const B2cConvDesc* d, const void* x, const void* w_fprop, void* y, const void* add_src,
                   float* part_sum, float* part_sq, int* stat_rows, void* stream) {
    int rc = check_conv(d);
    if (rc) return rc;
    if (!x || !w_fprop || !y || !aligned16(x) || !aligned16(w_fprop) || !aligned16(y) || !aligned16(add_src))
        return B2C_ERR_ARG;
    if ((part_sum == nullptr) != (part_sq == nullptr)) return B2C_ERR_ARG;
    const int ho = d->h / d->stride, wo = d->w / d->stride;
    if (!grid_tiles(ho, wo, 128)) return B2C_ERR_UNSUPPORTED;
    const long long m_tiles = (static_cast<long long>(d->n) * ho * wo + 127) / 128;
    const bool halo = use_halo(d);
    const int bn = pick_bn(d->block_n, d->cout, halo && d->cout <= 256 ? 128 : 256, m_tiles);
    if (d->cout % bn) return B2C_ERR_UNSUPPORTED;
    const int taps = d->ksize * d->ksize;
    if (halo)
        return launch_halo(d, x, d->cin, x_pitch(d), w_fprop, d->cout, y_pitch(d), bn, false, y, add_src, part_sum, part_sq, stat_rows,
                           as_stream(stream));

    TapGemmParams p = {};
    p.m_total = d->n * ho * wo;
    p.num_m_tiles = (p.m_total + 127) / 128;
    p.num_n_tiles = d->cout / bn;
    p.k_chunks = d->cin / 64;
    p.num_taps = taps;
    p.gH = ho;
    p.gW = wo;
    p.in_stride = d->stride;
    for (int t = 0; t < taps; ++t) {
        const int r = t / d->ksize, s = t % d->ksize;
        p.tap_dh[t] = static_cast<int8_t>(r - d->ksize / 2);
        p.tap_dw[t] = static_cast<int8_t>(s - d->ksize / 2);
        p.tap_a[t] = 0;
        p.tap_brow[t] = t * d->cout;
    }
    p.out = static_cast<__nv_bfloat16*>(y);
    p.add_src = static_cast<const __nv_bfloat16*>(add_src);
    p.out_H = ho;
    p.out_W = wo;
    p.out_ld = y_pitch(d);
    p.sy = p.sx = 1;
    single_phase(p);
    p.part_sum = part_sum;
    p.part_sq = part_sq;
    p.cout = d->cout;
    const bool acc = (d->flags & 512) != 0;       // statistics accumulated into one zero-filled row (fp32 atomics)
    p.stat_ld = acc ? 0 : d->cout;

    const int bh = (128 / wo) < ho ? (128 / wo) : ho;
    const int bnb = 128 / (wo * bh);
    CUtensorMap ma, mb;
    p.a_split_w = d->stride == 2 && !(d->flags & 4);      // flags & 4: traversal-stride map (A/B measurements)
    rc = p.a_split_w ? encode_nhwc_s2(&ma, x, d->n, d->h, d->w, d->cin, wo, bh, bnb, x_pitch(d))
                     : encode_nhwc(&ma, x, d->n, d->h, d->w, d->cin, wo, bh, bnb, d->stride, x_pitch(d));
    if (rc) return rc;
    const bool pair = (d->flags & 128) != 0;      // CTA-pair kernel (conv_gemm2.cu): each CTA loads half a weight tile
    rc = encode_2d(&mb, w_fprop, static_cast<long long>(taps) * d->cout, d->cin, pair ? bn / 2 : bn);
    if (rc) return rc;
    const int max_ctas = d->max_ctas > 0 ? d->max_ctas : 148;
    rc = pair ? cuda_rc(b2c::launch_tap_gemm2(bn, ma, ma, mb, p, max_ctas, stat_rows, as_stream(stream)))
              : cuda_rc(b2c::launch_tap_gemm(bn, ma, ma, mb, p, max_ctas, stat_rows, as_stream(stream)));
    if (acc && stat_rows != nullptr) *stat_rows = 1;
    return rc;
}

static int conv1x1_cat_impl(const B2cConvDesc* d, const void* const* xs, int n_src, const void* w_fprop, void* y,
                            float* part_sum, float* part_sq, int* stat_rows, void* stream, const int* cins = nullptr) {
    int rc = check_conv(d);
    if (rc) return rc;
    if (d->ksize != 1 || d->stride != 1) return B2C_ERR_UNSUPPORTED;
    if (n_src < 2 || n_src > 4) return B2C_ERR_UNSUPPORTED;
    if (!xs || !w_fprop || !y || !aligned16(w_fprop) || !aligned16(y)) return B2C_ERR_ARG;
    for (int t = 0; t < n_src; ++t)
        if (!xs[t] || !aligned16(xs[t])) return B2C_ERR_ARG;
    if ((part_sum == nullptr) != (part_sq == nullptr)) return B2C_ERR_ARG;
    if (!grid_tiles(d->h, d->w, 128)) return B2C_ERR_UNSUPPORTED;
    if (cins != nullptr)       // sources of different widths: each at most d->cin (the pitch of a weight slab)
        for (int t = 0; t < n_src; ++t)
            if (cins[t] < 64 || (cins[t] % 64) || cins[t] > d->cin || cins[t] / 64 > 255) return B2C_ERR_UNSUPPORTED;
    const long long m_tiles = (static_cast<long long>(d->n) * d->h * d->w + 127) / 128;
    const int bn = pick_bn(d->block_n, d->cout, 256, m_tiles);
    if (d->cout % bn) return B2C_ERR_UNSUPPORTED;
    // the sources are the "taps" of one GEMM: tap t reads activation map t and weight slab t ([t][cout][cin]),
    // all at pixel offset (0, 0) — the K axis of the concatenated input, split across the tensors
    TapGemmParams p = {};
    p.m_total = d->n * d->h * d->w;
    p.num_m_tiles = (p.m_total + 127) / 128;
    p.num_n_tiles = d->cout / bn;
    p.k_chunks = d->cin / 64;
    p.num_taps = n_src;
    p.gH = d->h;
    p.gW = d->w;
    p.in_stride = 1;
    for (int t = 0; t < n_src; ++t) {
        p.tap_dh[t] = 0;
        p.tap_dw[t] = 0;
        p.tap_a[t] = static_cast<uint8_t>(t);
        p.tap_brow[t] = t * d->cout;
        p.tap_kc[t] = cins != nullptr ? static_cast<uint8_t>(cins[t] / 64) : 0;
    }
    p.out = static_cast<__nv_bfloat16*>(y);
    p.out_H = d->h;
    p.out_W = d->w;
    p.out_ld = d->cout;
    p.sy = p.sx = 1;
    single_phase(p);
    p.part_sum = part_sum;
    p.part_sq = part_sq;
    p.cout = d->cout;
    const bool acc = (d->flags & 512) != 0;
    p.stat_ld = acc ? 0 : d->cout;
    const int bh = (128 / d->w) < d->h ? (128 / d->w) : d->h;
    const int bnb = 128 / (d->w * bh);
    CUtensorMap ma[4], mb;
    for (int t = 0; t < 4; ++t) {
        const int ts = t < n_src ? t : n_src - 1;
        rc = encode_nhwc(&ma[t], xs[ts], d->n, d->h, d->w, cins != nullptr ? cins[ts] : d->cin, d->w, bh, bnb, 1);
        if (rc) return rc;
    }
    rc = encode_2d(&mb, w_fprop, static_cast<long long>(n_src) * d->cout, d->cin, bn);
    if (rc) return rc;
    const int max_ctas = d->max_ctas > 0 ? d->max_ctas : 148;
    rc = cuda_rc(b2c::launch_tap_gemm4(bn, ma[0], ma[1], ma[2], ma[3], mb, p, max_ctas, stat_rows, as_stream(stream)));
    if (acc && stat_rows != nullptr) *stat_rows = 1;
    return rc;
}

int b2c_conv1x1_cat_fprop(const B2cConvDesc* d, const void* x0, const void* x1, const void* w_fprop2, void* y,
                          float* part_sum, float* part_sq, int* stat_rows, void* stream) {
    const void* xs[2] = {x0, x1};
    return conv1x1_cat_impl(d, xs, 2, w_fprop2, y, part_sum, part_sq, stat_rows, stream);
}

int b2c_conv1x1_catn_fprop(const B2cConvDesc* d, const void* const* xs, int n_src, const void* w_fprop_n, void* y,
                           float* part_sum, float* part_sq, int* stat_rows, void* stream) {
    return conv1x1_cat_impl(d, xs, n_src, w_fprop_n, y, part_sum, part_sq, stat_rows, stream);
}

int b2c_conv1x1_catr_fprop(const B2cConvDesc* d, const void* const* xs, const int* cins, int n_src,
                           const void* w_fprop_n, void* y, float* part_sum, float* part_sq, int* stat_rows,
                           void* stream) {
    if (!cins) return B2C_ERR_ARG;
    return conv1x1_cat_impl(d, xs, n_src, w_fprop_n, y, part_sum, part_sq, stat_rows, stream, cins);
}

static int conv_dgrad_impl(const B2cConvDesc* d, const void* dy, const void* dy_shortcut, const void* w_dgrad,
                           void* dx, const void* add_src, void* stream, const BnFuse* bn_fuse = nullptr) {
    int rc = check_conv(d);
    if (rc) return rc;
    if (!dy || !w_dgrad || !dx || !aligned16(dy) || !aligned16(w_dgrad) || !aligned16(dx) || !aligned16(add_src))
        return B2C_ERR_ARG;
    const int ho = d->h / d->stride, wo = d->w / d->stride;
    // GEMM: rows = dx pixels (per output phase for stride 2), N = cin, K = cout
    const long long m_tiles = (static_cast<long long>(d->n) * (d->stride == 1 ? d->h * d->w : ho * wo) + 127) / 128;
    const bool halo = use_halo(d);
    const int bn = pick_bn(d->block_n, d->cin, halo && d->cin <= 256 ? 128 : 256, m_tiles);
    if (d->cin % bn) return B2C_ERR_UNSUPPORTED;
    const int taps = d->ksize * d->ksize;
    const int max_ctas = d->max_ctas > 0 ? d->max_ctas : 148;
    cudaStream_t st = as_stream(stream);
    if (halo)
        return launch_halo(d, dy, d->cout, y_pitch(d), w_dgrad, d->cin, x_pitch(d), bn, true, dx, add_src, nullptr, nullptr, nullptr, st,
                           bn_fuse);

    TapGemmParams p = {};
    p.num_n_tiles = d->cin / bn;
    p.k_chunks = d->cout / 64;
    p.in_stride = 1;
    p.out = static_cast<__nv_bfloat16*>(dx);
    p.add_src = static_cast<const __nv_bfloat16*>(add_src);
    p.out_H = d->h;
    p.out_W = d->w;
    p.out_ld = x_pitch(d);
    p.part_sum = nullptr;
    p.part_sq = nullptr;
    p.cout = d->cin;
    p.stat_ld = d->cin;
    int* stat_rows = nullptr;
    if (bn_fuse != nullptr) {
        p.bwd_y = static_cast<const __nv_bfloat16*>(bn_fuse->y);
        p.bwd_scale = bn_fuse->scale;
        p.bwd_shift = bn_fuse->shift;
        p.part_sum = bn_fuse->partial;
        p.part_sq = bn_fuse->partial + d->cin;
        p.stat_ld = (d->flags & 512) ? 0 : 2 * d->cin;      // flags bit 9: one accumulated row
        stat_rows = bn_fuse->stat_rows;
    }

    CUtensorMap ma, mb;
    const bool pair = (d->flags & 128) != 0;   // conv_gemm2.cu (CTA pair), all phase / shortcut forms included
    rc = encode_2d(&mb, w_dgrad, static_cast<long long>(taps + (dy_shortcut != nullptr ? 1 : 0)) * d->cin, d->cout,
                   pair ? bn / 2 : bn);
    if (rc) return rc;

    if (d->stride == 1) {
        if (!grid_tiles(d->h, d->w, 128)) return B2C_ERR_UNSUPPORTED;
        const int bh = (128 / d->w) < d->h ? (128 / d->w) : d->h;
        const int bnb = 128 / (d->w * bh);
        rc = encode_nhwc(&ma, dy, d->n, ho, wo, d->cout, wo, bh, bnb, 1, y_pitch(d));
        if (rc) return rc;
        p.m_total = d->n * d->h * d->w;
        p.num_m_tiles = (p.m_total + 127) / 128;
        p.gH = d->h;
        p.gW = d->w;
        p.num_taps = taps;
        for (int t = 0; t < taps; ++t) {
            const int r = t / d->ksize, s = t % d->ksize;
            p.tap_dh[t] = static_cast<int8_t>(d->ksize / 2 - r);
            p.tap_dw[t] = static_cast<int8_t>(d->ksize / 2 - s);
            p.tap_a[t] = 0;
            p.tap_brow[t] = t * d->cin;
        }
        p.sy = p.sx = 1;
        single_phase(p);
        rc = pair ? cuda_rc(b2c::launch_tap_gemm2(bn, ma, ma, mb, p, max_ctas, stat_rows, st))
                  : cuda_rc(b2c::launch_tap_gemm(bn, ma, ma, mb, p, max_ctas, stat_rows, st));
        if ((d->flags & 512) && stat_rows != nullptr) *stat_rows = 1;
        return rc;
    }

    // stride 2: one GEMM per output parity phase (ph, pw); dx[2i+ph, 2j+pw] only sees taps with
    // r = ph+1 (mod 2), s = pw+1 (mod 2), reading dy[i + (ph+1-r)/2, j + (pw+1-s)/2].
    if (!grid_tiles(ho, wo, 128)) return B2C_ERR_UNSUPPORTED;
    const int bh = (128 / wo) < ho ? (128 / wo) : ho;
    const int bnb = 128 / (wo * bh);
    rc = encode_nhwc(&ma, dy, d->n, ho, wo, d->cout, wo, bh, bnb, 1, y_pitch(d));
    if (rc) return rc;
    p.m_total = d->n * ho * wo;
    p.num_m_tiles = (p.m_total + 127) / 128;
    p.gH = ho;
    p.gW = wo;
    p.sy = p.sx = 2;
    if (d->ksize == 1) {
        if (bn_fuse != nullptr) return B2C_ERR_UNSUPPORTED;
        if (add_src != nullptr && add_src != dx) return B2C_ERR_UNSUPPORTED;
        if (add_src == nullptr) {
            if (x_pitch(d) != d->cin) return B2C_ERR_UNSUPPORTED;      // (the zero fill assumes a dense dx)
            cudaError_t e = cudaMemsetAsync(dx, 0, static_cast<size_t>(d->n) * d->h * d->w * d->cin * 2, st);
            if (e != cudaSuccess) return cuda_rc(e);
        }
        p.num_taps = 1;
        p.tap_dh[0] = 0;
        p.tap_dw[0] = 0;
        p.tap_a[0] = 0;
        p.tap_brow[0] = 0;
        single_phase(p);
        if (pair) return cuda_rc(b2c::launch_tap_gemm2(bn, ma, ma, mb, p, max_ctas, nullptr, st));
        return cuda_rc(b2c::launch_tap_gemm(bn, ma, ma, mb, p, max_ctas, nullptr, st));
    }
    // 3x3: all four parity phases in ONE launch (tiles = phase x pixel tile): dy is read from HBM once and the
    // phases, which differ in tap count (1, 2, 2, 4), balance across the persistent CTAs.
    CUtensorMap ma1 = ma;
    int nt = 0, np = 0;
    for (int ph = 0; ph < 2; ++ph) {
        for (int pw = 0; pw < 2; ++pw) {
            p.phase_begin[np] = static_cast<uint8_t>(nt);
            p.phase_py[np] = static_cast<uint8_t>(ph);
            p.phase_px[np] = static_cast<uint8_t>(pw);
            for (int r = 0; r < 3; ++r) {
                if (((ph + 1 - r) & 1) != 0) continue;
                for (int s = 0; s < 3; ++s) {
                    if (((pw + 1 - s) & 1) != 0) continue;
                    p.tap_dh[nt] = static_cast<int8_t>((ph + 1 - r) / 2);
                    p.tap_dw[nt] = static_cast<int8_t>((pw + 1 - s) / 2);
                    p.tap_a[nt] = 0;
                    p.tap_brow[nt] = (r * 3 + s) * d->cin;
                    ++nt;
                }
            }
            if (np == 0 && dy_shortcut != nullptr) {
                // the 1x1 stride-2 shortcut conv reads x[2i, 2j]: its dgrad lands on phase (0, 0) only, as one more
                // tap whose A operand is the shortcut's dy and whose weights are slab 9 of the B matrix
                rc = encode_nhwc(&ma1, dy_shortcut, d->n, ho, wo, d->cout, wo, bh, bnb, 1, y_pitch(d));
                if (rc) return rc;
                p.tap_dh[nt] = 0;
                p.tap_dw[nt] = 0;
                p.tap_a[nt] = 1;
                p.tap_brow[nt] = 9 * d->cin;
                ++nt;
            }
            ++np;
        }
    }
    p.phase_begin[4] = static_cast<uint8_t>(nt);
    p.num_taps = nt;
    p.num_phases = 4;
    rc = pair ? cuda_rc(b2c::launch_tap_gemm2(bn, ma, ma1, mb, p, max_ctas, stat_rows, st))
              : cuda_rc(b2c::launch_tap_gemm(bn, ma, ma1, mb, p, max_ctas, stat_rows, st));
    if ((d->flags & 512) && stat_rows != nullptr) *stat_rows = 1;
    return rc;
}

int b2c_conv_dgrad(const B2cConvDesc* d, const void* dy, const void* w_dgrad, void* dx, const void* add_src,
                   void* stream) {
    return conv_dgrad_impl(d, dy, nullptr, w_dgrad, dx, add_src, stream);
}

int b2c_conv_dgrad_bn(const B2cConvDesc* d, const void* dy, const void* w_dgrad, void* dz, const void* add_src,
                      const void* y_bn, const float* scale, const float* shift, float* partial, int* stat_rows,
                      void* stream) {
    if (!y_bn || !scale || !shift || !partial || !stat_rows || !aligned16(y_bn)) return B2C_ERR_ARG;
    BnFuse f = {y_bn, scale, shift, partial, stat_rows};
    return conv_dgrad_impl(d, dy, nullptr, w_dgrad, dz, add_src, stream, &f);
}

int b2c_conv_dgrad_with_shortcut(const B2cConvDesc* d, const void* dy, const void* dy_shortcut, const void* w_dgrad10,
                                 void* dx, const void* add_src, void* stream) {
    if (d == nullptr || d->ksize != 3 || d->stride != 2) return B2C_ERR_UNSUPPORTED;
    if (!dy_shortcut || !aligned16(dy_shortcut)) return B2C_ERR_ARG;
    return conv_dgrad_impl(d, dy, dy_shortcut, w_dgrad10, dx, add_src, stream);
}

int b2c_conv_wgrad(const B2cConvDesc* d, const void* x, const void* dy, float* dw, float* scratch,
                   long long scratch_floats, void* stream) {
    int rc = check_conv(d);
    if (rc) return rc;
    if (!x || !dy || !dw || !aligned16(x) || !aligned16(dy) || !aligned16(scratch)) return B2C_ERR_ARG;
    const int ho = d->h / d->stride, wo = d->w / d->stride;
    // halo-reuse kernel: 3x3 stride 1, square 16/32-pixel images, unpadded channel counts, scratch provided.
    // Measured (profiles/r1b_conv_layer_times.json): a win for 64x64 channel slabs (125 -> 103 us), not for 128x128
    // (95 -> 103 us, the N=64 MMAs are shared-memory-operand bound), so larger layers keep the atomic kernel unless
    // flags bit 1 asks for the halo kernel.
    if (scratch != nullptr && !(d->flags & 1) && ((d->cin == 64 && d->cout == 64) || (d->flags & 2)) && d->ksize == 3 && d->stride == 1 && d->h == d->w &&
        (d->w == 16 || d->w == 32) && (d->cin_valid == 0 || d->cin_valid == d->cin) &&
        (d->cout_valid == 0 || d->cout_valid == d->cout) && d->kflat_cin == 0) {
        b2c::WgradHaloParams hp = {};
        hp.cin_slabs = d->cin / 64;
        hp.cout_slabs = d->cout / 64;
        hp.tiles_h = d->h / 16;
        hp.tiles_w = d->w / 8;
        hp.num_tiles = d->n * hp.tiles_h * hp.tiles_w;
        const int units = hp.cin_slabs * hp.cout_slabs;
        const int max_ctas = d->max_ctas > 0 ? d->max_ctas : 148;
        long long splits = max_ctas / units;
        const long long fit = scratch_floats / (static_cast<long long>(units) * 5 * 128 * 64);
        if (splits > fit) splits = fit;
        if (splits > hp.num_tiles) splits = hp.num_tiles;
        if (splits >= 1) {
            hp.tiles_per_split = static_cast<int>((hp.num_tiles + splits - 1) / splits);
            hp.splits = (hp.num_tiles + hp.tiles_per_split - 1) / hp.tiles_per_split;
            hp.partial = scratch;
            CUtensorMap mx, mdy;
            rc = encode_nhwc(&mx, x, d->n, d->h, d->w, d->cin, 10, 18, 1, 1, x_pitch(d));
            if (rc) return rc;
            rc = encode_nhwc(&mdy, dy, d->n, d->h, d->w, d->cout, 8, 16, 1, 1, y_pitch(d));
            if (rc) return rc;
            hp.tco = (d->flags & 32) ? 1 : 0;
            return cuda_rc(b2c::launch_wgrad_halo(mx, mdy, hp, dw, as_stream(stream)));
        }
    }
    // row kernel (conv_wgrad_row.cu): 3x3 stride 1 on 16x16 / 8x8 images, channels in units of 128, [tap][cin][cout]
    // gradient layout. flags bit 6 keeps the per-tap kernel (A/B measurements).
    if ((d->flags & 32) && !(d->flags & (1 | 64)) && d->ksize == 3 && d->stride == 1 && d->h == d->w &&
        (d->w == 16 || d->w == 8) && (d->cin % 128) == 0 && (d->cout % 128) == 0 &&
        (d->cin_valid == 0 || d->cin_valid == d->cin) && (d->cout_valid == 0 || d->cout_valid == d->cout) &&
        d->kflat_cin == 0) {
        b2c::WgradRowParams rp = {};
        rp.cin_pairs = d->cin / 128;
        rp.cout_slabs = d->cout / 128;
        rp.cin = d->cin;
        rp.cout = d->cout;
        rp.dw = dw;
        const bool w16 = d->w == 16;
        rp.num_steps = w16 ? d->n * 4 : d->n;
        // CTA-pair kernel (conv_wgrad_row2.cu) when the (kernel row, cin pair) units pair up evenly: with an odd count
        // a quarter of the MMAs is wasted (measured 128->128: 69.6 us vs 63.5 single; 256->256: 57.3 vs 62.2)
        const bool pair = (d->flags & 128) != 0 && ((3 * rp.cin_pairs) % 2 == 0 || (d->flags & 256) != 0);
        const int units = (pair ? (3 * rp.cin_pairs + 1) / 2 : 3 * rp.cin_pairs) * rp.cout_slabs;
        const int max_ctas = (d->max_ctas > 0 ? d->max_ctas : 148) / (pair ? 2 : 1);
        int splits = max_ctas / units;
        if (splits < 1) splits = 1;
        if (splits > rp.num_steps) splits = rp.num_steps;
        rp.steps_per_split = (rp.num_steps + splits - 1) / splits;
        rp.splits = (rp.num_steps + rp.steps_per_split - 1) / rp.steps_per_split;
        CUtensorMap mx, mdy;
        rc = encode_nhwc(&mx, x, d->n, d->h, d->w, d->cin, d->w + 2, w16 ? 4 : 8, 1, 1, x_pitch(d));
        if (rc) return rc;
        rc = encode_nhwc(&mdy, dy, d->n, d->h, d->w, d->cout, d->w, w16 ? 4 : 8, 1, 1, y_pitch(d));
        if (rc) return rc;
        if (pair) return cuda_rc(b2c::launch_wgrad_row2(mx, mdy, rp, w16 ? 1 : 0, as_stream(stream)));
        return cuda_rc(b2c::launch_wgrad_row(mx, mdy, rp, w16 ? 1 : 0, as_stream(stream)));
    }
    const int bn = pick_bn(d->block_n, d->cout, 256, 1 << 20);
    const int kpx = (bn <= 128 && grid_tiles(ho, wo, 128)) ? 128 : 64;     // pixels per k-step
    if (!grid_tiles(ho, wo, kpx)) return B2C_ERR_UNSUPPORTED;
    if (d->cout % bn) return B2C_ERR_UNSUPPORTED;
    const int taps = d->ksize * d->ksize;
    const int cin_valid = d->cin_valid > 0 ? d->cin_valid : d->cin;
    const int cout_valid = d->cout_valid > 0 ? d->cout_valid : d->cout;
    if (cin_valid > d->cin || cout_valid > d->cout) return B2C_ERR_ARG;
    if (d->kflat_cin > 0 && (taps != 1 || (cin_valid % d->kflat_cin) != 0)) return B2C_ERR_ARG;

    WgradParams p = {};
    const long long pixels = static_cast<long long>(d->n) * ho * wo;
    p.k_pixels = kpx;
    p.k_chunks_total = static_cast<int>((pixels + kpx - 1) / kpx);
    p.cin_blocks = d->cin / 64;
    p.num_slabs = taps * p.cin_blocks;
    p.num_m_blocks = (p.num_slabs + 1) / 2;
    p.num_n_tiles = d->cout / bn;
    const int tiles = p.num_m_blocks * p.num_n_tiles;
    const int max_ctas = d->max_ctas > 0 ? d->max_ctas : 148;
    int splits = max_ctas / tiles;
    if (splits < 1) splits = 1;
    if (splits > p.k_chunks_total) splits = p.k_chunks_total;
    p.chunks_per_split = (p.k_chunks_total + splits - 1) / splits;
    p.splits = (p.k_chunks_total + p.chunks_per_split - 1) / p.chunks_per_split;
    p.gH = ho;
    p.gW = wo;
    p.in_stride = d->stride;
    p.num_taps = taps;
    for (int t = 0; t < taps; ++t) {
        const int r = t / d->ksize, s = t % d->ksize;
        p.tap_dh[t] = static_cast<int8_t>(r - d->ksize / 2);
        p.tap_dw[t] = static_cast<int8_t>(s - d->ksize / 2);
    }
    p.cin = d->cin;
    p.cout = d->cout;
    p.cin_valid = cin_valid;
    p.cout_valid = cout_valid;
    p.kflat_cin = d->kflat_cin;
    p.dw = dw;
    if (d->kflat_cin > 0) {
        // dw is [cout, kflat_cin, kk, kk] with kk*kk = cin_valid / kflat_cin flattened taps
        const int ftaps = cin_valid / d->kflat_cin;
        p.o_stride = static_cast<long long>(cin_valid);
        p.c_stride = ftaps;
        p.t_stride = 1;
    } else if (d->flags & 32) {      // dw is [tap][cin_valid][cout_valid] (coalesced vector reductions)
        p.o_stride = 1;
        p.c_stride = cout_valid;
        p.t_stride = static_cast<long long>(cin_valid) * cout_valid;
    } else {
        p.o_stride = static_cast<long long>(cin_valid) * taps;
        p.c_stride = taps;
        p.t_stride = 1;
    }

    const int bh = (kpx / wo) < ho ? (kpx / wo) : ho;
    const int bnb = kpx / (wo * bh);
    CUtensorMap mx, mdy;
    p.debug = d->flags & 0x18;
    p.x_split_w = d->stride == 2 && !(d->flags & 4);
    rc = p.x_split_w ? encode_nhwc_s2(&mx, x, d->n, d->h, d->w, d->cin, wo, bh, bnb, x_pitch(d))
                     : encode_nhwc(&mx, x, d->n, d->h, d->w, d->cin, wo, bh, bnb, d->stride, x_pitch(d));
    if (rc) return rc;
    rc = encode_nhwc(&mdy, dy, d->n, ho, wo, d->cout, wo, bh, bnb, 1, y_pitch(d));
    if (rc) return rc;
    return cuda_rc(b2c::launch_wgrad(bn, mx, mdy, p, as_stream(stream)));
}

static int check_dw(const void* a, const void* b, const void* c, int n, int h, int w, int channels, int c_valid,
                    int stride) {
    if (!a || !b || !c || n < 1 || h < 1 || w < 1) return B2C_ERR_ARG;
    if (channels < 8 || (channels & 7) || (channels >> 3) > 256 || c_valid < 1 || c_valid > channels)
        return B2C_ERR_UNSUPPORTED;
    if ((stride != 1 && stride != 2) || (stride == 2 && ((h | w) & 1))) return B2C_ERR_UNSUPPORTED;
    return 0;
}

int b2c_dwconv_stat_rows(long long out_pixels, int channels) {
    if (channels < 8 || (channels & 7) || (channels >> 3) > 256 || out_pixels < 1) return B2C_ERR_UNSUPPORTED;
    return b2c::dw_fprop_partial_rows(out_pixels, channels);
}

int b2c_dwconv_fprop(const void* x, const float* weight, void* y, int n, int h, int w, int channels, int c_valid,
                     int stride, float* part_sum, float* part_sq, int* stat_rows, void* stream) {
    int rc = check_dw(x, weight, y, n, h, w, channels, c_valid, stride);
    if (rc) return rc;
    if (!aligned16(x) || !aligned16(y) || (part_sum == nullptr) != (part_sq == nullptr)) return B2C_ERR_ARG;
    return cuda_rc(b2c::launch_dw_fprop(x, weight, y, n, h, w, channels, c_valid, stride, part_sum, part_sq, stat_rows,
                                        as_stream(stream)));
}

int b2c_dwconv_fprop_acc(const void* x, const float* weight, void* y, int n, int h, int w, int channels, int c_valid,
                         int stride, float* sum, float* sumsq, void* stream) {
    int rc = check_dw(x, weight, y, n, h, w, channels, c_valid, stride);
    if (rc) return rc;
    if (!aligned16(x) || !aligned16(y) || !sum || !sumsq) return B2C_ERR_ARG;
    return cuda_rc(b2c::launch_dw_fprop(x, weight, y, n, h, w, channels, c_valid, stride, sum, sumsq, nullptr,
                                        as_stream(stream), 1));
}

int b2c_dwconv_input_transform_ok(int h, int w, int channels, int stride) {
    if (channels < 8 || (channels & 7) || (channels >> 3) > 256) return 0;
    return b2c::dw_input_transform_ok(h, w, channels, stride) ? 1 : 0;
}

int b2c_dwconv_fprop_bnrelu(const void* y_prev, const float* in_scale, const float* in_shift, const float* weight,
                            void* y, int n, int h, int w, int channels, int c_valid, float* part_sum, float* part_sq,
                            int* stat_rows, int accumulate, void* stream) {
    int rc = check_dw(y_prev, weight, y, n, h, w, channels, c_valid, 1);
    if (rc) return rc;
    if (!in_scale || !in_shift || !aligned16(y_prev) || !aligned16(y) || (part_sum == nullptr) != (part_sq == nullptr))
        return B2C_ERR_ARG;
    if (!b2c::dw_input_transform_ok(h, w, channels, 1)) return B2C_ERR_UNSUPPORTED;
    return cuda_rc(b2c::launch_dw_fprop_xf(y_prev, in_scale, in_shift, weight, y, n, h, w, channels, c_valid, part_sum,
                                           part_sq, stat_rows, as_stream(stream), accumulate ? 1 : 0));
}

int b2c_dwconv_wgrad_bnrelu(const void* y_prev, const float* in_scale, const float* in_shift, const void* dy,
                            float* dweight, int n, int h, int w, int channels, int c_valid, void* stream) {
    int rc = check_dw(y_prev, dy, dweight, n, h, w, channels, c_valid, 1);
    if (rc) return rc;
    if (!in_scale || !in_shift || !aligned16(y_prev) || !aligned16(dy)) return B2C_ERR_ARG;
    if (!b2c::dw_input_transform_ok(h, w, channels, 1)) return B2C_ERR_UNSUPPORTED;
    return cuda_rc(b2c::launch_dw_wgrad_xf(y_prev, in_scale, in_shift, dy, dweight, n, h, w, channels, c_valid,
                                           as_stream(stream)));
}

int b2c_dwconv_dgrad(const void* dy, const float* weight, void* dx, int n, int h, int w, int channels, int c_valid,
                     int stride, void* stream) {
    int rc = check_dw(dy, weight, dx, n, h, w, channels, c_valid, stride);
    if (rc) return rc;
    if (!aligned16(dy) || !aligned16(dx)) return B2C_ERR_ARG;
    return cuda_rc(b2c::launch_dw_dgrad(dy, weight, dx, n, h, w, channels, c_valid, stride, as_stream(stream)));
}

int b2c_dwconv_wgrad(const void* x, const void* dy, float* dweight, int n, int h, int w, int channels, int c_valid,
                     int stride, void* stream) {
    int rc = check_dw(x, dy, dweight, n, h, w, channels, c_valid, stride);
    if (rc) return rc;
    if (!aligned16(x) || !aligned16(dy)) return B2C_ERR_ARG;
    return cuda_rc(b2c::launch_dw_wgrad(x, dy, dweight, n, h, w, channels, c_valid, stride, as_stream(stream)));
}

int b2c_bn_finalize(const float* part_sum, const float* part_sq, int rows, int channels, float count,
                    const float* gamma, const float* beta, const float* conv_bias, float eps, float momentum,
                    float* running_mean, float* running_var, float* scale, float* shift, float* save_mean,
                    float* save_rstd, void* stream) {
    if (!part_sum || !part_sq || !gamma || !beta || !scale || !shift || !save_mean || !save_rstd || rows < 1 ||
        channels < 1 || count <= 0.f)
        return B2C_ERR_ARG;
    if ((running_mean == nullptr) != (running_var == nullptr)) return B2C_ERR_ARG;
    return cuda_rc(b2c::launch_bn_finalize(part_sum, part_sq, rows, channels, count, gamma, beta, conv_bias, eps, momentum,
                                           running_mean, running_var, scale, shift, save_mean, save_rstd,
                                           as_stream(stream)));
}

int b2c_bn_eval_scale_shift(int channels, const float* gamma, const float* beta, const float* conv_bias,
                            const float* running_mean, const float* running_var, float eps, float* scale,
                            float* shift, void* stream) {
    if (!gamma || !beta || !running_mean || !running_var || !scale || !shift || channels < 1) return B2C_ERR_ARG;
    return cuda_rc(b2c::launch_bn_eval_scale_shift(channels, gamma, beta, conv_bias, running_mean, running_var, eps, scale,
                                                   shift, as_stream(stream)));
}

int b2c_bn_act_apply(const void* y, const float* scale, const float* shift, const void* res,
                     const float* res_scale, const float* res_shift, void* out, long long rows, int channels,
                     int relu, unsigned char* relu_bits, void* stream) {
    if (!y || !scale || !shift || !out || rows < 0 || channels < 8 || (channels & 7)) return B2C_ERR_ARG;
    if (!aligned16(y) || !aligned16(out) || !aligned16(res) || !aligned16(scale) || !aligned16(shift) ||
        !aligned16(res_scale) || !aligned16(res_shift))
        return B2C_ERR_ARG;
    if ((res_scale == nullptr) != (res_shift == nullptr)) return B2C_ERR_ARG;
    if (res_scale != nullptr && res == nullptr) return B2C_ERR_ARG;
    if (rows == 0) return 0;
    return cuda_rc(b2c::launch_bn_act_apply(y, scale, shift, res, res_scale, res_shift, out, rows, channels, relu,
                                            relu_bits, as_stream(stream)));
}

static bool bn_stats_ok(const B2cBnStats* b) {
    return b && b->sum && b->sumsq && b->gamma && b->beta && b->scale && b->shift && b->save_mean && b->save_rstd &&
           b->count > 0.f && (b->running_mean == nullptr) == (b->running_var == nullptr);
}

int b2c_bn_act_apply_stats(const void* y, const B2cBnStats* bn, const void* res, const B2cBnStats* res_bn, void* out,
                           long long rows, int channels, int relu, unsigned char* relu_bits, void* stream) {
    static_assert(sizeof(B2cBnStats) == sizeof(b2c::BnStatArgs), "ABI struct mismatch");
    if (!y || !out || rows < 0 || channels < 8 || (channels & 7) || channels > 4096) return B2C_ERR_ARG;
    if (!bn_stats_ok(bn) || (res_bn != nullptr && (!bn_stats_ok(res_bn) || res == nullptr))) return B2C_ERR_ARG;
    if (!aligned16(y) || !aligned16(out) || !aligned16(res)) return B2C_ERR_ARG;
    if (rows == 0) return 0;
    return cuda_rc(b2c::launch_bn_act_apply_stats(y, *reinterpret_cast<const b2c::BnStatArgs*>(bn), res,
                                                  reinterpret_cast<const b2c::BnStatArgs*>(res_bn), out, rows, channels,
                                                  relu, relu_bits, as_stream(stream)));
}

int b2c_bn_bwd_reduce_acc(const void* dout, const void* act, const void* y, void* dz_out, long long rows, int channels,
                          float* sums, const float* relu_scale, const float* relu_shift,
                          const unsigned char* relu_bits, void* stream) {
    if (!dout || !y || !sums || rows < 1) return B2C_ERR_ARG;
    if (relu_bits != nullptr && (act != nullptr || relu_scale != nullptr)) return B2C_ERR_ARG;
    if ((relu_scale == nullptr) != (relu_shift == nullptr) || (relu_scale != nullptr && act != nullptr))
        return B2C_ERR_ARG;
    if (channels < 8 || (channels & 7) || (channels >> 3) > 256) return B2C_ERR_UNSUPPORTED;
    if (!aligned16(dout) || !aligned16(act) || !aligned16(y) || !aligned16(dz_out)) return B2C_ERR_ARG;
    return cuda_rc(b2c::launch_bn_bwd_reduce(dout, act, y, dz_out, rows, channels, sums, relu_scale, relu_shift,
                                             relu_bits, as_stream(stream), 1));
}

int b2c_bn_bwd_apply_stats(const void* dz, const void* y, const float* sums, float count, const float* gamma,
                           const float* save_mean, const float* save_rstd, float* dgamma, float* dbeta,
                           const void* add, void* dy, long long rows, int channels, const float* relu_scale,
                           const float* relu_shift, void* stream) {
    if ((relu_scale == nullptr) != (relu_shift == nullptr)) return B2C_ERR_ARG;
    if (!dz || !y || !sums || !gamma || !save_mean || !save_rstd || !dgamma || !dbeta || !dy || rows < 0 ||
        channels < 8 || (channels & 7) || channels > 4096 || count <= 0.f)
        return B2C_ERR_ARG;
    if (!aligned16(dz) || !aligned16(y) || !aligned16(dy) || !aligned16(add)) return B2C_ERR_ARG;
    if (rows == 0) return 0;
    return cuda_rc(b2c::launch_bn_bwd_apply_stats(dz, y, sums, count, gamma, save_mean, save_rstd, dgamma, dbeta, add,
                                                  dy, rows, channels, relu_scale, relu_shift, as_stream(stream)));
}

int b2c_bn_bwd_partial_rows(long long rows, int channels) {
    if (channels < 8 || (channels & 7) || (channels >> 3) > 256) return B2C_ERR_UNSUPPORTED;
    return b2c::bn_bwd_partial_rows(rows, channels);
}

int b2c_bn_bwd_reduce(const void* dout, const void* act, const void* y, void* dz_out, long long rows,
                      int channels, float* partial, const float* relu_scale, const float* relu_shift,
                      const unsigned char* relu_bits, void* stream) {
    if (!dout || !y || !partial || rows < 1) return B2C_ERR_ARG;
    if (relu_bits != nullptr && (act != nullptr || relu_scale != nullptr)) return B2C_ERR_ARG;
    if ((relu_scale == nullptr) != (relu_shift == nullptr) || (relu_scale != nullptr && act != nullptr))
        return B2C_ERR_ARG;
    if (channels < 8 || (channels & 7) || (channels >> 3) > 256) return B2C_ERR_UNSUPPORTED;
    if (!aligned16(dout) || !aligned16(act) || !aligned16(y) || !aligned16(dz_out)) return B2C_ERR_ARG;
    return cuda_rc(b2c::launch_bn_bwd_reduce(dout, act, y, dz_out, rows, channels, partial, relu_scale, relu_shift,
                                             relu_bits, as_stream(stream)));
}

int b2c_bn_bwd_finalize(const float* partial, int partial_rows, int channels, float count, const float* gamma,
                        const float* save_mean, const float* save_rstd, float* dgamma, float* dbeta, float* coef_a,
                        float* coef_b, float* coef_c, void* stream) {
    if (!partial || !gamma || !save_mean || !save_rstd || !dgamma || !dbeta || !coef_a || !coef_b || !coef_c ||
        partial_rows < 1 || channels < 1 || count <= 0.f)
        return B2C_ERR_ARG;
    return cuda_rc(b2c::launch_bn_bwd_finalize(partial, partial_rows, channels, count, gamma, save_mean, save_rstd,
                                               dgamma, dbeta, coef_a, coef_b, coef_c, as_stream(stream)));
}

int b2c_bn_bwd_stats_noalloc(const void* dout, const void* act, const void* y, void* dz_out, long long rows,
                             int channels, float count, const float* gamma, const float* save_mean,
                             const float* save_rstd, float* dgamma, float* dbeta, float* coef_a, float* coef_b,
                             float* coef_c, void* stream) {
    static_assert(B2C_NOALLOC_MAX_CHANNELS == 512, "keep in step with kNoAllocMaxChannels (elementwise.cu)");
    if (!dout || !y || !gamma || !save_mean || !save_rstd || !dgamma || !dbeta || !coef_a || !coef_b || !coef_c ||
        rows < 1 || count <= 0.f)
        return B2C_ERR_ARG;
    if (channels > b2c::bn_bwd_noalloc_max_channels()) return B2C_ERR_TOO_LARGE;
    if (channels < 8 || (channels & 7)) return B2C_ERR_UNSUPPORTED;
    if (!aligned16(dout) || !aligned16(act) || !aligned16(y) || !aligned16(dz_out)) return B2C_ERR_ARG;
    return cuda_rc(b2c::launch_bn_bwd_stats_noalloc(dout, act, y, dz_out, rows, channels, count, gamma, save_mean,
                                                    save_rstd, dgamma, dbeta, coef_a, coef_b, coef_c,
                                                    as_stream(stream)));
}

int b2c_bn_bwd_apply(const void* dz, const void* y, const float* coef_a, const float* coef_b, const float* coef_c,
                     const void* add, void* dy, long long rows, int channels, const float* relu_scale,
                     const float* relu_shift, void* stream) {
    if ((relu_scale == nullptr) != (relu_shift == nullptr)) return B2C_ERR_ARG;
    if (!dz || !y || !coef_a || !coef_b || !coef_c || !dy || rows < 0 || channels < 8 || (channels & 7))
        return B2C_ERR_ARG;
    if (!aligned16(dz) || !aligned16(y) || !aligned16(dy) || !aligned16(add) || !aligned16(coef_a) ||
        !aligned16(coef_b) || !aligned16(coef_c))
        return B2C_ERR_ARG;
    if (rows == 0) return 0;
    return cuda_rc(b2c::launch_bn_bwd_apply(dz, y, coef_a, coef_b, coef_c, add, dy, rows, channels, relu_scale,
                                            relu_shift, as_stream(stream)));
}

int b2c_maxpool2x2_fwd(const void* x, void* y, int n, int h, int w, int channels, void* stream) {
    if (!x || !y || n < 1 || h < 2 || w < 2 || (h & 1) || (w & 1) || channels < 8 || (channels & 7) ||
        !aligned16(x) || !aligned16(y))
        return B2C_ERR_ARG;
    return cuda_rc(b2c::launch_maxpool_fwd(x, y, n, h, w, channels, as_stream(stream)));
}

int b2c_maxpool2x2_bwd(const void* x, const void* y, const void* dy, void* dx, int n, int h, int w, int channels,
                       void* stream) {
    if (!x || !y || !dy || !dx || n < 1 || h < 2 || w < 2 || (h & 1) || (w & 1) || channels < 8 || (channels & 7) ||
        !aligned16(x) || !aligned16(y) || !aligned16(dy) || !aligned16(dx))
        return B2C_ERR_ARG;
    return cuda_rc(b2c::launch_maxpool_bwd(x, y, dy, dx, n, h, w, channels, as_stream(stream)));
}

int b2c_head_fwd(const void* act, int n, int hw, int channels, const float* weight, const float* bias,
                 int classes, float* pooled, float* logits, void* stream) {
    if (!act || !weight || !bias || !pooled || !logits || n < 1 || hw < 1 || classes < 1) return B2C_ERR_ARG;
    if (channels > 4096 || (channels & 31)) return B2C_ERR_UNSUPPORTED;
    return cuda_rc(b2c::launch_head_fwd(act, n, hw, channels, weight, bias, classes, pooled, logits,
                                        as_stream(stream)));
}

int b2c_head_fwd_ld(const void* act, int n, int hw, int channels, int act_ld, const float* weight, const float* bias,
                    int classes, float* pooled, float* logits, void* stream) {
    if (!act || !weight || !bias || !pooled || !logits || n < 1 || hw < 1 || classes < 1 || channels < 1) return B2C_ERR_ARG;
    if (channels > 4096 || act_ld < channels) return B2C_ERR_UNSUPPORTED;
    return cuda_rc(b2c::launch_head_fwd(act, n, hw, channels, weight, bias, classes, pooled, logits,
                                        as_stream(stream), act_ld));
}

int b2c_ce_loss(const float* logits, const long long* target, int n, int classes, float* dlogits, float* loss,
                int* correct, void* stream) {
    if (!logits || !target || !dlogits || !loss || n < 1 || n > 65536 || classes < 1) return B2C_ERR_ARG;
    return cuda_rc(b2c::launch_ce_loss(logits, target, n, classes, dlogits, loss, correct, as_stream(stream)));
}

int b2c_head_bwd(const float* dlogits, const float* pooled, const float* weight, int n, int hw, int channels,
                 int classes, void* dact, float* dweight, float* dbias, void* stream) {
    if (!dlogits || !pooled || !weight || !dact || !dweight || !dbias || n < 1 || hw < 1 || classes < 1)
        return B2C_ERR_ARG;
    if (channels > 4096 || (channels & 31)) return B2C_ERR_UNSUPPORTED;
    return cuda_rc(b2c::launch_head_bwd(dlogits, pooled, weight, n, hw, channels, classes, dact, dweight, dbias,
                                        as_stream(stream)));
}

int b2c_head_bwd_ld(const float* dlogits, const float* pooled, const float* weight, int n, int hw, int channels,
                    int act_ld, int classes, void* dact, float* dweight, float* dbias, void* stream) {
    if (!dlogits || !pooled || !weight || !dact || !dweight || !dbias || n < 1 || hw < 1 || classes < 1 || channels < 1)
        return B2C_ERR_ARG;
    if (channels > 4096 || act_ld < channels) return B2C_ERR_UNSUPPORTED;
    return cuda_rc(b2c::launch_head_bwd(dlogits, pooled, weight, n, hw, channels, classes, dact, dweight, dbias,
                                        as_stream(stream), act_ld));
}

int b2c_se_fwd(const void* y, int n, int hw, int channels, int ld, int reduced, const float* w1, const float* b1,
               const float* w2, const float* b2, float* pooled, float* hidden, float* scale, const void* add, void* out,
               void* stream) {
    if (!y || !w1 || !b1 || !w2 || !b2 || !pooled || !hidden || !scale || !out || n < 1 || hw < 1 || channels < 1 ||
        reduced < 1 || !aligned16(y) || !aligned16(out) || !aligned16(add) || !aligned16(scale))
        return B2C_ERR_ARG;
    if (ld < channels || (ld & 7) || channels > 4096 || reduced > 1024) return B2C_ERR_UNSUPPORTED;
    cudaStream_t st = as_stream(stream);
    int rc = cuda_rc(b2c::launch_se_pool(y, nullptr, n, hw, ld, 1.f / static_cast<float>(hw), pooled, st));
    if (rc) return rc;
    rc = cuda_rc(b2c::launch_se_fc(pooled, n, channels, reduced, ld, w1, b1, w2, b2, hidden, scale, st));
    if (rc) return rc;
    return cuda_rc(b2c::launch_se_scale(y, scale, add, out, n, hw, ld, st));
}

int b2c_se_bwd(const void* dout, const void* y, int n, int hw, int channels, int ld, int reduced, const float* w1,
               const float* w2, const float* pooled, const float* hidden, const float* scale, float* ds_scratch,
               float* dz1_scratch, float* dp_scratch, float* dw1, float* db1, float* dw2, float* db2, void* dy,
               void* stream) {
    if (!dout || !y || !w1 || !w2 || !pooled || !hidden || !scale || !ds_scratch || !dz1_scratch || !dp_scratch ||
        !dw1 || !db1 || !dw2 || !db2 || !dy || n < 1 || hw < 1 || channels < 1 || reduced < 1 || !aligned16(dout) ||
        !aligned16(y) || !aligned16(dy))
        return B2C_ERR_ARG;
    if (ld < channels || (ld & 7) || channels > 4096 || reduced > 1024) return B2C_ERR_UNSUPPORTED;
    return cuda_rc(b2c::launch_se_bwd(dout, y, ds_scratch, scale, hidden, pooled, n, hw, channels, reduced, ld, w1, w2,
                                      dz1_scratch, dp_scratch, dw1, db1, dw2, db2, dy, as_stream(stream)));
}

int b2c_maxpool3x3_fwd(const void* x, void* y, void* argmax, int n, int h, int w, int channels, int stride,
                       void* stream) {
    if (!x || !y || !argmax || n < 0 || h < 1 || w < 1 || !aligned16(x) || !aligned16(y) || !aligned16(argmax))
        return B2C_ERR_ARG;
    if ((channels & 7) || channels < 8 || (stride != 1 && stride != 2) || (h % stride) || (w % stride))
        return B2C_ERR_UNSUPPORTED;
    return cuda_rc(b2c::launch_maxpool3_fwd(x, y, argmax, n, h, w, channels, stride, as_stream(stream)));
}

int b2c_maxpool3x3_bwd(const void* dy, const void* argmax, const void* add, void* dx, int n, int h, int w, int channels,
                       int stride, void* stream) {
    if (!dy || !argmax || !dx || n < 0 || h < 1 || w < 1 || !aligned16(dy) || !aligned16(dx) || !aligned16(add) ||
        !aligned16(argmax))
        return B2C_ERR_ARG;
    if ((channels & 7) || channels < 8 || (stride != 1 && stride != 2) || (h % stride) || (w % stride))
        return B2C_ERR_UNSUPPORTED;
    return cuda_rc(b2c::launch_maxpool3_bwd(dy, argmax, add, dx, n, h, w, channels, stride, as_stream(stream)));
}

int b2c_copy_channels(const void* src, int src_ld, int src_off, void* dst, int dst_ld, int dst_off, long long rows,
                      int channels, void* stream) {
    if (!src || !dst || rows < 0 || channels < 0 || !aligned16(src) || !aligned16(dst)) return B2C_ERR_ARG;
    if (((src_ld | src_off | dst_ld | dst_off | channels) & 7) || src_off + channels > src_ld ||
        dst_off + channels > dst_ld || src_off < 0 || dst_off < 0)
        return B2C_ERR_UNSUPPORTED;
    return cuda_rc(b2c::launch_copy_channels(src, src_ld, src_off, dst, dst_ld, dst_off, rows, channels,
                                             as_stream(stream)));
}

int b2c_augment_nchw(const uint8_t* data, const long long* idx, const int8_t* ijf, float* out_nchw_f32, int n,
                     const float* mean_host, const float* std_host, void* stream) {
    if (n == 0) return 0;   // empty batch: nothing to do (zero-size tensors have null pointers)
    if (!data || !idx || !ijf || !out_nchw_f32 || !mean_host || !std_host || n < 0 || !aligned16(data))
        return B2C_ERR_ARG;
    return cuda_rc(b2c::launch_augment_nchw(data, idx, ijf, out_nchw_f32, n, mean_host, std_host, as_stream(stream)));
}

int b2c_augment_im2col(const uint8_t* data, const long long* idx, const int8_t* ijf, void* out_im2col_bf16,
                       int n, const float* mean_host, const float* std_host, void* stream) {
    if (n == 0) return 0;   // empty batch: nothing to do (zero-size tensors have null pointers)
    if (!data || !idx || !ijf || !out_im2col_bf16 || !mean_host || !std_host || n < 0 || !aligned16(data) ||
        !aligned16(out_im2col_bf16))
        return B2C_ERR_ARG;
    return cuda_rc(
        b2c::launch_augment_im2col(data, idx, ijf, out_im2col_bf16, n, mean_host, std_host, as_stream(stream)));
}

int b2c_stem_im2col(const float* x_nchw_f32, void* out_im2col_bf16, int n, void* stream) {
    if (n == 0) return 0;   // empty batch: nothing to do (zero-size tensors have null pointers)
    if (!x_nchw_f32 || !out_im2col_bf16 || n < 0 || !aligned16(out_im2col_bf16)) return B2C_ERR_ARG;
    return cuda_rc(b2c::launch_stem_im2col(x_nchw_f32, out_im2col_bf16, n, as_stream(stream)));
}

int b2c_augment_nhwc64(const uint8_t* data, const long long* idx, const int8_t* ijf, void* out_nhwc64_bf16, int n,
                       const float* mean_host, const float* std_host, void* stream) {
    if (n == 0) return 0;
    if (!data || !idx || !ijf || !out_nhwc64_bf16 || !mean_host || !std_host || n < 0 || !aligned16(data) ||
        !aligned16(out_nhwc64_bf16))
        return B2C_ERR_ARG;
    return cuda_rc(
        b2c::launch_augment_im2col(data, idx, ijf, out_nhwc64_bf16, n, mean_host, std_host, as_stream(stream), 1));
}

int b2c_stem_nhwc64(const float* x_nchw_f32, void* out_nhwc64_bf16, int n, void* stream) {
    if (n == 0) return 0;
    if (!x_nchw_f32 || !out_nhwc64_bf16 || n < 0 || !aligned16(out_nhwc64_bf16)) return B2C_ERR_ARG;
    return cuda_rc(b2c::launch_stem_im2col(x_nchw_f32, out_nhwc64_bf16, n, as_stream(stream), 1));
}

int b2c_grouped_wgrad_extract(const float* dense_tco, float* dw_oihw, int cout, int cin, int taps, int groups,
                              void* stream) {
    if (!dense_tco || !dw_oihw || cout <= 0 || cin <= 0 || taps <= 0 || groups <= 0) return B2C_ERR_ARG;
    if (cout % groups || cin % groups) return B2C_ERR_ARG;
    return cuda_rc(b2c::launch_grouped_wgrad_extract(dense_tco, dw_oihw, cout, cin, taps, groups, as_stream(stream)));
}

int b2c_sgd_step(float* param, const float* grad, float* momentum_buf, long long n, const float* hyper,
                 void* stream) {
    if (!param || !grad || !momentum_buf || !hyper || n < 0 || (n & 3) || !aligned16(param) || !aligned16(grad) ||
        !aligned16(momentum_buf))
        return B2C_ERR_ARG;
    if (n == 0) return 0;
    return cuda_rc(b2c::launch_sgd(param, grad, momentum_buf, n, hyper, as_stream(stream)));
}

int b2c_repack_weights(const float* master, void* w_fprop_arena, void* w_dgrad_arena, const B2cRepackEntry* table,
                       int entries, void* stream) {
    static_assert(sizeof(B2cRepackEntry) == sizeof(b2c::RepackEntry), "ABI struct mismatch");
    if (!master || !w_fprop_arena || !table || entries < 0) return B2C_ERR_ARG;
    return cuda_rc(b2c::launch_repack(master, w_fprop_arena, w_dgrad_arena,
                                      reinterpret_cast<const b2c::RepackEntry*>(table), entries,
                                      as_stream(stream)));
}

}  // extern "C"
