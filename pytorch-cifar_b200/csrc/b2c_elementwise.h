// b2c_elementwise.h — internal launch interface for the HBM-bound kernels (elementwise.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b2c {

// One conv weight tensor in the flat fp32 master arena and where its bf16 operand copies go.
// Mirrors B2cRepackEntry in include/b2c.h (same field order and types).
struct RepackEntry {
    long long src_off;  // element offset of the OIHW fp32 tensor in the master arena
    long long wf_off;   // element offset of the fprop operand [tap][cout][cin_pad] in the bf16 arena
    long long wd_off;   // element offset of the dgrad operand [tap][cin][cout]; < 0: none
    int cout, cin, taps, cin_pad;
    int kflat;          // stem: fprop operand is [cout][cin_pad] with k = tap*cin + c
    int pad_;           // cout_pad (0: cout)
};

// One BatchNorm whose batch statistics arrive as ONE accumulated row (sum, sum of squares) and are finalized inside
// the kernel that applies it (bn_act_apply_stats). Mirrors B2cBnStats in include/b2c.h (same field order and types).
struct BnStatArgs {
    const float* sum;
    const float* sumsq;
    const float* gamma;
    const float* beta;
    const float* conv_bias;
    float* running_mean;
    float* running_var;
    float* scale;
    float* shift;
    float* save_mean;
    float* save_rstd;
    float count, eps, momentum;
};

cudaError_t launch_bn_finalize(const float* part_sum, const float* part_sq, int rows, int C, float count,
                               const float* gamma, const float* beta, const float* conv_bias, float eps,
                               float momentum, float* running_mean, float* running_var, float* scale, float* shift,
                               float* mean, float* rstd, cudaStream_t s);
cudaError_t launch_bn_eval_scale_shift(int C, const float* gamma, const float* beta, const float* conv_bias,
                                       const float* rm,
                                       const float* rv, float eps, float* scale, float* shift, cudaStream_t s);
cudaError_t launch_bn_act_apply(const void* y, const float* scale, const float* shift, const void* res,
                                const float* res_scale, const float* res_shift, void* out, long long rows, int C,
                                int relu, uint8_t* relu_bits,
                                cudaStream_t s);
int bn_bwd_partial_rows(long long rows, int C);
cudaError_t launch_bn_bwd_reduce(const void* dout, const void* act, const void* y, void* dz_out, long long rows,
                                 int C, float* partial, const float* relu_scale, const float* relu_shift,
                                 const uint8_t* relu_bits, cudaStream_t s, int accumulate = 0);
cudaError_t launch_bn_act_apply_stats(const void* y, const BnStatArgs& st, const void* res, const BnStatArgs* rst,
                                      void* out, long long rows, int C, int relu, uint8_t* relu_bits, cudaStream_t s);
cudaError_t launch_bn_bwd_apply_stats(const void* dz, const void* y, const float* sums, float count,
                                      const float* gamma, const float* mean, const float* rstd, float* dgamma,
                                      float* dbeta, const void* add, void* dy, long long rows, int C,
                                      const float* relu_scale, const float* relu_shift, cudaStream_t s);
cudaError_t launch_bn_bwd_finalize(const float* partial, int prow, int C, float count, const float* gamma,
                                   const float* mean, const float* rstd, float* dgamma, float* dbeta, float* coefA,
                                   float* coefB, float* coefC, cudaStream_t s);
cudaError_t launch_bn_bwd_apply(const void* dz, const void* y, const float* coefA, const float* coefB,
                                const float* coefC, const void* add, void* dy, long long rows, int C,
                                const float* relu_scale, const float* relu_shift, cudaStream_t s);
cudaError_t launch_maxpool_fwd(const void* x, void* y, int n, int h, int w, int C, cudaStream_t s);
cudaError_t launch_maxpool_bwd(const void* x, const void* y, const void* dy, void* dx, int n, int h, int w, int C,
                               cudaStream_t s);
// ld: channel pitch of act / dact (0 = C)
cudaError_t launch_head_fwd(const void* act, int N, int hw, int C, const float* W, const float* bias, int K,
                            float* pooled, float* logits, cudaStream_t s, int ld = 0);
cudaError_t launch_ce_loss(const float* logits, const long long* target, int N, int K, float* dlogits, float* loss,
                           int* correct, cudaStream_t s);
cudaError_t launch_head_bwd(const float* dlogits, const float* pooled, const float* W, int N, int hw, int C, int K,
                            void* dact, float* dW, float* db, cudaStream_t s, int ld = 0);
cudaError_t launch_augment_nchw(const uint8_t* data, const long long* idx, const int8_t* ijf, float* out, int N,
                                const float* mean, const float* stdv, cudaStream_t s);
// taps: 9 = im2col operand of a 3x3 stem (k = tap*3 + c), 1 = operand of a 1x1 stem (k = c)
cudaError_t launch_augment_im2col(const uint8_t* data, const long long* idx, const int8_t* ijf, void* out, int N,
                                  const float* mean, const float* stdv, cudaStream_t s, int taps = 9);
cudaError_t launch_stem_im2col(const float* x, void* out, int N, cudaStream_t s, int taps = 9);
cudaError_t launch_grouped_wgrad_extract(const float* dense, float* dw, int cout, int cin, int taps, int groups,
                                         cudaStream_t s);
cudaError_t launch_sgd(float* p, const float* g, float* buf, long long n, const float* hyper, cudaStream_t s);
cudaError_t launch_repack(const float* master, void* wf, void* wd, const RepackEntry* table_dev, int entries,
                          cudaStream_t s);
int dw_fprop_partial_rows(long long out_pixels, int C);
cudaError_t launch_dw_fprop(const void* x, const float* w, void* y, int N, int H, int W, int C, int c_valid, int stride,
                            float* part_sum, float* part_sq, int* stat_rows, cudaStream_t s, int stat_acc = 0);
// Depthwise fprop / wgrad whose input is the preceding conv's raw output, BatchNorm + ReLU applied on load.
bool dw_input_transform_ok(int H, int W, int C, int stride);
cudaError_t launch_dw_fprop_xf(const void* y_prev, const float* in_scale, const float* in_shift, const float* w, void* y,
                               int N, int H, int W, int C, int c_valid, float* part_sum, float* part_sq, int* stat_rows,
                               cudaStream_t s, int stat_acc);
cudaError_t launch_dw_wgrad_xf(const void* y_prev, const float* in_scale, const float* in_shift, const void* dy,
                               float* dw, int N, int H, int W, int C, int c_valid, cudaStream_t s);
cudaError_t launch_dw_dgrad(const void* dy, const float* w, void* dx, int N, int H, int W, int C, int c_valid,
                            int stride, cudaStream_t s);
cudaError_t launch_dw_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int C, int c_valid,
                            int stride, cudaStream_t s);
// Squeeze-and-excitation (se.cu). ld = channel pitch of the activations and of the [N][ld] fp32 per-image vectors.
cudaError_t launch_se_pool(const void* a, const void* b, int N, int hw, int ld, float mult, float* out, cudaStream_t s);
cudaError_t launch_se_fc(const float* p, int N, int C, int R, int ld, const float* W1, const float* b1,
                         const float* W2, const float* b2, float* h, float* sc, cudaStream_t s);
cudaError_t launch_se_scale(const void* y, const float* sc, const void* add, void* out, int N, int hw, int ld,
                            cudaStream_t s);
cudaError_t launch_se_bwd(const void* dout, const void* y, float* ds, const float* sc, const float* h, const float* p,
                          int N, int hw, int C, int R, int ld, const float* W1, const float* W2, float* dz1, float* dp,
                          float* dW1, float* db1, float* dW2, float* db2, void* dy, cudaStream_t s);
// nn.MaxPool2d(3, stride, padding=1) and the channel-slice copy (pool3.cu)
cudaError_t launch_maxpool3_fwd(const void* x, void* y, void* arg, int n, int h, int w, int C, int stride,
                                cudaStream_t s);
cudaError_t launch_maxpool3_bwd(const void* dy, const void* arg, const void* add, void* dx, int n, int h, int w, int C,
                                int stride, cudaStream_t s);
cudaError_t launch_copy_channels(const void* src, int src_ld, int src_off, void* dst, int dst_ld, int dst_off,
                                 long long rows, int channels, cudaStream_t s);
int bn_bwd_noalloc_max_channels();
cudaError_t launch_bn_bwd_stats_noalloc(const void* dout, const void* act, const void* y, void* dz_out,
                                        long long rows, int C, float count, const float* gamma, const float* mean,
                                        const float* rstd, float* dgamma, float* dbeta, float* coefA, float* coefB,
                                        float* coefC, cudaStream_t s);

}  // namespace b2c
