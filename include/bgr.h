/* libbgr.so — C ABI of the B200-native Bias-GAN hot path.
 *
 * Every entry point takes plain pointers, sizes and a cudaStream_t passed as void*; it returns
 * BGR_OK (0) or a negative status, and bgr_last_error() gives the thread-local message.  No torch
 * types cross this boundary.  Device pointers are bf16 NHWC unless stated; nothing here
 * allocates device memory (workspaces are passed in) except the reader's own IO buffers.
 *
 * Reference interfaces replaced (paths relative to the reference tree):
 *   reader    src/numpy_reader/cpp/numpy_reader.cpp:501-536 (pybind class numpy_reader)
 *   stage     src/deepCam/data/gpsro_dataset.py:169-170, data/cam_numpy_singlefile_dataset.py:200
 *   convs     src/deepCam/architecture/gpsro/deeplab.py (nn.Conv2d / BatchNorm2d / LeakyReLU /
 *             F.interpolate call sites, see each function)
 *   losses    src/deepCam/utils/losses.py:129-172, gpsro_train/train_gan.py:146,281
 */
#ifndef BGR_H_
#define BGR_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BGR_OK 0
#define BGR_ERR_RUNTIME -1   /* std::runtime_error in the reference -> RuntimeError */
#define BGR_ERR_INDEX -2     /* std::out_of_range in the reference  -> IndexError   */
#define BGR_ERR_CUDA -3      /* CUDA runtime / driver failure        -> RuntimeError */
#define BGR_ERR_ARG -4       /* invalid argument                     -> ValueError   */

/* dtype codes of .npy payloads (numpy_reader.cpp:94-118 supports exactly these four) */
#define BGR_I64 0
#define BGR_I32 1
#define BGR_F64 2
#define BGR_F32 3

/* layouts accepted by the stage kernel */
#define BGR_NCHW 0
#define BGR_NHWC 1

/* Reductions across thread blocks never use floating-point atomics: a kernel writes one partial row per
 * block (or split) into a caller-provided workspace and the rows are added in index order, so every
 * result of this library is independent of block scheduling (bit-identical from run to run).  Workspace
 * capacities the caller must provide: */
#define BGR_BN_MAX_ROWS 448            /* BatchNorm partial rows: fp32 [BGR_BN_MAX_ROWS][2*C] */
#define BGR_DW_WGRAD_WS_FLOATS (1 << 20) /* depthwise weight-gradient partial rows (fp32 elements) */
#define BGR_GAP_MAX_ROWS 64            /* global-average-pool partial rows: fp32 [64][N][C] */
#define BGR_HEAD_MAX_ROWS 296          /* linear-head partial rows: fp32 [296][N] */
#define BGR_LOSS_MAX_ROWS 2368         /* loss partial rows: fp32 [2368] */

const char* bgr_last_error(void);
int bgr_version(void);
/* number of CUDA kernels this library has launched in this process */
unsigned long long bgr_launch_count(void);

/* ---- .npy reader (numpy_reader.cpp:501-536) ---------------------------------------------- */

/* numpy_reader(split_axis, device) constructor (numpy_reader.h:54-95); device -1 = host. */
int bgr_reader_create(int split_axis, int device, void** handle);
int bgr_reader_destroy(void* handle);
/* NumpyReader::ParseFile (numpy_reader.cpp:58-185): header -> dtype/order/shape/strides. */
int bgr_reader_parse(void* handle, const char* filename);
/* NumpyReader::InitFile / FinalizeFile (numpy_reader.cpp:342-380). */
int bgr_reader_init_file(void* handle, const char* filename);
int bgr_reader_finalize_file(void* handle);
/* NumpyReader::SetBatchsize (numpy_reader.cpp:228-247); range errors -> BGR_ERR_INDEX. */
int bgr_reader_set_batchsize(void* handle, int64_t batch_size);
/* num_inter_threads / num_intra_threads setters (numpy_reader.h:135-159). */
int bgr_reader_set_threads(void* handle, int inter, int intra);
/* shape/strides/num_samples properties; shape and stride hold up to 8 entries. */
int bgr_reader_info(void* handle, int* ndim, int64_t* shape, int64_t* stride, int* dtype,
                    int64_t* num_samples, int* fortran_order, int64_t* payload_offset,
                    int64_t* batchsize, int* inter, int* intra);
/* getSample / getBatch (numpy_reader.cpp:434-497): read n == batchsize samples (negative ids
 * wrap) into dst, sample-major.  dst_device < 0: dst is host memory; else device memory on
 * cuda:dst_device, filled by chunked pinned-staging copies enqueued on `stream`. */
int bgr_reader_read(void* handle, const int64_t* ids, int n, void* dst, int dst_device,
                    void* stream);
int64_t bgr_reader_bytes_read(void* handle);
/* EnableP2P / DisableP2P (numpy_reader.cpp:17-56). */
int bgr_reader_enable_p2p(void);
int bgr_reader_disable_p2p(void);

/* ---- decode / standardise / crop / flip / stage (gpsro_dataset.py:169-170,
 *      cam_numpy_singlefile_dataset.py:200) ------------------------------------------------ */

/* dst[b, y, x, i] = bf16_rne(scale[i] * (float(src[b, channels[i], y0+y', x0+x']) - shift[i]))
 * with y' = flip_v ? h-1-y : y and x' = flip_h ? w-1-x : x.
 * src: device payload of B samples, dtype BGR_*, layout BGR_NCHW [B,Cf,H,W] or BGR_NHWC
 * [B,H,W,Cf].  channels: HOST array of C source-channel indices (NULL = identity); shift/scale:
 * DEVICE fp32 [C].  dst: device bf16 [B,h,w,C], or with s2d != 0 the 2x2 space-to-depth layout
 * [B,h/2,w/2,4C] with channel (py*2+px)*C + i.  C multiple of 8 (<= 64), w multiple of 4. */
int bgr_stage_normalize(const void* src, int src_dtype, int src_layout, int B, int Cf, int H, int W,
                        const int* channels, int C, const float* shift, const float* scale, int y0,
                        int x0, int h, int w, int flip_h, int flip_v, void* dst, int s2d,
                        void* stream);

/* ---- per-channel dataset statistics (src/utils/summarize_data_gpsro.py:31-50 create_token) ---- */

/* out[b][k][c] (fp64, k = 0 mean, 1 mean-square, 2 min, 3 max) over the H*W pixels of channel c of
 * sample b: one reference "token" per sample (the reference's np.load(...).astype(float64) followed
 * by np.average / np.amin / np.amax over the spatial axes).  src: device payload of B samples, dtype
 * BGR_*, layout BGR_NCHW [B,Cf,H,W] or BGR_NHWC [B,H,W,Cf].  partials: device fp64 workspace of
 * B*Cf*splits*4 values; splits (>= 1) = chunks each plane is cut into for parallelism; the partials
 * are combined in a fixed order, so the result is independent of scheduling.  B*Cf <= 65535. */
int bgr_channel_stats(const void* src, int src_dtype, int src_layout, int B, int Cf, int H, int W,
                      int splits, double* partials, double* out, void* stream);

/* ---- implicit-GEMM convolution (tcgen05 / TMEM / TMA) ------------------------------------ */

/* y[n,h,w,co] = bias[co] + sum_{t,ci} x[n, h+dy[t], w+dx[t], ci] * w[co,t,ci]   (zero outside)
 * x: bf16 NHWC view [N,H,W,Cin] with pixel stride ldx elements; w: bf16 [Cout][ntaps][ldw];
 * y: bf16 NHWC view, pixel stride ldy.  Cin, Cout, ldx, ldy multiples of 8; pointers 16-byte
 * aligned.  Covers nn.Conv2d 1x1 / 3x3 / atrous stride 1 (deeplab.py:58,81,95,182,331,363,366,
 * 622,626,630) and, with transposed+flipped weights, their data gradients.
 * The stride-2 3x3 entry conv (deeplab.py:178) runs through the same entry point as a 2x2-tap
 * stride-1 conv over fields stored as 2x2 space-to-depth blocks (see bgr_stage_normalize). */
int bgr_conv_fprop(const void* x, int N, int H, int W, int Cin, int ldx, const void* w, int ldw,
                   int Cout, int ntaps, const int* dy, const int* dx, void* y, int ldy,
                   const float* bias, float* bn_part, int* bn_rows, void* stream);
/* ldw: elements between consecutive (cout, tap) rows of w (>= Cin, multiple of 8).  TMA moves whole
 * 128-byte lines: keep ldx, ldy and ldw multiples of 64 elements where the channel count is not
 * (728 -> 768) so that no pixel row straddles a line — measured 1.65x on the 728-channel layers. */
/* Wide, deep layers (Cout tile 256 with >= 8 k-blocks, or >= 4 and several Cout tiles) run as CTA pairs
 * (tcgen05.mma.cta_group::2, clusters of two CTAs); results are bit-identical to the 1-CTA kernel.  The
 * environment variable BGR_CONV_CTAS=1 keeps every layer on the 1-CTA kernels (2 forces pairs wherever the pair
 * kernel exists); same for bgr_conv_wgrad. */
/* bn_part (optional, NULL to skip; excludes bias): fp32 [BGR_BN_MAX_ROWS][2*Cout] workspace; the conv
 * epilogue writes *bn_rows (HOST int) partial rows of [sum_m y | sum_m y^2] of the bf16-rounded outputs,
 * one per group of thread blocks, for bgr_bn_finalize to add in row order — the statistics pass of
 * the nn.BatchNorm2d that follows every such conv in the reference (deeplab.py:179,183,334,...)
 * without re-reading y (same contract as bgr_bn_stats). */

/* Launch plan of the implicit-GEMM kernels for one layer signature (host only, no device work; for tests and for
 * sizing workspaces; no counterpart in the reference).  kind 0 (bgr_conv_fprop): out8 = {Cout tile width, CTAs per
 * tile (2 = cta_group::2 pair), Cout tiles, CTA groups, pixel tiles, rounds of tiles, BatchNorm partial rows (0
 * without with_stats), grid}.  kind 1 (bgr_conv_wgrad): out8 = {Cin tile width, CTAs per tile, tiles, pixel splits,
 * chunks per split, 64-pixel chunks, grid, 1 if the slab workspace is needed}. */
int bgr_conv_plan(int kind, int N, int H, int W, int Cin, int Cout, int ntaps, int with_stats, int* out8);

/* fp32 master weight [Cout][Cin][taps] (nn.Conv2d.weight) -> the bf16 operand bgr_conv_fprop reads:
 * [Cout][taps][pitch] (transpose = 0) or, for the data gradient, [Cin][taps][pitch] with Cout
 * innermost (transpose = 1); pitch >= inner dimension, padding written as zeros. */
int bgr_weight_to_gemm(const float* w, int Cout, int Cin, int taps, int transpose, void* out, int pitch,
                       void* stream);

/* dw[co,t,ci] = sum_{n,h,w} dy_[n,h,w,co] * x[n, h+dy[t], w+dx[t], ci]   (fp32, overwritten).
 * Same views as bgr_conv_fprop.  The pixel reduction is split over thread blocks; each split writes its
 * own slab of `ws` (device, 16-byte aligned, at least bgr_conv_wgrad_workspace(...) bytes — 0 when one
 * split suffices and ws may be NULL) and the slabs are added in split order. */
size_t bgr_conv_wgrad_workspace(int N, int H, int W, int Cin, int Cout, int ntaps);
int bgr_conv_wgrad(const void* x, int N, int H, int W, int Cin, int ldx, const void* dy_, int Cout,
                   int ldy, int ntaps, const int* dy, const int* dx, float* dw, float* ws, size_t ws_bytes,
                   void* stream);

/* ---- BatchNorm (training mode) + LeakyReLU + residual (deeplab.py:100,141,179-183,334-338) --
 * Tensors are bf16 [M pixels][C] with pixel strides ld*; C multiple of 8. */

/* part: fp32 [BGR_BN_MAX_ROWS][2*C] workspace; writes *rows (HOST int) partial rows of
 * [sum_m x | sum_m x^2], one per thread block. */
int bgr_bn_stats(const void* x, long M, int C, int ld, float* part, int* rows, void* stream);
/* NO COUNTERPART IN THE REFERENCE.  bgr_bn_stats for callers that
 * cannot allocate device memory at all (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc): the partial rows go to a file-scope device array of the library (fp32
 * [BGR_BN_MAX_ROWS][2*BGR_BN_STATS_STATIC_MAX_C] = 7.0 MiB, held by every GPU that loads libbgr.so until exit) and the
 * result is the finished statistics sums[0..C) = sum_m x, sums[C..2C) = sum_m x^2 (DEVICE fp64 [2*C], rows added in
 * row order).  The call allocates nothing (no cudaMalloc / cudaFree, no library allocator).  C > the maximum is
 * rejected with BGR_ERR_ARG.  All calls share the one array: calls that overlap on the same GPU race (serialise
 * them on one stream); no state is carried from one call to the next. */
#define BGR_BN_STATS_STATIC_MAX_C 2048
int bgr_bn_stats_static(const void* x, long M, int C, int ld, double* sums, void* stream);
/* The `rows` partial rows of `part` are added in row order in fp64, then:
 * mean, invstd = 1/sqrt(var_biased+eps), a = gamma*invstd, b = beta - mean*a; running stats
 * updated with `momentum` and the unbiased variance (nn.BatchNorm2d semantics); gamma/beta/
 * running_* may be NULL; num_batches_tracked (device int64, may be NULL) is incremented by one. */
int bgr_bn_finalize(const float* part, int rows, long M, int C, const float* gamma, const float* beta, float eps,
                    float momentum, float* mean, float* invstd, float* a, float* b,
                    float* running_mean, float* running_var, long long* num_batches_tracked, void* stream);
/* out = act(a*x + b [+ res]); act != 0: LeakyReLU(slope). */
int bgr_bn_apply(const void* x, int ldx, const float* a, const float* b, const void* res, int ldr,
                 void* out, int ldo, long M, int C, int act, float slope, void* stream);
/* Backward of bgr_bn_apply o batch statistics.  g: dL/dout; g2 (optional, NULL): a second gradient
 * of the same output, added on load (the block input feeds both the separable convs and the skip /
 * residual path, deeplab.py:137-141); dyh ([M][C], pixel stride ldd) receives
 * g*act'(out), which is also dL/dres; dx (same stride) receives dL/dx; part: fp32
 * [BGR_BN_MAX_ROWS][2*C] workspace for the per-block partial rows of sum dyh and sum dyh*xhat (added in
 * row order); coef_ws: fp32 [7][C] workspace whose rows 5 and 6 return dL/dbeta and dL/dgamma. */
int bgr_bn_backward(const void* g, int ldg, const void* g2, int ldg2, const void* out, int ldo,
                    const void* x, int ldx,
                    const float* mean, const float* invstd, const float* gamma, int act, float slope,
                    long M, int C, void* dyh, void* dx, int ldd, float* part, float* coef_ws,
                    void* stream);

/* Same, for a BatchNorm(+LeakyReLU) whose activated output was never stored because the depthwise
 * conv that follows applied it while loading (bgr_dw_fwd_bn): act' comes from sign(a*x + b).
 * dyh may be NULL: pass 2 then recomputes it from g and x (which it reads anyway) instead of pass 1
 * writing and pass 2 re-reading it -- one tensor pass less in each. */
int bgr_bn_backward_recompute(const void* g, int ldg, const float* a, const float* b, const void* x, int ldx,
                              const float* mean, const float* invstd, const float* gamma, int act,
                              float slope, long M, int C, void* dyh, void* dx, int ldd, float* part,
                              float* coef_ws, void* stream);

/* bgr_bn_backward_recompute with a second gradient g2 of the same output (NULL = none), added on load by
 * both passes.  Serves every BatchNorm(+LeakyReLU) without a residual input: the stored output is not
 * read (act' from sign(a*x + b)) and, with dyh = NULL, no dyh tensor is written or re-read. */
int bgr_bn_backward_recompute_dual(const void* g, int ldg, const void* g2, int ldg2, const float* a,
                                   const float* b, const void* x, int ldx, const float* mean,
                                   const float* invstd, const float* gamma, int act, float slope, long M,
                                   int C, void* dyh, void* dx, int ldd, float* part, float* coef_ws,
                                   void* stream);

/* Backward of out = act(a*x + b [+ res]) with FROZEN statistics: nn.BatchNorm2d in eval mode (validation,
 * train_gan.py:330-398; freeze_bn) or an Identity normaliser (--disable_batchnorm, train_gan.py:125 — mean 0,
 * invstd 1, gamma NULL).  dyh = (g [+ g2])*act'(out) is also dL/dres; dx = gamma*invstd*dyh (NULL to skip:
 * dyh then is dL/dx); part / coef_ws as bgr_bn_backward (rows 5, 6 of coef_ws = dL/dbeta, dL/dgamma). */
int bgr_bn_backward_frozen(const void* g, int ldg, const void* g2, int ldg2, const void* out, int ldo,
                           const void* x, int ldx, const float* mean, const float* invstd, const float* gamma,
                           int act, float slope, long M, int C, void* dyh, void* dx, int ldd, float* part,
                           float* coef_ws, void* stream);

/* ---- depthwise 3x3, "fixed padding" (deeplab.py:56,66-72,79), stride 1|2, dilation d --------
 * w: fp32 [9][C] (tap-major); dw: fp32 [C][9] (the parameter's own [C,1,3,3] layout).  Output size ceil(H/stride) x ceil(W/stride). */
int bgr_dw_fwd(const void* x, int N, int H, int W, int C, int ldx, const float* w, int stride, int dil,
               void* y, int ldy, void* stream);
int bgr_dw_dgrad(const void* dy, int N, int H, int W, int C, int ldy, const float* w, int stride, int dil,
                 void* dx, int ldx, void* stream);
/* dw is overwritten; ws: fp32 workspace of BGR_DW_WGRAD_WS_FLOATS elements for the per-block partial rows. */
int bgr_dw_wgrad(const void* x, int N, int H, int W, int C, int ldx, const void* dy, int ldy, int stride,
                 int dil, float* dw, float* ws, void* stream);
/* Stride-1 dilation-1 depthwise conv of act(a[c]*x + b[c]) (act: 0 affine, 1 LeakyReLU(slope)) with
 * the transform applied on load and zero padding applied after it: BatchNorm2d (+LeakyReLU) followed
 * by SeparableConv2d_same.conv1 (deeplab.py:100-118,75-87) without storing the activated tensor;
 * and its weight gradient (x = the stored pre-BatchNorm tensor). */
int bgr_dw_fwd_bn(const void* x, int N, int H, int W, int C, int ldx, const float* w, const float* a,
                  const float* b, int act, float slope, void* y, int ldy, void* stream);
int bgr_dw_wgrad_bn(const void* x, int N, int H, int W, int C, int ldx, const float* a, const float* b,
                    int act, float slope, const void* dy, int ldy, float* dw, float* ws, void* stream);
/* y[n,oy,ox,:] = x[n,2oy,2ox,:] (input of the stride-2 1x1 skip convs, deeplab.py:95) + adjoint. */
int bgr_subsample2(const void* x, int N, int H, int W, int C, int ldx, void* y, int ldy, void* stream);
int bgr_subsample2_bwd(const void* dy, int N, int H, int W, int C, int ldy, void* dx, int ldx, void* stream);

/* ---- decoder / head / loss helpers -------------------------------------------------------- */
/* F.interpolate(mode='bilinear', align_corners=True) (deeplab.py:375-379,663): [N,h,w,C] ->
 * [N,H,W,C] (pixel stride ldy) or, s2d != 0, the field layout [N,H/2,W/2,4C]. */
int bgr_bilinear_fwd(const void* x, int N, int h, int w, int C, int ldx, void* y, int H, int W, int ldy,
                     int s2d, void* stream);
/* adjoint, gather form (no atomics): gx fp32 dense [N,h,w,C] is overwritten. */
int bgr_bilinear_bwd(const void* gy, int N, int h, int w, int C, int H, int W, int ldy, int s2d, float* gx,
                     void* stream);
/* nn.AdaptiveAvgPool2d(1) (deeplab.py:621): out_sums fp32 [N][C] = sum over P pixels (caller scales
 * by 1/P); ws: fp32 [BGR_GAP_MAX_ROWS][N][C].  bwd: dx[n,p,:] = g[n,:]/P (g bf16 [N][C], dx dense). */
int bgr_gap_fwd(const void* x, int N, long P, int C, int ld, float* out_sums, float* ws, void* stream);
int bgr_gap_bwd(const void* g, int N, long P, int C, void* dx, void* stream);
/* flatten + nn.Linear(K,1) (deeplab_gan.py:32-36): logits[n] += x[n,:].w (logits pre-set to
 * the bias; ws: fp32 [BGR_HEAD_MAX_ROWS][N]); bwd: dx[n,:] = g[n]*w (NULL to skip),
 * dw = sum_n g[n]*x[n,:] (NULL to skip). */
int bgr_head_fwd(const void* x, const void* w, int N, long K, float* logits, float* ws, void* stream);
int bgr_head_bwd(const float* g, const void* x, const void* w, int N, long K, void* dx, float* dw,
                 void* stream);
/* nn.L1Loss (train_gan.py:146,281) fused with its gradient: *sum = sum|a-b| (fp64),
 * grad = sign(a-b)*gscale (bf16, NULL to skip); ws: fp32 [BGR_LOSS_MAX_ROWS]. */
int bgr_l1_loss_grad(const void* a, const void* b, long n, float gscale, void* grad, double* sum,
                     float* ws, void* stream);
/* The script's regression-loss choices (train_gan.py:141-152,278-281) fused with their gradient: mode 0 =
 * nn.L1Loss, 1 = nn.SmoothL1Loss(beta), 2 = nn.MSELoss; w (optional, NULL = none): bf16 per-element weights
 * in the tensors' layout = the masks of --enable_masks (utils/losses.py:101-112 L1LossWeighted).
 * *sum = sum w*l(a-b) (fp64), grad = w*l'(a-b)*gscale[* *gscale_dev] (bf16, NULL to skip; gscale_dev: optional
 * DEVICE scalar factor, e.g. 1/(sum(w)+eps) of the normalised variant); ws: fp32 [BGR_LOSS_MAX_ROWS]. */
int bgr_reg_loss_grad(const void* a, const void* b, const void* w, long n, int mode, float beta, float gscale,
                      const float* gscale_dev, void* grad, double* sum, float* ws, void* stream);
/* field layout (bf16 [N,H/2,W/2,(py,px,c)]) <-> fp32 NCHW, the tensors the training script sees. */
int bgr_field_to_nchw(const void* field, float* nchw, int N, int C, int H, int W, void* stream);
int bgr_nchw_to_field(const float* nchw, void* field, int N, int C, int H, int W, void* stream);
/* Generator.forward's noise concat (deeplab_gan.py:88-90) in the field layout: out [N,H/2,W/2,4*Cp] =
 * per 2x2 sub-pixel the C channels of `field` [N,H/2,W/2,4*C], then the nd planes of `noise` (fp32
 * NCHW [N,nd,H,W], rounded to bf16), then zeros up to Cp (C, Cp multiples of 8, Cp >= C+nd). */
int bgr_field_append_noise(const void* field, const float* noise, void* out, int N, int C, int nd, int Cp,
                           int H, int W, void* stream);

/* ---- Deconv / Deconv1x decoders and odd-sized fields (deeplab.py:398-444,466-500; the reference's native
 *      45-channel 19x37 GPSRO data) ------------------------------------------------------------------------
 * nn.ConvTranspose2d(k3, s2, padding p, output_padding) = bgr_conv_fprop over the zero-stuffed input
 * (y[n,2i,2j,:] = x[n,i,j,:], zeros elsewhere, output size H x W >= 2h-1 x 2w-1) with tap offsets p - k;
 * bgr_zero_stuff2_bwd is the adjoint gx[n,i,j,:] = gy[n,2i,2j,:]. */
int bgr_zero_stuff2(const void* x, int N, int h, int w, int C, int ldx, void* y, int H, int W, int ldy, void* stream);
int bgr_zero_stuff2_bwd(const void* gy, int N, int h, int w, int C, int H, int W, int ldy, void* gx, int ldx,
                        void* stream);
/* nn.AvgPool2d(2, stride=1, padding=pad in {0,1}) (count_include_pad: always / 4) followed by LeakyReLU(slope)
 * when act != 0 (deeplab.py:409-411,478-479,658): [N,H,W,C] -> [N,H+2pad-1,W+2pad-1,C]; bwd is the adjoint
 * (yout = the stored forward output, needed for act'). */
int bgr_avgpool2_fwd(const void* x, int N, int H, int W, int C, int ldx, int pad, int act, float slope, void* y,
                     int ldy, void* stream);
int bgr_avgpool2_bwd(const void* gy, int ldg, const void* yout, int ldo, int N, int H, int W, int C, int pad, int act,
                     float slope, void* gx, int ldx, void* stream);
/* bf16 NHWC [N,H,W,C] <-> field [N,ceil(H/2),ceil(W/2),4C]; odd H / W are padded with a zero row / column (the
 * stride-2 pad-1 entry conv over the padded field equals the conv over the odd-sized image). */
int bgr_nhwc_to_field(const void* x, int N, int H, int W, int C, int ldx, void* field, void* stream);
int bgr_field_to_nhwc(const void* field, int N, int H, int W, int C, void* y, int ldy, void* stream);
/* fp32 NCHW payload [B,Cf,H,W] of any size -> field with Cp >= C channels per sub-pixel (multiple of 8, <= 64;
 * zero beyond C, beyond H, W): bf16(scale[i]*(src[channels[i]] - shift[i])) (shift = scale = NULL: plain cast;
 * channels HOST array or NULL) — the stage step for the 45 x 19 x 37 GPSRO files; and the way back to the
 * fp32 NCHW tensor the training script sees (first C of Cp channels, H x W real pixels). */
int bgr_stage_normalize_padded(const float* src, int B, int Cf, int H, int W, const int* channels, int C,
                               const float* shift, const float* scale, void* field, int Cp, void* stream);
int bgr_field_to_nchw_padded(const void* field, float* nchw, int B, int C, int Cp, int H, int W, void* stream);

/* Every derived weight copy of a model in one launch (what layouts.py / kernels.weight_to_gemm build one
 * parameter at a time after each optimizer step).  descs / first_block are DEVICE arrays: descriptor d owns
 * blocks [first_block[d], first_block[d+1]) of the 256-thread grid of total_blocks blocks.
 *   mode 0: bf16 dst[Cout][taps][pitch] = src[Cout][Cin][taps] (fprop operand), gx = ceil(pitch/256),
 *           blocks = Cout*gx, taps <= 9;
 *   mode 1: bf16 dst[Cin][taps][pitch] = src transposed (data-gradient operand), gx = ceil(Cin*taps/32),
 *           blocks = gx*ceil(pitch/32);
 *   mode 2: fp32 dst[9][Cout] = src[Cout][9] (depthwise taps), blocks = ceil(9*Cout/256).
 * Padding columns up to pitch are written as zeros. */
typedef struct BgrWeightDesc {
  const float* src;
  void* dst;
  int Cout, Cin, taps, pitch, mode, gx;
} BgrWeightDesc;
int bgr_weights_prepare(const BgrWeightDesc* descs, const int* first_block, int ndesc, int total_blocks,
                        void* stream);

#ifdef __cplusplus
}
#endif
#endif /* BGR_H_ */
