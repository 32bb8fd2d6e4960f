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

const char* bgr_last_error(void);
int bgr_version(void);

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

/* ---- implicit-GEMM convolution (tcgen05 / TMEM / TMA) ------------------------------------ */

/* y[n,h,w,co] = bias[co] + sum_{t,ci} x[n, h+dy[t], w+dx[t], ci] * w[co,t,ci]   (zero outside)
 * x: bf16 NHWC view [N,H,W,Cin] with pixel stride ldx elements; w: bf16 [Cout][ntaps][Cin];
 * y: bf16 NHWC view, pixel stride ldy.  Cin, Cout, ldx, ldy multiples of 8; pointers 16-byte
 * aligned.  Covers nn.Conv2d 1x1 / 3x3 / atrous stride 1 (deeplab.py:58,81,95,182,331,363,366,
 * 622,626,630) and, with transposed+flipped weights, their data gradients.
 * The stride-2 3x3 entry conv (deeplab.py:178) runs through the same entry point as a 2x2-tap
 * stride-1 conv over fields stored as 2x2 space-to-depth blocks (see bgr_stage_normalize). */
int bgr_conv_fprop(const void* x, int N, int H, int W, int Cin, int ldx, const void* w, int Cout,
                   int ntaps, const int* dy, const int* dx, void* y, int ldy, const float* bias,
                   void* stream);

/* dw[co,t,ci] += sum_{n,h,w} dy_[n,h,w,co] * x[n, h+dy[t], w+dx[t], ci]   (fp32, accumulated:
 * zero dw first).  Same views as bgr_conv_fprop. */
int bgr_conv_wgrad(const void* x, int N, int H, int W, int Cin, int ldx, const void* dy_, int Cout,
                   int ldy, int ntaps, const int* dy, const int* dx, float* dw, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* BGR_H_ */
