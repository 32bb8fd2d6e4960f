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

/* ---- implicit-GEMM convolution (tcgen05 / TMEM / TMA) ------------------------------------ */

/* y[n,h,w,co] = bias[co] + sum_{t,ci} x[n, h+dy[t], w+dx[t], ci] * w[co,t,ci]   (zero outside)
 * x: bf16 NHWC view [N,H,W,Cin] with pixel stride ldx elements; w: bf16 [Cout][ntaps][Cin];
 * y: bf16 NHWC view, pixel stride ldy.  Cin, Cout, ldx, ldy multiples of 8; pointers 16-byte
 * aligned.  Covers nn.Conv2d 1x1 / 3x3 / atrous stride 1 (deeplab.py:58,81,95,182,331,363,366,
 * 622,626,630) and, with transposed+flipped weights, their data gradients.
 * s2d != 0: x is the full-resolution [N,2H,2W,Cin/4] tensor read as 2x2 space-to-depth blocks
 * (Cin must be 64): the stride-2 3x3 entry conv (deeplab.py:178) as a 2x2-tap conv. */
int bgr_conv_fprop(const void* x, int N, int H, int W, int Cin, int ldx, const void* w, int Cout,
                   int ntaps, const int* dy, const int* dx, void* y, int ldy, const float* bias,
                   int s2d, void* stream);

/* dw[co,t,ci] += sum_{n,h,w} dy_[n,h,w,co] * x[n, h+dy[t], w+dx[t], ci]   (fp32, accumulated:
 * zero dw first).  Same views as bgr_conv_fprop. */
int bgr_conv_wgrad(const void* x, int N, int H, int W, int Cin, int ldx, const void* dy_, int Cout,
                   int ldy, int ntaps, const int* dy, const int* dx, float* dw, int s2d,
                   void* stream);

#ifdef __cplusplus
}
#endif
#endif /* BGR_H_ */
