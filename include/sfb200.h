/*
 * sfb200.h -- C ABI of libsfb200.so: the B200 (sm_100a) kernels behind the drop-in
 * replacement of zc402/video-classification's SlowFast hot path.
 *
 * The reference has no FFI of its own: its hot path is `init_my_slowfast`
 * (reference model/my_slowfast.py:44-126) -> torch.nn modules -> cuDNN/cuBLAS/ATen.
 * Every entry point below names the reference module call it replaces.  A reference
 * maintainer binds these with ctypes; no torch types appear here.
 *
 * Conventions
 *   - activations: bf16, NDHWC ("channels last 3d").  A tensor argument is a
 *     (pointer, ld, coff) triple: `ld` = channels per pixel of the underlying buffer,
 *     `coff` = first channel of the view.  This is how the Fast->Slow lateral output and
 *     the Slow stage output share ONE buffer (torch.cat at my_slowfast.py:343 becomes a no-op).
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*); no hidden syncs.
 *   - return value: 0 on success, negative SfbStatus otherwise; sfb_strerror() explains.
 *   - `impl`: SFB_IMPL_TC = tcgen05/TMEM tensor-core kernels (persistent TMA-fed kernel where the geometry
 *     allows, else the register-gather kernel); SFB_IMPL_TC_GATHER = force the register-gather tcgen05 kernel;
 *     SFB_IMPL_SIMT = plain CUDA-core kernel with the same contract (on-device cross-check; never a CPU path).
 */
#ifndef SFB200_H
#define SFB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum SfbStatus {
  SFB_OK = 0,
  SFB_ERR_BAD_ARG = -1,      /* shape / alignment / null pointer rejected before any launch */
  SFB_ERR_CUDA = -2,         /* a CUDA runtime call or launch failed (see sfb_last_cuda_error) */
  SFB_ERR_UNSUPPORTED = -3,  /* valid request outside what the kernels implement */
  SFB_ERR_DRIVER = -4        /* could not resolve cuTensorMapEncodeTiled from the driver */
} SfbStatus;

enum { SFB_IMPL_TC = 0, SFB_IMPL_SIMT = 1, SFB_IMPL_TC_GATHER = 2 };

/* Geometry of one implicit-GEMM convolution pass.
 *   fprop   (transposed=0): out[m, co] = sum_{tap,c} G[in(m,tap), c] * Wp[co][tap*Cg + c]
 *                           in = out*s - p + k                      (nn.Conv3d forward)
 *   dgrad   (transposed=1): same GEMM, G = dY, in = (out + p - k)/s where divisible
 *                           (autograd of nn.Conv3d w.r.t. its input)
 * G is the gathered tensor (N,Ti,Hi,Wi,Cg), the result is (N,To,Ho,Wo,Co). */
typedef struct SfbConvGeom {
  int32_t N, Ti, Hi, Wi, Cg;
  int32_t To, Ho, Wo, Co;
  int32_t kT, kH, kW;
  int32_t sT, sH, sW;
  int32_t pT, pH, pW;
  int32_t g_ld, g_coff;
  int32_t o_ld, o_coff;
  int32_t transposed;
} SfbConvGeom;

/* Optional per-channel affine(+ReLU) applied to the gathered operand while it is staged
 * ("consumer prologue": BatchNorm3d+ReLU of the producer layer, reference
 * pytorchvideo BottleneckBlock norm_a/act_a etc.).  Padding stays exactly zero. */
typedef struct SfbPrologue {
  const float* scale;   /* [Cg] or NULL = identity */
  const float* shift;   /* [Cg] */
  int32_t relu;
} SfbPrologue;

/* Optional fused epilogue of the GEMM. */
typedef struct SfbEpilogue {
  const float* scale;    /* [Co] or NULL: y = acc*scale + shift (inference-folded BatchNorm3d) */
  const float* shift;
  const void* residual;  /* bf16 or NULL: y += residual[m, co] (ResBlock branch_fusion / grad accumulation) */
  int32_t res_ld, res_coff;
  int32_t relu;
  float* stats;          /* [2*Co] fp32 or NULL: stats[co] += sum_m y, stats[Co+co] += sum_m y*y
                            (training BatchNorm3d batch statistics, fused) */
} SfbEpilogue;

const char* sfb_strerror(int status);
const char* sfb_last_cuda_error(void);
int sfb_version(void);
/* number of kernel launches issued by this library since load (bench.py's gpu_launches) */
int64_t sfb_launch_count(void);
/* A CUDA graph captured from these entry points launches the same kernels again at every replay without passing
 * through the library: the host reports the captured graph's kernel count here per replay so the counter keeps
 * meaning "kernels of this library that ran".  Returns the new total. */
int64_t sfb_launch_count_add(int64_t n);

/* sizes of the packed bf16 weight matrix for a conv with `taps`=kT*kH*kW over Cg gathered
 * channels producing Co channels: rows are padded to the N tile, K to 64. */
int sfb_packed_weight_dims(int32_t Co, int32_t Cg, int32_t taps, int32_t* rows_pad, int32_t* k_pad);

/* replaces: nn.Conv3d.weight as consumed by cuDNN.  fp32 OIDHW (Co, Ci, taps) ->
 * bf16 [rows_pad][k_pad], k = tap*Cg + c.  transposed=0: rows=Co, c runs over Ci (zero padded to Cg).
 * transposed=1 (dgrad operand): rows = Ci, k = tap*Cg + co with Cg = padded Co. */
int sfb_pack_weight(const float* w, int32_t Co, int32_t Ci, int32_t taps, int32_t Cg,
                    int32_t transposed, void* wp, int32_t rows_pad, int32_t k_pad, void* stream);


/* Operand of the input-gradient GEMM (sfb_conv_gemm with transposed=1, impl SFB_IMPL_TC).  A convolution with stride
 * s splits its input gradient into s_T*s_H*s_W parity classes, each a stride-1 transposed conv over a subset of the
 * taps (csrc/elementwise.cuh, DgradClass); the packed matrix is [rows_pad >= Ci][k_pad] with one contiguous,
 * 64-padded K range per class (column = k0(class) + tap_in_class*Cg + co).  For stride 1 this is the layout of
 * sfb_pack_weight(transposed=1).  SFB_IMPL_SIMT / SFB_IMPL_TC_GATHER keep taking sfb_pack_weight's tap-major layout. */
int sfb_packed_weight_dims_dgrad(int32_t Ci, int32_t Cg, int32_t kT, int32_t kH, int32_t kW, int32_t sT, int32_t sH,
                                 int32_t sW, int32_t* rows_pad, int32_t* k_pad);
int sfb_pack_weight_dgrad(const float* w, int32_t Co, int32_t Ci, int32_t kT, int32_t kH, int32_t kW, int32_t sT,
                          int32_t sH, int32_t sW, int32_t Cg, void* wp, int32_t rows_pad, int32_t k_pad, void* stream);

/* replaces: train.py:136-140 (_prepare_slowfast_data permute + channel slices) feeding Conv3d.
 * fp32 tensor with arbitrary element strides (n,c,t,h,w) -> bf16 NDHWC, channels zero-padded to ld. */
int sfb_layout_input(const float* x, int64_t sn, int64_t sc, int64_t st, int64_t sh, int64_t sw,
                     int32_t N, int32_t C, int32_t T, int32_t H, int32_t W,
                     void* out, int32_t out_ld, void* stream);

/* replaces: nn.Conv3d forward (cuDNN fprop) and its input-gradient (cuDNN bwd-data), with the
 * BatchNorm3d/ReLU/residual work of pytorchvideo ResBlock fused as prologue/epilogue. */
int sfb_conv_gemm(const SfbConvGeom* geom, const void* g, const SfbPrologue* pro,
                  const void* wp, int32_t rows_pad, int32_t k_pad,
                  void* out, const SfbEpilogue* epi, int32_t impl, void* stream);

/* Input-gradient GEMM whose result is the gradient dz reaching a conv -> BatchNorm3d -> ReLU layer's activation
 * (pytorchvideo BottleneckBlock norm_a/act_a, norm_b/act_b; reference model/my_slowfast.py:116-122), with the FIRST
 * pass of that layer's BatchNorm backward fused into the epilogue.  Replaces cuDNN bwd-data followed by autograd's
 * threshold_backward and the reduction half of native_batch_norm_backward:
 *     dz        = conv_transpose(g, W)                         (as sfb_conv_gemm with geom->transposed = 1)
 *     out[m,c]  = dz[m,c] * [ y[m,c]*scale[c] + shift[c] > 0 ]  -- the MASKED gradient (bf16)
 *     dbeta[c] += sum_m out[m,c]
 *     dgamma[c]+= invstd[c] * sum_m out[m,c] * (y[m,c] - mean[c])
 * y is the raw output of the layer's convolution (same pixels and view layout as `out`); the y tile of every output
 * tile is fetched by TMA while the accumulator is converted.  sfb_bn_bwd_apply (relu = 0) then turns `out` into the
 * gradient of y.  Runs on the persistent TMA kernel only (SFB_IMPL_TC geometry rules); anything else is
 * SFB_ERR_UNSUPPORTED -- there is no fallback inside this entry point. */
typedef struct SfbBnBwdFuse {
  const void* y;            /* bf16 (ptr, ld, coff) view, dims of `out` */
  int32_t y_ld, y_coff;
  const float* scale;       /* [C] forward coefficients gamma*invstd, beta - mean*gamma*invstd (ReLU mask recompute) */
  const float* shift;
  const float* mean;        /* [C] batch mean, 1/sqrt(var + eps) */
  const float* invstd;
  float* dgamma;            /* [C] fp32, accumulated */
  float* dbeta;
  int32_t C;                /* real channels (multiple of 8, <= geom->Co: zeroed pad lanes / pixel folding) */
  int32_t fold_ld;          /* 0, or the per-pixel channel stride of a pixel-folded launch: channel = column % fold_ld */
} SfbBnBwdFuse;
int sfb_conv_dgrad_bn(const SfbConvGeom* geom, const void* g, const void* wp, int32_t rows_pad, int32_t k_pad,
                      void* out, const SfbBnBwdFuse* bn, void* stream);

/* replaces: nn.Conv3d weight gradient (cuDNN bwd-filter).
 * dw[co][ci][tap] += sum_m dy[m, co] * pro(x)[in(m,tap), ci]   (fp32, OIDHW, accumulated) */
int sfb_conv_wgrad(const SfbConvGeom* geom, const void* x, const SfbPrologue* pro,
                   const void* dy, int32_t dy_ld, int32_t dy_coff,
                   float* dw, int32_t Ci_real, int32_t impl, void* stream);



/* ---- pixel folding for narrow-channel layers (Fast pathway: 8..32 channels) ------------------------------------
 * A stride-1 conv over a dense NDHWC tensor may be run on the SAME memory viewed as (N,T,H,W/f,f*ld): f neighbouring
 * pixels become one pixel with f*ld channels (128-byte rows for TMA instead of 32/64-byte ones) and the weight
 * becomes block-banded (csrc/elementwise.cuh).  These translate between the reference layout and the folded one:
 *   sfb_fold_weight : fp32 (Co,Ci,kT,kH,kW) -> fp32 (f*ldo, f*ldi, kT, kH, kW') , kW' = 1 (kW = 1) or 3 (kW = 3);
 *                     feed the result to sfb_pack_weight / sfb_pack_weight_dgrad with Co' = f*ldo, Ci' = f*ldi
 *   sfb_unfold_wgrad: dw (Co,Ci,kT,kH,kW) += band sums of the folded weight gradient produced by sfb_conv_wgrad
 *   sfb_unfold_stats: stats[2*C] += column-group sums of the folded BatchNorm statistics [2*f*ld] */
int sfb_fold_weight(const float* w, int32_t Co, int32_t Ci, int32_t kT, int32_t kH, int32_t kW, int32_t f, int32_t ldo,
                    int32_t ldi, float* wf, void* stream);
/* Weight gradient of a pixel-folded conv written straight into the reference-layout gradient: `geom` describes the
 * FOLDED conv (what sfb_conv_wgrad would get, with dy / x viewed as f*ld channels), dw is the fp32 (Co,Ci,kT,kH,kW)
 * gradient of the real conv; the band map of sfb_fold_weight is applied in the kernel's epilogue (replaces
 * sfb_conv_wgrad into a folded buffer + sfb_unfold_wgrad). */
int sfb_conv_wgrad_folded(const SfbConvGeom* geom, const void* x, const void* dy, int32_t dy_ld, int32_t dy_coff, float* dw,
                          int32_t Co, int32_t Ci, int32_t kW, int32_t f, int32_t ldo, int32_t ldi, void* stream);
int sfb_unfold_wgrad(const float* dwf, int32_t Co, int32_t Ci, int32_t kT, int32_t kH, int32_t kW, int32_t f, int32_t ldo,
                     int32_t ldi, float* dw, void* stream);
int sfb_unfold_stats(const float* stats_folded, int32_t f, int32_t ld, int32_t C, float* stats, void* stream);

/* ---- stem convolution (1,7,7)/s(1,2,2)/p(0,3,3) in space-to-depth form -------------------------------------------
 * replaces: ResNetBasicStem.conv of both pathways (reference model/my_slowfast.py:63-68: kernel (1,7,7), stride
 * (1,2,2), padding (0,3,3); pytorchvideo create_res_basic_stem) and its weight gradient.  The clip is laid out
 * "s2d": [N][T][H/2][W/2][Cs] bf16 with channel (sh*2+sw)*C + c = input channel c of sub-pixel (sh, sw) of the 2x2
 * block, zero padded to Cs = sfb_stem_s2d_channels(C) (32 or 64).  The conv is then a stride-1 4x4 conv whose input
 * rows are staged once in shared memory and reused by every tap (csrc/stem_win.cuh).  H and W must be even,
 * W/2 <= 125, 4*C <= 64, Co <= 64; anything else: use sfb_layout_input + sfb_conv_gemm / sfb_conv_wgrad. */
int sfb_stem_s2d_channels(int32_t C);
/* replaces: train.py:136-140 (_prepare_slowfast_data) feeding the stem: fp32 strided NCTHW -> s2d bf16 */
int sfb_layout_input_s2d(const float* x, int64_t sn, int64_t sc, int64_t st, int64_t sh, int64_t sw,
                         int32_t N, int32_t C, int32_t T, int32_t H, int32_t W, void* out, int32_t Cs, void* stream);
/* ---- raw clip staging ---------------------------------------------------------------------
 * replaces, in one device pass over the uint8 frames: dataset/chalearn_dataset.py:41-46,113-118 (ToTensor: HWC uint8
 * -> fp32 / 255, Normalize: (x - mean) / std), dataset/chalearn_dataset.py:79-88 (RandomCrop(size, padding) of the
 * whole clip, zero padding in the normalised domain), train.py:136-140 (permute + channel split 0:5 / 5:20) and
 * sfb_layout_input_s2d.  frames: uint8 [N][T][H][W][Ct] (cv2 HWC frames stacked).  crop_yx: int32 device [N][2] or
 * NULL; output pixel (y, x) of clip n takes source pixel (y + crop_yx[2n], x + crop_yx[2n+1]), zero outside the
 * frame (RandomCrop's (i, j) window of the padded clip = offsets (i - padding, j - padding)).  Every pathway takes
 * channels [c0, c0+C) of frames 0, t_stride, 2*t_stride ... and receives bf16 s2d [N][T/t_stride][H/2][W/2][Cs].
 * Values are bf16_rn(((float)v / 255 - mean) / std) in the reference's fp32 operation order: bit-identical to
 * sfb_layout_input_s2d of the reference's fp32 clip. */
typedef struct {
  int32_t c0, C;     /* channel range of the raw frames */
  int32_t t_stride;  /* temporal subsampling of this pathway (T % t_stride == 0) */
  int32_t Cs;        /* s2d channels of the destination (>= 4*C) */
  void* out;
} SfbStagePathway;
int sfb_stage_clips_u8(const uint8_t* frames, int32_t N, int32_t T, int32_t H, int32_t W, int32_t Ct,
                       const int32_t* crop_yx, float mean, float std, const SfbStagePathway* pw, int32_t n_pw,
                       void* stream);
/* fp32 (Co, Ci, 1, 7, 7) -> bf16 [rows_pad][16*Cs]; rows_pad = 16 / 32 / 64 (Co rounded up) */
int sfb_pack_weight_stem(const float* w, int32_t Co, int32_t Ci, int32_t Cs, void* wp, int32_t rows_pad, void* stream);

/* Batched form of the four packing entry points above (nn.Conv3d.weight -> the bf16 operands of every conv of the
 * network, re-derived after each optimizer step): the caller describes the jobs once, sfb_pack_plan turns them into a
 * device-resident table (synchronous copy; pointers must stay valid), sfb_pack_run executes the whole table in ONE
 * launch.  kind: 0 = sfb_pack_weight (kT*kH*kW taps, Cg, transposed), 1 = sfb_pack_weight_dgrad (k, s, Cg),
 * 2 = sfb_fold_weight (k, f, ldo, ldi), 3 = sfb_pack_weight_stem (Cg = Cs).  Jobs of one table must not depend on
 * each other (pack folded weights from a second table run after the table holding the fold jobs). */
typedef struct SfbPackJob {
  int32_t kind;
  int32_t Co, Ci, kT, kH, kW, sT, sH, sW, Cg, transposed, f, ldo, ldi, rows_pad, k_pad;
  const float* src;
  void* dst;
} SfbPackJob;
int64_t sfb_pack_table_bytes(int32_t n_jobs);
int sfb_pack_plan(const SfbPackJob* jobs, int32_t n_jobs, void* table_dev, int64_t table_bytes, int32_t* total_blocks);
int sfb_pack_run(const void* table_dev, int32_t n_jobs, int32_t total_blocks, void* stream);
/* y[N,T,H2,W2,Co] (view o_ld/o_coff) = conv(x) ; stats as in SfbEpilogue.stats (may be NULL) */
int sfb_stem_conv_fprop(const void* xs2d, int32_t N, int32_t T, int32_t H2, int32_t W2, int32_t Ci, int32_t Cs,
                        const void* wp, int32_t rows_pad, void* out, int32_t o_ld, int32_t o_coff, int32_t Co,
                        float* stats, void* stream);
/* dw (fp32, (Co, Ci, 1, 7, 7), accumulated) += wgrad */
int sfb_stem_conv_wgrad(const void* xs2d, int32_t N, int32_t T, int32_t H2, int32_t W2, int32_t Ci, int32_t Cs,
                        const void* dy, int32_t dy_ld, int32_t dy_coff, int32_t Co, float* dw, void* stream);

/* replaces: nn.BatchNorm3d training statistics finalisation + running-stat update.
 * stats = [sum(C), sumsq(C)] over `count` elements per channel.
 * scale = gamma*invstd, shift = beta - mean*scale; save_mean/save_invstd kept for backward;
 * running = (1-momentum)*running + momentum*{mean, unbiased var}.  running_* may be NULL. */
int sfb_bn_finalize(const float* stats, int32_t C, int64_t count, const float* gamma, const float* beta,
                    float eps, float momentum, float* running_mean, float* running_var,
                    float* scale, float* shift, float* save_mean, float* save_invstd, void* stream);

/* eval-mode BatchNorm3d fold: scale = gamma/sqrt(running_var+eps), shift = beta - running_mean*scale */
int sfb_bn_fold(const float* gamma, const float* beta, const float* running_mean, const float* running_var,
                float eps, int32_t C, float* scale, float* shift, void* stream);

/* replaces: ResBlock tail  out = ReLU( BN_c(y) + shortcut )  (branch_fusion + activation).
 * shortcut = r*r_scale + r_shift when r_scale != NULL (branch1_conv+branch1_norm), else r. */
int sfb_bn_add_relu(const void* y, int32_t y_ld, int32_t y_coff, const float* scale, const float* shift,
                    const void* r, int32_t r_ld, int32_t r_coff, const float* r_scale, const float* r_shift,
                    void* out, int32_t o_ld, int32_t o_coff, uint8_t* mask_bits, int64_t M, int32_t C, int32_t relu,
                    void* stream);

/* Training-mode form of sfb_bn_add_relu that finalises the BatchNorm statistics itself: the coefficients are derived
 * in-kernel from the raw sums the producing conv accumulated (sfb_conv_gemm's `stats`), and scale / shift /
 * save_mean / save_invstd are stored and the running statistics updated exactly as sfb_bn_finalize would -- without
 * the extra launch per BatchNorm.  fold_f > 1: `stats` is the pixel-folded layout [2*fold_f*fold_ld] that
 * sfb_unfold_stats would otherwise translate.  r_bn == NULL: the shortcut r (if any) is added as is. */
typedef struct SfbBnRaw {
  const float* stats;
  int64_t count;
  const float* gamma;
  const float* beta;
  float eps, momentum;
  float* running_mean;   /* may be NULL */
  float* running_var;
  float* scale;          /* outputs, [C] each */
  float* shift;
  float* save_mean;
  float* save_invstd;
  int32_t fold_f, fold_ld;
} SfbBnRaw;
int sfb_bn_add_relu_raw(const void* y, int32_t y_ld, int32_t y_coff, const SfbBnRaw* y_bn,
                        const void* r, int32_t r_ld, int32_t r_coff, const SfbBnRaw* r_bn,
                        void* out, int32_t o_ld, int32_t o_coff, uint8_t* mask_bits, int64_t M, int32_t C, int32_t relu,
                        void* stream);

/* BatchNorm3d(+ReLU) backward, phase 1: per-channel reductions accumulated (atomically) into
 * dgamma/dbeta:  g = dz * [relu ? (y*scale+shift > 0) : 1] ;  dbeta += sum g ; dgamma += sum g*xhat.
 * mask_src: optional bf16 tensor whose sign gives the ReLU mask instead (block output O).
 * mask_bits: optional packed mask [M][C/8] written by sfb_bn_add_relu (bit i of byte cv = channel 8*cv+i); wins over
 * mask_src: the residual tail's backward then reads one byte per 8 channels instead of the block output. */
int sfb_bn_bwd_reduce(const void* dz, int32_t dz_ld, int32_t dz_coff,
                      const void* y, int32_t y_ld, int32_t y_coff,
                      const void* mask_src, int32_t m_ld, int32_t m_coff, const uint8_t* mask_bits,
                      const float* scale, const float* shift, const float* mean, const float* invstd,
                      int32_t relu, int64_t M, int32_t C, float* dgamma, float* dbeta, void* stream);

/* phase 2:  dy = gamma*invstd * (g - dbeta/M - xhat*dgamma/M)  -> bf16.
 * If g_out != NULL also writes g (masked dz) there (identity-shortcut gradient). */
int sfb_bn_bwd_apply(const void* dz, int32_t dz_ld, int32_t dz_coff,
                     const void* y, int32_t y_ld, int32_t y_coff,
                     const void* mask_src, int32_t m_ld, int32_t m_coff, const uint8_t* mask_bits,
                     const float* scale, const float* shift, const float* mean, const float* invstd,
                     const float* gamma, const float* dgamma, const float* dbeta,
                     int32_t relu, int64_t M, int32_t C,
                     void* dy, int32_t dy_ld, int32_t dy_coff,
                     void* g_out, int32_t g_ld, int32_t g_coff, int32_t g_accumulate, void* stream);

/* replaces: ResNetBasicStem norm+activation+pool: out = MaxPool3d((1,3,3),s(1,2,2),p(0,1,1))(ReLU(y*scale+shift)).
 * idx (uint8, may be NULL) records the arg-max tap for backward. */
int sfb_maxpool_fwd(const void* y, int32_t y_ld, int32_t y_coff, const float* scale, const float* shift,
                    int32_t N, int32_t T, int32_t H, int32_t W, int32_t C,
                    void* out, int32_t o_ld, int32_t o_coff, uint8_t* idx, void* stream);
/* dz[n,t,h,w,c] = sum over windows whose arg-max is (h,w) of dout[window]  (gradient w.r.t. the ReLU output) */
int sfb_maxpool_bwd(const void* dout, int32_t d_ld, int32_t d_coff, const uint8_t* idx,
                    int32_t N, int32_t T, int32_t H, int32_t W, int32_t C,
                    void* dz, int32_t dz_ld, int32_t dz_coff, void* stream);

/* Backward of a ResBlock tail with a projection shortcut, O = ReLU(BN_c(y_c) + BN_s(y_s)) (first block of every stage;
 * pytorchvideo ResBlock with branch1_conv / branch1_norm, reference model/my_slowfast.py:116-122): both BatchNorm
 * backward passes share the masked gradient g = dO * mask(O), so ONE reduction and ONE apply pass serve both layers
 * (replaces two sfb_bn_bwd_reduce + two sfb_bn_bwd_apply, which read dO and the mask four times).
 *     dbeta_x += sum g ; dgamma_x += invstd_x * sum g*(y_x - mean_x) ;  dy_x = gamma_x*invstd_x*(g - dbeta_x/M - xhat_x*dgamma_x/M)
 * mask_bits: packed ReLU mask of O as written by sfb_bn_add_relu(_raw).  dgamma / dbeta accumulate (zero them first). */
typedef struct SfbBnTailSide {
  const void* y;           /* raw conv output the BatchNorm normalised, bf16 (ptr, ld, coff) */
  int32_t y_ld, y_coff;
  const float* mean;       /* [C] batch mean, 1/sqrt(var + eps), BatchNorm weight */
  const float* invstd;
  const float* gamma;
  float* dgamma;           /* [C] fp32 */
  float* dbeta;
  void* dy;                /* gradient of y, bf16 (ptr, ld, coff) */
  int32_t dy_ld, dy_coff;
} SfbBnTailSide;
int sfb_bn_bwd_tail2(const void* dz, int32_t dz_ld, int32_t dz_coff, const uint8_t* mask_bits,
                     const SfbBnTailSide* c, const SfbBnTailSide* s, int64_t M, int32_t C, void* stream);

/* replaces: PoolConcatPathway AvgPool3d(k, stride 1) + ResNetBasicHead Dropout + AdaptiveAvgPool3d(1)
 * (the Linear commutes with the mean).  feat[n, f_off + c] = mean_p drop(avgpool_k(x))[n, p, c].
 * dropout_p = 0 disables dropout; otherwise a counter-based mask from `seed` (recomputed in backward).
 * seed_dev (may be NULL): device address of the 64-bit seed; when given it replaces `seed`, so that a launch captured
 * in a CUDA graph draws a fresh mask every replay (the caller rewrites the word between replays; nn.Dropout in the
 * reference, ResNetBasicHead, draws from the process's CUDA generator at every call). */
int sfb_head_pool_fwd(const void* x, int32_t x_ld, int32_t x_coff, int32_t N, int32_t T, int32_t H, int32_t W, int32_t C,
                      int32_t kT, int32_t kH, int32_t kW, float dropout_p, uint64_t seed, const uint64_t* seed_dev,
                      float* feat, int32_t f_ld, int32_t f_off, void* stream);
int sfb_head_pool_bwd(const float* dfeat, int32_t f_ld, int32_t f_off, int32_t N, int32_t T, int32_t H, int32_t W, int32_t C,
                      int32_t kT, int32_t kH, int32_t kW, float dropout_p, uint64_t seed, const uint64_t* seed_dev,
                      void* dx, int32_t dx_ld, int32_t dx_coff, void* stream);

/* replaces: ResNetBasicHead.proj (nn.Linear 2304 -> NUM_CLASS), fp32. */
int sfb_fc_fwd(const float* feat, const float* w, const float* b, int32_t N, int32_t K, int32_t C,
               float* logits, void* stream);
/* dfeat = dlogits @ w ; dw += dlogits^T @ feat ; db += sum_n dlogits */
int sfb_fc_bwd(const float* dlogits, const float* feat, const float* w, int32_t N, int32_t K, int32_t C,
               float* dfeat, float* dw, float* db, void* stream);

/* ---- after the network: evaluation drivers ------------------------------------------
 * replaces: Trainer.run_eval's host tail, reference train.py:337-360 (softmax = np.exp(x)/np.sum(np.exp(x)) per clip,
 * np.mean over the clips of a video, np.argmax) and SparseFusionTrainer.test, train_sparse.py:214-232 (softmax = 0).
 * scores: fp32 [n_clips][C]; seg: int32 device [n_videos + 1], clips seg[v] .. seg[v+1]-1 belong to video v
 * (the running sum of samples_per_video).  Outputs (each may be NULL): clip_probs [n_clips][C], video_mean
 * [n_videos][C], top1 [n_videos] (lowest index on ties, -1 for a video without clips). */
int sfb_video_scores(const float* scores, int32_t n_clips, int32_t C, const int32_t* seg, int32_t n_videos,
                     int32_t softmax, float* clip_probs, float* video_mean, int32_t* top1, void* stream);
/* replaces: SparseModel.forward, reference train_sparse.py:89-105 (num_class nn.Linear(num_part, 1) in a Python loop +
 * concat): y[n][c] = sum_p w[c][p] * x[n][p][c] + b[c].  x: [N][P][C] part scores, w: [C][P] (fcs[c].weight), b: [C]. */
int sfb_sparse_fuse(const float* x, const float* w, const float* b, int32_t N, int32_t P, int32_t C, float* y,
                    void* stream);
/* its backward (the reference trains SparseModel through autograd, train_sparse.py:160-190): dw [C][P], db [C] are
 * overwritten; dx [N][P][C] may be NULL. */
int sfb_sparse_fuse_bwd(const float* dy, const float* x, const float* w, int32_t N, int32_t P, int32_t C, float* dw,
                        float* db, float* dx, void* stream);

/* replaces: optim.Adam(model.parameters(), lr=cfg.MODEL.LR).step(), reference train.py:182 (construction) and
 * train.py:231 (the call), i.e. torch.optim.Adam with amsgrad=False, maximize=False: per parameter tensor
 *   g = grad * grad_scale (+ weight_decay * param);  m = lerp(m, g, 1-beta1);  v = beta2*v + (1-beta2)*g*g;
 *   param -= lr/(1-beta1^t) * m / (sqrt(v)/sqrt(1-beta2^t) + eps)
 * in torch's operation order and roundings (oracle/adam_oracle.py).  The caller describes every live parameter once
 * (fp32 device pointers, element counts; they must stay valid), sfb_adam_plan turns the list into a device-resident
 * table (synchronous copy) and sfb_adam_step updates ALL of them in ONE launch.  step = 1-based count of this update
 * (the bias corrections are computed on the host in double, as torch does); grad_scale = 1 for an averaged gradient,
 * 1/world_size for a summed all-reduce. */
typedef struct SfbAdamJob {
  float* param;
  const float* grad;
  float* exp_avg;
  float* exp_avg_sq;
  int64_t n;
} SfbAdamJob;
int64_t sfb_adam_table_bytes(int32_t n_jobs);
int sfb_adam_plan(const SfbAdamJob* jobs, int32_t n_jobs, void* table_dev, int64_t table_bytes, int32_t* total_blocks);
int sfb_adam_step(const void* table_dev, int32_t n_jobs, int32_t total_blocks, int64_t step, double lr, double beta1,
                  double beta2, double eps, double weight_decay, double grad_scale, void* stream);

/* No counterpart in the reference: added for callers that cannot allocate
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * BatchNorm3d(+ReLU) backward in ONE call that allocates nothing: the same two kernels as sfb_bn_bwd_reduce +
 * sfb_bn_bwd_apply (ReLU mask recomputed from y*scale+shift, no mask tensor, no shortcut-gradient output), with the
 * per-channel sums the two phases share (sum g, sum g*xhat: 2 * SFB_NOALLOC_MAX_C floats = 16 KB) in a file-scope
 * `static __device__` array of the library instead of caller memory.  Every GPU that loads the library holds that array
 * until the process exits; its address is looked up once per device (cudaGetSymbolAddress with that device current: the
 * device that owns the buffers must be current when this is called).  The call issues one cudaMemsetAsync (the array keeps
 * the previous call's sums), the two kernels and up to two device-to-device cudaMemcpyAsync on `stream`; it never calls
 * cudaMalloc / cudaFree or an allocating library routine.
 *   dy = gamma*invstd*(g - sum(g)/M - xhat*sum(g*xhat)/M),  g = dz * [relu ? (y*scale+shift > 0) : 1]   -> bf16
 *   dgamma_out / dbeta_out (fp32 [C], each may be NULL): sum g*xhat / sum g, OVERWRITTEN (not accumulated).
 * Limit: C <= SFB_NOALLOC_MAX_C (and C % 8 == 0 as everywhere); a larger C returns SFB_ERR_UNSUPPORTED and
 * sfb_last_cuda_error() names the limit.  All calls on one GPU share the array: calls that overlap on the GPU (different
 * streams, or graph branches that may run concurrently) race; order them on one stream or with events. */
#define SFB_NOALLOC_MAX_C 2048
int sfb_bn_bwd_noalloc(const void* dz, int32_t dz_ld, int32_t dz_coff, const void* y, int32_t y_ld, int32_t y_coff,
                       const float* scale, const float* shift, const float* mean, const float* invstd,
                       const float* gamma, int32_t relu, int64_t M, int32_t C, void* dy, int32_t dy_ld, int32_t dy_coff,
                       float* dgamma_out, float* dbeta_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SFB200_H */
