/* stereoformer_b200 -- C ABI of the B200-native stereo-matching hot path.
 *
 * Plain pointers and sizes only; every pointer is a DEVICE pointer unless the name starts
 * with `h_` or the function name ends in `_host`.  All tensors are dense, row-major NCHW,
 * fp32 unless noted.  `stream` is a cudaStream_t passed as void*.  Every function returns
 * SF_OK (0) or an error code; sf_last_error() returns the message (thread-local).
 * Nothing here allocates device memory except the sf_ctx_* host-buffer front end.
 *
 * Each entry point names the reference interface it replaces (paths relative to the
 * Magicboomliu/StereoFormer repository root).
 */
#ifndef STEREOFORMER_B200_H
#define STEREOFORMER_B200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SF_ABI_VERSION 1
#define SF_API __attribute__((visibility("default")))

/* algo selector of sf_corr_volume_f32 */
#define SF_ALGO_AUTO 0
#define SF_ALGO_TMA 1     /* force the TMA + FFMA (cost volume) / TMA + tcgen05 (attention) kernel; errors if the shape does not fit */
#define SF_ALGO_GENERIC 2 /* shape-agnostic CUDA kernel (any W / alignment) */
#define SF_ALGO_EXACT 4   /* fp32 FFMA arithmetic (no bf16 operand split), the fastest such kernel that fits the shape: what the
                             drop-ins use under stereoformer_b200.set_precision("exact") */
#define SF_ALGO_TC 3      /* cost volume on tcgen05 (bf16 hi/mid split, fp32 TMEM accumulation); errors if the shape does not fit */

/* window modes of sf_local_window_f32 */
#define SF_WIN_FIXED 0    /* fixed_local_cost_volume.py: bounds = disp -/+ radius                  */
#define SF_WIN_VARIANCE 1 /* variance_local_cost_volume.py: bounds = disp -/+ gamma * sigma        */
#define SF_WIN_OFFSETS 2  /* Attempts/update.py:279-282: bounds = disp - lo, disp + up (tensors)   */
#define SF_WIN_ABSOLUTE 3 /* pure_learning_method.py:112: lo/up ARE the bounds                     */
#define SF_WIN_VARIANCE_BOUNDS 5 /* lo/up ARE the bounds, under the variance window's validity / clamp rules (variance_local_cost_volume.py:44-53); forward only: separates the index arithmetic from sigma's exp() rounding in the parity tests */
#define SF_WIN_HEAD 4     /* as OFFSETS, lo = OffsetPredictionHead conv2 output [B,2,H,W] before relu2 (update.py:96-105): ReLU + chunk fused; up unused */

/* scalar-division rounding (tensor / python scalar): PyTorch CPU divides, PyTorch CUDA multiplies
 * by the reciprocal.  Affects the last bit of sampling coordinates. */
#define SF_DIV_TRUE 0
#define SF_DIV_RECIP 1

SF_API const char* sf_last_error(void);
SF_API int sf_abi_version(void);
SF_API int sf_device_sm_count(void);

/* a1/a2: CostVolume('correlation').forward  models/utils/build_cost_volume.py:39-47 (groups = 1, out [B,D,H,W]);
 * build_gwc_volume  models/CrossAttentionCostVolume/GWC_Cost_Volume.py:14-24 (groups = G, out [B,G,D,H,W]). */
SF_API int sf_corr_volume_f32(const float* left, const float* right, float* out, int B, int C, int H, int W, int D,
                       int groups, int algo, void* stream);

/* a1': CostVolume('difference' | 'concat').forward  models/utils/build_cost_volume.py:21-37.
 * out [B, C or 2C, D, H, W]; bit-exact. */
SF_API int sf_shift_volume_f32(const float* left, const float* right, float* out, int B, int C, int H, int W, int D,
                        int concat, void* stream);

/* a3: DisparityEstimation(.WithProb).forward  models/utils/estimation.py:14-62 (disparity_regression,
 * models/utils/submodule.py:23-27).  disp [B,H,W]; prob [B,D,H,W] or NULL. */
SF_API int sf_soft_argmin_f32(const float* cv, float* disp, float* prob, int B, int D, int H, int W,
                              int match_similarity, void* stream);

/* a4/a5/a6: LocalCostVolume.forward  models/LocalCostVolume/Attempts/fixed_local_cost_volume.py:18-58;
 * Variance_Based_Local_Cost_Volume.forward  variance_local_cost_volume.py:34-78; the window tail of
 * DisparityUpdateDLC(WithMask).forward  Attempts/update.py:279-313 and PureLearningLocalCostVolume.forward
 * pure_learning_method.py:112-143.  `mode` is one of SF_WIN_*.  disp may be NULL: the current disparity is then
 * the soft-argmin of cv (fused a3+a4, LowCNN baseline.py:118-127) and is written to disp0_out when given.
 * Optional debug outputs [B,n_points+1,H,W]: candidates (fp32), floor / ceil indices (int32), sampled volume. */
SF_API int sf_local_window_f32(const float* cv, const float* disp, const float* lo, const float* up, float* out,
                               float* disp0_out, float* dbg_cand, int* dbg_idx_floor, int* dbg_idx_ceil,
                               float* dbg_local, int B, int D, int H, int W, int mode, float radius, float gamma,
                               int n_points, int consider_valid, int div_mode, void* stream);

/* a4 + the multi-GPU disparity gather in one kernel: sf_local_window_f32 whose result is ALSO stored
 * into every rank's gather buffer through NVLink peer mappings: h_peers is a HOST array of n_peers (<= SF_MAX_PEERS)
 * device pointers (this rank's own buffer included, e.g. torch symmetric memory), element [peer_offset + b*H*W + y*W
 * + x] of each receives the pixel.  Replaces the NCCL all-gather of trainer_DLC_GRU.py:102-106's DataParallel
 * gather for the inference path; completion is ordered by the caller (stream sync + barrier) like any P2P store. */
#define SF_MAX_PEERS 8
SF_API int sf_local_window_gather_f32(const float* cv, const float* disp, const float* lo, const float* up, float* out,
                                      int B, int D, int H, int W, int mode, float radius, float gamma, int n_points,
                                      int consider_valid, int div_mode, float* const* h_peers, int n_peers,
                                      long long peer_offset, void* stream);

/* The same with COMPLETION SEMANTICS (what an NCCL gather gives and plain peer stores do not), as a low-latency
 * flag-in-data protocol (the scheme of NCCL's LL protocol): no fences and no per-step host call on the data path.
 *
 *   slots    : on every destination rank, SF_GATHER_SLOTS slots of `slot_stride` 8-byte words {fp32 value, uint32 flag};
 *              step s of a producer stores pixel i as ONE 8-byte store {value, (uint32)(s + 1)} at
 *              peer_data[r] + (s % SF_GATHER_SLOTS) * slot_stride + peer_offset + i.  8-byte stores are single transactions, so a
 *              consumer that reads flag == s + 1 has the value of step s.
 *   signals  : every rank owns SF_SIG_WORDS 64-bit words in peer-mapped memory (zero-initialised):
 *     [SF_SIG_TICKET16 + r]     CTA tickets of this rank's window kernel, r = blockIdx % 16: a CTA's step is its ticket /
 *                               (number of CTAs with that r) -- launches of one gather all have the same grid and
 *                               kernels of a stream run one after the other
 *     [SF_SIG_WAIT_TICKET]      the same for this rank's wait kernel
 *     [SF_SIG_WAIT_CTA]         last-CTA election of the wait kernel
 *     [SF_SIG_ERROR]            set to 1 when a spin timed out (a peer never arrived / never consumed)
 *     [SF_SIG_CONSUMED + dst]   steps consumer `dst` has finished reading (st.release.sys by dst's wait kernel into
 *                               every producer's block): back-pressure, a producer reuses a slot for step s only once
 *                               every destination has consumed step s - SF_GATHER_SLOTS
 * h_peer_data / h_peer_sig / h_peer_ranks: HOST arrays over the n_peers destinations (rank 0 only = the reference's
 * DataParallel gather to device 0, trainer_DLC_GRU.py:102-106; all ranks = all-gather). */
#define SF_GATHER_SLOTS 4
#define SF_SIG_WORDS 64
#define SF_SIG_TICKET16 32
#define SF_SIG_WAIT_TICKET 2
#define SF_SIG_WAIT_CTA 3
#define SF_SIG_ERROR 4
#define SF_SIG_CONSUMED 24
SF_API int sf_local_window_gather_signal_f32(const float* cv, const float* disp, const float* lo, const float* up,
                                             float* out, int B, int D, int H, int W, int mode, float radius,
                                             float gamma, int n_points, int consider_valid, int div_mode,
                                             unsigned long long* const* h_peer_data,
                                             unsigned long long* const* h_peer_sig, const int* h_peer_ranks,
                                             int n_peers, long long peer_offset, long long slot_stride,
                                             unsigned long long* sig_local, void* stream);

/* Consumer side of the protocol above, one launch per step on a consuming rank: for step s (from its own ticket
 * counter) it polls every word of slot (s % SF_GATHER_SLOTS) until the flag is s + 1, writes the n_elems values to `out` (the
 * gathered [N*B,1,H,W] tensor, like an all-gather output), then publishes CONSUMED = s + 1 to every source
 * (h_src_sig: HOST array of the n_src sources' signal blocks).  When this kernel retires, step s is complete. */
SF_API int sf_gather_wait_f32(unsigned long long* sig_local, unsigned long long* const* h_src_sig, int n_src,
                              int my_rank, const unsigned long long* slots, long long slot_stride, float* out,
                              long long n_elems, void* stream);

/* build_cost_volume_from_volume(old_volume, sampling_candidates)  fixed_local_cost_volume.py:61-80 (defined
 * identically in 5 reference files).  cand/out [B,K,H,W]; bit-exact; idx_* optional int32 [B,K,H,W]. */
SF_API int sf_sample_volume_f32(const float* cv, const float* cand, float* out, int* idx_floor, int* idx_ceil, int B,
                                int D, int K, int H, int W, void* stream);

/* a7: disp_warp(img, disp, padding_mode)  utils/disparity_warper.py:83-106.  border != 0 -> 'border', else 'zeros'.
 * mask [B,C,H,W] or NULL.  dbg_grid / dbg_src: optional [B,H,W,2] normalised grid / pixel-space source coordinate.
 * neg_flag: optional device int set to 1 when any disparity is negative (the reference asserts, :93). */
SF_API int sf_disp_warp_f32(const float* img, const float* disp, float* warped, float* mask, float* dbg_grid,
                            float* dbg_src, int B, int C, int H, int W, int border, int div_mode, int* neg_flag,
                            void* stream);

/* a7 on a row band (partition 2: one large pair split by epipolar row bands).  The output covers rows
 * [y0, y0 + rows) of an image of full_height rows; img holds rows [src_y0, src_y0 + src_rows) and must include the
 * 1-row halo the align_corners quirk needs (rows y0-1 .. y0+rows, clipped to the image; SF_ERR_ARG otherwise).
 * Coordinates are computed from the global row index and height: the band equals the same rows of the whole-image
 * call bit for bit.  disp/warped/mask/dbg_* are [B,*,rows,W]. */
SF_API int sf_disp_warp_band_f32(const float* img, const float* disp, float* warped, float* mask, float* dbg_grid,
                                 float* dbg_src, int B, int C, int rows, int W, int full_height, int y0, int src_y0,
                                 int src_rows, int border, int div_mode, int* neg_flag, void* stream);

/* a6 (i),(iii),(iv): BasicGuidanceNet.forward up to the conv encoders  Attempts/update.py:68-86.
 * left/right [B,C,H,W] already at the disparity resolution; err [B,C,H,W]; unc [B,D,H,W]. */
SF_API int sf_guidance_f32(const float* cv, const float* disp, const float* left, const float* right, float* err,
                           float* unc, int B, int D, int C, int H, int W, int div_mode, int* neg_flag, void* stream);

/* The same on a row band: cv/disp/left/err/unc hold the band's rows, right holds [src_y0, src_y0 + src_rows)
 * (1-row halo as for sf_disp_warp_band_f32). */
SF_API int sf_guidance_band_f32(const float* cv, const float* disp, const float* left, const float* right, float* err,
                                float* unc, int B, int D, int C, int rows, int W, int full_height, int y0, int src_y0,
                                int src_rows, int div_mode, int* neg_flag, void* stream);

/* a8: CorrBlock1D.__init__ / corr  models/IterativeLocalCostVolume/corr/corr.py:55-76,103-111.
 * levels: HOST array of n_out device pointers, level l is [B*H*W1, W2 >> l]. */
SF_API int sf_corr1d_pyramid_f32(const float* f1, const float* f2, float* const* levels, int n_out, int B, int C,
                                 int H, int W1, int W2, int algo, void* stream);

/* a9: CorrBlock1D.__call__  corr.py:78-101 + bilinear_sampler  utils/utils.py:59-76.
 * levels / widths: HOST arrays; coords [B,2,H,W1]; out [B,num_levels*(2r+1),H,W1]; dbg_* optional, same shape. */
SF_API int sf_corr1d_lookup_f32(const float* const* levels, const int* widths, int num_levels, int radius,
                                const float* coords, float* out, float* dbg_x0, float* dbg_src, int B, int H, int W1,
                                int div_mode, void* stream);

/* a9 alternative: PytorchAlternateCorrBlock1D.__call__  corr/corr.py:8-51 (on-the-fly correlation, no all-pairs
 * volume): f1 [B,C,H,W1], f2 [B,C,H,W2], coords [B,2,H,W1] (x, y) -> out [B,num_levels*(2*radius+1),H,W1].
 * workspace: as many bytes as the workspace query below returns, for the x-pooled right feature maps; dbg_src:
 * optional [B,L*(2r+1),H,W1,2] un-normalised sampling coordinates.  num_levels <= 4, radius <= 4, H > 1. */
SF_API size_t sf_corr1d_alt_workspace_bytes(int B, int C, int H, int W2, int num_levels);
SF_API int sf_corr1d_alt_f32(const float* f1, const float* f2, const float* coords, float* out, float* dbg_src,
                             void* workspace, size_t workspace_bytes, int B, int C, int H, int W1, int W2,
                             int num_levels, int radius, int div_mode, void* stream);

/* a10: DCN.modulated_deform_conv_forward  deform/src/vision.cpp:8 -> deform/src/modulated_deform_conv.h:10-44 ->
 * deform/src/cuda/modulated_deform_conv_cuda.cu:19-121; mask == NULL gives DCN.deform_conv_forward
 * (deform/src/deform_conv.h:10-25).  out [B,Cout,Ho,Wo] with the reference's output-size formula. */
SF_API int sf_dcn_forward_f32(const float* input, const float* weight, const float* bias, const float* offset,
                              const float* mask, float* out, int B, int Cin, int H, int W, int Cout, int kh, int kw,
                              int stride_h, int stride_w, int pad_h, int pad_w, int dil_h, int dil_w, int group,
                              int deformable_group, void* stream);

/* a10 on the tensor cores: same operator with a caller-provided scratch buffer for the bf16 hi/mid copy of the
 * weights (sf_dcn_workspace_bytes bytes, 128-byte aligned; 0 = no tensor-core path for this shape).
 * algo: SF_ALGO_AUTO (tcgen05 implicit GEMM when >= 32 channels per group on both sides, else the exact FFMA
 * kernel), SF_ALGO_GENERIC (always the exact FFMA kernel), SF_ALGO_TC (force; errors if the shape does not fit). */
SF_API size_t sf_dcn_workspace_bytes(int Cin, int Cout, int kh, int kw, int group);
/* Workspace that also enables the channels-last gather kernel (csrc/dcn_tc2.cu: 32 or a multiple of 64 channels per group): the weights' bf16 copy + a [B][H*W][Cin] fp32 copy of the input.  sf_dcn_forward_ws_f32 picks
 * the kernel from the workspace it is given: >= this size -> channels-last gathers, >= sf_dcn_workspace_bytes -> gathers
 * from NCHW, less -> the exact FFMA kernel (SF_ALGO_AUTO) or an error (SF_ALGO_TC). */
SF_API size_t sf_dcn_workspace_bytes_ex(int B, int Cin, int H, int W, int Cout, int kh, int kw, int group,
                                        int deformable_group);
SF_API int sf_dcn_forward_ws_f32(const float* input, const float* weight, const float* bias, const float* offset,
                                 const float* mask, float* out, int B, int Cin, int H, int W, int Cout, int kh, int kw,
                                 int stride_h, int stride_w, int pad_h, int pad_w, int dil_h, int dil_w, int group,
                                 int deformable_group, void* workspace, size_t workspace_bytes, int algo, void* stream);

/* a10 without any allocation: NO COUNTERPART IN THE REFERENCE; added for
 * callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid it).  Same
 * operator, arguments and algo values as sf_dcn_forward_ws_f32, but the scratch for the bf16 hi/mid copy of the weights
 * is a file-scope device array of SF_DCN_STATIC_SCRATCH_BYTES inside the library (every GPU that loads the library
 * holds it until exit; csrc/dcn_static.cu).  Shapes whose sf_dcn_workspace_bytes exceeds that limit are rejected with
 * SF_ERR_ARG (use sf_dcn_forward_ws_f32).  Calls no cudaMalloc / cudaFree / allocating library routine.  All calls on a
 * device share the array: each call rewrites the region it uses, so nothing carries over, but calls that OVERLAP on the
 * GPU (other streams, concurrent graph replays) race on it and must be serialised by the caller. */
#define SF_DCN_STATIC_SCRATCH_BYTES (4 * 1024 * 1024)   /* 256 -> 256 channels 3x3 needs 2.25 MiB */
SF_API size_t sf_dcn_static_scratch_bytes(void);
SF_API int sf_dcn_forward_static_f32(const float* input, const float* weight, const float* bias, const float* offset,
                                     const float* mask, float* out, int B, int Cin, int H, int W, int Cout, int kh,
                                     int kw, int stride_h, int stride_w, int pad_h, int pad_w, int dil_h, int dil_w,
                                     int group, int deformable_group, int algo, void* stream);

/* f3: backward passes of the pure-PyTorch operators = the gradients autograd derives for the reference modules.
 * Any grad_* output may be NULL (skipped).
 *   sf_corr_volume_backward_f32   a1/a2  build_cost_volume.py:39-47, GWC_Cost_Volume.py:14-24 (groups as the forward)
 *   sf_soft_argmin_backward_f32   a3     estimation.py:14-62
 *   sf_local_window_backward_f32  a4/a6  fixed_local_cost_volume.py:18-80, Attempts/update.py:275-313; modes FIXED,
 *                                        OFFSETS, ABSOLUTE, HEAD (grad_lo is then [B,2,H,W], ReLU backward included);
 *                                        straight-through ceil/floor end in .long(): no gradient through indices
 *   sf_sample_volume_backward_f32 build_cost_volume_from_volume alone (cand [B,K,H,W])
 *   sf_disp_warp_backward_f32     a7     disparity_warper.py:83-106 (grad_img is zero-filled here, atomics) */
SF_API int sf_corr_volume_backward_f32(const float* left, const float* right, const float* grad_cv, float* grad_left,
                                       float* grad_right, int B, int C, int H, int W, int D, int groups, void* stream);
/* same with a kernel choice: SF_ALGO_AUTO (tcgen05 band product with the bf16 hi/mid split from 64 channels per group
 * when channels per group % 32 == 0, W % 4 == 0, D <= 65; else the fp32 FFMA kernel), SF_ALGO_GENERIC / SF_ALGO_EXACT
 * (always the FFMA kernel = sf_corr_volume_backward_f32), SF_ALGO_TC (force; errors if the shape does not fit). */
SF_API int sf_corr_volume_backward_ex_f32(const float* left, const float* right, const float* grad_cv,
                                          float* grad_left, float* grad_right, int B, int C, int H, int W, int D,
                                          int groups, int algo, void* stream);
SF_API int sf_soft_argmin_backward_f32(const float* cv, const float* grad_disp, float* grad_cv, int B, int D, int H,
                                       int W, int match_similarity, void* stream);
/* modes FIXED (radius), VARIANCE (radius = gamma: the chain through sigma = sqrt(sum_d p_d (d - disp)^2) is included),
 * OFFSETS / ABSOLUTE (lo, up given; grad_lo / grad_up produced), HEAD (lo = pre-ReLU head output [B,2,H,W]). */
SF_API int sf_local_window_backward_f32(const float* cv, const float* disp, const float* lo, const float* up,
                                        const float* grad_out, float* grad_cv, float* grad_disp, float* grad_lo,
                                        float* grad_up, int B, int D, int H, int W, int mode, float radius,
                                        int n_points, int consider_valid, int div_mode, void* stream);
SF_API int sf_sample_volume_backward_f32(const float* cv, const float* cand, const float* grad_local, float* grad_cv,
                                         float* grad_cand, int B, int D, int K, int H, int W, void* stream);
SF_API int sf_disp_warp_backward_f32(const float* img, const float* disp, const float* grad_warped, float* grad_img,
                                     float* grad_disp, int B, int C, int H, int W, int border, int div_mode,
                                     void* stream);

/* f3: backward of the f1 / f2 / a6-guidance kernels (what makes the DLC-GRU step trainable through the drop-ins).
 * Any grad_* output may be NULL.  sf_upsample_convex_backward_f32 zero-fills grad_disp (atomics). */
SF_API int sf_gru_reset_backward_f32(const float* b_pre, long long b_bstride, const float* prev_h, const float* grad_s,
                                     float* grad_b_pre, float* grad_prev_h, float* grad_input, int B, int C, int Cx,
                                     int H, int W, void* stream);
SF_API int sf_gru_blend_backward_f32(const float* z_pre, long long z_bstride, const float* g_pre, const float* prev_h,
                                     const float* grad_h, float* grad_z_pre, float* grad_g_pre, float* grad_prev_h,
                                     int B, int C, int H, int W, int softsign, void* stream);
SF_API int sf_guidance_backward_f32(const float* cv, const float* disp, const float* right, const float* grad_err,
                                    const float* grad_unc, float* grad_cv, float* grad_disp, int B, int D, int C,
                                    int H, int W, int div_mode, void* stream);
SF_API int sf_upsample_convex_backward_f32(const float* disp, const float* mask, float mask_scale,
                                           const float* grad_out, float* grad_disp, float* grad_mask, int N, int H,
                                           int W, int factor, void* stream);

/* f3: DCN backward  deform/src/cuda/modulated_deform_conv_cuda.cu:124-280 (DCNv1: deform_conv_cuda.cu, mask = NULL).
 * Every grad_* pointer may be NULL (that gradient is skipped); grad_input and grad_weight are zero-filled here and
 * accumulated with atomics (as the reference's col2im does), grad_offset / grad_mask / grad_bias are plain stores.
 * Kernels up to 36 taps (6x6). */
SF_API int sf_dcn_backward_f32(const float* input, const float* weight, const float* offset, const float* mask,
                               const float* grad_out, float* grad_input, float* grad_weight, float* grad_bias,
                               float* grad_offset, float* grad_mask, int B, int Cin, int H, int W, int Cout, int kh,
                               int kw, int stride_h, int stride_w, int pad_h, int pad_w, int dil_h, int dil_w,
                               int group, int deformable_group, void* stream);
/* The same over channels-last copies where the forward's channels-last kernel applies (csrc/dcn_bwd_tc.cu: weight
 * gradient on the tensor cores; csrc/dcn_bwd_data_tc.cu: column gradient on the tensor cores + a scatter kernel with
 * vector reductions for grad_input / grad_offset / grad_mask).  workspace = sf_dcn_backward_workspace_bytes bytes
 * (channels-last input and grad_input copies, the split weights and the column-gradient buffer [pixels][Cin/g * kh*kw];
 * 0 = shape not taken), algo = SF_ALGO_AUTO (whatever the workspace allows), SF_ALGO_GENERIC / SF_ALGO_EXACT
 * (= sf_dcn_backward_f32), SF_ALGO_TC (error if the weight gradient cannot take the tensor-core path). */
SF_API size_t sf_dcn_backward_workspace_bytes(int B, int Cin, int H, int W, int Cout, int kh, int kw, int stride_h,
                                              int stride_w, int pad_h, int pad_w, int dil_h, int dil_w, int group,
                                              int deformable_group);
SF_API int sf_dcn_backward_ws_f32(const float* input, const float* weight, const float* offset, const float* mask,
                                  const float* grad_out, float* grad_input, float* grad_weight, float* grad_bias,
                                  float* grad_offset, float* grad_mask, int B, int Cin, int H, int W, int Cout, int kh,
                                  int kw, int stride_h, int stride_w, int pad_h, int pad_w, int dil_h, int dil_w,
                                  int group, int deformable_group, void* workspace, size_t workspace_bytes, int algo,
                                  void* stream);

/* a10 / f3 in double precision: the reference dispatches its DCN over float and double
 * (modulated_deform_conv_cuda.cu:93) and its gradient tests run in double (deform/test.py:376-403).  Same arguments
 * as sf_dcn_forward_f32 / sf_dcn_backward_f32 with double pointers; direct (unfused) kernels, a verification dtype. */
SF_API int sf_dcn_forward_f64(const double* input, const double* weight, const double* bias, const double* offset,
                              const double* mask, double* out, int B, int Cin, int H, int W, int Cout, int kh, int kw,
                              int stride_h, int stride_w, int pad_h, int pad_w, int dil_h, int dil_w, int group,
                              int deformable_group, void* stream);
SF_API int sf_dcn_backward_f64(const double* input, const double* weight, const double* offset, const double* mask,
                               const double* grad_out, double* grad_input, double* grad_weight, double* grad_bias,
                               double* grad_offset, double* grad_mask, int B, int Cin, int H, int W, int Cout, int kh,
                               int kw, int stride_h, int stride_w, int pad_h, int pad_w, int dil_h, int dil_w,
                               int group, int deformable_group, void* stream);

/* a11: epipolar row cross-attention (reference file is empty; SPEC-DERIVED from north_star).  q,k,v,out are
 * bf16 [B,C,H,W]; scores optional fp32 [B,band,H,W] (-inf outside the band). */
SF_API int sf_epipolar_attn_bf16(const void* q, const void* k, const void* v, void* out, float* scores, int B, int C,
                                 int H, int W, int band, int algo, void* stream);

/* f1: upsample_convex8  models/utils/upsample.py:22-38 (factor 8) / RAFT upsample_flow (factor 4).
 * disp [N,1,H,W], mask [N,9*F*F,H,W], out [N,1,F*H,F*W]. */
SF_API int sf_upsample_convex_f32(const float* disp, const float* mask, float* out, int N, int H, int W, int factor,
                                  void* stream);
/* same with the mask multiplied by mask_scale first: `mask = .25 * self.mask(hidden_state)` of
 * Attempts/update.py:273 folded into the load (saves a read+write pass over the model's largest tensor). */
SF_API int sf_upsample_convex_scaled_f32(const float* disp, const float* mask, float mask_scale, float* out, int N,
                                         int H, int W, int factor, void* stream);

/* f1 on a row band, with the multi-GPU gather of the full-resolution rows fused into the store (SURVEY 8e partition
 * 2).  disp [N,1,H,W] / mask [N,9*F*F,H,W] hold H SOURCE rows (band + halo, >= 1 row unless at the image border);
 * only the low-resolution rows [y_first, y_first + rows) are computed; out is [N,1,rows*F,W*F].  Every result is also
 * stored to each of the n_peers (<= SF_MAX_PEERS, may be 0) device pointers of the HOST array h_peers at element
 * peer_offset + n*peer_batch_stride + (row-major index inside the band): with whole-map buffers [N,1,Hfull*F,W*F]
 * that is peer_offset = y0*F*W*F, peer_batch_stride = Hfull*F*W*F.  Completion is ordered by the caller. */
SF_API int sf_upsample_convex_gather_f32(const float* disp, const float* mask, float mask_scale, float* out, int N,
                                         int H, int W, int factor, int y_first, int rows, float* const* h_peers,
                                         int n_peers, long long peer_offset, long long peer_batch_stride, void* stream);

/* f1: upsample_simple8(disp, mode='bilinear')  models/utils/upsample.py:40-43 (the LocalCostVolume baseline's output
 * when no mask head is used, baseline.py:138): factor * F.interpolate(disp, bilinear, align_corners=True).
 * disp [N,1,H,W] -> out [N,1,factor*H,factor*W]; ATen's index / weight arithmetic op for op, no FMA contraction.
 * The backward is a deterministic gather (no atomics). */
SF_API int sf_upsample_simple_f32(const float* disp, float* out, int N, int H, int W, int factor, void* stream);
SF_API int sf_upsample_simple_backward_f32(const float* grad_out, float* grad_disp, int N, int H, int W, int factor,
                                           void* stream);

/* f2: ConvGRU.forward gate arithmetic  models/LocalCostVolume/Attempts/update.py:36-49, around the three
 * convolutions (which stay cuDNN).  b_pre / z_pre: conv_b / conv_z outputs BEFORE the sigmoid, [B,C,H,W] with batch
 * stride *_bstride elements (>= C*H*W; lets both gates live in one [B,2C,H,W] tensor); prev_h [B,C,H,W] or NULL
 * (= zeros, update.py:39); input [B,Cx,H,W].
 *   sf_gru_reset_f32:  s_out [B,C+Cx,H,W] = cat(sigmoid(b_pre) * prev_h, input)              (update.py:45-46)
 *   sf_gru_blend_f32:  h_out [B,C,H,W] = (1 - z) * prev_h + z * act(g_pre), z = sigmoid(z_pre),
 *                      act = tanh, or softsign when softsign != 0                             (update.py:31-34,47-48) */
SF_API int sf_gru_reset_f32(const float* b_pre, long long b_bstride, const float* prev_h, const float* input,
                            float* s_out, int B, int C, int Cx, int H, int W, void* stream);
SF_API int sf_gru_blend_f32(const float* z_pre, long long z_bstride, const float* g_pre, const float* prev_h,
                            float* h_out, int B, int C, int H, int W, int softsign, void* stream);

#ifdef __cplusplus
}
#endif
#endif
