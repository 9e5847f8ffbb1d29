// Parameter block + launcher of the tcgen05 DCN kernel (dcn_tc.cu), shared with the C-ABI entry in dcn.cu.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

namespace sf {

struct DcnTcParams {
  const float* in;      // [B,Cin,H,W]
  const float* bias;    // [Cout]
  const float* offset;  // [B,dg*2*T,Ho,Wo]
  const float* mask;    // [B,dg*T,Ho,Wo] or null (DCNv1)
  float* out;           // [B,Cout,Ho,Wo]
  int B, Cin, H, W, Cout, Ho, Wo;
  int kh, kw, sh, sw, ph, pw, dh, dw, G, dg;
  int K, nchunk;        // Cig*T, ceil(K/64)
  int N;                // MMA N: Cog rounded up to 16
  int NST;              // stages (B in shared memory, A in tensor memory)
  int nptile, ntiles;
  int CB, ntg;          // dcn_tc2 only: channels per chunk slot (16, 32 or 64; 64 / CB taps share a chunk), tap groups per channel block
  int NA;               // dcn_tc2 only: stages of the A (column tile) ring
  int compact;          // dcn_tc2 only: 20-byte sampling descriptors (base + step bits) instead of 32-byte ones (4 offsets)
  int tiles_x;          // dcn_tc2 only: its 128-pixel tiles are 8 rows x 16 columns, tiles_x of them per output row band
};

size_t dcn_tc_workspace_bytes(int Cin, int Cout, int kh, int kw, int group);
// Runs the tensor-core path when the shape fits and the workspace is large enough (*handled = true);
// otherwise leaves *handled = false and launches nothing.
int launch_dcn_tc(DcnTcParams p, const float* weight, void* workspace, size_t workspace_bytes, cudaStream_t stream,
                  bool* handled);

// Channels-last variant (dcn_tc2.cu): channels per group and per deformable group multiples of 64.  Its workspace also
// holds the [B][H*W][Cin] copy of the input; 0 = the shape does not take this path.
size_t dcn_tc2_workspace_bytes(int B, int Cin, int H, int W, int Cout, int kh, int kw, int group, int deformable_group);
int launch_dcn_tc2(DcnTcParams p, const float* weight, void* workspace, size_t workspace_bytes, cudaStream_t stream,
                   bool* handled);

// shared with the tensor-core weight gradient (dcn_bwd_tc.cu)
bool dcn_tc2_plan(DcnTcParams& p, size_t* smem_out, bool allow16 = false);
int launch_dcn_nhwc(const float* in, float* out, int B, int C, int HW, cudaStream_t stream);
// grad_weight += grad_out x column tile on tcgen05 (grad_weight zero-filled by the caller); *handled = false when the
// shape does not fit or the workspace (dcn_bwd_tc_workspace_bytes: the channels-last input copy) is too small
size_t dcn_bwd_tc_workspace_bytes(int B, int Cin, int H, int W, int Cout, int kh, int kw, int group, int deformable_group);
int launch_dcn_bwd_weight_tc(DcnTcParams p, const float* grad_out, float* grad_weight, void* workspace,
                             size_t workspace_bytes, cudaStream_t stream, bool have_input_copy, bool* handled);
// grad_input / grad_offset / grad_mask over channels-last copies (dcn_bwd_data_tc.cu): column gradient on tcgen05 into the
// workspace, then a scatter kernel with vector reductions; the workspace starts with the channels-last input copy
size_t dcn_bwd_data_tc_workspace_bytes(int B, int Cin, int H, int W, int Cout, int kh, int kw, int stride_h, int stride_w,
                                       int pad_h, int pad_w, int dil_h, int dil_w, int group, int deformable_group);
int launch_dcn_bwd_data_tc(DcnTcParams p, const float* weight, const float* grad_out, float* grad_input, float* grad_offset,
                           float* grad_mask, void* workspace, size_t workspace_bytes, cudaStream_t stream, bool* handled);

}  // namespace sf
