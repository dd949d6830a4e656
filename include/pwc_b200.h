/* pwc_b200.h -- C ABI of libpwc_b200.so: the B200 (sm_100a) PWC-Net inference hot path.
 *
 * Drop-in boundary.  Every entry point takes plain device pointers, sizes and a CUDA stream
 * handle (void* == cudaStream_t), returns a cudaError_t as int (0 == success), never throws,
 * never allocates, never synchronises.  No torch types appear here.  The reference interfaces
 * each one replaces are cited as /root/reference paths (file:line).
 *
 * Activation layout inside the library ("NHWC + pitch"): element (n,y,x,c) of a tensor lives at
 *     base[((n*H + y)*W + x) * pitch + coff + c]
 * so a layer can read/write a channel slice of a wider buffer (the DenseNet concat buffer of
 * PWCNet.py:202-264 is ONE such buffer per level; no torch.cat ever runs).
 * dtype codes: PWC_F32 = 0, PWC_BF16 = 1.
 */
#ifndef PWC_B200_H_
#define PWC_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PWC_F32 0
#define PWC_BF16 1

#define PWC_LAYOUT_NHWC 0
#define PWC_LAYOUT_NCHW 1

/* Library / build identification. Returns e.g. 100 for an sm_100a build. */
int pwc_b200_arch(void);
const char* pwc_b200_version(void);

/* ------------------------------------------------------------------------------------------
 * (1) API-compatible cost volume: replaces correlation_cuda.forward()
 *     models/correlation_package/correlation_cuda.cc:10-87 (binding),
 *     models/correlation_package/correlation_cuda_kernel.cu:46-147 (channels_first + correlation_forward)
 *     for the PWC-Net configuration pad_size = max_displacement = md, kernel_size = 1,
 *     stride1 = stride2 = 1 (PWCNet.py:71).
 *   f1, f2 : [B,C,H,W] contiguous NCHW, dtype fp32
 *   out    : [B,(2md+1)^2,H,W] contiguous NCHW fp32;
 *            out[n,(dy+md)*(2md+1)+(dx+md),y,x] = (1/C) * sum_c f1[n,c,y,x]*f2[n,c,y+dy,x+dx], 0 outside.
 *   No rInput1/rInput2 scratch is needed (the padded NHWC copies of correlation_cuda.cc:36-41 are gone).
 * ------------------------------------------------------------------------------------------ */
int pwc_corr_fwd(const void* f1, const void* f2, void* out,
                 int B, int C, int H, int W, int md, void* stream);

/* (1b) Correlation backward: replaces correlation_cuda.backward()
 *     models/correlation_package/correlation_cuda.cc:89-170 (binding),
 *     models/correlation_package/correlation_cuda_kernel.cu:150-334 (correlation_backward_input1 / _input2), 430-564 (host),
 *     same configuration as (1).  All tensors contiguous NCHW fp32; grad_out [B,(2md+1)^2,H,W];
 *       grad_f1[n,c,y,x] = (1/C) * sum_{dy,dx} grad_out[n,tc,y,x]       * f2[n,c,y+dy,x+dx]
 *       grad_f2[n,c,y,x] = (1/C) * sum_{dy,dx} grad_out[n,tc,y-dy,x-dx] * f1[n,c,y-dy,x-dx]
 *     grad_f1 or grad_f2 may be NULL (not needed).  Deterministic (no atomics), no scratch tensors. */
int pwc_corr_bwd(const float* f1, const float* f2, const float* grad_out, float* grad_f1, float* grad_f2,
                 int B, int C, int H, int W, int md, void* stream);

/* (1c) Standalone warp: replaces PWCDCNet.warp (models/PWCNet.py:141-177) for callers of that method:
 *     out[n,c,y,x] = grid_sample(x, grid + flow*flow_scale, bilinear, zeros)[n,c,y,x] * (grid_sample(ones) >= mask_thresh)
 *   x, out : [B,C,H,W] contiguous NCHW fp32; flow : [B,2,H,W] (u,v in pixels). */
int pwc_warp_fwd(const float* x, const float* flow, float* out, int B, int C, int H, int W,
                 float flow_scale, int align_corners, float mask_thresh, void* stream);

/* ------------------------------------------------------------------------------------------
 * (2) Fused warp + validity mask + cost volume + LeakyReLU: replaces, in one kernel,
 *     PWCDCNet.warp (PWCNet.py:141-177), self.corr (correlation_cuda_kernel.cu:46-147 + the three
 *     fill_ of correlation_cuda.cc:40-42) and self.leakyRELU (PWCNet.py:199,214,228,242,258).
 *   f1   : features of image 1, element (n,y,x,c) at f1[((n*H+y)*W+x)*f1_pitch + c]   (NHWC)
 *          or f1[((n*C+c)*H+y)*W+x] when layout == PWC_LAYOUT_NCHW (pitches ignored)
 *   f2   : features of image 2, same convention with f2_pitch
 *   flow : NULL (level 6: no warp, PWCNet.py:198) or the upsampled flow, element (n,y,x,k),
 *          k=0:u(x) k=1:v(y), at flow[((n*H+y)*W+x)*flow_pitch + k] (NHWC) / [((n*2+k)*H+y)*W+x] (NCHW)
 *   flow_scale  : 20/2^L (PWCNet.py:212,226,240,256)
 *   align_corners : 0 = how the reference runs under torch>=1.3 (grid_sample default); 1 = 2018 semantics
 *   mask_thresh : 0.9999f (PWCNet.py:174);  slope : 0.1f LeakyReLU (1.0f = none)
 *   out  : 81 channels, element (n,y,x,k) at out[((n*H+y)*W+x)*out_pitch + k] (NHWC)
 *          or out[((n*81+k)*H+y)*W+x] (NCHW)
 *   dtype: PWC_F32 or PWC_BF16 for f1/f2/out (flow is always fp32); arithmetic is fp32.
 * ------------------------------------------------------------------------------------------ */
int pwc_warp_corr_lrelu_fwd(const void* f1, long f1_pitch, const void* f2, long f2_pitch,
                            const float* flow, long flow_pitch, float flow_scale,
                            int align_corners, float mask_thresh, float slope,
                            void* out, long out_pitch,
                            int B, int C, int H, int W, int md,
                            int layout, int dtype, void* stream);

/* ------------------------------------------------------------------------------------------
 * (3) fp32 convolution on CUDA cores (exact-fp32 mode and the thin layers): 3x3, stride 1|2,
 *     dilation d, padding d, fused bias + LeakyReLU(slope); replaces nn.Conv2d + nn.LeakyReLU
 *     of conv() (PWCNet.py:26-30) and predict_flow() (PWCNet.py:32-33).
 *   in   : NHWC+pitch fp32; reads channels [in_coff, in_coff+Cin) (Cin, in_coff, in_pitch % 4 == 0);
 *          in2 (may be NULL) is the source for images n >= n_split (same split as the outputs below)
 *   w    : packed [9][CinP][CoutP] fp32, CinP = roundup(Cin,16), CoutP = roundup(Cout,64), zero padded
 *          (pwc_pack_conv_weight_f32 below produces it from the torch [Cout,Cin,3,3] tensor)
 *   bias : [Cout] fp32
 *   out  : NHWC+pitch fp32; images n < n_split go to out (pitch/coff), images n >= n_split go to
 *          out2 (index n - n_split) -- lets the siamese pyramid write image-1 features straight into
 *          the concat buffer and image-2 features into their own tensor. n_split >= N disables it.
 *   slope: 0.1f for conv(), 1.0f for predict_flow (no activation)
 * ------------------------------------------------------------------------------------------ */
int pwc_conv3x3_f32_fwd(const float* in, long in_pitch, int in_coff,
                        const float* in2, long in2_pitch, int in2_coff, int Cin,
                        const float* w, const float* bias,
                        float* out, long out_pitch, int out_coff,
                        float* out2, long out2_pitch, int out2_coff, int n_split,
                        int Cout, int N, int Hin, int Win, int stride, int dil, float slope,
                        const float* residual, long res_pitch,
                        void* stream);

/* Host-side helper (runs on the given stream, device pointers): torch conv weight [Cout,Cin,3,3]
 * -> [9][CinP][CoutP]; input channel i of the layer lands at packed row chan_map[i] (NULL = identity). */
int pwc_pack_conv_weight_f32(const float* w_torch, float* w_packed, const int* chan_map,
                             int Cout, int Cin, int CinP, int CoutP, void* stream);

/* ------------------------------------------------------------------------------------------
 * (4) Two-channel transposed convolution 4x4, stride 2, padding 1 (+bias, no activation):
 *     replaces deconv() (PWCNet.py:35-36) for deconvL (2->2) and upfeatL (od+448 -> 2),
 *     PWCNet.py:208-209,222-223,236-237,252-253.
 *   in  : NHWC+pitch fp32, channels [in_coff, in_coff+Cin), spatial Hin x Win
 *   w   : torch ConvTranspose2d layout [Cin,2,4,4] fp32 with chan_map applied by the caller (see packer)
 *         packed as [16 taps][2][CinP]
 *   out : NHWC+pitch fp32, 2 channels at out_coff, spatial 2Hin x 2Win; out_f32b (optional, may be NULL)
 *         receives a second copy with pitch 2 (the flow tensor the next level's warp reads).
 * ------------------------------------------------------------------------------------------ */
int pwc_deconv4x4s2_c2_f32_fwd(const float* in, long in_pitch, int in_coff, int Cin,
                               const float* w, const float* bias,
                               float* out, long out_pitch, int out_coff,
                               float* out_b, long out_b_pitch,
                               int N, int Hin, int Win, void* stream);

int pwc_pack_deconv_weight_f32(const float* w_torch, float* w_packed, const int* chan_map,
                               int Cin, int CinP, void* stream);

/* ------------------------------------------------------------------------------------------
 * (5) Input / output format kernels (the tensor the drivers build: [B,6,H,W] fp32 NCHW,
 *     inference.py:217-223, pwc_extract_flow_video.py:204-206; flow2 returned as [B,2,H/4,W/4]).
 *   pwc_pack_input : x[B,6,H,W] -> NHWC [2B,H,W,4] (channel 3 = 0); images 0..B-1 = im1, B..2B-1 = im2
 *   pwc_unpack_flow: NHWC [B,h,w,pitch] channels [coff,coff+2) -> NCHW [B,2,h,w], times `scale`
 * ------------------------------------------------------------------------------------------ */
int pwc_pack_input_f32(const float* x, float* out, int B, int H, int W, void* stream);
int pwc_unpack_flow_f32(const float* in, long in_pitch, int in_coff, float* out,
                        int B, int H, int W, float scale, void* stream);

/* ------------------------------------------------------------------------------------------
 * (6) bf16 tensor-core convolution (tcgen05.mma + TMEM + TMA implicit GEMM), 3x3, stride 1|2, dilation d,
 *     padding d, fused bias + LeakyReLU; same role as (3) for the bf16-conv mode (conv(), PWCNet.py:26-30;
 *     decoder PWCNet.py:202-264; context net PWCNet.py:126-132,267-268; predict_flow PWCNet.py:32-33).
 *   A descriptor object holds the TMA tensor maps (they embed the buffer addresses), so it is created once per
 *   (layer, buffers, shape) -- host memory only -- and launched many times.  create/destroy are host calls;
 *   launch enqueues one kernel on `stream`.
 *   in/in2 : bf16 NHWC+pitch, channels [coff, coff+cin_span); cin_span % kc == 0, kc in {16,32,64};
 *            pitch and coff multiples of 8.  in2 (may be NULL) serves images n >= n_split.
 *   w      : packed bf16 [9][cin_span/kc][coutp][kc] (pwc_pack_conv_weight_bf16), coutp % 16 == 0, <= 256
 *   out    : out_f32 == 0: bf16 NHWC+pitch slice (all coutp channels are written; pad channels get 0);
 *            out_f32 == 1: fp32, channels < cout only, plus optional fp32 residual (dc_conv7 + flow2, PWCNet.py:268)
 *            out_f32 == 2: fused predict_flowL + upfeatL (PWCNet.py:207-209...): cout = 10; channels 0,1 -> out (fp32 flow),
 *                          channel 2+2*(2a+b)+j -> out2 (bf16) at pixel (2y+a, 2x+b), channel out2_coff + j, spatial 2H x 2W
 *            out_f32 | 16: tap-stacked weights for layers with 9*cout <= 128 (predict_flowL, dc_conv7): coutp = 128 and
 *                          w_packed comes from pwc_pack_conv_weight_bf16 applied to the [128,Cin,3,3] tensor whose row
 *                          tap*cout + co holds W[co,:,tap/3,tap%3] at kernel position (0,0) (zeros elsewhere); the kernel
 *                          runs ONE MMA group per 64-channel chunk on the unshifted halo tile and sums the 9 taps in the
 *                          epilogue.  stride 1, dilation 1, fp32 output modes only.
 * ------------------------------------------------------------------------------------------ */
typedef struct pwc_tc_conv pwc_tc_conv;
int pwc_tc_conv_create(pwc_tc_conv** handle,
                       const void* in, long in_pitch, int in_coff,
                       const void* in2, long in2_pitch, int in2_coff, int n_split, int cin_span, int kc,
                       const void* w_packed, const float* bias, int cout, int coutp,
                       void* out, long out_pitch, int out_coff,
                       void* out2, long out2_pitch, int out2_coff, int out_f32,
                       const float* residual, long res_pitch,
                       int N, int Hin, int Win, int stride, int dil, float slope);
int pwc_tc_conv_launch(const pwc_tc_conv* handle, void* stream);
/* Optional tuning of a created handle (space-to-depth execution of Cout = 32 / 64 layers: g = 128/Cout adjacent pixels
 * of a dense NHWC buffer are one super-pixel, so all 128 MMA rows carry outputs):
 *   set_skip       : bit ch of masks[tap] (tap = 3*ky + kx, 9 entries) set <=> the packed weight block of that tap and
 *                    64-channel chunk ch is all zero; its TMA load and MMAs are dropped.
 *   set_out_groups : the CoutP output channels are g groups of CoutP/g; group i goes to channel out_coff + i*group_pitch.
 *   set_kk_skip    : single-chunk layers (Cin span <= 64): bit kk (0..3) of masks[tap] set <=> the 16-channel K step kk of
 *                    that tap's weight tile is all zero; its MMA is dropped. */
int pwc_tc_conv_set_skip(pwc_tc_conv* handle, const unsigned long long* masks);
int pwc_tc_conv_set_kk_skip(pwc_tc_conv* handle, const int* masks);
int pwc_tc_conv_set_out_groups(pwc_tc_conv* handle, int g, long group_pitch);
void pwc_tc_conv_destroy(pwc_tc_conv* handle);
/* kernels ONE pwc_tc_conv_launch of the handle enqueues: 1, or 2 for small layers that run as deterministic two-pass split-K
 * (conv kernel storing per-CTA fp32 partial tiles to handle-owned scratch + a reduce kernel); for launch accounting. */
/* Two layers that read the same input in one pass.  A three-lane tap-lane layer (Cout <= 32, out_f32 | 32) leaves MMA rows 96..127
 * idle: pwc_tc_conv_set_head_rows makes them `nhead` (<= 10) output channels of another layer -- packed weight rows 96 + ky*nhead + h =
 * kernel row ky of channel h -- whose raw sums (no bias, no activation) are stored as fp32 [N*H*W][nhead] to head_out;
 * pwc_tc_conv_set_partial makes a tap-stacked layer (out_f32 | 16) add such a tensor to its outputs before the activation.  Used for
 * predict_flow2 and predict_flow3 + upfeat3 (PWCNet.py:264-266, 252-255): their 533 shared input channels ride along with conv2_4 /
 * conv3_4, the head itself reads the remaining 32.  No counterpart in the reference (it runs every layer as its own cuDNN call). */
int pwc_tc_conv_set_head_rows(pwc_tc_conv* h, float* head_out, int nhead);
int pwc_tc_conv_set_partial(pwc_tc_conv* h, const float* partial, int pitch);
int pwc_tc_conv_num_launches(const pwc_tc_conv* handle);
/* Split-K launch of a created handle that ALLOCATES NOTHING.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  The split-K form of the tcgen05 kernel (`ksplit` CTAs share one pixel tile, each a range of
 * 64-channel chunks; fp32 partial sums + per-tile arrival counters in scratch memory) normally gets its scratch from
 * cudaMalloc inside pwc_tc_conv_create (opt-in, PWC_TC_WM_KSPLIT=1).  This entry point runs the same kernel on ANY
 * eligible handle with the scratch in a file-scope `static __device__` array of 16 MiB (held by every GPU that loads the
 * library until exit; device address from cudaGetSymbolAddress, once per device).  Inputs, weights and outputs are the
 * ones the handle was created with; the handle is not modified.
 *   ksplit : CTAs per pixel tile, 2..8 (clamped to the chunk count); 0 = as many as fill the SMs.
 *   eligible: stride 1, coutp <= 128 and % 32 == 0, bf16 output (out_f32 == 0), >= 2 channel chunks, no tap stacking.
 *   errors : cudaErrorInvalidValue (and a message on stderr) for an ineligible handle or when
 *            tiles * tile_pixels * 128 * 4 + tiles * 4 bytes exceed the 16 MiB array.
 *   Calls no cudaMalloc / cudaFree / allocating library routine; enqueues one cudaMemsetAsync of the used prefix of the
 *   array (sums and counters start at zero) and one kernel on `stream`.  Every call on a GPU shares the array: calls
 *   that overlap on that GPU (different streams, no ordering) race.  The atomic summation order varies, so results may
 *   differ from pwc_tc_conv_launch in the last bit of the bf16 output. */
int pwc_tc_conv_launch_static_splitk(const pwc_tc_conv* handle, int ksplit, void* stream);
/* ------------------------------------------------------------------------------------------
 * Split precision (conv_mode 'mixed').  An fp32-grade activation tensor is stored as TWO 16-bit planes of the same pixel
 * row: hi = RN16(v) at channel c, lo = RN16(v - hi) at channel c + lo_off.  A convolution over such a tensor issues three
 * tensor-core products per K step into the same fp32 accumulator, x_hi*W_hi + x_hi*W_lo + x_lo*W_hi (W_hi = RN16(W),
 * W_lo = RN16(W - W_hi); the dropped x_lo*W_lo term is 2^-22 relative), i.e. the convolution of the fp32 reference
 * (conv(), PWCNet.py:26-30) to ~2^-21 -- the same technique conv1a uses on its fp32 input.  No counterpart in the reference.
 *   pwc_tc_conv_create_split : pwc_tc_conv_create plus
 *       in_lo_off  != 0 : the input is split; w_packed then holds THREE 64-channel weight chunks per logical chunk ch, in
 *                         the order 3ch: W_hi (for x_hi), 3ch+1: W_lo (for x_hi), 3ch+2: W_hi (for x_lo): pack a virtual
 *                         [Cout, 3*Cin] tensor with cin_span = 3*kc*ceil(span/kc).  cin_span here stays the logical span.
 *       in_lo_off  <  0 : split ACTIVATIONS only, lo plane at -in_lo_off: two chunks per logical chunk, 2ch: W_hi (for x_hi),
 *                         2ch+1: W_hi (for x_lo) -- weights in single precision (the level 5/4 decoders: costs 0.002 px).
 *       out_lo_off != 0 : the 16-bit output is stored as both planes (out_coff + c and out_coff + out_lo_off + c).
 *       up_lo_off / up_dup_off (out_f32 == 2 only): the fused upfeat outputs also go to out2_coff + up_lo_off + j (lo plane)
 *                         and out2_coff + up_dup_off + j (second copy of hi; see pwc_deconv4x4s2_c2_split_fwd).
 *     stride-1 dilation-1 layers with coutp <= 128 and the stride-2 layers (kc | cin_span) are supported.
 *   pwc_conv3x3_thin_split_fwd : the thin pyramid layers (cin, cout in {16,32}, stride 1); in / out pixels are
 *       [hi(C) | lo(C)]; w_packed rows are [W_hi(cin) | W_lo(cin)] (pack a [Cout, 2*cin] tensor with kc = span = 2*cin).
 *   pwc_conv1a_split_fwd : conv1a (PWCNet.py:52,184) with out [N, H/2, W/2, 32] = [hi 16 | lo 16].
 *   pwc_deconv4x4s2_c2_split_fwd : deconvL with hi / lo (/ second hi) copies of its 2 outputs in a 16-bit slice.  A
 *       single-precision consumer gets split accuracy for these channels by treating lo and the second hi copy as extra
 *       input channels with weights W_hi and W_lo ("K padding": the concat buffers have 11 spare channels).
 *   pwc_copy_channels_h16 : 16-bit channel-slice copy (hi plane of the image-1 features -> c1 slot of the concat buffer).
 * ------------------------------------------------------------------------------------------ */
int pwc_tc_conv_create_split(pwc_tc_conv** handle,
                             const void* in, long in_pitch, int in_coff,
                             const void* in2, long in2_pitch, int in2_coff, int n_split, int cin_span, int kc,
                             const void* w_packed, const float* bias, int cout, int coutp,
                             void* out, long out_pitch, int out_coff,
                             void* out2, long out2_pitch, int out2_coff, int out_f32,
                             const float* residual, long res_pitch,
                             int N, int Hin, int Win, int stride, int dil, float slope,
                             int in_lo_off, int out_lo_off, int up_lo_off, int up_dup_off);
int pwc_conv3x3_thin_split_fwd(const void* in, long in_pitch, int in_coff, int cin,
                               const void* w_packed, const float* bias, int cout,
                               void* out, long out_pitch, int out_coff,
                               int N, int H, int W, float slope, void* stream);
/* the same layer at stride 1 or 2 (stride 2: cin = 16 only -- conv2a of the 'mixed' plan, 16 -> 32 channels, PWCNet.py:54,186 --
 * wider pixels do not fit two input patches in shared memory: cudaErrorInvalidValue; H, W are the INPUT dims) */
int pwc_conv3x3_thin_split_strided_fwd(const void* in, long in_pitch, int in_coff, int cin,
                                       const void* w_packed, const float* bias, int cout,
                                       void* out, long out_pitch, int out_coff,
                                       int N, int H, int W, int stride, float slope, void* stream);
int pwc_conv1a_split_fwd(const float* x, const float* w, const float* bias, void* out,
                         int B, int H, int W, float slope, int frames, void* stream);
int pwc_deconv4x4s2_c2_split_fwd(const void* in, int in_dtype, long in_pitch, int in_coff, int Cin,
                                 const float* w, const float* bias,
                                 void* out, int out_dtype, long out_pitch, int out_coff,
                                 float* out_b, long out_b_pitch,
                                 int N, int Hin, int Win, int lo_off, int dup_off, void* stream);
int pwc_copy_channels_h16(const void* src, long src_pitch, int src_coff, void* dst, long dst_pitch, int dst_coff,
                          int C, long npix, void* stream);

/* 16-bit storage format of the tensor-core conv mode.  Every entry point that takes or produces 16-bit tensors (the
 * pwc_tc_conv_* family, pwc_conv3x3_mma_fwd, pwc_conv3x3_thin_fwd, pwc_conv3x3_flow_head_fwd, pwc_conv1a_*_bf16_fwd,
 * pwc_pack_conv_weight_bf16, and dtype == PWC_BF16 of pwc_warp_corr_lrelu_fwd / pwc_deconv4x4s2_c2_fwd) reads it at
 * create / launch time: 0 = bfloat16 (default), 1 = IEEE half (conv_mode 'f16': 3 more mantissa bits; same kernels, same speed;
 * ~7x lower end-to-end flow error).  Thread-local; a created pwc_tc_conv handle keeps the format it was created
 * with; captured CUDA graphs keep what was current at capture.  No counterpart in the reference (it computes in fp32). */
int pwc_set_half_format(int fmt);
int pwc_get_half_format(void);
/* debug aid: timeline of CTA 0 of the halo kernel (enabled by env PWC_TC_HALO_BO bit 16); host memory destination */
int pwc_debug_halo_trace(unsigned long long* host_dst, int count);
int pwc_debug_corr_trace(unsigned long long* host_dst, int count);
int pwc_debug_wm_trace(unsigned long long* host_dst, int count);   /* PWC_TC_HALO_BO bit 5 (32) enables the marks */
int pwc_pack_conv_weight_bf16(const float* w_torch, void* w_packed, const int* chan_map,
                              int Cout, int Cin, int cin_span, int kc, int coutp, void* stream);

/* ------------------------------------------------------------------------------------------
 * (7) bf16-mode companions (fp32 arithmetic, bf16 storage).
 *   pwc_conv1a_bf16_fwd : first pyramid layer conv1a (3->16, stride 2, +LeakyReLU; PWCNet.py:52,184,190) fused with
 *       the im1/im2 split of PWCNet.py:181-182.  x [B,6,H,W] fp32 NCHW (the driver tensor), w torch [16,3,3,3] fp32,
 *       out bf16 NHWC [2B,H/2,W/2,16]; images 0..B-1 = im1 of each pair, B..2B-1 = im2.
 *   pwc_deconv4x4s2_c2_fwd : (4) with selectable input / output dtype (PWC_F32 | PWC_BF16); out_b stays fp32.
 * ------------------------------------------------------------------------------------------ */
int pwc_conv1a_bf16_fwd(const float* x, const float* w, const float* bias, void* out,
                        int B, int H, int W, float slope, void* stream);
int pwc_deconv4x4s2_c2_fwd(const void* in, int in_dtype, long in_pitch, int in_coff, int Cin,
                           const float* w, const float* bias,
                           void* out, int out_dtype, long out_pitch, int out_coff,
                           float* out_b, long out_b_pitch,
                           int N, int Hin, int Win, void* stream);

/* ------------------------------------------------------------------------------------------
 * (7b) 3x3 convolution for the small pyramid levels (5, 6): same operator as (6) -- conv(), PWCNet.py:26-30, call sites
 *      PWCNet.py:186-195 (conv5a..conv6b) and PWCNet.py:202-206,217-221 (conv6_0..4, conv5_0..4) -- cut three ways
 *      (pixel tile x output-channel slice x K split inside the CTA) on warp-level bf16 MMA so that layers of a few
 *      thousand pixels fill the machine; stateless (no handle, no tensor maps).
 *   in / in2 : bf16 NHWC + pitch; images >= n_split_in read in2 (NULL: one input).  Channels [coff, coff + cin_span).
 *   w_packed : pwc_pack_conv_weight_bf16(..., kc = 64, coutp): [9][cin_span/64][coutp][64]; coutp % 16 == 0
 *   out/out2 : bf16 NHWC + pitch, channels [coff, coff + coutp); images >= n_split go to out2 (NULL: one output)
 *   stride 1 | 2, padding 1; bias + LeakyReLU(slope) fused, 0 <= slope <= 1.
 * ------------------------------------------------------------------------------------------ */
int pwc_conv3x3_mma_fwd(const void* in, long in_pitch, int in_coff,
                        const void* in2, long in2_pitch, int in2_coff, int n_split_in, int cin_span,
                        const void* w_packed, const float* bias, int cout, int coutp,
                        void* out, long out_pitch, int out_coff,
                        void* out2, long out2_pitch, int out2_coff, int n_split,
                        int N, int H, int W, int stride, float slope, void* stream);

/* ------------------------------------------------------------------------------------------
 * (7c) 3x3 convolution for the thin feature-pyramid layers (Cin, Cout in {16, 32, 64}): conv1aa, conv1b, conv2a, conv2aa,
 *      conv2b, conv3a, conv3aa, conv3b (PWCNet.py:52-60, call sites PWCNet.py:184-195).  HBM bound: persistent CTAs,
 *      weights resident in shared memory, double-buffered cp.async input patches, warp-level bf16 MMA, 16-byte stores.
 *   in / in2 : bf16 NHWC + pitch, channels [coff, coff + cin); images >= n_split_in read in2 (NULL: one input)
 *   w_packed : pwc_pack_conv_weight_bf16(..., cin_span = cin, kc = cin, coutp = cout): [9][cout][cin] bf16
 *   out/out2 : bf16 NHWC + pitch, channels [coff, coff + cout); images >= n_split go to out2 (NULL: one output);
 *              pitches and offsets multiples of 8 channels, bases 16-byte aligned
 *   stride 1 | 2, padding 1; bias + LeakyReLU(slope) fused, 0 <= slope <= 1.
 * ------------------------------------------------------------------------------------------ */
int pwc_conv3x3_thin_fwd(const void* in, long in_pitch, int in_coff,
                         const void* in2, long in2_pitch, int in2_coff, int n_split_in, int cin,
                         const void* w_packed, const float* bias, int cout,
                         void* out, long out_pitch, int out_coff,
                         void* out2, long out2_pitch, int out2_coff, int n_split,
                         int N, int H, int W, int stride, float slope, void* stream);

/* Flow head on the same kernel: dc_conv7 (3x3, Cin -> 2, no activation; PWCNet.py:132) fused with the residual add
 * `flow2 + dc_conv7(x)` (PWCNet.py:268) and the NHWC -> NCHW unpack: out_nchw[N,2,H,W] fp32 = conv + bias2 + residual[...,0:2].
 *   w_packed16 : pwc_pack_conv_weight_bf16 of the [2,Cin,3,3] weight zero-padded to 16 output rows (kc = cin, coutp = 16)
 *   residual   : fp32 [N,H,W,res_pitch], channels 0,1 */
int pwc_conv3x3_flow_head_fwd(const void* in, long in_pitch, int in_coff, int cin,
                              const void* w_packed16, const float* bias2, const float* residual, long res_pitch,
                              float* out_nchw, int N, int H, int W, void* stream);

/* ------------------------------------------------------------------------------------------
 * (8) Driver-side pre/post-processing and metrics on the device.
 *   pwc_preprocess_u8_fwd : frame_to_tensor + pad_to_multiple_of_64 + torch.cat([t1p, t2p], 1)
 *       (pwc_extract_flow_video.py:27-42,197-205; inference_kitti.py:53-63,216-217; script_pwc.py:61-71).
 *       im1, im2 : B uint8 frames [H][W][3] each (interleaved, as cv2 / imageio deliver them), `row_stride` bytes
 *                  between rows, `img_stride` bytes between frames; swap_rb = 1 reverses the channel order
 *                  (cv2.COLOR_BGR2RGB / the [:, :, ::-1] of script_pwc.py:63).
 *       out      : [B,6,Hp,Wp] fp32 NCHW = value / 255 (correctly rounded fp32 division), channels 0-2 frame 1,
 *                  3-5 frame 2, replicate-padded at the bottom / right to Hp x Wp (Wp % 4 == 0).
 *   pwc_flow_postprocess_fwd : unpad + flow_resize + scale (inference_kitti.py:66-91,219-223; script_pwc.py:72-81).
 *       flow : [B,2,h,w] fp32 NCHW; its top-left crop_h x crop_w window (the "unpadded" flow) is resized to H x W with
 *              bilinear interpolation, align_corners = True (F.interpolate), u scaled by W/crop_w, v by H/crop_h, both by
 *              `mult` (20.0 for pwc_net.pth.tar flows, script_pwc.py:72).
 *       out  : [B,2,H,W] NCHW, or [B,H,W,2] when out_hwc = 1 (the [H,W,2] array the .flo / KITTI writers take).
 *   pwc_flow_metrics_fwd : EPE and Fl-all sums over valid pixels (inference_kitti.py:94-128, inference.py:105-159).
 *       pred, gt : [B,2,H,W] (layout_hwc = 0) or [B,H,W,2] (1) fp32; valid : [B,H,W] uint8 or NULL (all valid).
 *       acc      : [B][3] fp64 (zeroed here): sum of EPE over valid pixels, #valid, #outliers with
 *                  EPE > max(abs_thr, rel_thr * |gt|) (3 px / 5% for KITTI).  EPE = acc[0]/acc[1], Fl = 100*acc[2]/acc[1].
 * ------------------------------------------------------------------------------------------ */
int pwc_preprocess_u8_fwd(const uint8_t* im1, const uint8_t* im2, long img_stride, long row_stride,
                          int B, int H, int W, int Hp, int Wp, int swap_rb, float* out, void* stream);
/* frame-stack variants for consecutive pairs of one video (pair k = frames k, k+1: every frame is converted and run through
 * the feature pyramid ONCE): N uint8 frames -> [N,3,Hp,Wp] fp32, and conv1a on such a stack (N frames -> [N,H/2,W/2,16] bf16). */
/* ImageNet-normalised variant of the KITTI drivers (T.ToTensor + T.Normalize, inference_kitti.py:175-179, inference.py:305-310):
 * out = (byte / 255 - mean[c]) / std[c] per output channel c (mean3 / std3: 3 HOST floats), then replicate pad + cat. */
int pwc_preprocess_u8_norm_fwd(const uint8_t* im1, const uint8_t* im2, long img_stride, long row_stride,
                               int B, int H, int W, int Hp, int Wp, int swap_rb, const float* mean3, const float* std3,
                               float* out, void* stream);
int pwc_preprocess_frames_u8_fwd(const uint8_t* frames, long img_stride, long row_stride,
                                 int N, int H, int W, int Hp, int Wp, int swap_rb, float* out, void* stream);
int pwc_conv1a_frames_bf16_fwd(const float* frames, const float* w, const float* bias, void* out,
                               int N, int H, int W, float slope, void* stream);
int pwc_flow_postprocess_fwd(const float* flow, int B, int h, int w, int crop_h, int crop_w,
                             float* out, int H, int W, float mult, int out_hwc, void* stream);
int pwc_flow_metrics_fwd(const float* pred, const float* gt, const uint8_t* valid, int B, int H, int W,
                         int layout_hwc, float abs_thr, float rel_thr, double* acc, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PWC_B200_H_ */
