// Internal (C++) launch wrappers of the irb200 kernels. All take channels-last activations with an explicit
// row stride (ld, in elements) so that zero-copy concatenation / channel slicing works, and a cudaStream_t.
// Return 0 on success; on failure the message is available through irb_last_error().
#pragma once
#include "common.cuh"
#include <cuda.h>

namespace irb {

// ---- gemm_tc.cu : tcgen05 (bf16) -------------------------------------------------------------
int make_tmap_bf16(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                   const uint32_t* box);
int make_tmap_bf16_sw(CUtensorMap* m, const void* base, int rank, const uint64_t* dims,
                      const uint64_t* strides_bytes, const uint32_t* box, int swizzle_bytes);
// true when tc_gemm / tc_conv3x3 can also emit the LayerNorm statistics of their output rows (Epi::stats_out)
bool tc_can_fuse_stats(int N, const Epi& epi);
int tc_gemm(const bf16* A, long long lda, int M, int K, int batch, long long a_batch_stride, const bf16* Wt, int ldw,
            int N, int w_rows_total, int w_rows_per_batch, const Epi& epi, cudaStream_t stream);
int tc_conv3x3(const bf16* A, long long lda, int Nimg, int H, int W, int C, const bf16* Wt, int kpad, int N,
               int w_rows_total, const Epi& epi, cudaStream_t stream);

// ---- conv_small.cu : 3x3 convolutions with N <= 4 output channels as a shared-memory stencil (the reconstruction
// convolutions at the end of both networks); conv3x3_bf16 picks between it and tc_conv3x3 ----
bool conv3x3_small_n_ok(int C, int N, const Epi& epi);
int conv3x3_small_n(const bf16* x, long long ldx, int Nimg, int H, int W, int C, const bf16* Wt, int kpad, int N,
                    const Epi& epi, cudaStream_t stream);
int conv3x3_bf16(const bf16* x, long long ldx, int Nimg, int H, int W, int C, const bf16* Wt, int kpad, int N,
                 int w_rows_total, const Epi& epi, cudaStream_t stream);

// ---- gemm_simt.cu : fp32 FMA path (fp32 precision mode, and tiny-channel convolutions) ------
// A element type a_dt (DT_F32/DT_BF16); weights always float [Nrows, taps*kpad].
int simt_gemm(const void* A, int a_dt, long long lda, int M, int K, int batch, long long a_batch_stride,
              const float* Wt, int ldw, int N, int w_rows_per_batch, const Epi& epi, cudaStream_t stream);
int simt_conv3x3(const void* A, int a_dt, long long lda, int Nimg, int H, int W, int C, const float* Wt, int kpad,
                 int N, const Epi& epi, cudaStream_t stream);

// ---- pointwise.cu ----------------------------------------------------------------------------
// LayerNorm over channels per pixel. mode 0: WithBias ((x-mu)*rstd*w+b), mode 1: BiasFree (x*rstd*w).
int layernorm(const void* x, int x_dt, long long ldx, void* y, int y_dt, long long ldy, long long rows, int C,
              const float* w, const float* b, int mode, float eps, cudaStream_t stream);
// (mu, rstd) per row for a LayerNorm folded into the next GEMM (Epi::ln_stats / ln_colsum).
int row_stats(const void* x, int x_dt, long long ldx, float2* stats, long long rows, int C, float eps,
              cudaStream_t stream);
// depthwise 3x3 (zero pad 1). wt is [9][C] float (tap-major). If sumsq != nullptr, accumulates per (image, channel)
// sum of squares of the fp32 outputs for channels < sumsq_channels into sumsq[img*sumsq_channels + c].
int dwconv3x3(const void* x, int dt, long long ldx, void* y, long long ldy, int Nimg, int H, int W, int C,
              const float* wt, float* sumsq, int sumsq_channels, cudaStream_t stream);
// GDFN gate: y[., c] = gelu(dw(x)[., c]) * dw(x)[., hp + c] for c < hp ; x has 2*hp channels, wt is [9][2*hp].
int dwconv3x3_gate(const void* x, int dt, long long ldx, void* y, long long ldy, int Nimg, int H, int W, int hp,
                   const float* wt, cudaStream_t stream);
// bf16 TMA-fed versions (dw_tma.cu); x rows must be 16-byte aligned
int dwconv3x3_bf16_tma(const bf16* x, long long ldx, bf16* y, long long ldy, int Nimg, int H, int W, int C,
                       const float* wt, cudaStream_t stream);
int dwconv3x3_gate_bf16_tma(const bf16* x, long long ldx, bf16* y, long long ldy, int Nimg, int H, int W, int hp,
                            const float* wt, cudaStream_t stream);
// ---- mlp_fused.cu : SwinIR fc1 -> GELU -> fc2 -> + x (norm2 folded, in place) as one kernel ----
bool mlp_fused_shape_ok(int C, int hid);
int mlp_fused_bf16(bf16* x, long long ldx, long long P, int C, int hid, const float2* stats_in, const bf16* w1,
                   int ldw1, const float* cs1, const float* b1, const bf16* w2, int ldw2, const float* b2,
                   float2* stats_out, float eps, cudaStream_t stream);

// ---- ocab.cu : X-Restormer overlapping cross-attention (xrestormer_arch.py:190-237) ----
int ocab_attention(const void* qkv, int dt, long long ld, void* out, long long ldo, int Nimg, int H, int W, int heads,
                   int dim_head, int ws, int ows, const float* rel_h, const float* rel_w, cudaStream_t stream);

// ---- hat.cu : HAT's window-16 self-attention / overlapping cross-attention and CAB channel attention ----
// ows == ws: (shifted-)window self-attention; ows > ws: OCAB (keys from the zero-padded ows x ows window, no shift)
int hat_attention(const void* qkv, int dt, long long ld, void* out, long long ldo, int n, int H, int W, int heads, int d,
                  int dstride, int ws, int ows, int shift, const float* table, cudaStream_t stream);
int channel_attention_slabs(int n, long long hw);
int channel_attention_scale(const void* y, int dt, long long ldy, int n, long long hw, int C, int Cs, const float* w1,
                            const float* b1, const float* w2, const float* b2, float out_scale, float* sums,
                            float* scale, cudaStream_t stream);
int add_scaled(const void* x, long long ldx, const void* y, long long ldy, void* out, long long ldo, int dt, long long rows,
               long long hw, int C, const float* scale, cudaStream_t stream);

// ---- pwdw_fused.cu : 1x1 convolution (folded LayerNorm) -> depthwise 3x3 [-> GELU gate] without the HBM round trip ----
bool pwdw_fused_shape_ok(int C, int nrows, bool gate, int hp);
int pwdw_fused_bf16(bool gate, const bf16* x, long long ldx, const float2* stats, const bf16* w1, int nrows, int hp,
                    const float* cs, const float* tb, const float* dww, int dw_ld, bf16* y, long long ldy, int Nimg,
                    int H, int W, int C, cudaStream_t stream);

// fp32 [rows][cin] (cin <= 4) -> bf16 [rows][8]: channel c = bf16(x), channel cin + c = bf16(x - hi), rest zero
int split_hi_lo_bf16(const float* x, int cin, bf16* y, long long rows, cudaStream_t stream);
// fp32 [rows][K] -> bf16 [rows][6K] (h | h | h | m | m | l, x = h + m + l): the A operand of the fp32-mode tensor-core GEMMs
int split3_bf16(const float* x, long long ldx, bf16* y, long long ldy, long long rows, int K, cudaStream_t stream);
int nearest_up2(const void* x, int dt, long long ldx, void* y, long long ldy, int Nimg, int H, int W, int C,
                cudaStream_t stream);

// ---- mdta.cu ---------------------------------------------------------------------------------
// Gram over pixels of channels-last q,k (inside a qkv tensor of row stride ld): G[b,h,i,j] += sum_p q[p,hc+i]*k[p,hc+j]
// The pixel axis is split into nchunks = mdta_nchunks(P, heads, Nimg) chunks; every chunk writes its own slot:
//   Gpart [Nimg][heads][nchunks][c][c], ssq_part [Nimg][heads][nchunks][2][c] (|q_i|^2 then |k_j|^2). No atomics.
int mdta_nchunks(int P, int heads, int Nimg);
int mdta_gram_tc(const bf16* qkv, long long ld, int Nimg, int P, int C, int heads, int nchunks, float* Gpart,
                 float* ssq_part, cudaStream_t stream);
int mdta_gram_simt(const void* qkv, int dt, long long ld, int Nimg, int P, int C, int heads, int nchunks,
                   float* Gpart, float* ssq_part, cudaStream_t stream);
// sums the slots in order; attn = softmax_j(G[i,j] / (max(|q_i|,eps) max(|k_j|,eps)) * temp[h]);
// Weff[b][o][h*c+j] = sum_i Wp[o][h*c+i] attn[i][j]
int mdta_softmax_weff(const float* Gpart, const float* ssq_part, int nchunks, const float* temperature,
                      const float* Wp, int Nimg, int C, int heads, void* Weff, int w_dt,
                      float* attn /*[Nimg][heads][c][c], required*/, cudaStream_t stream);

// ---- swin_attn.cu ----------------------------------------------------------------------------
int window_attention(const void* qkv, int dt, long long ld, void* out, long long ldo, int n, int H, int W, int C,
                     int heads, int dstride, int ws, int shift, const float* bias_table, cudaStream_t stream);

// ---- tiles.cu --------------------------------------------------------------------------------
// Gather tiles from a planar float image [c][h][w] into channels-last float tiles [n][t][t][c],
// applying (x - mean[c]) * scale.
int tiles_gather(const float* img, int c, int h, int w, const int* ys, const int* xs, int ny, int nx, int th, int tw,
                 float* tiles, const float* mean, float scale, cudaStream_t stream);
// Overlap blend (reference order h-major, w-minor), divide by the analytic count, crop, clamp, round-half-even.
int tiles_stitch_rows(const float* tiles, int c, int th, int tw, int sf, const int* ys, const int* xs, int ny, int nx,
                      int h, int w, int crop_h, int crop_w, int row0, int rows, float* out_f32, uint8_t* out_u8,
                      int u8_reverse_channels, cudaStream_t stream);
int tiles_stitch(const float* tiles, int c, int th, int tw, int sf, const int* ys, const int* xs, int ny, int nx, int h,
                 int w, int crop_h, int crop_w, float* out_f32 /*[c][crop_h][crop_w] or null*/,
                 uint8_t* out_u8 /*[crop_h][crop_w][c] or null*/, int u8_reverse_channels, cudaStream_t stream);
// crop-with-pad tiling (HAT): paste the [ph][pw] rectangle at (oy, ox) of a channels-last tile to (Y0, X0) of the outputs
int tile_paste(const float* tile, int c, int TH, int TW, int oy, int ox, int ph, int pw, int Y0, int X0, int out_h,
               int out_w, float* out_f32, uint8_t* out_u8, int u8_reverse_channels, cudaStream_t stream);
// uint8 HWC (optionally BGR) -> planar float [c][h'][w'] in [0,1] with the reference's padding rule.
// pad_mode 0: reflect (Restormer / F.pad 'reflect'), 1: symmetric flip-and-concat (SwinIR scripts).
int image_prepare(const uint8_t* src, int h, int w, int c, int reverse_channels, float* dst, int hp, int wp,
                  int pad_mode, cudaStream_t stream);
int image_pad_f32(const float* src, int c, int h, int w, float* dst, int hp, int wp, int pad_mode, cudaStream_t stream);

}  // namespace irb
