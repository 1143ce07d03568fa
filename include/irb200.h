/* irb200 — C ABI of the B200-native tiled forward pass of the SwinIR / Restormer restoration tools.
 *
 * This is the drop-in boundary for the one hot path of blackbean001/Image-Restoration-Agent: everything a
 * reference tool does between "image decoded" and "uint8 image ready to encode":
 *     pad -> [tile loop: model(in_patch); E += out; W += 1] -> E/W -> crop -> clamp -> round -> uint8
 * Reference call sites replaced (paths relative to the reference root):
 *   AgenticIR/executor/denoising/tools/SwinIR/inference.py:69-90, 294-322          (main loop body, test())
 *   AgenticIR/executor/defocus_deblurring/tools/Restormer/demo.py:118-161           (per-file loop body)
 *   AgentApp/model_service/denoising/SwinIR/model_serving.py:131-175                (process_image)
 *   AgentApp/model_service/motion_deblurring/Restormer/model_serving.py:157-208     (restore_image)
 *   .../SwinIR/models/network_swinir.py:805-840  (SwinIR.forward)
 *   .../Restormer/basicsr/models/archs/restormer_arch.py:245-284 (Restormer.forward)
 *
 * Conventions
 *   - plain C types only; every pointer named d_* is DEVICE memory owned by the caller (e.g. tensor.data_ptr()).
 *   - `stream` is a cudaStream_t passed as void*; the library never synchronises it (irb_create excepted) and
 *     never allocates device memory during a forward call: scratch comes from the caller's workspace.
 *   - return 0 = ok; 1 = invalid argument; 2 = CUDA/driver error; 3 = bad weights blob. The message is
 *     available from irb_last_error() (thread-local).
 *   - there is no CPU fallback and no alternative backend: without an sm_100 device every compute entry fails.
 */
#ifndef IRB200_H
#define IRB200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IRB_ARCH_RESTORMER 1
#define IRB_ARCH_SWINIR 2
#define IRB_ARCH_XRESTORMER 3 /* X-Restormer: Restormer blocks + overlapping cross-attention */
#define IRB_ARCH_HAT 4         /* HAT: SwinIR skeleton + CAB branch + OCAB per group, window 16 */

#define IRB_F32 0  /* fp32 storage + fp32 FMA arithmetic (verification mode, <=1e-4 of the reference) */
#define IRB_BF16 1 /* bf16 storage, tcgen05 tensor-core contractions with fp32 accumulation */

#define IRB_UPSAMPLER_NONE 0         /* denoise / JPEG-CAR tail: x + conv_last(res)   network_swinir.py:832-836 */
#define IRB_UPSAMPLER_PIXELSHUFFLE 1 /* classical SR                                   network_swinir.py:812-817 */
#define IRB_UPSAMPLER_NEAREST_CONV 2 /* real-world SR                                  network_swinir.py:823-831 */
#define IRB_UPSAMPLER_PIXELSHUFFLE_DIRECT 3 /* lightweight SR: one conv + PixelShuffle(s), s in {2, 3, 4}   :818-822 */

typedef struct irb_handle irb_handle;

/* Mirrors the constructor arguments of the two reference nn.Modules. */
typedef struct irb_model_desc {
    int arch;      /* IRB_ARCH_* */
    int precision; /* IRB_F32 / IRB_BF16 */
    int in_chans;  /* 1 or 3; out_chans is the same in every reference configuration */
    /* Restormer(...)  restormer_arch.py:194-206, demo.py:92 */
    int dim;
    int num_blocks[4];
    int num_refinement_blocks;
    int heads[4];
    float ffn_expansion_factor;
    int ln_bias_free; /* 1: LayerNorm_type='BiasFree' (denoising tasks), 0: 'WithBias' */
    /* SwinIR(...)  network_swinir.py:646-654, inference.py:136-195 */
    int img_size;
    int window_size;
    int embed_dim;
    int num_layers;
    int depths[16];
    int num_heads[16];
    int mlp_ratio;
    int upscale;
    float img_range;
    int upsampler;  /* IRB_UPSAMPLER_* */
    int resi_3conv; /* 1: resi_connection='3conv', 0: '1conv' */
    /* XRestormer(...)  xrestormer_arch.py:324-341: dim, num_blocks, num_refinement_blocks, heads (= channel_heads),
     * ffn_expansion_factor, ln_bias_free as above, window_size as above, upscale = scale (1; 4 = the SR config: the input
     * is bilinearly up-sampled first, :383-384), plus: */
    int spatial_heads[4];
    int spatial_dim_head;
    float overlap_ratio; /* the same for every level in all agent configs */
    /* SwinIR, DiffBIR stage 1 (DiffBIR/model/swinir.py:716-724): 0 = plain conv_first; s > 1: conv_first =
     * PixelUnshuffle(s) + conv(in_chans * s^2 -> embed_dim), upscale must equal s (nearest+conv tail, x2 per stage), so
     * the model's output has the size of its input (irb_scale() = 1) */
    int unshuffle_scale;
    /* HAT(...)  hat_arch.py:738-766 (configs/hat.yml:21-37): the SwinIR fields above (img_size, window_size <= 16,
     * embed_dim, depths, num_heads, mlp_ratio, upscale, img_range, upsampler = pixelshuffle, resi_3conv = 0) and
     * overlap_ratio, plus: */
    int compress_ratio;  /* CAB: conv3x3 C -> C / compress_ratio -> C */
    int squeeze_factor;  /* ChannelAttention: C -> C / squeeze_factor -> C */
    float conv_scale;    /* weight of the CAB branch in HAB.forward (:306) */
} irb_model_desc;

const char* irb_last_error(void);
int irb_version(void);

/* Tile index list of the reference loop: list(range(0, n - tile, tile - overlap)) + [n - tile]
 * (inference.py:305-308, demo.py:139-141) with tile = min(tile, n) applied by the CALLER.
 * idx may be NULL to query the count only. Host-only, no GPU needed. */
int irb_plan_tiles(int n, int tile, int overlap, int* idx, int* count);

/* Padded size the reference scripts feed to the model: SwinIR (h / ws + 1) * ws (always pads,
 * inference.py:76-78); Restormer next multiple of 8 only when needed (demo.py:122-126); X-Restormer next multiple of
 * `window_size` (64 for the scale-1 configs, 16 for the scale-4 SR config; 0 = 64) only when needed
 * (xrestormer_model.py:56-66); HAT next multiple of window_size only when needed (hat_model.py:16-26). Host-only. */
int irb_padded_size(int arch, int window_size, int h, int w, int* hp, int* wp);

/* weights_blob: host memory, format documented in image_restoration_agent_b200/weights.py (IRBW v1: named fp32
 * tensors of the reference state_dict). Uploads + repacks weights on `device`; synchronises. */
int irb_create(const irb_model_desc* desc, const void* weights_blob, size_t nbytes, int device, irb_handle** out);
void irb_destroy(irb_handle* h);
/* options: "tile_batch" (max tiles per forward batch inside irb_restore_image; 0 = automatic);
 *          "fuse_pwdw" (default 0; 1: Restormer blocks with C <= 128 run each 1x1 -> depthwise pair as one kernel,
 *          irb_pwdw_fused — parity-equal but measured slower than the separate kernels);
 *          "fuse_mlp" (SwinIR: fc1 -> GELU -> fc2 -> + x as one kernel, irb_mlp_fused, where the shape allows);
 *          "fp32_tc" (default 1; SwinIR / HAT in IRB_F32 mode: linear layers and 3x3 convolutions run on the tensor cores
 *          with both operands split into three bf16 terms, six bf16 products per fp32 product, fp32 accumulation; 0: the
 *          fp32 FMA kernels);
 *          "strict_tile" (default 1: SwinIR tiles must be a multiple of window_size like inference.py:302; 0: the
 *          Flask service's process_image, model_serving.py:146-166, which has no such assert) */
int irb_set_option(irb_handle* h, const char* key, long long value);
int irb_scale(const irb_handle* h); /* upscale factor (1 for Restormer / denoise) */
/* Tiles per forward batch that irb_restore_image would use for n_tiles tiles of tile_h x tile_w (the "tile_batch"
 * option, else the largest batch whose workspace stays under 40 GiB). Callers that drive irb_forward_tiles_cl
 * themselves (multi-rank sharding) chunk their tile range by it. */
int irb_tile_batch(irb_handle* h, int n_tiles, int tile_h, int tile_w, int* out_batch);

/* model(x): x is [n, c, th, tw] fp32 NCHW (device), y is [n, c, th*s, tw*s] fp32 NCHW (device).
 * Restormer needs th, tw multiples of 8; SwinIR reflect-pads to the window internally like the reference. */
int irb_workspace_bytes(irb_handle* h, int n_tiles, int tile_h, int tile_w, size_t* out_bytes);
int irb_forward_tiles(irb_handle* h, const float* d_in, float* d_out, int n_tiles, int tile_h, int tile_w,
                      void* d_workspace, size_t workspace_bytes, void* stream);

/* The whole tool body for one image.
 *   in_fmt 0: d_in is uint8 [h][w][c] (as cv2.imread gives it; reverse_channels=1 means BGR on both sides)
 *   in_fmt 1: d_in is fp32 planar [c][h][w] in [0,1] (RGB / gray), reverse_channels ignored on input
 *   tile = 0 means "--tile None" (whole image); otherwise the reference's tile / tile_overlap arguments.
 *   d_out_u8  : uint8 [h*s][w*s][c] (NULL to skip)      d_out_f32 : fp32 planar [c][h*s][w*s], un-clamped E/W
 *   (NULL to skip). */
int irb_restore_workspace_bytes(irb_handle* h, int img_h, int img_w, int tile, int overlap, size_t* out_bytes);
int irb_restore_image(irb_handle* h, const void* d_in, int in_fmt, int reverse_channels, int img_h, int img_w,
                      int tile, int overlap, uint8_t* d_out_u8, float* d_out_f32, void* d_workspace,
                      size_t workspace_bytes, void* stream);

/* HATModel's tool body (hat/models/hat_model.py:16-109): reflect pad bottom / right to the window (pre_process), cut
 * tile_size tiles with a tile_pad border clipped to the image (tile_process; tile_size = 0: the whole image, process),
 * run each through the model, paste each tile's own area (no blending), crop (post_process). tile_size and tile_pad
 * must be multiples of the model's window. Formats as irb_restore_image. Works for every arch; the padded size follows
 * the arch's own rule (irb_padded_size). */
int irb_restore_crop_workspace_bytes(irb_handle* h, int img_h, int img_w, int tile_size, int tile_pad, size_t* out_bytes);
int irb_restore_image_crop(irb_handle* h, const void* d_in, int in_fmt, int reverse_channels, int img_h, int img_w,
                           int tile_size, int tile_pad, uint8_t* d_out_u8, float* d_out_f32, void* d_workspace,
                           size_t workspace_bytes, void* stream);

/* ---- the stages of irb_restore_image, exposed so that tiles can be sharded across ranks (one process per GPU):
 * plan (host) -> tiles_prepare -> forward_tiles_cl on this rank's tile range -> [NCCL gather] -> irb_stitch. ------
 * ys/xs must hold 128 ints each. d_padded: fp32 [c][hp][wp]; d_tiles: fp32 channels-last [ny*nx][th][tw][c]. */
int irb_restore_plan(irb_handle* h, int img_h, int img_w, int tile, int overlap, int* hp, int* wp, int* th, int* tw,
                     int* ys, int* ny, int* xs, int* nx);
int irb_tiles_prepare(irb_handle* h, const void* d_in, int in_fmt, int reverse_channels, int img_h, int img_w,
                      int tile, int overlap, float* d_padded, float* d_tiles, void* stream);
/* channels-last variant of irb_forward_tiles: in [n][th][tw][c] -> out [n][th*s][tw*s][c]; workspace size =
 * irb_workspace_bytes(). */
int irb_forward_tiles_cl(irb_handle* h, const float* d_in, float* d_out, int n_tiles, int tile_h, int tile_w,
                         void* d_workspace, size_t workspace_bytes, void* stream);

/* ---- single-kernel entry points (parity tests, ncu captures). dtype: IRB_F32 / IRB_BF16; activations are
 * channels-last with an explicit row stride ld (elements). -------------------------------------------------- */
int irb_layernorm(const void* d_x, int dtype, long long ldx, void* d_y, long long ldy, long long rows, int C,
                  const float* d_w, const float* d_b, int bias_free, float eps, void* stream);
/* y[b,m,:N] = act(x[b,m,:K] . W[b*w_rows_per_batch + n,:K] + bias) + res ; W has dtype `dtype`, row stride ldw.
 * act: 0 none, 1 exact GELU, 2 LeakyReLU(slope). bf16 -> tcgen05 kernel, fp32 -> FMA kernel. */
int irb_gemm1x1(const void* d_x, int dtype, long long ldx, int M, int K, int batch, const void* d_w, int ldw, int N,
                int w_rows_total, int w_rows_per_batch, const float* d_bias, const void* d_res, long long ldr,
                int act, float slope, void* d_y, long long ldy, void* stream);
/* 3x3 conv, zero pad 1, on [n,H,W,C]; W is [N_rows][9*kpad] (K index = (ky*3+kx)*kpad + c), dtype `dtype`.
 * shuffle: 0 none, 1 PixelShuffle(2) fused into the store, 2 PixelUnshuffle(2). */
int irb_conv3x3(const void* d_x, int dtype, long long ldx, int n, int H, int W, int C, const void* d_w, int kpad,
                int N, int w_rows_total, const float* d_bias, const void* d_res, long long ldr, int act, float slope,
                int shuffle, void* d_y, long long ldy, void* stream);
/* depthwise 3x3; d_w is fp32 [9][C] tap-major. d_sumsq (optional) [n][sumsq_channels] += sum_p y^2 */
int irb_dwconv3x3(const void* d_x, int dtype, long long ldx, void* d_y, long long ldy, int n, int H, int W, int C,
                  const float* d_w, float* d_sumsq, int sumsq_channels, void* stream);
/* GDFN: y[.,c] = gelu(dw(x)[.,c]) * dw(x)[.,hp+c], c < hp; d_w fp32 [9][2*hp] */
int irb_gdfn_dw_gate(const void* d_x, int dtype, long long ldx, void* d_y, long long ldy, int n, int H, int W, int hp,
                     const float* d_w, void* stream);
/* A Restormer "1x1 convolution -> depthwise 3x3" pair as ONE bf16 kernel: the tensor between the two convolutions
 * stays in shared memory (restormer_arch.py:115-116 qkv -> qkv_dwconv; :88-93 project_in -> dwconv -> gelu gate).
 *   gate 0: y[n,H,W,:nrows] = dw3x3(h),                      h = W1' . LN(x)   (nrows % 8 == 0)
 *   gate 1: y[n,H,W,:hp]    = gelu(dw3x3(h)[:hp]) * dw3x3(h)[hp:2hp]           (nrows == 2*hp, hp % 32 == 0)
 * d_x bf16 [n,H,W,C] (row stride ldx), C in {16..128} multiple of 16; d_stats fp32 [n*H*W][2] = (mu, rstd) of the
 * rows of x; d_w1 bf16 [nrows][C] = the 1x1 weights with the LayerNorm gain folded in; d_colsum / d_t fp32 [nrows]
 * (row sums of the rounded folded weights / W1 . ln_bias; both NULL for BiasFree: h = rstd * (W1' . x));
 * d_dw fp32 [nrows/2][9][2]: the 9 taps (ky*3+kx) of each channel PAIR contiguous (dw_ld is ignored).
 * h is zero outside the image (the depthwise convolution's zero padding). */
int irb_pwdw_fused(int gate, const void* d_x, long long ldx, const float* d_stats, const void* d_w1, int nrows, int hp,
                   const float* d_colsum, const float* d_t, const float* d_dw, int dw_ld, void* d_y, long long ldy, int n,
                   int H, int W, int C, void* stream);
/* SwinIR's MLP half-block in place, bf16 (network_swinir.py:24-30 with the block's norm2 and residual, :277):
 *   x[p,:] <- x[p,:] + W2 . gelu(W1' . LN(x[p,:]) + b1') + b2, and (mu, rstd) of the new rows into d_stats_out.
 * d_x bf16 [rows][ldx] (C valid columns, C <= 192, C % 4 == 0); d_stats_in fp32 [rows][2] = (mu, rstd) of the rows of
 * x; d_w1 bf16 [hid][ldw1] = fc1 with norm2's gain folded in, d_colsum1 / d_b1 fp32 [hid] (row sums of the rounded
 * folded weights / fc1 bias + W1 . ln_bias); d_w2 bf16 [C][ldw2], d_b2 fp32 [C]. The hidden tensor stays on the SM. */
int irb_mlp_fused(void* d_x, long long ldx, long long rows, int C, int hid, const float* d_stats_in, const void* d_w1,
                  int ldw1, const float* d_colsum1, const float* d_b1, const void* d_w2, int ldw2, const float* d_b2,
                  float* d_stats_out, float eps, void* stream);
/* X-Restormer overlapping cross-attention (xrestormer_arch.py:190-237) on a channels-last qkv tensor
 * [n, H, W, 3*heads*dim_head] (q | k | v, head-major): queries = the window x window blocks, keys / values = the
 * overlap_window-wide zero-padded neighbourhoods, logits += q . rel_width[dj] + q . rel_height[di] with q pre-scaled by
 * dim_head^-0.5 inside the kernel. d_rel_h / d_rel_w fp32 [2*overlap_window-1][dim_head]. dim_head must be 16. */
int irb_ocab(const void* d_qkv, int dtype, long long ld, void* d_out, long long ldo, int n, int H, int W, int heads,
             int dim_head, int window, int overlap_window, const float* d_rel_h, const float* d_rel_w, void* stream);
/* HAT window attention (hat_arch.py:165-197 WindowAttention with window <= 16 and the shift mask of :921-941 when
 * overlap_window == window; :392-432 OCAB when overlap_window > window: keys / values from the zero-padded
 * overlap_window-wide neighbourhood, shift must be 0) on a channels-last qkv tensor [n, H, W, ld] with sections q | k | v
 * of heads * dstride channels each (head-major, the first head_dim of every dstride real, q pre-scaled).
 * d_bias_table fp32 [(2*window-1)^2][heads] resp. [(window+overlap_window-1)^2][heads], indexed like
 * calculate_rpi_sa / calculate_rpi_oca (:882-919), negative OCA indices wrapping like PyTorch indexing.
 * out [n, H, W, ldo], channel head * head_dim + dd. */
int irb_hat_attention(const void* d_qkv, int dtype, long long ld, void* d_out, long long ldo, int n, int H, int W,
                      int heads, int head_dim, int dstride, int window, int overlap_window, int shift,
                      const float* d_bias_table, void* stream);
/* MDTA Gram over a channels-last qkv tensor (q | k | v): the P pixels are split into nchunks =
 * irb_mdta_nchunks(P, heads, n) chunks and every chunk writes its own slot (no atomics -> bit-reproducible):
 *   d_Gpart   [n][heads][nchunks][c][c]   partial sum_p q k^T
 *   d_ssq_part[n][heads][nchunks][2][c]   partial sum_p q^2 (then k^2): the F.normalize denominators */
int irb_mdta_nchunks(int P, int heads, int n);
int irb_mdta_gram(const void* d_qkv, int dtype, long long ld, int n, int P, int C, int heads, int nchunks,
                  float* d_Gpart, float* d_ssq_part, void* stream);
/* sums the slots in order, softmax(G / (|q||k|) * temperature), folded into project_out:
 * d_weff[n][C][C] (dtype); d_attn fp32 [n][heads][c][c] receives the attention matrix (required: it is also the
 * hand-off buffer between the softmax kernel and the fold kernel) */
int irb_mdta_attn(const float* d_Gpart, const float* d_ssq_part, int nchunks, const float* d_temperature,
                  const float* d_wproj, int n, int C, int heads, void* d_weff, int dtype, float* d_attn,
                  void* stream);
/* irb_mdta_gram + irb_mdta_attn as ONE call that allocates nothing (no cudaMalloc / cudaFree / allocating library
 * routine): the per-chunk partial sums live in a file-scope static __device__ array of
 * irb_mdta_attn_static_max_bytes() = 4 MiB, held by every GPU that loads the library until the process exits.
 * No counterpart in the reference: added for callers that cannot allocate
 * (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). The caller still passes the input
 * (d_qkv, channels-last q | k | v, row stride ld, n images of P pixels) and the outputs (d_weff [n][C][C] in `dtype`,
 * d_attn fp32 [n][heads][c][c]); results are bit-identical to the existing pair (same kernels).
 * Limit: n * heads * irb_mdta_nchunks(P, heads, n) * (c*c + 2*c) floats (c = C / heads) must fit in the array; larger
 * shapes are rejected with an error that names both sizes. Every call shares the array and overwrites it (nothing needs
 * to start at zero: the Gram kernel writes every slot it later reads): calls that OVERLAP on the GPU (two streams, or
 * graph branches without a dependency) race on it; serialise them on one stream or with events. */
size_t irb_mdta_attn_static_max_bytes(void);
int irb_mdta_attn_static(const void* d_qkv, int dtype, long long ld, int n, int P, int C, int heads,
                         const float* d_temperature, const float* d_wproj, void* d_weff, float* d_attn, void* stream);
/* overlap-blend stitch of channels-last fp32 tiles [ny*nx][th*sf][tw*sf][c] (host index lists) */
int irb_stitch(const float* d_tiles, int c, int th, int tw, int sf, const int* ys, const int* xs, int ny, int nx,
               int h, int w, int crop_h, int crop_w, float* d_out_f32, uint8_t* d_out_u8, int reverse_channels,
               void* stream);
/* The same blend restricted to output rows [row0, row0 + rows) of the crop — the strip one rank owns when the tiles
 * of an image are sharded. d_tiles is still indexed by the GLOBAL tile number iy*nx+ix; only the tile
 * rows that cover the strip are read, so a rank needs its own tiles plus its neighbours' seam rows in place.
 * Outputs are strip-local: d_out_u8 [rows][crop_w][c], d_out_f32 [c][rows][crop_w]. Per pixel the tiles are summed
 * in the reference's (h_idx, w_idx) order, so the strips concatenate to irb_stitch()'s result bit for bit. */
int irb_stitch_rows(const float* d_tiles, int c, int th, int tw, int sf, const int* ys, const int* xs, int ny, int nx,
                    int h, int w, int crop_h, int crop_w, int row0, int rows, float* d_out_f32, uint8_t* d_out_u8,
                    int reverse_channels, void* stream);
/* SwinIR (shifted-)window attention on a channels-last qkv tensor [n, H, W, 3*heads*dstride] (sections q | k | v,
 * head-major, dstride >= C/heads channels per head, q pre-multiplied by head_dim^-0.5):
 * out[n,H,W,C] = window_reverse(softmax(q k^T + rel_pos_bias + shift_mask) v), with the cyclic shift folded into
 * the addressing (network_swinir.py:114-145, 216-237, 239-279). d_bias_table is fp32 [(2ws-1)^2][heads]. */
int irb_window_attention(const void* d_qkv, int dtype, long long ld, void* d_out, long long ldo, int n, int H, int W,
                         int C, int heads, int dstride, int window, int shift, const float* d_bias_table,
                         void* stream);

/* ---- per-launch profiler (bench.py's live roofline): between begin and end every kernel launched by this
 * library on `stream` is bracketed by CUDA events; end synchronises the stream and aggregates by kernel name +
 * shape class. flops / bytes are ALGORITHMIC (operands read once, results written once). Not thread-safe. ---- */
int irb_profile_begin(void* stream);
int irb_profile_end(void* stream);
int irb_profile_count(void);
int irb_profile_entry(int i, char* name56, int* launches, double* ms, double* flops, double* bytes);

#ifdef __cplusplus
}
#endif
#endif /* IRB200_H */
