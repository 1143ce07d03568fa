// Restormer forward (restormer_arch.py:245-284) as a sequence of irb200 kernels on channels-last activations.
//
// Data layout in the caller's workspace (E = element size of the storage dtype):
//   cat1 [P1, 2d]   decoder-level-1 stream; the encoder-level-1 stream lives in its channels [d, 2d)  (zero-copy cat)
//   cat2 [P2, 4d]   channels [0,2d) <- up3_2 output, [2d,4d) = encoder-level-2 stream
//   cat3 [P3, 8d]   channels [0,4d) <- up4_3 output, [4d,8d) = encoder-level-3 stream
//   x2 [P2,2d], x3 [P3,4d], x4 [P4,8d]  decoder-2 / decoder-3 / latent streams
//   scratch: ln [P,C], qkv_pre / qkv [P,3C], hidden [P,2hp], gated [P,hp], G [n,heads,c,c] + sumsq [n,2C] (fp32),
//            weff [n,C,C]
// where P1 = n*H*W, P2 = P1/4, ... and d = dim (48).
#include "model.h"
#include <algorithm>
#include <string.h>

namespace irb {

struct RBlockW {
    int C, heads, h, hp;
    float *n1w, *n1b, *n2w, *n2b;
    void* qkv_w;     // [3C][C]
    float* qkv_dw;   // [9][3C]
    float *qkv_dwp, *ffn_dwp;  // the same taps as [ch/2][9][2] (pwdw_fused.cu: a channel pair's 9 taps contiguous)
    float* temp;     // [heads]
    float* proj_w;   // [C][C] fp32 (folded with the attention matrix per image)
    void* pin_w;     // [2hp][C]   rows [0,h) = x1, [hp,hp+h) = x2, others zero
    float* ffn_dw;   // [9][2hp]
    void* pout_w;    // [C][hp]
    // bf16 mode: norm1 / norm2 are folded into qkv_w / pin_w (W' = W diag(gain)); the GEMM epilogue applies
    // rstd * (acc - mu * colsum) + t with colsum[n] = sum_k W'[n,k], t[n] = sum_k W[n,k] * ln_bias[k]
    float *qkv_cs, *qkv_t, *pin_cs, *pin_t;
    // X-Restormer (xrestormer_arch.py:239-260): the block continues with x += OCAB(norm3(x)); x += GDFN(norm4(x))
    bool has_spatial = false;
    int sheads = 0, inner = 0, ows = 0;  // spatial heads, sheads * dim_head, overlapping window size
    float *n3w = nullptr, *n3b = nullptr;
    void* oc_qkv_w = nullptr;    // [3*inner][C] (norm3 folded in bf16 mode)
    float *oc_cs = nullptr, *oc_t = nullptr;
    void* oc_proj_w = nullptr;   // [C][inner]
    float *rel_h = nullptr, *rel_w = nullptr;  // [2*ows-1][dim_head]
    RBlockW* sp = nullptr;       // the second GDFN (norm4 + spatial_ffn), stored in the ffn fields of another RBlockW
};

// state_dict names of the pieces of a block: Restormer {norm1, attn, norm2, ffn}; X-Restormer's first half
// {norm1, channel_attn, norm2, channel_ffn} and its second GDFN {-, -, norm4, spatial_ffn}
struct BlockNames {
    const char *norm_a, *attn, *norm_f, *ffn;
};

// W [N][K] (row stride ldw) -> W' = W * diag(gain) in place; colsum of the bf16-ROUNDED W'; t = W . ln_bias
static void fold_ln(std::vector<float>& w, int N, int K, int ldw, const float* gain, const float* ln_bias,
                    std::vector<float>& colsum, std::vector<float>& t) {
    colsum.assign(N, 0.f);
    t.assign(N, 0.f);
    for (int n = 0; n < N; ++n) {
        double cs = 0.0, tt = 0.0;
        for (int k = 0; k < K; ++k) {
            float& e = w[size_t(n) * ldw + k];
            if (ln_bias) tt += double(e) * ln_bias[k];
            e *= gain[k];
            cs += double(__bfloat162float(__float2bfloat16_rn(e)));
        }
        colsum[n] = float(cs);
        t[n] = float(tt);
    }
}

// [9][n] tap-major -> [n/2][9][2]
static std::vector<float> pair_major(const std::vector<float>& w, int n) {
    std::vector<float> o(size_t(9) * n);
    for (int c = 0; c < n; ++c)
        for (int k = 0; k < 9; ++k) o[(size_t(c >> 1) * 9 + k) * 2 + (c & 1)] = w[size_t(k) * n + c];
    return o;
}

struct ConvW {
    void* w;       // storage dtype, [npad][9*kpad]
    float* wf32;   // fp32 copy (SIMT path for tiny channel counts)
    void* wsplit;  // bf16 mode, C <= 4: [npad][9*64] weights for the hi/lo-split tensor-core first conv
    int kpad, npad, N, C;
    int ps;        // PixelShuffle store mode when this conv is followed by PixelShuffle(2): 1 reference rows, 3 permuted
};

struct Restormer : Model {
    int dim;
    std::vector<RBlockW> enc1, enc2, enc3, latent, dec3, dec2, dec1, refine;
    ConvW patch, down12, down23, down34, up43, up32, up21, outc;
    void *reduce3, *reduce2;

    int tile_multiple() const override { return desc.arch == IRB_ARCH_XRESTORMER ? 64 / scale : 8; }
    size_t forward_workspace(int n, int H, int W) override;
    int forward(const float* in, float* out, int n, int H, int W, void* ws, size_t ws_bytes,
                cudaStream_t stream) override;
    int forward_core(const float* in, float* out, int n, int H, int W, void* ws, size_t ws_bytes, cudaStream_t stream);

    struct WS {
        void *cat1, *cat2, *cat3, *x2, *x3, *x4, *ln, *qkv_pre, *qkv, *hidden, *gated, *weff;
        float *G, *sumsq, *attn;
        bf16* in8;      // [P1][8] hi/lo split of the fp32 input (bf16 mode)
        float2* stats;  // [P] (mu, rstd) of the LayerNorm folded into the next GEMM (bf16 mode)
        size_t gs_bytes;
        bool stats_valid;  // ws.stats already holds the LayerNorm statistics of the current stream (written by the producing GEMM)
    };
    void layout(Bump& b, int n, int H, int W, WS& w);
    int load_block(const Blob& blob, const std::string& p, int C, int heads, RBlockW& o,
                   BlockNames nm = {"norm1", "attn", "norm2", "ffn"});
    int load_spatial(const Blob& blob, const std::string& p, int C, int level, RBlockW& o);
    int run_attn(const RBlockW& w, void* x, long long ldx, int n, int H, int W, WS& ws, cudaStream_t s);
    int run_ffn(const RBlockW& w, void* x, long long ldx, int n, int H, int W, WS& ws, cudaStream_t s);
    int run_ocab(const RBlockW& w, void* x, long long ldx, int n, int H, int W, WS& ws, cudaStream_t s);
    int load_conv(const Blob& blob, const std::string& name, int N, int C, ConvW& o, bool pixelshuffle = false);
    int run_block(const RBlockW& w, void* x, long long ldx, int n, int H, int W, WS& ws, cudaStream_t s);
    int run_conv(const ConvW& c, const void* x, int x_dt, long long ldx, int n, int H, int W, Epi e, cudaStream_t s);
    int run_gemm(const void* A, long long lda, long long M, int K, const void* Wt, int ldw, int N, int wrows, Epi e,
                 cudaStream_t s);
    char* at(void* p, long long elems) { return reinterpret_cast<char*>(p) + elems * (long long)esize(); }
};

int Restormer::load_block(const Blob& blob, const std::string& p, int C, int heads, RBlockW& o, BlockNames nm) {
    const bool has_attn = nm.attn != nullptr;
    const std::string NA = p + "." + (has_attn ? nm.norm_a : nm.norm_f), NF = p + "." + nm.norm_f;
    const std::string AT = p + "." + (has_attn ? nm.attn : ""), FF = p + "." + nm.ffn;
    o.C = C;
    o.heads = heads;
    o.h = int(C * desc.ffn_expansion_factor);  // restormer_arch.py:80 int(dim*ffn_expansion_factor)
    o.hp = (o.h + 63) / 64 * 64;
    BlobTensor t;
    IRB_TRY(blob.get(NA + ".body.weight", t, C));
    IRB_TRY(upload_f32(t.data, C, &o.n1w));
    IRB_TRY(blob.get(NF + ".body.weight", t, C));
    IRB_TRY(upload_f32(t.data, C, &o.n2w));
    o.n1b = o.n2b = nullptr;
    if (!desc.ln_bias_free) {
        IRB_TRY(blob.get(NA + ".body.bias", t, C));
        IRB_TRY(upload_f32(t.data, C, &o.n1b));
        IRB_TRY(blob.get(NF + ".body.bias", t, C));
        IRB_TRY(upload_f32(t.data, C, &o.n2b));
    }
    o.temp = nullptr;
    if (has_attn) {
        IRB_TRY(blob.get(AT + ".temperature", t, heads, 1, 1));
        IRB_TRY(upload_f32(t.data, heads, &o.temp));
    }
    const bool fold = dt == DT_BF16;
    BlobTensor g1, b1, g2, b2;
    IRB_TRY(blob.get(NA + ".body.weight", g1, C));
    IRB_TRY(blob.get(NF + ".body.weight", g2, C));
    if (!desc.ln_bias_free) {
        IRB_TRY(blob.get(NA + ".body.bias", b1, C));
        IRB_TRY(blob.get(NF + ".body.bias", b2, C));
    }
    o.qkv_cs = o.qkv_t = o.pin_cs = o.pin_t = nullptr;
    o.qkv_w = nullptr;
    o.qkv_dw = o.qkv_dwp = o.proj_w = nullptr;
    if (has_attn) {
    IRB_TRY(blob.get(AT + ".qkv.weight", t, 3 * C, C, 1, 1));
    {
        std::vector<float> w(t.data, t.data + size_t(3) * C * C), cs, tt;
        if (fold) {
            fold_ln(w, 3 * C, C, C, g1.data, desc.ln_bias_free ? nullptr : b1.data, cs, tt);
            if (!desc.ln_bias_free) {
                IRB_TRY(upload_f32(cs.data(), cs.size(), &o.qkv_cs));
                IRB_TRY(upload_f32(tt.data(), tt.size(), &o.qkv_t));
            }
        }
        IRB_TRY(upload_act(w.data(), w.size(), &o.qkv_w));
    }
    IRB_TRY(blob.get(AT + ".qkv_dwconv.weight", t, 3 * C, 1, 3, 3));
    {
        std::vector<float> d = pack_dw(t);
        IRB_TRY(upload_f32(d.data(), d.size(), &o.qkv_dw));
        std::vector<float> dp = pair_major(d, 3 * C);
        IRB_TRY(upload_f32(dp.data(), dp.size(), &o.qkv_dwp));
    }
    IRB_TRY(blob.get(AT + ".project_out.weight", t, C, C, 1, 1));
    IRB_TRY(upload_f32(t.data, size_t(C) * C, &o.proj_w));
    }
    const int h = o.h, hp = o.hp;
    IRB_TRY(blob.get(FF + ".project_in.weight", t, 2 * h, C, 1, 1));
    {
        std::vector<float> w(size_t(2) * hp * C, 0.f);
        for (int r = 0; r < 2 * h; ++r) {
            const int dst = r < h ? r : hp + (r - h);
            memcpy(&w[size_t(dst) * C], t.data + size_t(r) * C, size_t(C) * 4);
        }
        if (fold) {
            std::vector<float> cs, tt;
            fold_ln(w, 2 * hp, C, C, g2.data, desc.ln_bias_free ? nullptr : b2.data, cs, tt);
            if (!desc.ln_bias_free) {
                IRB_TRY(upload_f32(cs.data(), cs.size(), &o.pin_cs));
                IRB_TRY(upload_f32(tt.data(), tt.size(), &o.pin_t));
            }
        }
        IRB_TRY(upload_act(w.data(), w.size(), &o.pin_w));
    }
    IRB_TRY(blob.get(FF + ".dwconv.weight", t, 2 * h, 1, 3, 3));
    {
        std::vector<float> w(size_t(9) * 2 * hp, 0.f);
        for (int r = 0; r < 2 * h; ++r) {
            const int dst = r < h ? r : hp + (r - h);
            for (int k = 0; k < 9; ++k) w[size_t(k) * 2 * hp + dst] = t.data[size_t(r) * 9 + k];
        }
        IRB_TRY(upload_f32(w.data(), w.size(), &o.ffn_dw));
        std::vector<float> wp = pair_major(w, 2 * hp);
        IRB_TRY(upload_f32(wp.data(), wp.size(), &o.ffn_dwp));
    }
    IRB_TRY(blob.get(FF + ".project_out.weight", t, C, h, 1, 1));
    {
        std::vector<float> w(size_t(C) * hp, 0.f);
        for (int r = 0; r < C; ++r) memcpy(&w[size_t(r) * hp], t.data + size_t(r) * h, size_t(h) * 4);
        IRB_TRY(upload_act(w.data(), w.size(), &o.pout_w));
    }
    return 0;
}

// OCAB + norm3 + the second GDFN of an X-Restormer block (xrestormer_arch.py:190-260)
int Restormer::load_spatial(const Blob& blob, const std::string& p, int C, int level, RBlockW& o) {
    const int dh = desc.spatial_dim_head;
    o.has_spatial = true;
    o.sheads = desc.spatial_heads[level];
    o.inner = dh * o.sheads;
    o.ows = int(desc.window_size * desc.overlap_ratio) + desc.window_size;  // :196
    const int inner = o.inner, rel = 2 * o.ows - 1;
    BlobTensor t, g3, b3;
    IRB_TRY(blob.get(p + ".norm3.body.weight", g3, C));
    IRB_TRY(upload_f32(g3.data, C, &o.n3w));
    if (!desc.ln_bias_free) {
        IRB_TRY(blob.get(p + ".norm3.body.bias", b3, C));
        IRB_TRY(upload_f32(b3.data, C, &o.n3b));
    }
    IRB_TRY(blob.get(p + ".spatial_attn.qkv.weight", t, 3 * inner, C, 1, 1));
    {
        std::vector<float> w(t.data, t.data + size_t(3) * inner * C), cs, tt;
        if (dt == DT_BF16) {
            fold_ln(w, 3 * inner, C, C, g3.data, desc.ln_bias_free ? nullptr : b3.data, cs, tt);
            if (!desc.ln_bias_free) {
                IRB_TRY(upload_f32(cs.data(), cs.size(), &o.oc_cs));
                IRB_TRY(upload_f32(tt.data(), tt.size(), &o.oc_t));
            }
        }
        IRB_TRY(upload_act(w.data(), w.size(), &o.oc_qkv_w));
    }
    IRB_TRY(blob.get(p + ".spatial_attn.project_out.weight", t, C, inner, 1, 1));
    IRB_TRY(upload_act(t.data, size_t(C) * inner, &o.oc_proj_w));
    IRB_TRY(blob.get(p + ".spatial_attn.rel_pos_emb.rel_height", t, rel, dh));
    IRB_TRY(upload_f32(t.data, size_t(rel) * dh, &o.rel_h));
    IRB_TRY(blob.get(p + ".spatial_attn.rel_pos_emb.rel_width", t, rel, dh));
    IRB_TRY(upload_f32(t.data, size_t(rel) * dh, &o.rel_w));
    o.sp = new RBlockW();
    return load_block(blob, p, C, 1, *o.sp, BlockNames{nullptr, nullptr, "norm4", "spatial_ffn"});
}

int Restormer::load_conv(const Blob& blob, const std::string& name, int N, int C, ConvW& o, bool pixelshuffle) {
    BlobTensor t;
    IRB_TRY(blob.get(name, t, N, C, 3, 3));
    std::vector<float> w = pack_conv3x3(t, o.kpad, o.npad);
    o.N = N;
    o.C = C;
    o.ps = pixelshuffle ? 1 : 0;
    if (pixelshuffle && N % 4 == 0 && ((N / 4) % 64 == 0 || (N / 4) % 48 == 0)) {
        permute_rows_ps_major(w, N, 9 * o.kpad, nullptr);
        o.ps = 3;
    }
    IRB_TRY(upload_act(w.data(), w.size(), &o.w));
    o.wf32 = nullptr;
    if (dt == DT_F32)
        o.wf32 = reinterpret_cast<float*>(o.w);
    else if (C < 16)
        IRB_TRY(upload_f32(w.data(), w.size(), &o.wf32));
    o.wsplit = nullptr;
    if (dt == DT_BF16 && C <= 4) {
        int kp, np;
        std::vector<float> ws = pack_conv3x3_split(t, kp, np);
        IRB_TRY(upload_act(ws.data(), ws.size(), &o.wsplit));
    }
    return 0;
}

int build_restormer(const irb_model_desc& d, const Blob& blob, int device, Model** out) {
    Restormer* m = new Restormer();
    *out = m;
    m->desc = d;
    m->dt = d.precision == IRB_BF16 ? DT_BF16 : DT_F32;
    m->device = device;
    m->cin = d.in_chans;
    m->scale = 1;
    if (d.arch == IRB_ARCH_XRESTORMER && d.upscale > 1) {  // the SR config: scale 4 (executor/super_resolution/configs/xrestormer.yml)
        IRB_REQUIRE(d.upscale == 2 || d.upscale == 4, "XRestormer: scale must be 1, 2 or 4");
        m->scale = d.upscale;
    }
    m->dim = d.dim;
    IRB_REQUIRE(d.dim > 0 && d.dim % 16 == 0, "Restormer: dim must be a positive multiple of 16");
    IRB_REQUIRE(d.in_chans == 1 || d.in_chans == 3, "Restormer: in_chans must be 1 or 3");
    for (int l = 0; l < 4; ++l)
        IRB_REQUIRE(d.heads[l] > 0 && ((d.dim << l) % d.heads[l]) == 0 && (d.dim << l) / d.heads[l] <= 96,
                    "Restormer: channels per head must divide and be <= 96");
    IRB_REQUIRE((2 * d.dim) % d.heads[0] == 0 && (2 * d.dim) / d.heads[0] <= 96, "Restormer: decoder-1 head size");
    const int dim = d.dim;
    const bool xr = d.arch == IRB_ARCH_XRESTORMER;
    if (xr) {
        IRB_REQUIRE(d.spatial_dim_head == 16 && d.window_size >= 2 && d.window_size <= 8, "XRestormer: spatial_dim_head 16, window <= 8");
        IRB_REQUIRE(d.overlap_ratio >= 0.f && int(d.window_size * d.overlap_ratio) % 2 == 0 &&
                        int(d.window_size * d.overlap_ratio) + d.window_size <= 16,
                    "XRestormer: overlapping window must be <= 16 and symmetric");
        for (int l = 0; l < 4; ++l)
            IRB_REQUIRE(d.spatial_heads[l] >= 1 && 3 * 16 * d.spatial_heads[l] <= 3 * (d.dim << l),
                        "XRestormer: spatial heads x 16 must not exceed the level's channel count");
    }
    auto stage = [&](std::vector<RBlockW>& v, const char* name, int nblk, int C, int heads) -> int {
        v.resize(nblk);
        for (int i = 0; i < nblk; ++i) {
            const std::string p = std::string(name) + "." + std::to_string(i);
            if (!xr) {
                IRB_TRY(m->load_block(blob, p, C, heads, v[i]));
            } else {  // level of this stage: 0 for C = dim and for the 2*dim decoder-1 / refinement stages
                const bool lvl0 = strcmp(name, "encoder_level1") == 0 || strcmp(name, "decoder_level1") == 0 ||
                                  strcmp(name, "refinement") == 0;
                int level = 0;
                if (!lvl0) level = C == dim * 2 ? 1 : (C == dim * 4 ? 2 : 3);
                IRB_TRY(m->load_block(blob, p, C, heads, v[i], BlockNames{"norm1", "channel_attn", "norm2", "channel_ffn"}));
                IRB_TRY(m->load_spatial(blob, p, C, level, v[i]));
            }
        }
        return 0;
    };
    IRB_TRY(m->load_conv(blob, "patch_embed.proj.weight", dim, d.in_chans, m->patch));
    IRB_TRY(stage(m->enc1, "encoder_level1", d.num_blocks[0], dim, d.heads[0]));
    IRB_TRY(m->load_conv(blob, "down1_2.body.0.weight", dim / 2, dim, m->down12));
    IRB_TRY(stage(m->enc2, "encoder_level2", d.num_blocks[1], dim * 2, d.heads[1]));
    IRB_TRY(m->load_conv(blob, "down2_3.body.0.weight", dim, dim * 2, m->down23));
    IRB_TRY(stage(m->enc3, "encoder_level3", d.num_blocks[2], dim * 4, d.heads[2]));
    IRB_TRY(m->load_conv(blob, "down3_4.body.0.weight", dim * 2, dim * 4, m->down34));
    IRB_TRY(stage(m->latent, "latent", d.num_blocks[3], dim * 8, d.heads[3]));
    IRB_TRY(m->load_conv(blob, "up4_3.body.0.weight", dim * 16, dim * 8, m->up43, true));
    BlobTensor t;
    IRB_TRY(blob.get("reduce_chan_level3.weight", t, dim * 4, dim * 8, 1, 1));
    IRB_TRY(m->upload_act(t.data, size_t(dim) * 4 * dim * 8, &m->reduce3));
    IRB_TRY(stage(m->dec3, "decoder_level3", d.num_blocks[2], dim * 4, d.heads[2]));
    IRB_TRY(m->load_conv(blob, "up3_2.body.0.weight", dim * 8, dim * 4, m->up32, true));
    IRB_TRY(blob.get("reduce_chan_level2.weight", t, dim * 2, dim * 4, 1, 1));
    IRB_TRY(m->upload_act(t.data, size_t(dim) * 2 * dim * 4, &m->reduce2));
    IRB_TRY(stage(m->dec2, "decoder_level2", d.num_blocks[1], dim * 2, d.heads[1]));
    IRB_TRY(m->load_conv(blob, "up2_1.body.0.weight", dim * 4, dim * 2, m->up21, true));
    IRB_TRY(stage(m->dec1, "decoder_level1", d.num_blocks[0], dim * 2, d.heads[0]));
    IRB_TRY(stage(m->refine, "refinement", d.num_refinement_blocks, dim * 2, d.heads[0]));
    IRB_TRY(m->load_conv(blob, "output.weight", d.in_chans, dim * 2, m->outc));
    return 0;
}

void Restormer::layout(Bump& b, int n, int H, int W, WS& w) {
    const size_t E = esize();
    const size_t P1 = size_t(n) * H * W, P2 = P1 / 4, P3 = P1 / 16, P4 = P1 / 64;
    const int d = dim;
    w.cat1 = b.take(P1 * 2 * d * E);
    w.cat2 = b.take(P2 * 4 * d * E);
    w.cat3 = b.take(P3 * 8 * d * E);
    w.x2 = b.take(P2 * 2 * d * E);
    w.x3 = b.take(P3 * 4 * d * E);
    w.x4 = b.take(P4 * 8 * d * E);
    // scratch sized by the worst level: per-pixel channel counts shrink 2x while pixels shrink 4x, so the
    // full-resolution decoder level (C = 2d) dominates; take the max explicitly.
    size_t ln = 0, qkv = 0, qkv_pre = 0, hid = 0, gat = 0, weff = 0, gs = 0, at = 0;
    const size_t Ps[4] = {P1, P2, P3, P4};
    for (int l = 0; l < 4; ++l) {
        const int Cs[2] = {d << l, l == 0 ? 2 * d : d << l};
        for (int k = 0; k < 2; ++k) {
            const int C = Cs[k];
            const int heads = (k == 1 && l == 0) ? desc.heads[0] : desc.heads[l];
            const int hp = (int(C * desc.ffn_expansion_factor) + 63) / 64 * 64;
            ln = std::max(ln, Ps[l] * C * E);
            qkv = std::max(qkv, Ps[l] * 3 * C * E);
            // the tensors between the 1x1 and the depthwise convolutions exist only where the pair is not fused
            const bool fq = dt == DT_BF16 && fuse_pwdw && pwdw_fused_shape_ok(C, 3 * C, false, 0);
            const bool fg = dt == DT_BF16 && fuse_pwdw && pwdw_fused_shape_ok(C, 2 * hp, true, hp);
            if (!fq) qkv_pre = std::max(qkv_pre, Ps[l] * 3 * C * E);
            if (!fg) hid = std::max(hid, Ps[l] * 2 * hp * E);
            gat = std::max(gat, Ps[l] * hp * E);
            weff = std::max(weff, size_t(n) * C * C * E);
            at = std::max(at, size_t(n) * C * (C / heads) * 4);
            const int P = int(Ps[l] / n);
            gs = std::max(gs, size_t(n) * mdta_nchunks(P, heads, n) * (size_t(C) * (C / heads) + 2 * C) * 4);
        }
    }
    w.ln = b.take(ln);
    w.qkv_pre = b.take(qkv_pre);
    w.qkv = b.take(qkv);
    w.hidden = b.take(hid);
    w.gated = b.take(gat);
    w.weff = b.take(weff);
    w.stats = b.take<float2>(P1 * sizeof(float2));
    w.in8 = b.take<bf16>(P1 * 8 * sizeof(bf16));
    w.G = b.take<float>(gs);
    w.gs_bytes = gs;
    w.attn = b.take<float>(at);
    w.sumsq = nullptr;
}

size_t Restormer::forward_workspace(int n, int H, int W) {
    Bump b(nullptr);
    WS w;
    layout(b, n, H * scale, W * scale, w);
    const size_t up = scale == 1 ? 0 : ((size_t(n) * H * scale * W * scale * cin * 4 + 255) & ~size_t(255));
    return b.off + 256 + up;
}

int Restormer::run_gemm(const void* A, long long lda, long long M, int K, const void* Wt, int ldw, int N, int wrows,
                        Epi e, cudaStream_t s) {
    IRB_REQUIRE(M < (1ll << 31), "Restormer: too many pixels in one batch (lower tile_batch)");
    if (dt == DT_BF16)
        return tc_gemm((const bf16*)A, lda, int(M), K, 1, 0, (const bf16*)Wt, ldw, N, wrows, 0, e, s);
    return simt_gemm(A, DT_F32, lda, int(M), K, 1, 0, (const float*)Wt, ldw, N, 0, e, s);
}

int Restormer::run_conv(const ConvW& c, const void* x, int x_dt, long long ldx, int n, int H, int W, Epi e,
                        cudaStream_t s) {
    if (dt == DT_BF16 && x_dt == DT_BF16 && c.C >= 16)
        return conv3x3_bf16((const bf16*)x, ldx, n, H, W, c.C, (const bf16*)c.w, c.kpad, c.N, c.npad, e, s);
    IRB_REQUIRE(c.wf32 != nullptr, "Restormer: missing fp32 conv weights");
    return simt_conv3x3(x, x_dt, ldx, n, H, W, c.C, c.wf32, c.kpad, c.N, e, s);
}

// TransformerBlock.forward (restormer_arch.py:146-150) on the stream x [n*H*W, C] (row stride ldx), in place.
int Restormer::run_block(const RBlockW& w, void* x, long long ldx, int n, int H, int W, WS& ws, cudaStream_t s) {
    IRB_TRY(run_attn(w, x, ldx, n, H, W, ws, s));
    IRB_TRY(run_ffn(w, x, ldx, n, H, W, ws, s));
    if (w.has_spatial) {  // xrestormer_arch.py:259-260
        IRB_TRY(run_ocab(w, x, ldx, n, H, W, ws, s));
        IRB_TRY(run_ffn(*w.sp, x, ldx, n, H, W, ws, s));
    }
    return 0;
}

// x = x + project_out(attn(norm1(x)))   (restormer_arch.py:148)
int Restormer::run_attn(const RBlockW& w, void* x, long long ldx, int n, int H, int W, WS& ws, cudaStream_t s) {
    const int C = w.C, heads = w.heads, c = C / heads;
    const long long P = (long long)n * H * W;
    const int mode = desc.ln_bias_free ? 1 : 0;
    // --- x = x + project_out(attn(norm1(x))) ---
    // bf16, C <= 128: each "1x1 -> depthwise 3x3" pair is one kernel (pwdw_fused.cu): qkv_pre / hidden stay on the SM
    const bool fuse_qkv = dt == DT_BF16 && fuse_pwdw && pwdw_fused_shape_ok(C, 3 * C, false, 0);
    if (dt == DT_BF16) {  // norm1 folded into the qkv GEMM
        if (!ws.stats_valid) IRB_TRY(row_stats(x, dt, ldx, ws.stats, P, C, 1e-5f, s));
        if (fuse_qkv) {
            IRB_TRY(pwdw_fused_bf16(false, (const bf16*)x, ldx, ws.stats, (const bf16*)w.qkv_w, 3 * C, 0, w.qkv_cs,
                                    w.qkv_t, w.qkv_dwp, 3 * C, (bf16*)ws.qkv, 3 * C, n, H, W, C, s));
        } else {
            Epi e = make_epi(ws.qkv_pre, 3 * C, dt);
            e.ln_stats = ws.stats;
            e.ln_colsum = w.qkv_cs;
            e.bias = w.qkv_t;
            IRB_TRY(run_gemm(x, ldx, P, C, w.qkv_w, C, 3 * C, 3 * C, e, s));
        }
    } else {
        IRB_TRY(layernorm(x, dt, ldx, ws.ln, dt, C, P, C, w.n1w, w.n1b, mode, 1e-5f, s));
        IRB_TRY(run_gemm(ws.ln, C, P, C, w.qkv_w, C, 3 * C, 3 * C, make_epi(ws.qkv_pre, 3 * C, dt), s));
    }
    const int nch = mdta_nchunks(H * W, heads, n);
    float* G = ws.G;                                       // [n][heads][nch][c][c]
    float* ssq = ws.G + size_t(n) * heads * nch * c * c;   // [n][heads][nch][2][c]
    if (!fuse_qkv) IRB_TRY(dwconv3x3(ws.qkv_pre, dt, 3 * C, ws.qkv, 3 * C, n, H, W, 3 * C, w.qkv_dw, nullptr, 0, s));
    if (dt == DT_BF16)
        IRB_TRY(mdta_gram_tc((const bf16*)ws.qkv, 3 * C, n, H * W, C, heads, nch, G, ssq, s));
    else
        IRB_TRY(mdta_gram_simt(ws.qkv, DT_F32, 3 * C, n, H * W, C, heads, nch, G, ssq, s));
    IRB_TRY(mdta_softmax_weff(G, ssq, nch, w.temp, w.proj_w, n, C, heads, ws.weff, dt, ws.attn, s));
    {
        Epi e = make_epi(x, ldx, dt);
        e.res = x;
        e.ldr = ldx;
        e.res_dt = dt;
        const void* v = at(ws.qkv, 2 * C);
        ws.stats_valid = false;
        if (dt == DT_BF16 && tc_can_fuse_stats(C, e)) {  // norm2's statistics come out of this GEMM's epilogue
            e.stats_out = ws.stats;
            e.stats_eps = 1e-5f;
            ws.stats_valid = true;
        }
        if (dt == DT_BF16)
            IRB_TRY(tc_gemm((const bf16*)v, 3 * C, H * W, C, n, 0, (const bf16*)ws.weff, C, C, n * C, C, e, s));
        else
            IRB_TRY(simt_gemm(v, DT_F32, 3 * C, H * W, C, n, 0, (const float*)ws.weff, C, C, C, e, s));
    }
    return 0;
}

// x = x + project_out(gelu(dw(x1)) * dw(x2)), [x1 | x2] = project_in(norm2(x))   (restormer_arch.py:149, :88-93)
int Restormer::run_ffn(const RBlockW& w, void* x, long long ldx, int n, int H, int W, WS& ws, cudaStream_t s) {
    const int C = w.C;
    const long long P = (long long)n * H * W;
    const int mode = desc.ln_bias_free ? 1 : 0;
    const bool fuse_gate = dt == DT_BF16 && fuse_pwdw && pwdw_fused_shape_ok(C, 2 * w.hp, true, w.hp);
    // --- x = x + project_out(gelu(dw(x1)) * dw(x2)) ---
    if (dt == DT_BF16) {  // norm2 folded into project_in
        if (!ws.stats_valid) IRB_TRY(row_stats(x, dt, ldx, ws.stats, P, C, 1e-5f, s));
        if (fuse_gate) {
            IRB_TRY(pwdw_fused_bf16(true, (const bf16*)x, ldx, ws.stats, (const bf16*)w.pin_w, 2 * w.hp, w.hp, w.pin_cs,
                                    w.pin_t, w.ffn_dwp, 2 * w.hp, (bf16*)ws.gated, w.hp, n, H, W, C, s));
        } else {
            Epi e = make_epi(ws.hidden, 2 * w.hp, dt);
            e.ln_stats = ws.stats;
            e.ln_colsum = w.pin_cs;
            e.bias = w.pin_t;
            IRB_TRY(run_gemm(x, ldx, P, C, w.pin_w, C, 2 * w.hp, 2 * w.hp, e, s));
        }
    } else {
        IRB_TRY(layernorm(x, dt, ldx, ws.ln, dt, C, P, C, w.n2w, w.n2b, mode, 1e-5f, s));
        IRB_TRY(run_gemm(ws.ln, C, P, C, w.pin_w, C, 2 * w.hp, 2 * w.hp, make_epi(ws.hidden, 2 * w.hp, dt), s));
    }
    if (!fuse_gate) IRB_TRY(dwconv3x3_gate(ws.hidden, dt, 2 * w.hp, ws.gated, w.hp, n, H, W, w.hp, w.ffn_dw, s));
    {
        Epi e = make_epi(x, ldx, dt);
        e.res = x;
        e.ldr = ldx;
        e.res_dt = dt;
        ws.stats_valid = false;
        if (dt == DT_BF16 && tc_can_fuse_stats(C, e)) {  // the next block's norm1 statistics
            e.stats_out = ws.stats;
            e.stats_eps = 1e-5f;
            ws.stats_valid = true;
        }
        IRB_TRY(run_gemm(ws.gated, w.hp, P, w.hp, w.pout_w, w.hp, C, C, e, s));
    }
    return 0;
}

// x = x + project_out(OCA(qkv(norm3(x))))   (xrestormer_arch.py:259, :190-237)
int Restormer::run_ocab(const RBlockW& w, void* x, long long ldx, int n, int H, int W, WS& ws, cudaStream_t s) {
    const int C = w.C, inner = w.inner;
    const long long P = (long long)n * H * W;
    const int mode = desc.ln_bias_free ? 1 : 0;
    if (dt == DT_BF16) {  // norm3 folded into the qkv GEMM
        if (!ws.stats_valid) IRB_TRY(row_stats(x, dt, ldx, ws.stats, P, C, 1e-5f, s));
        Epi e = make_epi(ws.qkv, 3 * inner, dt);
        e.ln_stats = ws.stats;
        e.ln_colsum = w.oc_cs;
        e.bias = w.oc_t;
        IRB_TRY(run_gemm(x, ldx, P, C, w.oc_qkv_w, C, 3 * inner, 3 * inner, e, s));
    } else {
        IRB_TRY(layernorm(x, dt, ldx, ws.ln, dt, C, P, C, w.n3w, w.n3b, mode, 1e-5f, s));
        IRB_TRY(run_gemm(ws.ln, C, P, C, w.oc_qkv_w, C, 3 * inner, 3 * inner, make_epi(ws.qkv, 3 * inner, dt), s));
    }
    IRB_TRY(ocab_attention(ws.qkv, dt, 3 * inner, ws.ln, inner, n, H, W, w.sheads, desc.spatial_dim_head,
                           desc.window_size, w.ows, w.rel_h, w.rel_w, s));
    Epi e = make_epi(x, ldx, dt);
    e.res = x;
    e.ldr = ldx;
    e.res_dt = dt;
    ws.stats_valid = false;
    if (dt == DT_BF16 && tc_can_fuse_stats(C, e)) {  // norm4's statistics
        e.stats_out = ws.stats;
        e.stats_eps = 1e-5f;
        ws.stats_valid = true;
    }
    return run_gemm(ws.ln, inner, P, inner, w.oc_proj_w, inner, C, C, e, s);
}

// F.interpolate(x, scale_factor=S, mode='bilinear', align_corners=False) on channels-last fp32 (xrestormer_arch.py:383-384,
// the scale-4 SR variant of X-Restormer up-samples its INPUT and then runs the scale-1 network on it)
__global__ void bilinear_up_kernel(const float* __restrict__ in, float* __restrict__ out, int n, int H, int W, int c, int S) {
    const int Ho = H * S, Wo = W * S;
    const long long total = (long long)n * Ho * Wo;
    const float inv = 1.0f / float(S);
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int ox = int(i % Wo), oy = int((i / Wo) % Ho);
        const long long b = i / ((long long)Wo * Ho);
        // ATen area_pixel_compute_source_index (align_corners=False): src = max((dst + 0.5) * scale - 0.5, 0)
        const float sy = fmaxf((oy + 0.5f) * inv - 0.5f, 0.f), sx = fmaxf((ox + 0.5f) * inv - 0.5f, 0.f);
        const int y0 = int(sy), x0 = int(sx);
        const int y1 = y0 + (y0 < H - 1 ? 1 : 0), x1 = x0 + (x0 < W - 1 ? 1 : 0);
        const float ly = sy - y0, lx = sx - x0, hy = 1.f - ly, hx = 1.f - lx;
        const float* p00 = in + ((b * H + y0) * W + x0) * c;
        const float* p01 = in + ((b * H + y0) * W + x1) * c;
        const float* p10 = in + ((b * H + y1) * W + x0) * c;
        const float* p11 = in + ((b * H + y1) * W + x1) * c;
        for (int ch = 0; ch < c; ++ch)
            out[i * c + ch] = hy * (hx * p00[ch] + lx * p01[ch]) + ly * (hx * p10[ch] + lx * p11[ch]);
    }
}

int Restormer::forward(const float* in, float* out, int n, int H, int W, void* wsp, size_t ws_bytes, cudaStream_t s) {
    if (scale == 1) return forward_core(in, out, n, H, W, wsp, ws_bytes, s);
    // X-Restormer SR: bilinear x`scale` of the input, then the network at the up-sampled size (:383-412)
    IRB_REQUIRE(ws_bytes >= forward_workspace(n, H, W), "XRestormer: workspace too small");
    const size_t up_bytes = (size_t(n) * H * scale * W * scale * cin * 4 + 255) & ~size_t(255);
    float* up = reinterpret_cast<float*>(wsp);
    const long long total = (long long)n * H * scale * W * scale;
    const int blocks = int(ceil_div_ll(total, 256) > 148 * 16 ? 148 * 16 : ceil_div_ll(total, 256));
    bilinear_up_kernel<<<blocks, 256, 0, s>>>(in, up, n, H, W, cin, scale);
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("bilinear_up", cin, scale, 0.0, double(total) * cin * 4.0 * (1.0 + 1.0 / (scale * scale)), s);
    return forward_core(up, out, n, H * scale, W * scale, reinterpret_cast<char*>(wsp) + up_bytes, ws_bytes - up_bytes, s);
}

int Restormer::forward_core(const float* in, float* out, int n, int H, int W, void* wsp, size_t ws_bytes,
                            cudaStream_t s) {
    IRB_REQUIRE(H % 8 == 0 && W % 8 == 0, "Restormer: tile height/width must be multiples of 8");
    IRB_REQUIRE((long long)H * W < (1ll << 31) / 8, "Restormer: tile too large");
    Bump b(wsp);
    WS ws;
    layout(b, n, H, W, ws);
    IRB_REQUIRE(ws_bytes >= b.off, "Restormer: workspace too small");
    const int d = dim;
    auto blocks = [&](std::vector<RBlockW>& v, void* x, long long ldx, int h, int w) -> int {
        ws.stats_valid = false;  // the stream was just written by a convolution
        for (auto& blk : v) IRB_TRY(run_block(blk, x, ldx, n, h, w, ws, s));
        return 0;
    };
    auto shuffle_epi = [&](void* o, long long ldo, int ps, int h, int w) {
        Epi e = make_epi(o, ldo, dt);
        e.ps = ps;
        e.ps_h = h;
        e.ps_w = w;
        return e;
    };
    void* e1 = at(ws.cat1, d);       // encoder-1 stream: channels [d,2d) of cat1
    void* e2 = at(ws.cat2, 2 * d);   // encoder-2 stream: channels [2d,4d) of cat2
    void* e3 = at(ws.cat3, 4 * d);   // encoder-3 stream: channels [4d,8d) of cat3
    const int H2 = H / 2, W2 = W / 2, H3 = H / 4, W3 = W / 4, H4 = H / 8, W4 = W / 8;

    if (patch.wsplit) {  // :247 on the tensor core: x = hi + lo in bf16, both halves share the weights
        IRB_TRY(split_hi_lo_bf16(in, cin, ws.in8, (long long)n * H * W, s));
        IRB_TRY(tc_conv3x3(ws.in8, 8, n, H, W, 8, (const bf16*)patch.wsplit, 64, patch.N, patch.npad,
                           make_epi(e1, 2 * d, dt), s));
    } else {
        IRB_TRY(run_conv(patch, in, DT_F32, cin, n, H, W, make_epi(e1, 2 * d, dt), s));           // :247
    }
    IRB_TRY(blocks(enc1, e1, 2 * d, H, W));                                                        // :248
    IRB_TRY(run_conv(down12, e1, dt, 2 * d, n, H, W, shuffle_epi(e2, 4 * d, 2, H, W), s));         // :250
    IRB_TRY(blocks(enc2, e2, 4 * d, H2, W2));                                                      // :251
    IRB_TRY(run_conv(down23, e2, dt, 4 * d, n, H2, W2, shuffle_epi(e3, 8 * d, 2, H2, W2), s));     // :253
    IRB_TRY(blocks(enc3, e3, 8 * d, H3, W3));                                                      // :254
    IRB_TRY(run_conv(down34, e3, dt, 8 * d, n, H3, W3, shuffle_epi(ws.x4, 8 * d, 2, H3, W3), s));  // :256
    IRB_TRY(blocks(latent, ws.x4, 8 * d, H4, W4));                                                 // :257
    IRB_TRY(run_conv(up43, ws.x4, dt, 8 * d, n, H4, W4, shuffle_epi(ws.cat3, 8 * d, up43.ps, H4, W4), s));  // :259-260
    IRB_TRY(run_gemm(ws.cat3, 8 * d, (long long)n * H3 * W3, 8 * d, reduce3, 8 * d, 4 * d, 4 * d,
                     make_epi(ws.x3, 4 * d, dt), s));                                              // :261
    IRB_TRY(blocks(dec3, ws.x3, 4 * d, H3, W3));                                                   // :262
    IRB_TRY(run_conv(up32, ws.x3, dt, 4 * d, n, H3, W3, shuffle_epi(ws.cat2, 4 * d, up32.ps, H3, W3), s));  // :264-265
    IRB_TRY(run_gemm(ws.cat2, 4 * d, (long long)n * H2 * W2, 4 * d, reduce2, 4 * d, 2 * d, 2 * d,
                     make_epi(ws.x2, 2 * d, dt), s));                                              // :266
    IRB_TRY(blocks(dec2, ws.x2, 2 * d, H2, W2));                                                   // :267
    IRB_TRY(run_conv(up21, ws.x2, dt, 2 * d, n, H2, W2, shuffle_epi(ws.cat1, 2 * d, up21.ps, H2, W2), s));  // :269-270
    IRB_TRY(blocks(dec1, ws.cat1, 2 * d, H, W));                                                   // :271
    IRB_TRY(blocks(refine, ws.cat1, 2 * d, H, W));                                                 // :273
    {
        Epi e = make_epi(out, cin, DT_F32);  // output conv + inp_img (:281), fp32 out
        e.res = in;
        e.ldr = cin;
        e.res_dt = DT_F32;
        IRB_TRY(run_conv(outc, ws.cat1, dt, 2 * d, n, H, W, e, s));
    }
    return 0;
}

}  // namespace irb
