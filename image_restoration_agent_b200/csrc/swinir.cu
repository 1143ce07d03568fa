// SwinIR forward (network_swinir.py:805-840) as a sequence of irb200 kernels on channels-last tokens.
//
// Token stream x is [n*H*W, ldC] (ldC = C rounded up to 16 so every row is 16-byte aligned for TMA; the pad
// channels are never read: TMA tensor maps and LayerNorm are told C, not ldC). q/k/v use a head-padded layout
// [3][heads][dp] (dp = head_dim rounded up to 16, pad weights are zero, q weights pre-multiplied by head_dim^-0.5).
#include "model.h"
#include <algorithm>
#include <math.h>
#include <string.h>

namespace irb {

struct LinW {
    void* w;     // storage dtype [npad][kpad]
    float* b;    // [N] fp32 (nullptr = no bias)
    int N, K, kpad, npad;
};
struct Conv3W {
    void* w;       // storage dtype [npad][9*kpad]
    float* wf32;   // fp32 copy when the FMA kernel runs it
    float* b;
    int N, C, kpad, npad;
};
struct ResiW {  // '1conv' (c0 only) or '3conv' (c0 -> lrelu -> lin -> lrelu -> c2)
    Conv3W c0, c2;
    LinW lin;
};
struct STLW {
    float *n1w, *n1b, *n2w, *n2b, *bias_table;
    LinW qkv, proj, fc1, fc2;
    int shift, heads;
};

// reflect pad (network_swinir.py:783-788) + (x - mean) * img_range (:809-810) on channels-last fp32
__global__ void swin_prep_kernel(const float* __restrict__ in, float* __restrict__ out, int n, int H0, int W0, int H,
                                 int W, int c, float m0, float m1, float m2, float range) {
    const long long total = (long long)n * H * W;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int x = int(i % W), y = int((i / W) % H);
        const long long b = i / ((long long)W * H);
        const int sy = y < H0 ? y : 2 * (H0 - 1) - y, sx = x < W0 ? x : 2 * (W0 - 1) - x;
        const float* s = in + ((b * H0 + sy) * W0 + sx) * c;
        for (int ch = 0; ch < c; ++ch) {
            const float mean = ch == 0 ? m0 : (ch == 1 ? m1 : m2);
            out[i * c + ch] = (s[ch] - mean) * range;
        }
    }
}
__global__ void crop_nhwc_kernel(const float* __restrict__ in, float* __restrict__ out, int n, int H, int W, int Hc,
                                 int Wc, int c) {
    const long long total = (long long)n * Hc * Wc * c;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int ch = int(i % c);
        long long p = i / c;
        const int x = int(p % Wc);
        p /= Wc;
        const int y = int(p % Hc);
        const long long b = p / Hc;
        out[i] = in[((b * H + y) * W + x) * c + ch];
    }
}
static int grid_for(long long total) {
    const long long b = ceil_div_ll(total, 256);
    return int(b > 148 * 16 ? 148 * 16 : b);
}

struct SwinIR : Model {
    int C, ldC, ws, ws_eff, can_shift, hid, dp, Cq;
    float mean[3];
    Conv3W conv_first, conv_before_up, conv_up1, conv_up2, conv_hr, conv_last;
    std::vector<Conv3W> upsample;
    float *pen_w, *pen_b, *norm_w, *norm_b, *d_mean;
    std::vector<std::vector<STLW>> layers;
    std::vector<ResiW> layer_conv;
    ResiW conv_after_body;

    int tile_multiple() const override { return desc.window_size; }
    size_t forward_workspace(int n, int H, int W) override;
    int forward(const float* in, float* out, int n, int H, int W, void* ws, size_t ws_bytes,
                cudaStream_t stream) override;

    struct WS {
        float *xin, *out_tmp;
        void *xa, *xb, *ffirst, *ln, *qkv, *attn, *hidden, *q4a, *q4b, *t64, *u1, *u2, *u3;
    };
    void layout(Bump& b, int n, int H0, int W0, WS& w);
    void padded(int H0, int W0, int& H, int& W) const {
        const int m = desc.window_size;
        H = (H0 + m - 1) / m * m;
        W = (W0 + m - 1) / m * m;
    }
    int load_lin(const Blob& blob, const std::string& name, int N, int K, LinW& o);
    int load_conv(const Blob& blob, const std::string& name, int N, int Cin, Conv3W& o);
    int load_resi(const Blob& blob, const std::string& name, ResiW& o);
    int load_qkv(const Blob& blob, const std::string& name, int heads, LinW& o);
    int load_norm(const Blob& blob, const std::string& name, float** w, float** b);
    int gemm(const void* A, long long lda, long long M, const LinW& L, Epi e, cudaStream_t s);
    int conv(const Conv3W& c, const void* x, int x_dt, long long ldx, int n, int H, int W, Epi e, cudaStream_t s);
    int resi(const ResiW& r, const void* x, long long ldx, int n, int H, int W, Epi e, WS& w, cudaStream_t s);
    int stl(const STLW& L, const void* xin, void* xout, int n, int H, int W, WS& w, cudaStream_t s);
};

int SwinIR::load_norm(const Blob& blob, const std::string& name, float** w, float** b) {
    BlobTensor t;
    IRB_TRY(blob.get(name + ".weight", t, C));
    IRB_TRY(upload_f32(t.data, C, w));
    IRB_TRY(blob.get(name + ".bias", t, C));
    IRB_TRY(upload_f32(t.data, C, b));
    return 0;
}

int SwinIR::load_lin(const Blob& blob, const std::string& name, int N, int K, LinW& o) {
    BlobTensor t;
    const bool conv1x1 = blob.has(name + ".weight") && blob.t.at(name + ".weight").ndim == 4;
    if (conv1x1)
        IRB_TRY(blob.get(name + ".weight", t, N, K, 1, 1));
    else
        IRB_TRY(blob.get(name + ".weight", t, N, K));
    std::vector<float> w = pack_linear(t.data, N, K, o.kpad, o.npad);
    IRB_TRY(upload_act(w.data(), w.size(), &o.w));
    IRB_TRY(blob.get(name + ".bias", t, N));
    IRB_TRY(upload_f32(t.data, N, &o.b));
    o.N = N;
    o.K = K;
    return 0;
}

// qkv Linear [3C][C] (rows [3][heads][d], network_swinir.py:121-122) -> rows [3][heads][dp], q rows * d^-0.5
int SwinIR::load_qkv(const Blob& blob, const std::string& name, int heads, LinW& o) {
    BlobTensor tw, tb;
    IRB_TRY(blob.get(name + ".weight", tw, 3 * C, C));
    IRB_TRY(blob.get(name + ".bias", tb, 3 * C));
    const int d = C / heads;
    const float scale = 1.0f / sqrtf(float(d));  // head_dim ** -0.5 (:85-86)
    std::vector<float> w(size_t(3) * Cq * C, 0.f), b(size_t(3) * Cq, 0.f);
    for (int sec = 0; sec < 3; ++sec)
        for (int h = 0; h < heads; ++h)
            for (int dd = 0; dd < d; ++dd) {
                const int src = sec * C + h * d + dd, dst = sec * Cq + h * dp + dd;
                const float s = sec == 0 ? scale : 1.0f;
                for (int k = 0; k < C; ++k) w[size_t(dst) * C + k] = tw.data[size_t(src) * C + k] * s;
                b[dst] = tb.data[src] * s;
            }
    std::vector<float> wp = pack_linear(w.data(), 3 * Cq, C, o.kpad, o.npad);
    IRB_TRY(upload_act(wp.data(), wp.size(), &o.w));
    IRB_TRY(upload_f32(b.data(), b.size(), &o.b));
    o.N = 3 * Cq;
    o.K = C;
    return 0;
}

int SwinIR::load_conv(const Blob& blob, const std::string& name, int N, int Cin, Conv3W& o) {
    BlobTensor t;
    IRB_TRY(blob.get(name + ".weight", t, N, Cin, 3, 3));
    std::vector<float> w = pack_conv3x3(t, o.kpad, o.npad);
    o.N = N;
    o.C = Cin;
    IRB_TRY(upload_act(w.data(), w.size(), &o.w));
    o.wf32 = nullptr;
    if (dt == DT_F32)
        o.wf32 = reinterpret_cast<float*>(o.w);
    else if (Cin < 16)
        IRB_TRY(upload_f32(w.data(), w.size(), &o.wf32));
    IRB_TRY(blob.get(name + ".bias", t, N));
    IRB_TRY(upload_f32(t.data, N, &o.b));
    return 0;
}

int SwinIR::load_resi(const Blob& blob, const std::string& name, ResiW& o) {
    if (!desc.resi_3conv) return load_conv(blob, name, C, C, o.c0);
    IRB_TRY(load_conv(blob, name + ".0", C / 4, C, o.c0));
    IRB_TRY(load_lin(blob, name + ".2", C / 4, C / 4, o.lin));
    IRB_TRY(load_conv(blob, name + ".4", C, C / 4, o.c2));
    return 0;
}

int build_swinir(const irb_model_desc& d, const Blob& blob, int device, Model** out) {
    SwinIR* m = new SwinIR();
    *out = m;
    m->desc = d;
    m->dt = d.precision == IRB_BF16 ? DT_BF16 : DT_F32;
    m->device = device;
    m->cin = d.in_chans;
    m->scale = d.upscale;
    IRB_REQUIRE(d.in_chans == 1 || d.in_chans == 3, "SwinIR: in_chans must be 1 or 3");
    IRB_REQUIRE(d.embed_dim > 0 && d.embed_dim % 4 == 0 && d.embed_dim <= 512, "SwinIR: embed_dim must be %4, <= 512");
    IRB_REQUIRE(d.window_size >= 1 && d.window_size <= 8, "SwinIR: window_size must be <= 8");
    IRB_REQUIRE(d.num_layers >= 1 && d.num_layers <= 16, "SwinIR: 1..16 RSTB layers");
    IRB_REQUIRE(d.upscale >= 1 && (d.upscale & (d.upscale - 1)) == 0, "SwinIR: upscale must be a power of two");
    IRB_REQUIRE(d.upsampler == IRB_UPSAMPLER_NONE || d.upscale > 1 || d.upsampler == IRB_UPSAMPLER_PIXELSHUFFLE ||
                    d.upsampler == IRB_UPSAMPLER_NEAREST_CONV, "SwinIR: bad upsampler");
    if (d.upsampler == IRB_UPSAMPLER_NONE) IRB_REQUIRE(d.upscale == 1, "SwinIR: upsampler '' needs upscale 1");
    if (d.upsampler == IRB_UPSAMPLER_NEAREST_CONV)
        IRB_REQUIRE(d.upscale == 2 || d.upscale == 4, "SwinIR: nearest+conv supports x2 / x4");
    const int C = m->C = d.embed_dim;
    m->ldC = (C + 15) / 16 * 16;
    m->ws = d.window_size;
    m->can_shift = d.img_size > d.window_size;  // network_swinir.py:193-196 (constructor img_size)
    m->ws_eff = m->can_shift ? d.window_size : d.img_size;
    m->hid = C * d.mlp_ratio;
    IRB_REQUIRE(m->hid % 8 == 0, "SwinIR: mlp hidden dim must be a multiple of 8");
    const int heads0 = d.num_heads[0];
    for (int r = 0; r < d.num_layers; ++r)
        IRB_REQUIRE(d.num_heads[r] == heads0 && d.depths[r] >= 1, "SwinIR: all RSTBs must use the same head count");
    IRB_REQUIRE(heads0 > 0 && C % heads0 == 0 && C / heads0 <= 32, "SwinIR: head_dim must be <= 32");
    m->dp = (C / heads0 + 15) / 16 * 16;
    m->Cq = heads0 * m->dp;
    m->mean[0] = m->mean[1] = m->mean[2] = 0.f;
    if (d.in_chans == 3) {  // :656-657
        m->mean[0] = 0.4488f;
        m->mean[1] = 0.4371f;
        m->mean[2] = 0.4040f;
    }
    IRB_TRY(m->upload_f32(m->mean, 3, &m->d_mean));
    IRB_TRY(m->load_conv(blob, "conv_first", C, d.in_chans, m->conv_first));
    IRB_TRY(m->load_norm(blob, "patch_embed.norm", &m->pen_w, &m->pen_b));
    m->layers.resize(d.num_layers);
    m->layer_conv.resize(d.num_layers);
    for (int r = 0; r < d.num_layers; ++r) {
        m->layers[r].resize(d.depths[r]);
        for (int b = 0; b < d.depths[r]; ++b) {
            STLW& L = m->layers[r][b];
            const std::string p = "layers." + std::to_string(r) + ".residual_group.blocks." + std::to_string(b);
            L.heads = d.num_heads[r];
            L.shift = (b % 2 == 1 && m->can_shift) ? m->ws / 2 : 0;  // :383
            IRB_TRY(m->load_norm(blob, p + ".norm1", &L.n1w, &L.n1b));
            IRB_TRY(m->load_norm(blob, p + ".norm2", &L.n2w, &L.n2b));
            BlobTensor t;
            const int nb = (2 * m->ws_eff - 1) * (2 * m->ws_eff - 1);
            IRB_TRY(blob.get(p + ".attn.relative_position_bias_table", t, nb, L.heads));
            IRB_TRY(m->upload_f32(t.data, size_t(nb) * L.heads, &L.bias_table));
            IRB_TRY(m->load_qkv(blob, p + ".attn.qkv", L.heads, L.qkv));
            IRB_TRY(m->load_lin(blob, p + ".attn.proj", C, C, L.proj));
            IRB_TRY(m->load_lin(blob, p + ".mlp.fc1", m->hid, C, L.fc1));
            IRB_TRY(m->load_lin(blob, p + ".mlp.fc2", C, m->hid, L.fc2));
        }
        IRB_TRY(m->load_resi(blob, "layers." + std::to_string(r) + ".conv", m->layer_conv[r]));
    }
    IRB_TRY(m->load_norm(blob, "norm", &m->norm_w, &m->norm_b));
    IRB_TRY(m->load_resi(blob, "conv_after_body", m->conv_after_body));
    if (d.upsampler == IRB_UPSAMPLER_PIXELSHUFFLE) {
        IRB_TRY(m->load_conv(blob, "conv_before_upsample.0", 64, C, m->conv_before_up));
        int steps = 0;
        for (int s = d.upscale; s > 1; s >>= 1) ++steps;
        m->upsample.resize(steps);
        for (int i = 0; i < steps; ++i)
            IRB_TRY(m->load_conv(blob, "upsample." + std::to_string(2 * i), 256, 64, m->upsample[i]));
        IRB_TRY(m->load_conv(blob, "conv_last", d.in_chans, 64, m->conv_last));
    } else if (d.upsampler == IRB_UPSAMPLER_NEAREST_CONV) {
        IRB_TRY(m->load_conv(blob, "conv_before_upsample.0", 64, C, m->conv_before_up));
        IRB_TRY(m->load_conv(blob, "conv_up1", 64, 64, m->conv_up1));
        if (d.upscale == 4) IRB_TRY(m->load_conv(blob, "conv_up2", 64, 64, m->conv_up2));
        IRB_TRY(m->load_conv(blob, "conv_hr", 64, 64, m->conv_hr));
        IRB_TRY(m->load_conv(blob, "conv_last", d.in_chans, 64, m->conv_last));
    } else {
        IRB_TRY(m->load_conv(blob, "conv_last", d.in_chans, C, m->conv_last));
    }
    return 0;
}
