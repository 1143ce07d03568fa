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
    float* cs;   // bf16 mode, LayerNorm folded in: colsum[n] = sum_k bf16(W[n,k] * gain[k]); b then holds b + W.ln_bias
    int N, K, kpad, npad;
    void* w3 = nullptr;  // fp32 mode: bf16 [npad][k3pad] = Wh | Wm | Wl | Wh | Wm | Wh (W = Wh + Wm + Wl), see split3_bf16
    int k3pad = 0;
};

// the three bf16 terms of an fp32 value
static inline void split3_host(float v, float* t) {
    t[0] = __bfloat162float(__float2bfloat16_rn(v));
    const float r1 = v - t[0];
    t[1] = __bfloat162float(__float2bfloat16_rn(r1));
    t[2] = __bfloat162float(__float2bfloat16_rn(r1 - t[1]));
}
// W [N][K] (row stride ldw) -> [npad][k3pad] with segments Wh | Wm | Wl | Wh | Wm | Wh: pairs with the A segments
// h | h | h | m | m | l into h.Wh + h.Wm + h.Wl + m.Wh + m.Wm + l.Wh
static std::vector<float> pack_split3(const float* w, int N, int K, int ldw, int npad, int& k3pad) {
    static const int term[6] = {0, 1, 2, 0, 1, 0};
    k3pad = (6 * K + 7) / 8 * 8;
    std::vector<float> o(size_t(npad) * k3pad, 0.f);
    for (int n = 0; n < N; ++n)
        for (int k = 0; k < K; ++k) {
            float t[3];
            split3_host(w[size_t(n) * ldw + k], t);
            for (int sg = 0; sg < 6; ++sg) o[size_t(n) * k3pad + sg * K + k] = t[term[sg]];
        }
    return o;
}

// W [N][K] -> W * diag(gain) in place; colsum of the bf16-rounded result; bias += W . ln_bias
static void fold_ln(std::vector<float>& w, std::vector<float>& bias, int N, int K, const float* gain,
                    const float* ln_bias, std::vector<float>& colsum) {
    colsum.assign(N, 0.f);
    for (int n = 0; n < N; ++n) {
        double cs = 0.0, tt = 0.0;
        for (int k = 0; k < K; ++k) {
            float& e = w[size_t(n) * K + k];
            tt += double(e) * ln_bias[k];
            e *= gain[k];
            cs += double(__bfloat162float(__float2bfloat16_rn(e)));
        }
        colsum[n] = float(cs);
        bias[n] += float(tt);
    }
}
struct Conv3W {
    void* w;       // storage dtype [npad][9*kpad]
    float* wf32;   // fp32 copy when the FMA kernel runs it
    void* wsplit;  // bf16 mode, C <= 4: weights of the hi/lo-split tensor-core first conv
    float* b;
    int N, C, kpad, npad;
    int ps;        // PixelShuffle epilogue mode for this conv: 0 none, 1 reference row order, 3 sub-pixel-major rows
    void* w3 = nullptr;  // fp32 mode: bf16 [npad][9 * kpad3], per tap the six split segments of 6 * C input channels
    int kpad3 = 0;
};
struct ResiW {  // '1conv' (c0 only) or '3conv' (c0 -> lrelu -> lin -> lrelu -> c2)
    Conv3W c0, c2;
    LinW lin;
};
struct STLW {
    float *n1w, *n1b, *n2w, *n2b, *bias_table;
    LinW qkv, proj, fc1, fc2;
    int shift, heads;
    // HAT only (hat_arch.py:61-74 CAB): conv3x3 C -> C/r, GELU, conv3x3 C/r -> C, channel attention (two 1x1 convolutions)
    Conv3W cab0, cab2;
    float *ca_w1, *ca_b1, *ca_w2, *ca_b2;
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
// DiffBIR stage 1 (DiffBIR/model/swinir.py:716-724): (x - mean) * img_range, then PixelUnshuffle(us):
// out[b, y, x, c * us^2 + i * us + j] = in[b, y * us + i, x * us + j, c]
template <typename T>
__global__ void swin_prep_unshuffle_kernel(const float* __restrict__ in, T* __restrict__ out, int n, int H, int W, int c,
                                           int us, float m0, float m1, float m2, float range) {
    const int co = c * us * us;
    const long long total = (long long)n * H * W * co;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int oc = int(i % co);
        long long px = i / co;
        const int x = int(px % W);
        px /= W;
        const int y = int(px % H);
        const long long b = px / H;
        const int ch = oc / (us * us), sub = oc - ch * us * us, si = sub / us, sj = sub - si * us;
        const float mean = ch == 0 ? m0 : (ch == 1 ? m1 : m2);
        const float v = (in[((b * (H * us) + y * us + si) * (long long)(W * us) + x * us + sj) * c + ch] - mean) * range;
        if (sizeof(T) == 4)
            reinterpret_cast<float*>(out)[i] = v;
        else
            reinterpret_cast<bf16*>(out)[i] = __float2bfloat16_rn(v);
    }
}
// UpsampleOneStep's PixelShuffle(s) (network_swinir.py:594-612) + x / img_range + mean (:838):
// out[b, y*s + i, x*s + j, c] = t[b, y, x, c*s*s + i*s + j] / range + mean[c]
__global__ void pixel_shuffle_out_kernel(const float* __restrict__ t, float* __restrict__ out, int n, int H, int W, int c,
                                         int s, float m0, float m1, float m2, float inv_range) {
    const int Ho = H * s, Wo = W * s;
    const long long total = (long long)n * Ho * Wo * c;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int ch = int(i % c);
        long long px = i / c;
        const int xo = int(px % Wo);
        px /= Wo;
        const int yo = int(px % Ho);
        const long long b = px / Ho;
        const int y = yo / s, si = yo - y * s, x = xo / s, sj = xo - x * s;
        const float mean = ch == 0 ? m0 : (ch == 1 ? m1 : m2);
        out[i] = t[((b * H + y) * W + x) * (long long)(c * s * s) + ch * s * s + si * s + sj] * inv_range + mean;
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
    bool hat = false;  // HAT (hat_arch.py): HAB = STL + CAB branch, one OCAB per residual group, windows up to 16
    int ows = 0, c_cab = 0, ld_cab = 0, c_sq = 0;  // HAT: overlapping window, C / compress_ratio (+ row stride), C / squeeze_factor
    std::vector<STLW> ocab;                         // HAT: layers[r].residual_group.overlap_attn
    int us = 1;   // PixelUnshuffle factor of conv_first (DiffBIR stage 1: 8), 1 = none
    int tsc = 1;  // scale of the reconstruction tail (desc.upscale); the model's in -> out scale is tsc / us
    Conv3W conv_up3;
    float mean[3];
    Conv3W conv_first, conv_before_up, conv_up1, conv_up2, conv_hr, conv_last;
    std::vector<Conv3W> upsample;
    float *pen_w, *pen_b, *norm_w, *norm_b, *d_mean;
    std::vector<std::vector<STLW>> layers;
    std::vector<ResiW> layer_conv;
    ResiW conv_after_body;

    int tile_multiple() const override { return desc.window_size * us; }  // HAT: the network does not pad
    size_t forward_workspace(int n, int H, int W) override;
    int forward(const float* in, float* out, int n, int H, int W, void* ws, size_t ws_bytes,
                cudaStream_t stream) override;

    struct WS {
        float *xin, *out_tmp;
        void *xa, *xb, *ffirst, *ln, *qkv, *attn, *hidden, *q4a, *q4b, *t64, *u1, *u2, *u3;
        void* a3;                  // fp32 mode: the split A operand of the tensor-core GEMM / convolution in flight
        size_t a3_bytes;
        void *cab1, *cab;          // HAT: the two CAB convolution outputs
        float *ca_sums, *ca_scale; // HAT: per-slab channel sums, per-image channel scale
        float2* stats;
        bool stats_valid;  // stats already describe the stream the next block normalises (written by the producing GEMM)
    };
    WS* cur_ws = nullptr;  // the workspace of the forward in flight (fp32 mode: its split buffer)
    void layout(Bump& b, int n, int H0, int W0, WS& w);
    void padded(int H0, int W0, int& H, int& W) const {
        const int m = desc.window_size;
        H = (H0 + m - 1) / m * m;
        W = (W0 + m - 1) / m * m;
    }
    int load_lin(const Blob& blob, const std::string& name, int N, int K, LinW& o, const std::string& fold_norm = "");
    int load_conv(const Blob& blob, const std::string& name, int N, int Cin, Conv3W& o, bool pixelshuffle = false);
    int load_resi(const Blob& blob, const std::string& name, ResiW& o);
    int load_qkv(const Blob& blob, const std::string& name, int heads, LinW& o, const std::string& fold_norm);
    int load_norm(const Blob& blob, const std::string& name, float** w, float** b);
    int gemm(const void* A, long long lda, long long M, const LinW& L, Epi e, cudaStream_t s);
    int conv(const Conv3W& c, const void* x, int x_dt, long long ldx, int n, int H, int W, Epi e, cudaStream_t s);
    int resi(const ResiW& r, const void* x, long long ldx, int n, int H, int W, Epi e, WS& w, cudaStream_t s);
    int stl(const STLW& L, const void* xin, void* xout, int n, int H, int W, WS& w, cudaStream_t s);
    int mlp(const STLW& L, void* x, long long P, WS& w, cudaStream_t s);
    int hab(const STLW& L, const void* xin, void* xout, int n, int H, int W, WS& w, cudaStream_t s);
    int ocab_block(const STLW& L, void* x, int n, int H, int W, WS& w, cudaStream_t s);
    int load_block(const Blob& blob, const std::string& p, const std::string& attn, int heads, int table_rows, bool fold1,
                   STLW& L);
};

int SwinIR::load_norm(const Blob& blob, const std::string& name, float** w, float** b) {
    BlobTensor t;
    IRB_TRY(blob.get(name + ".weight", t, C));
    IRB_TRY(upload_f32(t.data, C, w));
    IRB_TRY(blob.get(name + ".bias", t, C));
    IRB_TRY(upload_f32(t.data, C, b));
    return 0;
}

int SwinIR::load_lin(const Blob& blob, const std::string& name, int N, int K, LinW& o, const std::string& fold_norm) {
    BlobTensor t, tb;
    const bool conv1x1 = blob.has(name + ".weight") && blob.t.at(name + ".weight").ndim == 4;
    if (conv1x1)
        IRB_TRY(blob.get(name + ".weight", t, N, K, 1, 1));
    else
        IRB_TRY(blob.get(name + ".weight", t, N, K));
    IRB_TRY(blob.get(name + ".bias", tb, N));
    std::vector<float> w0(t.data, t.data + size_t(N) * K), b0(tb.data, tb.data + N);
    o.cs = nullptr;
    if (!fold_norm.empty() && dt == DT_BF16) {
        BlobTensor g, lb;
        IRB_TRY(blob.get(fold_norm + ".weight", g, K));
        IRB_TRY(blob.get(fold_norm + ".bias", lb, K));
        std::vector<float> cs;
        fold_ln(w0, b0, N, K, g.data, lb.data, cs);
        IRB_TRY(upload_f32(cs.data(), cs.size(), &o.cs));
    }
    std::vector<float> w = pack_linear(w0.data(), N, K, o.kpad, o.npad);
    IRB_TRY(upload_act(w.data(), w.size(), &o.w));
    IRB_TRY(upload_f32(b0.data(), N, &o.b));
    o.N = N;
    o.K = K;
    if (dt == DT_F32 && K % 4 == 0) {
        std::vector<float> w3 = pack_split3(w0.data(), N, K, K, o.npad, o.k3pad);
        IRB_TRY(upload_bf16(w3.data(), w3.size(), &o.w3));
    }
    return 0;
}

// qkv Linear [3C][C] (rows [3][heads][d], network_swinir.py:121-122) -> rows [3][heads][dp], q rows * d^-0.5
int SwinIR::load_qkv(const Blob& blob, const std::string& name, int heads, LinW& o, const std::string& fold_norm) {
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
    o.cs = nullptr;
    if (dt == DT_BF16 && !fold_norm.empty()) {
        BlobTensor g, lb;
        IRB_TRY(blob.get(fold_norm + ".weight", g, C));
        IRB_TRY(blob.get(fold_norm + ".bias", lb, C));
        std::vector<float> cs;
        fold_ln(w, b, 3 * Cq, C, g.data, lb.data, cs);
        IRB_TRY(upload_f32(cs.data(), cs.size(), &o.cs));
    }
    std::vector<float> wp = pack_linear(w.data(), 3 * Cq, C, o.kpad, o.npad);
    IRB_TRY(upload_act(wp.data(), wp.size(), &o.w));
    IRB_TRY(upload_f32(b.data(), b.size(), &o.b));
    o.N = 3 * Cq;
    o.K = C;
    if (dt == DT_F32 && C % 4 == 0) {
        std::vector<float> w3 = pack_split3(w.data(), 3 * Cq, C, C, o.npad, o.k3pad);
        IRB_TRY(upload_bf16(w3.data(), w3.size(), &o.w3));
    }
    return 0;
}

int SwinIR::load_conv(const Blob& blob, const std::string& name, int N, int Cin, Conv3W& o, bool pixelshuffle) {
    BlobTensor t, tb;
    IRB_TRY(blob.get(name + ".weight", t, N, Cin, 3, 3));
    IRB_TRY(blob.get(name + ".bias", tb, N));
    std::vector<float> w = pack_conv3x3(t, o.kpad, o.npad);
    std::vector<float> bias(tb.data, tb.data + N);
    o.N = N;
    o.C = Cin;
    o.ps = pixelshuffle ? 1 : 0;
    if (pixelshuffle && N % 4 == 0 && ((N / 4) % 64 == 0 || (N / 4) % 48 == 0)) {
        permute_rows_ps_major(w, N, 9 * o.kpad, &bias);
        o.ps = 3;
    }
    IRB_TRY(upload_act(w.data(), w.size(), &o.w));
    o.wf32 = nullptr;
    if (dt == DT_F32)
        o.wf32 = reinterpret_cast<float*>(o.w);
    else if (Cin < 16)
        IRB_TRY(upload_f32(w.data(), w.size(), &o.wf32));
    o.wsplit = nullptr;
    if (dt == DT_BF16 && Cin <= 4) {
        int kp, np;
        std::vector<float> wsp = pack_conv3x3_split(t, kp, np);
        IRB_TRY(upload_act(wsp.data(), wsp.size(), &o.wsplit));
    }
    IRB_TRY(upload_f32(bias.data(), N, &o.b));
    if (dt == DT_F32 && Cin >= 16 && Cin % 4 == 0) {
        // w holds [npad][9][kpad] (rows possibly permuted for the PixelShuffle store): split every tap's Cin weights
        static const int term[6] = {0, 1, 2, 0, 1, 0};
        o.kpad3 = (6 * Cin + 63) / 64 * 64;
        std::vector<float> w3(size_t(o.npad) * 9 * o.kpad3, 0.f);
        for (int n = 0; n < o.npad; ++n)
            for (int tp = 0; tp < 9; ++tp)
                for (int c = 0; c < Cin; ++c) {
                    float t3[3];
                    split3_host(w[(size_t(n) * 9 + tp) * o.kpad + c], t3);
                    for (int sg = 0; sg < 6; ++sg) w3[(size_t(n) * 9 + tp) * o.kpad3 + sg * Cin + c] = t3[term[sg]];
                }
        IRB_TRY(upload_bf16(w3.data(), w3.size(), &o.w3));
    }
    return 0;
}

int SwinIR::load_resi(const Blob& blob, const std::string& name, ResiW& o) {
    if (!desc.resi_3conv) return load_conv(blob, name, C, C, o.c0);
    IRB_TRY(load_conv(blob, name + ".0", C / 4, C, o.c0));
    IRB_TRY(load_lin(blob, name + ".2", C / 4, C / 4, o.lin));
    IRB_TRY(load_conv(blob, name + ".4", C, C / 4, o.c2));
    return 0;
}

// norm1, attention (qkv / bias table / proj under `attn`), norm2, mlp of one block. fold1: LayerNorm 1 folded into the
// qkv GEMM (SwinIR, bf16); HAT materialises norm1(x) because its convolution branch reads it too.
int SwinIR::load_block(const Blob& blob, const std::string& p, const std::string& attn, int heads, int table_rows,
                       bool fold1, STLW& L) {
    L.heads = heads;
    IRB_TRY(load_norm(blob, p + ".norm1", &L.n1w, &L.n1b));
    IRB_TRY(load_norm(blob, p + ".norm2", &L.n2w, &L.n2b));
    BlobTensor t;
    IRB_TRY(blob.get(attn + "relative_position_bias_table", t, table_rows, heads));
    IRB_TRY(upload_f32(t.data, size_t(table_rows) * heads, &L.bias_table));
    IRB_TRY(load_qkv(blob, attn + "qkv", heads, L.qkv, fold1 ? p + ".norm1" : std::string()));
    IRB_TRY(load_lin(blob, attn + "proj", C, C, L.proj));
    IRB_TRY(load_lin(blob, p + ".mlp.fc1", hid, C, L.fc1, p + ".norm2"));
    IRB_TRY(load_lin(blob, p + ".mlp.fc2", C, hid, L.fc2));
    return 0;
}

int build_swinir(const irb_model_desc& d, const Blob& blob, int device, Model** out) {
    SwinIR* m = new SwinIR();
    *out = m;
    m->desc = d;
    m->dt = d.precision == IRB_BF16 ? DT_BF16 : DT_F32;
    m->device = device;
    m->cin = d.in_chans;
    m->tsc = d.upscale;
    m->us = d.unshuffle_scale > 1 ? d.unshuffle_scale : 1;
    if (m->us > 1)  // DiffBIR/model/swinir.py:716-724: the tail undoes the head's PixelUnshuffle
        IRB_REQUIRE(d.upsampler == IRB_UPSAMPLER_NEAREST_CONV && d.upscale == m->us && (m->us == 2 || m->us == 4 || m->us == 8),
                    "SwinIR: unshuffle needs the nearest+conv tail with upscale == unshuffle_scale in {2, 4, 8}");
    m->scale = d.upscale / m->us;
    IRB_REQUIRE(d.in_chans == 1 || d.in_chans == 3, "SwinIR: in_chans must be 1 or 3");
    IRB_REQUIRE(d.embed_dim > 0 && d.embed_dim % 4 == 0 && d.embed_dim <= 512, "SwinIR: embed_dim must be %4, <= 512");
    m->hat = d.arch == IRB_ARCH_HAT;
    if (m->hat) {
        IRB_REQUIRE(d.window_size >= 2 && d.window_size <= 16 && d.window_size % 2 == 0, "HAT: window_size must be even, <= 16");
        IRB_REQUIRE(d.img_size > d.window_size, "HAT: img_size must exceed window_size");
        IRB_REQUIRE(d.upsampler == IRB_UPSAMPLER_PIXELSHUFFLE && !d.resi_3conv && m->us == 1,
                    "HAT: only upsampler 'pixelshuffle' with resi_connection '1conv' (hat.yml)");
        IRB_REQUIRE(d.compress_ratio >= 1 && d.squeeze_factor >= 1 && d.embed_dim / d.compress_ratio >= 1 &&
                        d.embed_dim / d.squeeze_factor >= 1, "HAT: bad compress_ratio / squeeze_factor");
        m->ows = int(d.window_size * d.overlap_ratio) + d.window_size;  // hat_arch.py:373
        IRB_REQUIRE((m->ows - d.window_size) % 2 == 0 && m->ows <= 64, "HAT: bad overlap_ratio");
        m->c_cab = d.embed_dim / d.compress_ratio;
        m->ld_cab = (m->c_cab + 15) / 16 * 16;
        m->c_sq = d.embed_dim / d.squeeze_factor;
    } else {
        IRB_REQUIRE(d.window_size >= 1 && d.window_size <= 8, "SwinIR: window_size must be <= 8");
    }
    IRB_REQUIRE(d.num_layers >= 1 && d.num_layers <= 16, "SwinIR: 1..16 RSTB layers");
    if (d.upsampler == IRB_UPSAMPLER_PIXELSHUFFLE_DIRECT)
        IRB_REQUIRE(d.upscale >= 2 && d.upscale <= 4 && m->us == 1 && !m->hat, "SwinIR: pixelshuffledirect supports x2 / x3 / x4");
    else
        IRB_REQUIRE(d.upscale >= 1 && (d.upscale & (d.upscale - 1)) == 0, "SwinIR: upscale must be a power of two");
    IRB_REQUIRE(d.upsampler >= IRB_UPSAMPLER_NONE && d.upsampler <= IRB_UPSAMPLER_PIXELSHUFFLE_DIRECT, "SwinIR: bad upsampler");
    if (d.upsampler == IRB_UPSAMPLER_NONE) IRB_REQUIRE(d.upscale == 1, "SwinIR: upsampler '' needs upscale 1");
    if (d.upsampler == IRB_UPSAMPLER_NEAREST_CONV)
        IRB_REQUIRE(d.upscale == 2 || d.upscale == 4 || (d.upscale == 8 && m->us == 8),
                    "SwinIR: nearest+conv supports x2 / x4 (x8 with the unshuffle head)");
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
    if (m->us > 1)  // Sequential(PixelUnshuffle, Conv2d): the convolution is entry 1
        IRB_TRY(m->load_conv(blob, "conv_first.1", C, d.in_chans * m->us * m->us, m->conv_first));
    else
        IRB_TRY(m->load_conv(blob, "conv_first", C, d.in_chans, m->conv_first));
    IRB_TRY(m->load_norm(blob, "patch_embed.norm", &m->pen_w, &m->pen_b));
    m->layers.resize(d.num_layers);
    m->layer_conv.resize(d.num_layers);
    for (int r = 0; r < d.num_layers; ++r) {
        m->layers[r].resize(d.depths[r]);
        for (int b = 0; b < d.depths[r]; ++b) {
            STLW& L = m->layers[r][b];
            const std::string p = "layers." + std::to_string(r) + ".residual_group.blocks." + std::to_string(b);
            L.shift = (b % 2 == 1 && m->can_shift) ? m->ws / 2 : 0;  // :383 (hat_arch.py:495)
            const int nb = (2 * m->ws_eff - 1) * (2 * m->ws_eff - 1);
            IRB_TRY(m->load_block(blob, p, p + ".attn.", d.num_heads[r], nb, !m->hat, L));
            if (m->hat) {  // conv_block = CAB (hat_arch.py:61-74)
                IRB_TRY(m->load_conv(blob, p + ".conv_block.cab.0", m->c_cab, C, L.cab0));
                IRB_TRY(m->load_conv(blob, p + ".conv_block.cab.2", C, m->c_cab, L.cab2));
                BlobTensor t;
                IRB_TRY(blob.get(p + ".conv_block.cab.3.attention.1.weight", t, m->c_sq, C, 1, 1));
                IRB_TRY(m->upload_f32(t.data, size_t(m->c_sq) * C, &L.ca_w1));
                IRB_TRY(blob.get(p + ".conv_block.cab.3.attention.1.bias", t, m->c_sq));
                IRB_TRY(m->upload_f32(t.data, m->c_sq, &L.ca_b1));
                IRB_TRY(blob.get(p + ".conv_block.cab.3.attention.3.weight", t, C, m->c_sq, 1, 1));
                IRB_TRY(m->upload_f32(t.data, size_t(C) * m->c_sq, &L.ca_w2));
                IRB_TRY(blob.get(p + ".conv_block.cab.3.attention.3.bias", t, C));
                IRB_TRY(m->upload_f32(t.data, C, &L.ca_b2));
            }
        }
        if (m->hat) {  // overlap_attn = OCAB (hat_arch.py:352-438); table (ws + ows - 1)^2 x heads
            if (m->ocab.empty()) m->ocab.resize(d.num_layers);
            const std::string p = "layers." + std::to_string(r) + ".residual_group.overlap_attn";
            const int nb = (m->ws + m->ows - 1) * (m->ws + m->ows - 1);
            m->ocab[r].shift = 0;
            IRB_TRY(m->load_block(blob, p, p + ".", d.num_heads[r], nb, false, m->ocab[r]));
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
            IRB_TRY(m->load_conv(blob, "upsample." + std::to_string(2 * i), 256, 64, m->upsample[i], true));
        IRB_TRY(m->load_conv(blob, "conv_last", d.in_chans, 64, m->conv_last));
    } else if (d.upsampler == IRB_UPSAMPLER_NEAREST_CONV) {
        IRB_TRY(m->load_conv(blob, "conv_before_upsample.0", 64, C, m->conv_before_up));
        IRB_TRY(m->load_conv(blob, "conv_up1", 64, 64, m->conv_up1));
        if (d.upscale >= 4) IRB_TRY(m->load_conv(blob, "conv_up2", 64, 64, m->conv_up2));
        if (d.upscale == 8) IRB_TRY(m->load_conv(blob, "conv_up3", 64, 64, m->conv_up3));
        IRB_TRY(m->load_conv(blob, "conv_hr", 64, 64, m->conv_hr));
        IRB_TRY(m->load_conv(blob, "conv_last", d.in_chans, 64, m->conv_last));
    } else if (d.upsampler == IRB_UPSAMPLER_PIXELSHUFFLE_DIRECT) {
        IRB_TRY(m->load_conv(blob, "upsample.0", d.upscale * d.upscale * d.in_chans, C, m->conv_last));  // UpsampleOneStep :594-612
    } else {
        IRB_TRY(m->load_conv(blob, "conv_last", d.in_chans, C, m->conv_last));
    }
    return 0;
}

void SwinIR::layout(Bump& b, int n, int Hin, int Win, WS& w) {
    const int H0 = Hin / us, W0 = Win / us;  // the body runs at 1 / us of the input resolution (us = 1 except DiffBIR)
    int H, W;
    padded(H0, W0, H, W);
    const size_t E = esize();
    const size_t P = size_t(n) * H * W;
    const int s = tsc;
    w.xin = b.take<float>(P * cin * us * us * 4);
    w.out_tmp = (H != H0 || W != W0) ? b.take<float>(P * s * s * cin * 4) : nullptr;
    w.xa = b.take(P * ldC * E);
    w.xb = b.take(P * ldC * E);
    w.ffirst = b.take(P * ldC * E);
    w.ln = b.take(P * ldC * E);
    w.stats = b.take<float2>(P * sizeof(float2));
    w.qkv = b.take(P * 3 * Cq * E);
    w.attn = b.take(P * ldC * E);
    w.hidden = b.take(P * hid * E);
    w.a3 = nullptr;
    w.a3_bytes = 0;
    if (dt == DT_F32) {
        // widest split operand: 6 * hidden per token in the body, 6 * 64 per up-sampled pixel in the tails
        size_t cols = size_t(6) * (hid > C ? hid : C);
        size_t rows = P;
        if (desc.upsampler == IRB_UPSAMPLER_PIXELSHUFFLE || desc.upsampler == IRB_UPSAMPLER_NEAREST_CONV) {
            const size_t tail = P * s * s * 6 * 64;
            if (tail > rows * cols) {
                rows = P * s * s;
                cols = 6 * 64;
            }
        }
        w.a3_bytes = rows * cols * 2;
        w.a3 = b.take(w.a3_bytes);
    }
    w.cab1 = w.cab = nullptr;
    w.ca_sums = w.ca_scale = nullptr;
    if (hat) {
        w.cab1 = b.take(P * ld_cab * E);
        w.cab = b.take(P * ldC * E);
        w.ca_sums = b.take<float>(size_t(n) * channel_attention_slabs(n, (long long)H * W) * C * 4);
        w.ca_scale = b.take<float>(size_t(n) * C * 4);
    }
    w.q4a = w.q4b = nullptr;
    if (desc.resi_3conv) {
        w.q4a = b.take(P * 64 * E);
        w.q4b = b.take(P * 64 * E);
    }
    w.t64 = w.u1 = w.u2 = w.u3 = nullptr;
    if (desc.upsampler == IRB_UPSAMPLER_PIXELSHUFFLE) {
        // ping-pong 64-channel maps: LR, then x2, x4, ... (largest = s*s*P)
        w.t64 = b.take(P * 64 * E);
        w.u1 = b.take(P * s * s * 64 * E);
        w.u2 = b.take(P * (s * s / 4 > 0 ? s * s / 4 : 1) * 64 * E);
    } else if (desc.upsampler == IRB_UPSAMPLER_NEAREST_CONV) {
        w.t64 = b.take(P * 64 * E);
        w.u1 = b.take(P * s * s * 64 * E);
        w.u2 = b.take(P * s * s * 64 * E);
    } else if (desc.upsampler == IRB_UPSAMPLER_PIXELSHUFFLE_DIRECT) {
        w.u1 = b.take(P * s * s * cin * 4);  // fp32 [P][s*s*cin] before the shuffle
    }
}

size_t SwinIR::forward_workspace(int n, int H, int W) {
    Bump b(nullptr);
    WS w;
    layout(b, n, H, W, w);
    return b.off + 256;
}

int SwinIR::gemm(const void* A, long long lda, long long M, const LinW& L, Epi e, cudaStream_t s) {
    IRB_REQUIRE(M < (1ll << 31), "SwinIR: too many tokens in one batch (lower tile_batch)");
    e.bias = L.b;
    if (dt == DT_BF16)
        return tc_gemm((const bf16*)A, lda, int(M), L.K, 1, 0, (const bf16*)L.w, L.kpad, L.N, L.npad, 0, e, s);
    if (fp32_tc && L.w3 && cur_ws && size_t(M) * 6 * L.K * 2 <= cur_ws->a3_bytes) {
        // fp32 mode on the tensor cores: x = h + m + l in bf16, six bf16 products per fp32 product (split3_bf16)
        IRB_TRY(split3_bf16((const float*)A, lda, (bf16*)cur_ws->a3, 6ll * L.K, M, L.K, s));
        return tc_gemm((const bf16*)cur_ws->a3, 6ll * L.K, int(M), 6 * L.K, 1, 0, (const bf16*)L.w3, L.k3pad, L.N, L.npad, 0,
                       e, s);
    }
    return simt_gemm(A, DT_F32, lda, int(M), L.K, 1, 0, (const float*)L.w, L.kpad, L.N, 0, e, s);
}

int SwinIR::conv(const Conv3W& c, const void* x, int x_dt, long long ldx, int n, int H, int W, Epi e,
                 cudaStream_t s) {
    e.bias = c.b;
    if (dt == DT_BF16 && x_dt == DT_BF16 && c.C >= 16)
        return conv3x3_bf16((const bf16*)x, ldx, n, H, W, c.C, (const bf16*)c.w, c.kpad, c.N, c.npad, e, s);
    const long long Px = (long long)n * H * W;
    if (dt == DT_F32 && x_dt == DT_F32 && fp32_tc && c.w3 && cur_ws && size_t(Px) * 6 * c.C * 2 <= cur_ws->a3_bytes) {
        IRB_TRY(split3_bf16((const float*)x, ldx, (bf16*)cur_ws->a3, 6ll * c.C, Px, c.C, s));
        return tc_conv3x3((const bf16*)cur_ws->a3, 6ll * c.C, n, H, W, 6 * c.C, (const bf16*)c.w3, c.kpad3, c.N, c.npad, e, s);
    }
    IRB_REQUIRE(c.wf32 != nullptr, "SwinIR: missing fp32 conv weights");
    return simt_conv3x3(x, x_dt, ldx, n, H, W, c.C, c.wf32, c.kpad, c.N, e, s);
}

// '1conv' / '3conv' residual convolution (:464-471, :728-736); `e` is the epilogue of the LAST convolution.
int SwinIR::resi(const ResiW& r, const void* x, long long ldx, int n, int H, int W, Epi e, WS& w, cudaStream_t s) {
    if (!desc.resi_3conv) return conv(r.c0, x, dt, ldx, n, H, W, e, s);
    Epi e0 = make_epi(w.q4a, 64, dt);
    e0.act = ACT_LRELU;
    e0.slope = 0.2f;
    IRB_TRY(conv(r.c0, x, dt, ldx, n, H, W, e0, s));
    Epi e1 = make_epi(w.q4b, 64, dt);
    e1.act = ACT_LRELU;
    e1.slope = 0.2f;
    IRB_TRY(gemm(w.q4a, 64, (long long)n * H * W, r.lin, e1, s));
    return conv(r.c2, w.q4b, dt, 64, n, H, W, e, s);
}

// SwinTransformerBlock.forward (:239-279): xout = xin + attn(LN1(xin)); xout += mlp(LN2(xout))
int SwinIR::stl(const STLW& L, const void* xin, void* xout, int n, int H, int W, WS& w, cudaStream_t s) {
    const long long P = (long long)n * H * W;
    if (dt == DT_BF16) {  // norm1 folded into the qkv GEMM
        if (!w.stats_valid) IRB_TRY(row_stats(xin, dt, ldC, w.stats, P, C, 1e-5f, s));
        Epi e = make_epi(w.qkv, 3 * Cq, dt);
        e.ln_stats = w.stats;
        e.ln_colsum = L.qkv.cs;
        IRB_TRY(gemm(xin, ldC, P, L.qkv, e, s));
    } else {
        IRB_TRY(layernorm(xin, dt, ldC, w.ln, dt, ldC, P, C, L.n1w, L.n1b, 0, 1e-5f, s));
        IRB_TRY(gemm(w.ln, ldC, P, L.qkv, make_epi(w.qkv, 3 * Cq, dt), s));
    }
    IRB_TRY(window_attention(w.qkv, dt, 3 * Cq, w.attn, ldC, n, H, W, C, L.heads, dp, ws_eff, L.shift, L.bias_table, s));
    {
        Epi e = make_epi(xout, ldC, dt);
        e.res = xin;
        e.ldr = ldC;
        e.res_dt = dt;
        e.bias = L.proj.b;
        w.stats_valid = false;
        if (dt == DT_BF16 && tc_can_fuse_stats(C, e)) {  // norm2's statistics come out of this GEMM's epilogue
            e.stats_out = w.stats;
            e.stats_eps = 1e-5f;
            w.stats_valid = true;
        }
        IRB_TRY(gemm(w.attn, ldC, P, L.proj, e, s));
    }
    return mlp(L, xout, P, w, s);
}

// x += mlp(norm2(x)) (:276 / hat_arch.py:307, :437); norm2's statistics are in w.stats when w.stats_valid
int SwinIR::mlp(const STLW& L, void* xout, long long P, WS& w, cudaStream_t s) {
    if (dt == DT_BF16 && fuse_mlp && mlp_fused_shape_ok(C, hid)) {
        // fc1 -> GELU -> fc2 -> + x in one kernel (mlp_fused.cu): the hidden tensor stays on the SM; the kernel also
        // writes the next block's norm1 statistics (it reads norm2's from w.stats and overwrites them row by row)
        if (!w.stats_valid) IRB_TRY(row_stats(xout, dt, ldC, w.stats, P, C, 1e-5f, s));
        IRB_TRY(mlp_fused_bf16((bf16*)xout, ldC, P, C, hid, w.stats, (const bf16*)L.fc1.w, L.fc1.kpad, L.fc1.cs, L.fc1.b,
                               (const bf16*)L.fc2.w, L.fc2.kpad, L.fc2.b, w.stats, 1e-5f, s));
        w.stats_valid = true;
        return 0;
    }
    if (dt == DT_BF16) {  // norm2 folded into fc1
        if (!w.stats_valid) IRB_TRY(row_stats(xout, dt, ldC, w.stats, P, C, 1e-5f, s));
        Epi e = make_epi(w.hidden, hid, dt);
        e.act = ACT_GELU;
        e.ln_stats = w.stats;
        e.ln_colsum = L.fc1.cs;
        IRB_TRY(gemm(xout, ldC, P, L.fc1, e, s));
    } else {
        IRB_TRY(layernorm(xout, dt, ldC, w.ln, dt, ldC, P, C, L.n2w, L.n2b, 0, 1e-5f, s));
        Epi e = make_epi(w.hidden, hid, dt);
        e.act = ACT_GELU;
        IRB_TRY(gemm(w.ln, ldC, P, L.fc1, e, s));
    }
    {
        Epi e = make_epi(xout, ldC, dt);
        e.res = xout;
        e.ldr = ldC;
        e.res_dt = dt;
        e.bias = L.fc2.b;
        w.stats_valid = false;
        if (dt == DT_BF16 && tc_can_fuse_stats(C, e)) {  // the next block's norm1 statistics
            e.stats_out = w.stats;
            e.stats_eps = 1e-5f;
            w.stats_valid = true;
        }
        IRB_TRY(gemm(w.hidden, hid, P, L.fc2, e, s));
    }
    return 0;
}

// HAB.forward (hat_arch.py:266-309): x = shortcut + attn(norm1(x)) + conv_scale * CAB(norm1(x)); x += mlp(norm2(x)).
// norm1(x) is materialised (w.ln) because two consumers read it; the channel-attention scale and conv_scale are
// applied when the convolution branch is added to the shortcut, before the projection GEMM adds the attention branch.
int SwinIR::hab(const STLW& L, const void* xin, void* xout, int n, int H, int W, WS& w, cudaStream_t s) {
    const long long P = (long long)n * H * W;
    IRB_TRY(layernorm(xin, dt, ldC, w.ln, dt, ldC, P, C, L.n1w, L.n1b, 0, 1e-5f, s));
    IRB_TRY(gemm(w.ln, ldC, P, L.qkv, make_epi(w.qkv, 3 * Cq, dt), s));
    IRB_TRY(hat_attention(w.qkv, dt, 3 * Cq, w.attn, ldC, n, H, W, L.heads, C / L.heads, dp, ws, ws, L.shift, L.bias_table, s));
    {
        Epi e = make_epi(w.cab1, ld_cab, dt);
        e.act = ACT_GELU;
        IRB_TRY(conv(L.cab0, w.ln, dt, ldC, n, H, W, e, s));
        IRB_TRY(conv(L.cab2, w.cab1, dt, ld_cab, n, H, W, make_epi(w.cab, ldC, dt), s));
    }
    IRB_TRY(channel_attention_scale(w.cab, dt, ldC, n, (long long)H * W, C, c_sq, L.ca_w1, L.ca_b1, L.ca_w2, L.ca_b2,
                                    desc.conv_scale, w.ca_sums, w.ca_scale, s));
    IRB_TRY(add_scaled(xin, ldC, w.cab, ldC, xout, ldC, dt, P, (long long)H * W, C, w.ca_scale, s));
    {
        Epi e = make_epi(xout, ldC, dt);
        e.res = xout;
        e.ldr = ldC;
        e.res_dt = dt;
        w.stats_valid = false;
        if (dt == DT_BF16 && tc_can_fuse_stats(C, e)) {
            e.stats_out = w.stats;
            e.stats_eps = 1e-5f;
            w.stats_valid = true;
        }
        IRB_TRY(gemm(w.attn, ldC, P, L.proj, e, s));
    }
    return mlp(L, xout, P, w, s);
}

// OCAB.forward (hat_arch.py:392-438), in place on x
int SwinIR::ocab_block(const STLW& L, void* x, int n, int H, int W, WS& w, cudaStream_t s) {
    const long long P = (long long)n * H * W;
    IRB_TRY(layernorm(x, dt, ldC, w.ln, dt, ldC, P, C, L.n1w, L.n1b, 0, 1e-5f, s));
    IRB_TRY(gemm(w.ln, ldC, P, L.qkv, make_epi(w.qkv, 3 * Cq, dt), s));
    IRB_TRY(hat_attention(w.qkv, dt, 3 * Cq, w.attn, ldC, n, H, W, L.heads, C / L.heads, dp, ws, ows, 0, L.bias_table, s));
    {
        Epi e = make_epi(x, ldC, dt);
        e.res = x;
        e.ldr = ldC;
        e.res_dt = dt;
        w.stats_valid = false;
        if (dt == DT_BF16 && tc_can_fuse_stats(C, e)) {
            e.stats_out = w.stats;
            e.stats_eps = 1e-5f;
            w.stats_valid = true;
        }
        IRB_TRY(gemm(w.attn, ldC, P, L.proj, e, s));
    }
    return mlp(L, x, P, w, s);
}

int SwinIR::forward(const float* in, float* out, int n, int Hin, int Win, void* wsp, size_t ws_bytes, cudaStream_t s) {
    IRB_REQUIRE(ws_bytes >= forward_workspace(n, Hin, Win), "SwinIR: workspace too small");
    // DiffBIR stage 1: the body runs on the PixelUnshuffle'd input; the reference pads only to the window at full
    // resolution (:844-849), so inputs that are not multiples of us * window fail there too
    IRB_REQUIRE(us == 1 || (Hin % (us * desc.window_size) == 0 && Win % (us * desc.window_size) == 0),
                "SwinIR (unshuffle head): height / width must be multiples of unshuffle_scale * window_size");
    const int H0 = Hin / us, W0 = Win / us;
    int H, W;
    padded(H0, W0, H, W);
    IRB_REQUIRE(!hat || (H == H0 && W == W0), "HAT: height / width must be multiples of window_size (HATModel.pre_process pads)");
    IRB_REQUIRE(H - H0 < H0 && W - W0 < W0, "SwinIR: image smaller than the reflect padding");
    IRB_REQUIRE(H % ws_eff == 0 && W % ws_eff == 0, "SwinIR: padded size must be a multiple of the window");
    if (desc.resi_3conv) IRB_REQUIRE(C / 4 <= 64, "SwinIR: 3conv supports embed_dim <= 256");
    Bump b(wsp);
    WS w;
    layout(b, n, Hin, Win, w);
    // calls on one model are serialised by the caller (Engine's lock): the pointer is per forward and cleared on every
    // return path
    struct WsScope {
        WS*& slot;
        WsScope(WS*& s_, WS* v) : slot(s_) { slot = v; }
        ~WsScope() { slot = nullptr; }
    } ws_scope(cur_ws, &w);
    const long long P = (long long)n * H * W;
    const int sc = tsc;
    if (us > 1) {
        const int co = cin * us * us;
        void* xu = dt == DT_BF16 ? w.hidden : (void*)w.xin;  // w.hidden is free here (P * hid >= P * co elements)
        IRB_REQUIRE(dt != DT_BF16 || hid >= co, "SwinIR (unshuffle head): hidden buffer too small for the unshuffled input");
        if (dt == DT_BF16)
            swin_prep_unshuffle_kernel<bf16><<<grid_for(P * co), 256, 0, s>>>(in, (bf16*)xu, n, H, W, cin, us, mean[0],
                                                                              mean[1], mean[2], desc.img_range);
        else
            swin_prep_unshuffle_kernel<float><<<grid_for(P * co), 256, 0, s>>>(in, (float*)xu, n, H, W, cin, us, mean[0],
                                                                               mean[1], mean[2], desc.img_range);
        IRB_CHECK_CUDA(cudaGetLastError());
        prof_mark("swin_prep_unshuffle", cin, us, 0.0, double(P) * co * (4.0 + esize()), s);
        IRB_TRY(conv(conv_first, xu, dt, co, n, H, W, make_epi(w.ffirst, ldC, dt), s));
    } else {
    // (x - mean) * img_range on the reflect-padded tile (:806-810)
    swin_prep_kernel<<<grid_for(P), 256, 0, s>>>(in, w.xin, n, H0, W0, H, W, cin, mean[0], mean[1], mean[2],
                                                 desc.img_range);
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("swin_prep", cin, 0, 0.0, 8.0 * P * cin, s);
    }
    if (us > 1) {
        // conv_first already ran on the unshuffled input
    } else if (conv_first.wsplit) {  // :813 / :833 on the tensor core: x = hi + lo in bf16 (w.hidden is free here)
        IRB_TRY(split_hi_lo_bf16(w.xin, cin, (bf16*)w.hidden, P, s));
        Epi e = make_epi(w.ffirst, ldC, dt);
        e.bias = conv_first.b;
        IRB_TRY(tc_conv3x3((const bf16*)w.hidden, 8, n, H, W, 8, (const bf16*)conv_first.wsplit, 64, conv_first.N,
                           conv_first.npad, e, s));
    } else {
        IRB_TRY(conv(conv_first, w.xin, DT_F32, cin, n, H, W, make_epi(w.ffirst, ldC, dt), s));  // :813 / :833
    }
    // forward_features (:790-803)
    IRB_TRY(layernorm(w.ffirst, dt, ldC, w.xa, dt, ldC, P, C, pen_w, pen_b, 0, 1e-5f, s));  // patch_embed.norm
    w.stats_valid = false;
    for (size_t r = 0; r < layers.size(); ++r) {
        // RSTB (:481-482): xa = conv(blocks(xa)) + xa ; the blocks run in xb
        const void* src = w.xa;
        for (auto& L : layers[r]) {
            IRB_TRY(hat ? hab(L, src, w.xb, n, H, W, w, s) : stl(L, src, w.xb, n, H, W, w, s));
            src = w.xb;
        }
        if (hat) IRB_TRY(ocab_block(ocab[r], w.xb, n, H, W, w, s));  // AttenBlocks.forward, hat_arch.py:526-534
        Epi e = make_epi(w.xa, ldC, dt);
        e.res = w.xa;
        e.ldr = ldC;
        e.res_dt = dt;
        w.stats_valid = false;
        const int last_cin = desc.resi_3conv ? layer_conv[r].c2.C : layer_conv[r].c0.C;  // >= 16: tensor-core conv
        if (dt == DT_BF16 && r + 1 < layers.size() && last_cin >= 16 && tc_can_fuse_stats(C, e)) {  // norm1 of the next RSTB's first block
            e.stats_out = w.stats;
            e.stats_eps = 1e-5f;
            w.stats_valid = true;
        }
        IRB_TRY(resi(layer_conv[r], w.xb, ldC, n, H, W, e, w, s));
    }
    IRB_TRY(layernorm(w.xa, dt, ldC, w.ln, dt, ldC, P, C, norm_w, norm_b, 0, 1e-5f, s));  // :800
    {
        // conv_after_body(forward_features(x)) + x  -> xb
        Epi e = make_epi(w.xb, ldC, dt);
        e.res = w.ffirst;
        e.ldr = ldC;
        e.res_dt = dt;
        IRB_TRY(resi(conv_after_body, w.ln, ldC, n, H, W, e, w, s));
    }
    const bool crop = (H != H0) || (W != W0);
    float* dst = crop ? w.out_tmp : out;
    auto last_epi = [&]() {
        Epi e = make_epi(dst, cin, DT_F32);  // x / img_range + mean (:838)
        e.post_scale = 1.0f / desc.img_range;
        e.post_shift = d_mean;
        return e;
    };
    auto lrelu_epi = [&](void* o, float slope) {
        Epi e = make_epi(o, 64, dt);
        e.act = ACT_LRELU;
        e.slope = slope;
        return e;
    };
    if (desc.upsampler == IRB_UPSAMPLER_NONE) {
        Epi e = last_epi();  // x + conv_last(res) on the normalised input (:834-836)
        e.res = w.xin;
        e.ldr = cin;
        e.res_dt = DT_F32;
        IRB_TRY(conv(conv_last, w.xb, dt, ldC, n, H, W, e, s));
    } else if (desc.upsampler == IRB_UPSAMPLER_PIXELSHUFFLE_DIRECT) {
        // x = upsample(x): conv C -> s*s*cin, PixelShuffle(s) (:818-822); the shuffle kernel also applies / img_range + mean
        IRB_TRY(conv(conv_last, w.xb, dt, ldC, n, H, W, make_epi(w.u1, sc * sc * cin, DT_F32), s));
        const long long total = P * sc * sc * cin;
        pixel_shuffle_out_kernel<<<grid_for(total), 256, 0, s>>>(reinterpret_cast<const float*>(w.u1), dst, n, H, W, cin, sc,
                                                                 mean[0], mean[1], mean[2], 1.0f / desc.img_range);
        IRB_CHECK_CUDA(cudaGetLastError());
        prof_mark("pixel_shuffle_out", cin, sc, 0.0, 8.0 * total, s);
    } else if (desc.upsampler == IRB_UPSAMPLER_PIXELSHUFFLE) {
        IRB_TRY(conv(conv_before_up, w.xb, dt, ldC, n, H, W, lrelu_epi(w.t64, 0.01f), s));  // nn.LeakyReLU() :742-743
        const void* cur = w.t64;
        int h = H, wd = W;
        void* bufs[2] = {w.u1, w.u2};
        // keep the LAST stage in u1 (the largest buffer)
        int which = (upsample.size() % 2 == 1) ? 0 : 1;
        for (size_t i = 0; i < upsample.size(); ++i) {
            Epi e = make_epi(bufs[which], 64, dt);  // conv 64->256 + PixelShuffle(2) fused into the store (:572-591)
            e.ps = upsample[i].ps;
            e.ps_h = h;
            e.ps_w = wd;
            IRB_TRY(conv(upsample[i], cur, dt, 64, n, h, wd, e, s));
            cur = bufs[which];
            which ^= 1;
            h *= 2;
            wd *= 2;
        }
        IRB_TRY(conv(conv_last, cur, dt, 64, n, h, wd, last_epi(), s));
    } else {
        IRB_TRY(conv(conv_before_up, w.xb, dt, ldC, n, H, W, lrelu_epi(w.t64, 0.01f), s));
        // lrelu(conv_up1(nearest x2)) [, lrelu(conv_up2(nearest x2))], conv_last(lrelu(conv_hr(x)))  (:826-831)
        IRB_TRY(nearest_up2(w.t64, dt, 64, w.u1, 64, n, H, W, 64, s));
        IRB_TRY(conv(conv_up1, w.u1, dt, 64, n, 2 * H, 2 * W, lrelu_epi(w.u2, 0.2f), s));
        int h = 2 * H, wd = 2 * W;
        void* cur = w.u2;
        void* other = w.u1;
        if (sc >= 4) {
            IRB_TRY(nearest_up2(w.u2, dt, 64, w.u1, 64, n, h, wd, 64, s));
            h *= 2;
            wd *= 2;
            IRB_TRY(conv(conv_up2, w.u1, dt, 64, n, h, wd, lrelu_epi(w.u2, 0.2f), s));
        }
        if (sc == 8) {  // DiffBIR stage 1 (:891)
            IRB_TRY(nearest_up2(w.u2, dt, 64, w.u1, 64, n, h, wd, 64, s));
            h *= 2;
            wd *= 2;
            IRB_TRY(conv(conv_up3, w.u1, dt, 64, n, h, wd, lrelu_epi(w.u2, 0.2f), s));
        }
        IRB_TRY(conv(conv_hr, cur, dt, 64, n, h, wd, lrelu_epi(other, 0.2f), s));
        IRB_TRY(conv(conv_last, other, dt, 64, n, h, wd, last_epi(), s));
    }
    if (crop) {
        const long long total = (long long)n * H0 * sc * W0 * sc * cin;
        crop_nhwc_kernel<<<grid_for(total), 256, 0, s>>>(w.out_tmp, out, n, H * sc, W * sc, H0 * sc, W0 * sc, cin);
        IRB_CHECK_CUDA(cudaGetLastError());
        prof_mark("crop_nhwc", cin, 0, 0.0, 8.0 * total, s);
    }
    return 0;
}

}  // namespace irb
