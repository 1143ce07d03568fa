// C ABI: error state, host-only planners, and the single-kernel entry points declared in include/irb200.h.
#include "../../include/irb200.h"
#include "common.cuh"
#include "kernels.h"
#include <string.h>

namespace irb {
static thread_local std::string g_last_error;
void set_error(const std::string& msg) { g_last_error = msg; }
}  // namespace irb

using namespace irb;

extern "C" {

const char* irb_last_error(void) { return g_last_error.c_str(); }
int irb_version(void) { return 100; }

// inference.py:305-308 / demo.py:139-141
int irb_plan_tiles(int n, int tile, int overlap, int* idx, int* count) {
    IRB_REQUIRE(count != nullptr, "irb_plan_tiles: count is NULL");
    IRB_REQUIRE(n > 0 && tile > 0 && tile <= n, "irb_plan_tiles: need 0 < tile <= n (apply tile = min(tile, h, w) first)");
    IRB_REQUIRE(overlap >= 0 && overlap < tile, "irb_plan_tiles: need 0 <= overlap < tile");
    const int stride = tile - overlap;
    int k = 0;
    for (int i = 0; i < n - tile; i += stride) {
        if (idx) idx[k] = i;
        ++k;
    }
    if (idx) idx[k] = n - tile;
    ++k;
    *count = k;
    return 0;
}

int irb_padded_size(int arch, int window_size, int h, int w, int* hp, int* wp) {
    IRB_REQUIRE(hp && wp && h > 0 && w > 0, "irb_padded_size: bad arguments");
    if (arch == IRB_ARCH_SWINIR) {
        IRB_REQUIRE(window_size > 0, "irb_padded_size: window_size");
        *hp = (h / window_size + 1) * window_size;  // inference.py:76-78: always pads
        *wp = (w / window_size + 1) * window_size;
    } else if (arch == IRB_ARCH_RESTORMER) {
        *hp = (h % 8) ? ((h + 8) / 8) * 8 : h;  // demo.py:122-126
        *wp = (w % 8) ? ((w + 8) / 8) * 8 : w;
    } else {
        set_error("irb_padded_size: unknown arch");
        return 1;
    }
    return 0;
}

static int require_device() {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
        set_error("no CUDA device: irb200 has no CPU fallback");
        return 2;
    }
    return 0;
}

static Epi epi_from(void* y, long long ldy, int dtype, const float* bias, const void* res, long long ldr, int act,
                    float slope) {
    Epi e = make_epi(y, ldy, dtype);
    e.bias = bias;
    e.res = res;
    e.ldr = ldr;
    e.res_dt = dtype;
    e.act = act;
    e.slope = slope;
    return e;
}

int irb_layernorm(const void* d_x, int dtype, long long ldx, void* d_y, long long ldy, long long rows, int C,
                  const float* d_w, const float* d_b, int bias_free, float eps, void* stream) {
    IRB_TRY(require_device());
    return layernorm(d_x, dtype, ldx, d_y, dtype, ldy, rows, C, d_w, d_b, bias_free ? 1 : 0, eps,
                     (cudaStream_t)stream);
}

int irb_gemm1x1(const void* d_x, int dtype, long long ldx, int M, int K, int batch, const void* d_w, int ldw, int N,
                int w_rows_total, int w_rows_per_batch, const float* d_bias, const void* d_res, long long ldr,
                int act, float slope, void* d_y, long long ldy, void* stream) {
    IRB_TRY(require_device());
    Epi e = epi_from(d_y, ldy, dtype, d_bias, d_res, ldr, act, slope);
    if (dtype == IRB_BF16)
        return tc_gemm((const bf16*)d_x, ldx, M, K, batch, 0, (const bf16*)d_w, ldw, N, w_rows_total,
                       w_rows_per_batch, e, (cudaStream_t)stream);
    return simt_gemm(d_x, DT_F32, ldx, M, K, batch, 0, (const float*)d_w, ldw, N, w_rows_per_batch, e,
                     (cudaStream_t)stream);
}

int irb_conv3x3(const void* d_x, int dtype, long long ldx, int n, int H, int W, int C, const void* d_w, int kpad,
                int N, int w_rows_total, const float* d_bias, const void* d_res, long long ldr, int act, float slope,
                int shuffle, void* d_y, long long ldy, void* stream) {
    IRB_TRY(require_device());
    Epi e = epi_from(d_y, ldy, dtype, d_bias, d_res, ldr, act, slope);
    e.ps = shuffle;
    e.ps_h = H;
    e.ps_w = W;
    if (shuffle == 1) IRB_REQUIRE(N % 4 == 0, "irb_conv3x3: PixelShuffle needs N % 4 == 0");
    if (shuffle == 2) IRB_REQUIRE(H % 2 == 0 && W % 2 == 0, "irb_conv3x3: PixelUnshuffle needs even H, W");
    if (dtype == IRB_BF16)
        return tc_conv3x3((const bf16*)d_x, ldx, n, H, W, C, (const bf16*)d_w, kpad, N, w_rows_total, e,
                          (cudaStream_t)stream);
    return simt_conv3x3(d_x, DT_F32, ldx, n, H, W, C, (const float*)d_w, kpad, N, e, (cudaStream_t)stream);
}

int irb_dwconv3x3(const void* d_x, int dtype, long long ldx, void* d_y, long long ldy, int n, int H, int W, int C,
                  const float* d_w, float* d_sumsq, int sumsq_channels, void* stream) {
    IRB_TRY(require_device());
    return dwconv3x3(d_x, dtype, ldx, d_y, ldy, n, H, W, C, d_w, d_sumsq, sumsq_channels, (cudaStream_t)stream);
}

int irb_gdfn_dw_gate(const void* d_x, int dtype, long long ldx, void* d_y, long long ldy, int n, int H, int W, int hp,
                     const float* d_w, void* stream) {
    IRB_TRY(require_device());
    return dwconv3x3_gate(d_x, dtype, ldx, d_y, ldy, n, H, W, hp, d_w, (cudaStream_t)stream);
}

int irb_mdta_gram(const void* d_qkv, int dtype, long long ld, int n, int P, int C, int heads, float* d_G,
                  void* stream) {
    IRB_TRY(require_device());
    if (dtype == IRB_BF16) return mdta_gram_tc((const bf16*)d_qkv, ld, n, P, C, heads, d_G, (cudaStream_t)stream);
    return mdta_gram_simt(d_qkv, DT_F32, ld, n, P, C, heads, d_G, (cudaStream_t)stream);
}

int irb_mdta_attn(const float* d_G, const float* d_sumsq, const float* d_temperature, const float* d_wproj, int n,
                  int C, int heads, void* d_weff, int dtype, float* d_attn, void* stream) {
    IRB_TRY(require_device());
    return mdta_softmax_weff(d_G, d_sumsq, d_temperature, d_wproj, n, C, heads, d_weff, dtype, d_attn,
                             (cudaStream_t)stream);
}

int irb_stitch(const float* d_tiles, int c, int th, int tw, int sf, const int* ys, const int* xs, int ny, int nx,
               int h, int w, int crop_h, int crop_w, float* d_out_f32, uint8_t* d_out_u8, int reverse_channels,
               void* stream) {
    IRB_TRY(require_device());
    return tiles_stitch(d_tiles, c, th, tw, sf, ys, xs, ny, nx, h, w, crop_h, crop_w, d_out_f32, d_out_u8,
                        reverse_channels, (cudaStream_t)stream);
}

}  // extern "C"
