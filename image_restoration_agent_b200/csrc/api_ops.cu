// C ABI: error state, host-only planners, and the single-kernel entry points declared in include/irb200.h.
#include "../../include/irb200.h"
#include "common.cuh"
#include "kernels.h"
#include <stdlib.h>
#include <string.h>
#include <mutex>
#include <set>
#include <vector>

namespace irb {
static thread_local std::string g_last_error;
void set_error(const std::string& msg) { g_last_error = msg; }

// ---- per-launch profiler -----------------------------------------------------------------------------
bool g_prof_on = false;

bool pdl_enabled() {
    static int on = -1;
    if (on < 0) {
        on = IRB_DIAG_ENV("IRB_PDL", 1);
    }
    return on != 0;
}

#ifdef IRB_DIAG
int diag_env(const char* name, int dflt) {
    const char* e = getenv(name);
    return e ? atoi(e) : dflt;
}
#endif

cudaError_t smem_optin(const void* fn, int bytes) {
    static std::mutex mu;
    static std::set<std::pair<const void*, int>> done;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lk(mu);
    if (done.count({fn, dev})) return cudaSuccess;
    e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    if (e == cudaSuccess) done.insert({fn, dev});
    return e;
}
struct ProfRec {
    char name[56];
    double flops, bytes;
    cudaEvent_t ev;
};
struct ProfAgg {
    char name[56];
    int launches;
    double ms, flops, bytes;
};
static std::vector<ProfRec> g_prof_recs;
static std::vector<cudaEvent_t> g_prof_pool;
static std::vector<ProfAgg> g_prof_agg;
static cudaEvent_t g_prof_start = nullptr;
static size_t g_prof_pool_used = 0;

static cudaEvent_t prof_event() {
    if (g_prof_pool_used == g_prof_pool.size()) {
        cudaEvent_t e = nullptr;
        cudaEventCreate(&e);
        g_prof_pool.push_back(e);
    }
    return g_prof_pool[g_prof_pool_used++];
}

void prof_mark_slow(const char* name, int p0, int p1, double flops, double bytes, cudaStream_t s) {
    ProfRec r{};
    snprintf(r.name, sizeof r.name, "%s[%d,%d]", name, p0, p1);
    r.flops = flops;
    r.bytes = bytes;
    r.ev = prof_event();
    cudaEventRecord(r.ev, s);
    g_prof_recs.push_back(r);
}
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
    } else if (arch == IRB_ARCH_XRESTORMER) {
        // xrestormer_model.py:56-66: padding_size 64 (scale-1 configs) / 16 (the scale-4 SR config); the caller passes it
        // in `window_size` (0 = 64)
        const int m = window_size > 0 ? window_size : 64;
        *hp = (h % m) ? (h / m + 1) * m : h;
        *wp = (w % m) ? (w / m + 1) * m : w;
    } else if (arch == IRB_ARCH_HAT) {
        IRB_REQUIRE(window_size > 0, "irb_padded_size: window_size");
        *hp = (h % window_size) ? h + window_size - h % window_size : h;  // hat_model.py:16-26 pre_process
        *wp = (w % window_size) ? w + window_size - w % window_size : w;
    } else {
        set_error("irb_padded_size: unknown arch");
        return 1;
    }
    return 0;
}

// ---- profiler API: begin -> run any irb_* calls on `stream` -> end (synchronises) -> read entries ---------
int irb_profile_begin(void* stream) {
    g_prof_recs.clear();
    g_prof_agg.clear();
    g_prof_pool_used = 0;
    g_prof_start = prof_event();
    IRB_CHECK_CUDA(cudaEventRecord(g_prof_start, (cudaStream_t)stream));
    g_prof_on = true;
    return 0;
}

int irb_profile_end(void* stream) {
    g_prof_on = false;
    IRB_CHECK_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    cudaEvent_t prev = g_prof_start;
    for (const ProfRec& r : g_prof_recs) {
        float ms = 0.f;
        IRB_CHECK_CUDA(cudaEventElapsedTime(&ms, prev, r.ev));
        prev = r.ev;
        ProfAgg* a = nullptr;
        for (ProfAgg& x : g_prof_agg)
            if (strcmp(x.name, r.name) == 0) a = &x;
        if (!a) {
            ProfAgg n{};
            memcpy(n.name, r.name, sizeof n.name);
            g_prof_agg.push_back(n);
            a = &g_prof_agg.back();
        }
        a->launches += 1;
        a->ms += ms;
        a->flops += r.flops;
        a->bytes += r.bytes;
    }
    return 0;
}

int irb_profile_count(void) { return int(g_prof_agg.size()); }

int irb_profile_entry(int i, char* name56, int* launches, double* ms, double* flops, double* bytes) {
    IRB_REQUIRE(i >= 0 && i < int(g_prof_agg.size()) && name56 && launches && ms && flops && bytes,
                "irb_profile_entry: bad arguments");
    const ProfAgg& a = g_prof_agg[i];
    memcpy(name56, a.name, sizeof a.name);
    *launches = a.launches;
    *ms = a.ms;
    *flops = a.flops;
    *bytes = a.bytes;
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
        return conv3x3_bf16((const bf16*)d_x, ldx, n, H, W, C, (const bf16*)d_w, kpad, N, w_rows_total, e,
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

int irb_pwdw_fused(int gate, const void* d_x, long long ldx, const float* d_stats, const void* d_w1, int nrows, int hp,
                   const float* d_colsum, const float* d_t, const float* d_dw, int dw_ld, void* d_y, long long ldy, int n,
                   int H, int W, int C, void* stream) {
    IRB_TRY(require_device());
    return pwdw_fused_bf16(gate != 0, (const bf16*)d_x, ldx, reinterpret_cast<const float2*>(d_stats), (const bf16*)d_w1,
                           nrows, hp, d_colsum, d_t, d_dw, dw_ld, (bf16*)d_y, ldy, n, H, W, C, (cudaStream_t)stream);
}

int irb_mlp_fused(void* d_x, long long ldx, long long rows, int C, int hid, const float* d_stats_in, const void* d_w1,
                  int ldw1, const float* d_colsum1, const float* d_b1, const void* d_w2, int ldw2, const float* d_b2,
                  float* d_stats_out, float eps, void* stream) {
    IRB_TRY(require_device());
    return mlp_fused_bf16((bf16*)d_x, ldx, rows, C, hid, reinterpret_cast<const float2*>(d_stats_in), (const bf16*)d_w1,
                          ldw1, d_colsum1, d_b1, (const bf16*)d_w2, ldw2, d_b2, reinterpret_cast<float2*>(d_stats_out), eps,
                          (cudaStream_t)stream);
}

int irb_ocab(const void* d_qkv, int dtype, long long ld, void* d_out, long long ldo, int n, int H, int W, int heads,
             int dim_head, int window, int overlap_window, const float* d_rel_h, const float* d_rel_w, void* stream) {
    IRB_TRY(require_device());
    return ocab_attention(d_qkv, dtype == IRB_BF16 ? DT_BF16 : DT_F32, ld, d_out, ldo, n, H, W, heads, dim_head, window,
                          overlap_window, d_rel_h, d_rel_w, (cudaStream_t)stream);
}

int irb_mdta_nchunks(int P, int heads, int n) { return mdta_nchunks(P, heads, n); }

int irb_mdta_gram(const void* d_qkv, int dtype, long long ld, int n, int P, int C, int heads, int nchunks,
                  float* d_Gpart, float* d_ssq_part, void* stream) {
    IRB_TRY(require_device());
    if (dtype == IRB_BF16)
        return mdta_gram_tc((const bf16*)d_qkv, ld, n, P, C, heads, nchunks, d_Gpart, d_ssq_part, (cudaStream_t)stream);
    return mdta_gram_simt(d_qkv, DT_F32, ld, n, P, C, heads, nchunks, d_Gpart, d_ssq_part, (cudaStream_t)stream);
}

int irb_mdta_attn(const float* d_Gpart, const float* d_ssq_part, int nchunks, const float* d_temperature,
                  const float* d_wproj, int n, int C, int heads, void* d_weff, int dtype, float* d_attn,
                  void* stream) {
    IRB_TRY(require_device());
    return mdta_softmax_weff(d_Gpart, d_ssq_part, nchunks, d_temperature, d_wproj, n, C, heads, d_weff, dtype, d_attn,
                             (cudaStream_t)stream);
}

int irb_stitch(const float* d_tiles, int c, int th, int tw, int sf, const int* ys, const int* xs, int ny, int nx,
               int h, int w, int crop_h, int crop_w, float* d_out_f32, uint8_t* d_out_u8, int reverse_channels,
               void* stream) {
    IRB_TRY(require_device());
    return tiles_stitch(d_tiles, c, th, tw, sf, ys, xs, ny, nx, h, w, crop_h, crop_w, d_out_f32, d_out_u8,
                        reverse_channels, (cudaStream_t)stream);
}

int irb_stitch_rows(const float* d_tiles, int c, int th, int tw, int sf, const int* ys, const int* xs, int ny, int nx,
                    int h, int w, int crop_h, int crop_w, int row0, int rows, float* d_out_f32, uint8_t* d_out_u8,
                    int reverse_channels, void* stream) {
    IRB_TRY(require_device());
    return tiles_stitch_rows(d_tiles, c, th, tw, sf, ys, xs, ny, nx, h, w, crop_h, crop_w, row0, rows, d_out_f32,
                             d_out_u8, reverse_channels, (cudaStream_t)stream);
}

}  // extern "C"
