// Shared types for the irb200 kernels: storage dtypes, the fused GEMM/conv epilogue, error plumbing.
#pragma once
#include <utility>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

namespace irb {

typedef __nv_bfloat16 bf16;

enum { DT_F32 = 0, DT_BF16 = 1 };
enum { ACT_NONE = 0, ACT_GELU = 1, ACT_LRELU = 2 };

// ---- error plumbing (host) --------------------------------------------------------------------
void set_error(const std::string& msg);
#define IRB_CHECK_CUDA(expr)                                                                          \
    do {                                                                                              \
        cudaError_t _e = (expr);                                                                      \
        if (_e != cudaSuccess) {                                                                      \
            ::irb::set_error(std::string(#expr) + ": " + cudaGetErrorString(_e) + " @" + __FILE__ + ":" + \
                             std::to_string(__LINE__));                                               \
            return 2;                                                                                 \
        }                                                                                             \
    } while (0)
#define IRB_REQUIRE(cond, msg)                                                        \
    do {                                                                              \
        if (!(cond)) {                                                                \
            ::irb::set_error(std::string("invalid argument: ") + (msg) + " (" #cond ")"); \
            return 1;                                                                 \
        }                                                                             \
    } while (0)
#define IRB_TRY(expr)          \
    do {                       \
        int _rc = (expr);      \
        if (_rc != 0) return _rc; \
    } while (0)

// ---- diagnosis switches -------------------------------------------------------------------------
// The A/B switches of DESIGN.md §5 (IRB_GEMM_DBG, IRB_GEMM_SPEC, ...) exist only in a library built with -DIRB_DIAG
// (build.py --diag -> libirb200_diag.so). The shipped library evaluates every switch to its default at compile
// time: no environment variable can skip stores, MMAs or epilogues, or change numerics, at run time.
#ifdef IRB_DIAG
int diag_env(const char* name, int dflt);
#define IRB_DIAG_ENV(name, dflt) ::irb::diag_env(name, dflt)
#else
#define IRB_DIAG_ENV(name, dflt) (dflt)
#endif

// cudaFuncAttributeMaxDynamicSharedMemorySize is a PER-DEVICE attribute: the opt-in is remembered per
// (function, current device), so a second GPU driven from the same process gets it too. Returns a cudaError_t.
cudaError_t smem_optin(const void* fn, int bytes);
#define IRB_SMEM_OPTIN(fn, bytes) IRB_CHECK_CUDA(::irb::smem_optin(reinterpret_cast<const void*>(fn), bytes))

// ---- optional per-launch profiling (irb_profile_begin/end; bench.py's live roofline) ---------------
// Every launch wrapper calls prof_mark() right after its kernel launch. When profiling is off this is one
// predictable branch; when on, a CUDA event is recorded on the launching stream, so the device time between
// two consecutive marks is the duration of the kernel in between. flops / bytes are the ALGORITHMIC work of
// the launch (each operand read once, each result written once).
extern bool g_prof_on;
void prof_mark_slow(const char* name, int p0, int p1, double flops, double bytes, cudaStream_t s);
static inline void prof_mark(const char* name, int p0, int p1, double flops, double bytes, cudaStream_t s) {
    if (g_prof_on) prof_mark_slow(name, p0, p1, flops, bytes, s);
}

// ---- fused epilogue of every GEMM / implicit-GEMM convolution ---------------------------------
//   y = post_scale * ( act(acc + bias[col]) + res[row, col] ) + post_shift[col]
// `row` is the global output pixel/token index, `col` the output channel.
struct Epi {
    void* out;
    long long ldo;
    int out_dt;
    const float* bias;        // [N] or nullptr
    const void* res;          // residual, indexed like out (ldr) or nullptr
    long long ldr;
    int res_dt;
    int act;
    float slope;              // LeakyReLU slope
    float post_scale;         // 1.0f when unused
    const float* post_shift;  // [N] or nullptr
    // optional fused PixelShuffle(2) store (ps=1): column n = c*4 + i*2 + j of LR pixel (y,x) goes to
    // channel c of HR pixel (2y+i, 2x+j); out is [B,2H,2W,N/4]. ps=2 is PixelUnshuffle(2): out is
    // [B,H/2,W/2,4N]. ps_w / ps_h are the width / height the convolution runs at. ps=3 is PixelShuffle(2) with the
    // weight rows permuted sub-pixel major (column n = (i*2+j) * N/4 + c): every 64-column block is then one
    // contiguous 128-byte run of an HR pixel and goes out as a strided 5-D TMA store.
    int ps;
    int ps_h, ps_w;
    // optional fused LayerNorm of the GEMM's A operand (weights pre-multiplied by the LN gain, see row_stats()):
    //   acc <- rstd[row] * (acc - mu[row] * ln_colsum[col])      (WithBias / nn.LayerNorm; the LN bias is folded
    //   acc <- rstd[row] * acc                                   (BiasFree: ln_colsum == nullptr)     into `bias`)
    const float2* ln_stats;   // [rows] (mu, rstd) or nullptr
    // optional: LayerNorm statistics of the OUTPUT rows (mu, 1/sqrt(var + stats_eps)) for the LayerNorm folded into
    // the next GEMM; only the tcgen05 kernel with one n-tile (N <= 256), a bf16 residual and the TMA-store epilogue
    float2* stats_out;
    float stats_eps;
    const float* ln_colsum;   // [N] sum_k W'[n,k] of the (rounded) folded weights
};

__device__ __forceinline__ float gelu_erf(float x) { return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f)); }

__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float ex2_approx(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// exact-erf GELU with erf from Abramowitz-Stegun 7.1.26 (|abs err| <= 1.5e-7 + MUFU round-off, i.e. far below a
// bf16 ulp); used where the result is stored as bf16. One MUFU.RCP, one MUFU.EX2, 9 FMA-pipe instructions.
__device__ __forceinline__ float gelu_as(float x) {
    const float h = 0.5f * x;
    const float t = rcp_approx(fmaf(0.3275911f * 1.41421356237f, fabsf(h), 1.0f));  // 1 / (1 + p |x| / sqrt2)
    float poly = fmaf(-1.061405429f, t, 1.453152027f);
    poly = fmaf(poly, t, -1.421413741f);
    poly = fmaf(poly, t, 0.284496736f);
    poly = fmaf(poly, t, -0.254829592f);
    poly *= t;                                                   // -(a1 t + ... + a5 t^5)
    const float ez = ex2_approx(h * h * -2.88539008178f);       // exp(-x^2 / 2) = 2^(-2 log2(e) h^2)
    const float e = fmaf(poly, ez, 1.0f);                        // erf(|x| / sqrt2)
    return fmaf(fabsf(h), e, h);                                 // 0.5 x (1 + sign(x) erf(|x| / sqrt2))
}

__device__ __forceinline__ float epi_act(const Epi& e, float x) {
    if (e.act == ACT_GELU) return gelu_erf(x);
    if (e.act == ACT_LRELU) return x > 0.f ? x : x * e.slope;
    return x;
}

__device__ __forceinline__ float ld_as_float(const void* p, long long idx, int dt) {
    return dt == DT_F32 ? reinterpret_cast<const float*>(p)[idx]
                        : __bfloat162float(reinterpret_cast<const bf16*>(p)[idx]);
}
__device__ __forceinline__ void st_from_float(void* p, long long idx, int dt, float v) {
    if (dt == DT_F32)
        reinterpret_cast<float*>(p)[idx] = v;
    else
        reinterpret_cast<bf16*>(p)[idx] = __float2bfloat16_rn(v);
}

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ void unpack_bf16x2(uint32_t u, float& a, float& b) {
    a = __uint_as_float(u << 16);
    b = __uint_as_float(u & 0xFFFF0000u);
}

// Apply the epilogue to NV consecutive columns [col0, col0+NV) of output row `row` and store them.
// Columns >= ncols_total are skipped. (y, x, b) are only used by the PixelShuffle store.
template <int NV>
__device__ __forceinline__ void epi_apply_store(const Epi& e, long long row, int col0, int ncols_total, float* v,
                                                int b = 0, int y = 0, int x = 0) {
    const int nvalid = min(NV, ncols_total - col0);
    if (nvalid <= 0) return;
    if (e.ln_stats) {
        const float2 st = __ldg(e.ln_stats + row);
#pragma unroll
        for (int i = 0; i < NV; ++i)
            if (i < nvalid)
                v[i] = e.ln_colsum ? st.y * (v[i] - st.x * __ldg(e.ln_colsum + col0 + i)) : st.y * v[i];
    }
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        if (i < nvalid) {
            float t = v[i];
            if (e.bias) t += __ldg(e.bias + col0 + i);
            t = epi_act(e, t);
            v[i] = t;
        }
    }
    if (e.res) {
        const long long rbase = row * e.ldr + col0;
        if (e.res_dt == DT_BF16 && nvalid == NV && (NV % 8) == 0 && ((rbase & 7) == 0)) {
            const uint4* rp = reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(e.res) + rbase);
#pragma unroll
            for (int q = 0; q < NV / 8; ++q) {
                uint4 u = __ldg(rp + q);
                float a, c;
                unpack_bf16x2(u.x, a, c); v[q * 8 + 0] += a; v[q * 8 + 1] += c;
                unpack_bf16x2(u.y, a, c); v[q * 8 + 2] += a; v[q * 8 + 3] += c;
                unpack_bf16x2(u.z, a, c); v[q * 8 + 4] += a; v[q * 8 + 5] += c;
                unpack_bf16x2(u.w, a, c); v[q * 8 + 6] += a; v[q * 8 + 7] += c;
            }
        } else {
#pragma unroll
            for (int i = 0; i < NV; ++i)
                if (i < nvalid) v[i] += ld_as_float(e.res, rbase + i, e.res_dt);
        }
    }
    if (e.post_scale != 1.0f || e.post_shift) {
#pragma unroll
        for (int i = 0; i < NV; ++i)
            if (i < nvalid) v[i] = v[i] * e.post_scale + (e.post_shift ? __ldg(e.post_shift + col0 + i) : 0.f);
    }
    if (e.ps == 2) {
        // PixelUnshuffle(2) scatter (restormer_arch.py:175-176): conv pixel (y,x), channel n goes to
        // channel n*4 + (y&1)*2 + (x&1) of pixel (y/2, x/2).
        const long long orow = ((long long)b * (e.ps_h >> 1) + (y >> 1)) * (e.ps_w >> 1) + (x >> 1);
        const int sub = ((y & 1) << 1) | (x & 1);
#pragma unroll
        for (int i = 0; i < NV; ++i)
            if (i < nvalid) st_from_float(e.out, orow * e.ldo + (col0 + i) * 4 + sub, e.out_dt, v[i]);
        return;
    }
    if (e.ps == 3) {
        const int cout = ncols_total >> 2;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            if (i < nvalid) {
                const int n = col0 + i;
                const int sub = n / cout, c = n - sub * cout;
                const long long orow = ((long long)b * (2 * e.ps_h) + (2 * y + (sub >> 1))) * (2 * e.ps_w) + (2 * x + (sub & 1));
                st_from_float(e.out, orow * e.ldo + c, e.out_dt, v[i]);
            }
        }
        return;
    }
    if (e.ps == 1) {
        // PixelShuffle(2) scatter (network_swinir.py:585 / restormer_arch.py:185-186 semantics).
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            if (i < nvalid) {
                const int n = col0 + i;
                const int c = n >> 2, si = (n >> 1) & 1, sj = n & 1;
                const long long orow = ((long long)b * (2 * e.ps_h) + (2 * y + si)) * (2 * e.ps_w) + (2 * x + sj);
                st_from_float(e.out, orow * e.ldo + c, e.out_dt, v[i]);
            }
        }
        return;
    }
    const long long obase = row * e.ldo + col0;
    if (e.out_dt == DT_BF16 && nvalid == NV && (NV % 8) == 0 && ((obase & 7) == 0)) {
        uint4* op = reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(e.out) + obase);
#pragma unroll
        for (int q = 0; q < NV / 8; ++q) {
            uint4 u;
            u.x = pack_bf16x2(v[q * 8 + 0], v[q * 8 + 1]);
            u.y = pack_bf16x2(v[q * 8 + 2], v[q * 8 + 3]);
            u.z = pack_bf16x2(v[q * 8 + 4], v[q * 8 + 5]);
            u.w = pack_bf16x2(v[q * 8 + 6], v[q * 8 + 7]);
            op[q] = u;
        }
    } else if (e.out_dt == DT_F32 && nvalid == NV && (NV % 4) == 0 && ((obase & 3) == 0)) {
        float4* op = reinterpret_cast<float4*>(reinterpret_cast<float*>(e.out) + obase);
#pragma unroll
        for (int q = 0; q < NV / 4; ++q) op[q] = make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
    } else {
#pragma unroll
        for (int i = 0; i < NV; ++i)
            if (i < nvalid) st_from_float(e.out, obase + i, e.out_dt, v[i]);
    }
}

static inline Epi make_epi(void* out, long long ldo, int out_dt) {
    Epi e{};
    e.out = out;
    e.ldo = ldo;
    e.out_dt = out_dt;
    e.post_scale = 1.0f;
    return e;
}

// algorithmic bytes of an epilogue: result written once (+ residual read once) for `rows` x `cols`
// Launch with the PDL attribute (see griddep_wait() in ptx.cuh); IRB_PDL=0 launches the ordinary way (diagnosis).
bool pdl_enabled();
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                     Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

static inline double epi_bytes(const Epi& e, double rows, double cols) {
    const double eo = e.out_dt == DT_BF16 ? 2 : 4, er = e.res_dt == DT_BF16 ? 2 : 4;
    return rows * cols * (eo + (e.res ? er : 0));
}

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline long long ceil_div_ll(long long a, long long b) { return (a + b - 1) / b; }

}  // namespace irb
