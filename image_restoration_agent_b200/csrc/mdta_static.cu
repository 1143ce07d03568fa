// irb_mdta_attn_static: the MDTA channel attention (Gram over pixels -> softmax -> fold into project_out) in ONE
// entry point that allocates nothing. The existing pair irb_mdta_gram + irb_mdta_attn hands the per-chunk partial sums
// (Gpart / ssq_part) from the Gram kernel to the softmax kernel through scratch buffers that the CALLER allocates; some
// callers cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc). Here
// that scratch lives in a file-scope __device__ array of a fixed, modest size: every GPU that loads the library holds
// kStaticScratchBytes of it until the process exits.
//
// No counterpart in the reference. The kernels are the unchanged ones of
// mdta.cu (mdta_gram_tc / mdta_gram_simt / mdta_softmax_weff), so the results are bit-identical to the existing pair.
//
// Rules of the shared array:
//  * every call uses the SAME array, which keeps the previous call's contents. Nothing has to start at zero: the Gram
//    kernel writes every slot [img][head][chunk] it later reads (one slot per chunk, no atomics), so no clearing pass is
//    needed and none is done;
//  * two calls that overlap on the GPU (different streams, or graph branches without a dependency) RACE on it: issue
//    them on one stream or order them with events;
//  * the array's name is a host-side stand-in in host code, not a device address: the device address is looked up once
//    per device with cudaGetSymbolAddress (that device current) and cached.
#include "../../include/irb200.h"
#include "common.cuh"
#include "kernels.h"
#include <mutex>

namespace irb {

static constexpr size_t kStaticScratchFloats = size_t(1) << 20;                     // 4 MiB
static constexpr size_t kStaticScratchBytes = kStaticScratchFloats * sizeof(float);
static __device__ __align__(256) float g_mdta_scratch[kStaticScratchFloats];

static constexpr int kMaxDevices = 64;
static float* g_scratch_ptr[kMaxDevices] = {};
static std::mutex g_scratch_mu;

// device address of g_mdta_scratch on the CURRENT device (looked up once per device)
static int static_scratch(float** out) {
    int dev = -1;
    IRB_CHECK_CUDA(cudaGetDevice(&dev));
    IRB_REQUIRE(dev >= 0 && dev < kMaxDevices, "irb_mdta_attn_static: device index out of range");
    std::lock_guard<std::mutex> lock(g_scratch_mu);
    if (!g_scratch_ptr[dev]) {
        float* p = nullptr;
        IRB_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_mdta_scratch));
        IRB_REQUIRE(p != nullptr && (reinterpret_cast<uintptr_t>(p) & 255) == 0, "irb_mdta_attn_static: bad scratch address");
        g_scratch_ptr[dev] = p;
    }
    *out = g_scratch_ptr[dev];
    return 0;
}

}  // namespace irb

using namespace irb;

extern "C" size_t irb_mdta_attn_static_max_bytes(void) { return kStaticScratchBytes; }

extern "C" int irb_mdta_attn_static(const void* d_qkv, int dtype, long long ld, int n, int P, int C, int heads,
                                    const float* d_temperature, const float* d_wproj, void* d_weff, float* d_attn,
                                    void* stream) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        set_error("no CUDA device: irb200 has no CPU fallback");
        return 2;
    }
    IRB_REQUIRE(d_qkv && d_temperature && d_wproj && d_weff && d_attn, "irb_mdta_attn_static: NULL argument");
    IRB_REQUIRE(dtype == IRB_BF16 || dtype == IRB_F32, "irb_mdta_attn_static: bad dtype");
    IRB_REQUIRE(n > 0 && P > 0 && C > 0 && heads > 0 && C % heads == 0, "irb_mdta_attn_static: bad shape");
    const int c = C / heads;
    const int nchunks = mdta_nchunks(P, heads, n);
    // Gpart [n][heads][nchunks][c][c], then ssq_part [n][heads][nchunks][2][c] (256-byte aligned start)
    const size_t g_floats = (size_t(n) * heads * nchunks * c * c + 63) & ~size_t(63);
    const size_t s_floats = size_t(n) * heads * nchunks * 2 * c;
    if (g_floats + s_floats > kStaticScratchFloats) {
        set_error("irb_mdta_attn_static: this shape needs " + std::to_string((g_floats + s_floats) * sizeof(float)) +
                  " bytes of scratch, the static array holds " + std::to_string(kStaticScratchBytes) +
                  " (n * heads * nchunks * (c*c + 2*c) floats): use irb_mdta_gram + irb_mdta_attn with caller scratch");
        return 1;
    }
    float* scratch = nullptr;
    IRB_TRY(static_scratch(&scratch));
    float* Gpart = scratch;
    float* ssq = scratch + g_floats;
    cudaStream_t s = (cudaStream_t)stream;
    if (dtype == IRB_BF16)
        IRB_TRY(mdta_gram_tc((const bf16*)d_qkv, ld, n, P, C, heads, nchunks, Gpart, ssq, s));
    else
        IRB_TRY(mdta_gram_simt(d_qkv, DT_F32, ld, n, P, C, heads, nchunks, Gpart, ssq, s));
    return mdta_softmax_weff(Gpart, ssq, nchunks, d_temperature, d_wproj, n, C, heads, d_weff,
                             dtype == IRB_BF16 ? DT_BF16 : DT_F32, d_attn, s);
}
