// 3x3 convolution (stride 1, zero pad 1) with a handful of output channels (N <= 4): the reconstruction convolutions
// that end both networks (Restormer `output` 96 -> 3, restormer_arch.py:281; SwinIR `conv_last` 64 / 180 -> 3,
// network_swinir.py:836-838), bf16 channels-last in, generic fused epilogue (bias, fp32 residual image, post scale /
// shift, fp32 or bf16 out).
//
// Why not the tcgen05 implicit GEMM: with N = 3 the tensor core has nothing to do, and the implicit GEMM re-reads
// the activation once per tap (9 x 128-pixel boxes per tile through L2): tc_conv3x3[64,3] ran at 0.09 of HBM peak.
// This is a memory-bound stencil: the halo tile is staged ONCE in shared memory (16-byte chunks, pixel stride padded
// to an odd number of chunks so that a quarter-warp of consecutive pixels hits distinct banks), every thread owns PX
// pixels of one column and walks taps x 8-channel chunks with packed FFMA2; the weights are read as broadcast
// 16-byte shared loads and shared by the thread's PX pixels.
#include "common.cuh"
#include "ptx.cuh"
#include "kernels.h"
#include <stdlib.h>

namespace irb {

static constexpr int kCsTW = 32;       // tile width  (= lanes of a warp)
#ifndef IRB_CS_WARPS
#define IRB_CS_WARPS 8
#endif
static constexpr int kCsWarps = IRB_CS_WARPS;     // rows of threads (8: sixteen warps per SM at two CTAs, PX = 2 / 1)
static constexpr int kCsPxBig = 16 / kCsWarps, kCsPxSmall = 8 / kCsWarps;   // pixels per thread of the 16- / 8-row tile
static constexpr int kCsThreads = kCsTW * kCsWarps;

struct ConvSmallParams {
    const bf16* x;
    long long ldx;
    const bf16* wt;  // [npad][9 * kpad], K index = tap * kpad + channel
    int kpad;
    int Nimg, H, W, C, N;
    int nchunk;      // ceil(C / 8)
    int pstride;     // shared-memory pixel stride in 16-byte units (odd)
    uint32_t nchunk_magic;  // ceil(2^32 / nchunk)
    int tiles_x, tiles_y;
    Epi epi;
};

template <int PX, int NOUT>
__global__ void __launch_bounds__(kCsThreads) conv3x3_small_n_kernel(const ConvSmallParams p) {
    constexpr int TH = PX * kCsWarps;
    extern __shared__ uint4 cs_smem[];
    uint4* halo = cs_smem;                                             // [(TH+2)][(TW+2)][pstride]
    float* wsm = reinterpret_cast<float*>(cs_smem + (TH + 2) * (kCsTW + 2) * p.pstride);  // [9][nchunk][NOUT][8]
    const int tid = threadIdx.x;
    const int img = blockIdx.z;
    const int x0 = blockIdx.x * kCsTW, y0 = blockIdx.y * TH;

    // ---- weights: bf16 [n][tap * kpad + c] -> fp32 [tap][chunk][n][8], zero for c >= C ----
    // one item = (tap, chunk, n): a 16-byte bf16 vector of the packed weights -> 8 floats (kpad % 8 == 0)
    const int witems = 9 * p.nchunk * NOUT;
    for (int i = tid; i < witems; i += kCsThreads) {
        const int n = i % NOUT;
        const int tc = i / NOUT;  // tap * nchunk + chunk
        const int tap = tc / p.nchunk, ck = tc - tap * p.nchunk;
        uint4 u = make_uint4(0u, 0u, 0u, 0u);
        if (n < p.N) u = __ldg(reinterpret_cast<const uint4*>(p.wt + (long long)n * 9 * p.kpad + tap * p.kpad + ck * 8));
        float4 lo, hi;  // channels >= C are zero in the packed weights (pack_conv3x3 pads K with zeros)
        unpack_bf16x2(u.x, lo.x, lo.y);
        unpack_bf16x2(u.y, lo.z, lo.w);
        unpack_bf16x2(u.z, hi.x, hi.y);
        unpack_bf16x2(u.w, hi.z, hi.w);
        reinterpret_cast<float4*>(wsm)[2 * i] = lo;
        reinterpret_cast<float4*>(wsm)[2 * i + 1] = hi;
    }
    // ---- halo tile: (TH+2) x (TW+2) pixels x nchunk chunks; out-of-image pixels and channels >= C are zero ----
    const int hw = kCsTW + 2;
    const int total = (TH + 2) * hw * p.nchunk;
    // cp.async: every thread's ~30 chunk copies are in flight together (a load -> store loop was one DRAM latency per
    // iteration: 32 us per CTA, slower than the implicit GEMM it replaces)
    const uint32_t halo_s0 = smem_u32(halo);
    for (int i = tid; i < total; i += kCsThreads) {
        const int px = int(__umulhi(uint32_t(i), p.nchunk_magic));  // i / nchunk (exact for i < 2^16)
        const int ck = i - px * p.nchunk;
        const int hy = px / hw, hx = px - hy * hw;
        const int gy = y0 + hy - 1, gx = x0 + hx - 1;
        const bool inside = gy >= 0 && gy < p.H && gx >= 0 && gx < p.W;
        // one zero-filling cp.async for every item: whole chunks copy 16 bytes, the ragged last chunk (C % 8 != 0) its
        // valid channels, out-of-image pixels nothing (src_bytes 0 reads no memory). The first version took a
        // load -> pack -> store path for ragged chunks: with C = 180 every warp holds one (23 chunks per pixel), so
        // every pass of every warp waited a DRAM round trip: 44 % of the kernel's stall samples on that one
        // instruction, 0.037 of the HBM peak (ncu profiles/r02_s85).
        const int valid = inside ? min(8, p.C - ck * 8) * 2 : 0;
        const bf16* src = inside ? p.x + (((long long)img * p.H + gy) * p.W + gx) * p.ldx + ck * 8 : p.x;
        cp_async16_zfill(halo_s0 + (px * p.pstride + ck) * 16, src, valid);
    }
    cp_async_commit();
    cp_async_wait<0>();
    __syncthreads();

    const int tx = tid & 31, ty = tid >> 5;
    f32x2 acc[PX][NOUT];
#pragma unroll
    for (int i = 0; i < PX; ++i)
#pragma unroll
        for (int n = 0; n < NOUT; ++n) acc[i][n] = 0ull;

    const uint32_t halo_s = smem_u32(halo);
    const uint32_t w_s = smem_u32(wsm);
#pragma unroll 1
    for (int tap = 0; tap < 9; ++tap) {
        const int dy = tap / 3, dx = tap - dy * 3;  // halo offsets (already +1)
        // pixel i of this thread: output row ty + i * kCsWarps
        uint32_t pa[PX];
#pragma unroll
        for (int i = 0; i < PX; ++i) pa[i] = halo_s + (((ty + i * kCsWarps + dy) * hw + tx + dx) * p.pstride) * 16;
        uint32_t wa = w_s + tap * p.nchunk * NOUT * 32;
#pragma unroll 4
        for (int ck = 0; ck < p.nchunk; ++ck) {
            f32x2 w[NOUT][4];
#pragma unroll
            for (int n = 0; n < NOUT; ++n) {
                lds_f32x4(wa + n * 32, w[n][0], w[n][1]);
                lds_f32x4(wa + n * 32 + 16, w[n][2], w[n][3]);
            }
            wa += NOUT * 32;
#pragma unroll
            for (int i = 0; i < PX; ++i) {
                const uint4 d = lds_v4(pa[i] + ck * 16);
                const f32x2 x0v = bf16x2_to_f32x2_alu(d.x), x1v = bf16x2_to_f32x2_alu(d.y);
                const f32x2 x2v = bf16x2_to_f32x2_alu(d.z), x3v = bf16x2_to_f32x2_alu(d.w);
#pragma unroll
                for (int n = 0; n < NOUT; ++n) {
                    ffma2(acc[i][n], x0v, w[n][0]);
                    ffma2(acc[i][n], x1v, w[n][1]);
                    ffma2(acc[i][n], x2v, w[n][2]);
                    ffma2(acc[i][n], x3v, w[n][3]);
                }
            }
        }
    }
    const int gx = x0 + tx;
#pragma unroll
    for (int i = 0; i < PX; ++i) {
        const int gy = y0 + ty + i * kCsWarps;
        if (gx < p.W && gy < p.H) {
            float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int n = 0; n < NOUT; ++n) {
                float lo, hi;
                unpack2(acc[i][n], lo, hi);
                v[n] = lo + hi;
            }
            const long long row = ((long long)img * p.H + gy) * p.W + gx;
            epi_apply_store<4>(p.epi, row, 0, p.N, v, img, gy, gx);
        }
    }
}

bool conv3x3_small_n_ok(int C, int N, const Epi& e) {
    if (!(N >= 1 && N <= 4 && C >= 16 && C <= 256 && e.ps == 0 && e.ln_stats == nullptr && e.stats_out == nullptr))
        return false;
    const int nchunk = (C + 7) / 8, pstride = nchunk | 1;
    const size_t smem = size_t(8 + 2) * (kCsTW + 2) * pstride * 16 + size_t(9) * nchunk * 4 * 32;
    return smem <= 200 * 1024;
}

template <int PX, int NOUT>
static int launch_cs(const ConvSmallParams& p, size_t smem, cudaStream_t stream) {
    IRB_SMEM_OPTIN((conv3x3_small_n_kernel<PX, NOUT>), 200 * 1024);
    dim3 grid(p.tiles_x, p.tiles_y, p.Nimg);
    conv3x3_small_n_kernel<PX, NOUT><<<grid, kCsThreads, smem, stream>>>(p);
    IRB_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int conv3x3_small_n(const bf16* x, long long ldx, int Nimg, int H, int W, int C, const bf16* Wt, int kpad, int N,
                    const Epi& epi, cudaStream_t stream) {
    IRB_REQUIRE(conv3x3_small_n_ok(C, N, epi), "conv3x3_small_n: unsupported shape / epilogue");
    IRB_REQUIRE(ldx % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0, "conv3x3_small_n: rows must be 16-byte aligned");
    IRB_REQUIRE(Nimg <= 65535, "conv3x3_small_n: too many images in one launch");
    IRB_REQUIRE(kpad % 8 == 0 && kpad >= (C + 7) / 8 * 8 && (reinterpret_cast<uintptr_t>(Wt) & 15) == 0, "conv3x3_small_n: packed weights must be 16-byte aligned, K padded to 8");
    ConvSmallParams p{};
    p.x = x;
    p.ldx = ldx;
    p.wt = Wt;
    p.kpad = kpad;
    p.Nimg = Nimg;
    p.H = H;
    p.W = W;
    p.C = C;
    p.N = N;
    p.nchunk = (C + 7) / 8;
    p.pstride = p.nchunk | 1;
    p.nchunk_magic = uint32_t(((1ull << 32) + p.nchunk - 1) / p.nchunk);
    p.epi = epi;
    // 16-row tiles (kCsPxBig pixels per thread) when two CTAs of that size still fit an SM, else 8-row tiles. Eight warps
    // per CTA instead of four (same tiles, half the pixels per thread): [64 -> 3] 2.58 -> 2.26 ms (profiles/r02_s92)
    const size_t wbytes = size_t(9) * p.nchunk * (N <= 1 ? 1 : (N <= 3 ? 3 : 4)) * 32;
    const size_t smem4 = size_t(16 + 2) * (kCsTW + 2) * p.pstride * 16 + wbytes;
    const size_t smem2 = size_t(8 + 2) * (kCsTW + 2) * p.pstride * 16 + wbytes;
    const bool px4 = smem4 <= 100 * 1024 && H >= 16;
    const int TH = px4 ? 16 : 8;
    p.tiles_x = ceil_div(W, kCsTW);
    p.tiles_y = ceil_div(H, TH);
    const size_t smem = px4 ? smem4 : smem2;
    int rc;
    if (N <= 1)
        rc = px4 ? launch_cs<kCsPxBig, 1>(p, smem, stream) : launch_cs<kCsPxSmall, 1>(p, smem, stream);
    else if (N <= 3)
        rc = px4 ? launch_cs<kCsPxBig, 3>(p, smem, stream) : launch_cs<kCsPxSmall, 3>(p, smem, stream);
    else
        rc = px4 ? launch_cs<kCsPxBig, 4>(p, smem, stream) : launch_cs<kCsPxSmall, 4>(p, smem, stream);
    IRB_TRY(rc);
    const double rows = double(H) * W * Nimg;
    prof_mark("conv3x3_small_n", C, N, 18.0 * rows * C * N, rows * C * 2 + 9.0 * N * C * 2 + epi_bytes(epi, rows, N), stream);
    return 0;
}

// bf16 3x3 convolution: the stencil kernel above for N <= 4 outputs, the tcgen05 implicit GEMM otherwise
// (IRB_CONV_SMALLN=0 forces the implicit GEMM; diagnosis).
int conv3x3_bf16(const bf16* x, long long ldx, int Nimg, int H, int W, int C, const bf16* Wt, int kpad, int N,
                 int w_rows_total, const Epi& epi, cudaStream_t stream) {
    static int on = -1;
    if (on < 0) {
        on = IRB_DIAG_ENV("IRB_CONV_SMALLN", 1);
    }
    // With two or more 64-channel K slabs the implicit GEMM's dy-reuse path (gemm_tc.cu, one 16 x 10 pixel A box per
    // (dx, K slab), N padded to 16) now beats the stencil: [180 -> 3] 0.09 vs 0.19 ms, [96 -> 3] 0.35 vs 0.37 ms; with one
    // slab ([64 -> 3] at HR resolution) the stencil wins, 1.32 vs 1.63 ms (profiles/r02_s87, same box).
    const bool stencil = C <= 64 || W < 16 || H < 8;
    if (on && stencil && conv3x3_small_n_ok(C, N, epi) && ldx % 8 == 0 && Nimg <= 65535)
        return conv3x3_small_n(x, ldx, Nimg, H, W, C, Wt, kpad, N, epi, stream);
    return tc_conv3x3(x, ldx, Nimg, H, W, C, Wt, kpad, N, w_rows_total, epi, stream);
}

}  // namespace irb
