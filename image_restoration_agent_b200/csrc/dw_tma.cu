// Depthwise 3x3 (zero pad 1) on channels-last bf16 activations, fed by TMA, for sm_100a.
//   plain : y[p, c]  = dw(x)[p, c]                                  (MDTA qkv_dwconv, restormer_arch.py:106,116)
//   gate  : y[p, c]  = gelu(dw(x)[p, c]) * dw(x)[p, hp + c]          (GDFN, restormer_arch.py:88-93)
//
// Work item = (image, 32x8 pixel tile, 64-channel chunk). A persistent CTA walks its items with a 2-stage
// shared-memory ring: one producer thread issues a 4-D TMA box {64 ch, 34, 10, 1} per stream (the halo'd tile;
// rows / columns outside the image are ZERO-FILLED by TMA, which is exactly the convolution's zero padding), 16
// consumer warps compute from shared memory. Consumer warp = two adjacent output columns, lane = one channel PAIR
// (a warp reads the 128 contiguous bytes of one pixel: conflict-free, one wavefront). The warp walks down the 10
// halo rows; every input row is loaded and converted ONCE (4 columns) and scattered into the three output rows it
// touches (input-stationary, 3 live accumulator rows), so each element costs one LDS.32 + two ALU ops per 2 output
// columns. Arithmetic is fp32 with packed fma.rn.f32x2 (both channels of the pair in one instruction); the GELU is
// evaluated packed as well, in its sigmoid form (gelu2_fast: the kernel is bound by the FP32 pipe, 2 MUFU per element).
#include <cstdio>
#include "common.cuh"
#include "ptx.cuh"
#include "kernels.h"
#include <stdlib.h>

#ifndef IRB_DW_GATE_TH
#define IRB_DW_GATE_TH 8
#endif

namespace irb {

static constexpr int kDwTW = 32;                               // output tile width (pixels)
static constexpr int kDwHW = kDwTW + 2;                        // halo'd tile width
static constexpr int kDwCh = 64;                               // channels per chunk = one 128-byte row
// Output tile height and ring depth. plain: 32x8 tiles, 4 stages of 43.5 KB. gate (two streams per stage): 32x8 tiles,
// 2 stages of 87 KB. ncu (profiles/r02_s08_ncu_dw_gate.txt) shows 15 % of the gate kernel's stall samples on the
// TMA-full barrier, so 32x6 tiles with 3 stages of 69.6 KB were tried (-DIRB_DW_GATE_TH=6): 6.49-6.64 ms against
// 6.34-6.40 ms per 20 launches in a same-box A/B (profiles/r02_s09_ab_*): the extra halo rows (8/6 instead of 10/8 per
// output row) cost more than the third stage gains. The default stays 8.
template <int GATE> struct DwGeom {
    static constexpr int TH = GATE != 0 ? IRB_DW_GATE_TH : 8;
    static constexpr int HH = TH + 2;
    static constexpr int kStream = HH * kDwHW * kDwCh * 2;    // bytes of one stream of one stage
    static constexpr int kStages = GATE != 0 ? (IRB_DW_GATE_TH == 8 ? 2 : 3) : 4;
};
static constexpr int kDwConsumerWarps = kDwTW / 2;             // 16
static constexpr int kDwThreads = (kDwConsumerWarps + 1) * 32; // + producer warp

struct alignas(64) DwParams {
    CUtensorMap tm;  // {Ctot, W, H, N}, box {64, 34, 10, 1}, no swizzle
    bf16* y;
    long long ldy;
    int H, W, Cout, hp_off;
    const float* wt;  // [9][wt_ld] fp32 tap-major
    int wt_ld;
    int tiles_x, tiles_y, chunks, total;
};

// Work order: CTA b owns channel chunk (b % chunks) for the whole launch (its 9 x 2 weight pairs stay in registers)
// and walks the spatial tiles (image, tile row, tile column) b / chunks, + grid / chunks, ...; the grid is a
// multiple of `chunks`, so at any moment the CTAs cover ALL chunks of one neighbourhood of tiles: DRAM rows are
// consumed whole and halo re-reads hit L2. The producer decodes the tile once and hands (img, ty, tx) to the
// consumers through the stage's shared-memory slot, so the 16 consumer warps do no integer division.
template <int GATE>  // 0: plain; 1: GELU gate, sigmoid form (ex2 + rcp); 2: GELU gate, tanh form (MUFU.TANH)
__global__ void __launch_bounds__(kDwThreads, 1) dw3x3_tma_kernel(const __grid_constant__ DwParams p) {
    constexpr int NS = GATE != 0 ? 2 : 1;
    constexpr int kDwTH = DwGeom<GATE>::TH, kDwHH = DwGeom<GATE>::HH;
    constexpr int kDwStreamBytes = DwGeom<GATE>::kStream;
    constexpr int stage_bytes = NS * kDwStreamBytes;
    constexpr int kDwStages = DwGeom<GATE>::kStages;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + kDwStages * stage_bytes);
    uint64_t* empty = full + kDwStages;
    int4* item = reinterpret_cast<int4*>(empty + kDwStages);  // [stages] (img, ty, tx, chunk)
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < kDwStages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], kDwConsumerWarps);
        }
        fence_mbar_init();
    }
    __syncthreads();
    griddep_wait();  // PDL: the barrier setup above overlapped the previous kernel's tail
    griddep_launch();
    const int per_chunk = p.total / p.chunks;  // images * tiles
    const int per_img = p.tiles_x * p.tiles_y;
    const int chunk = blockIdx.x % p.chunks;
    const int sp0 = blockIdx.x / p.chunks, sp_step = gridDim.x / p.chunks;

    if (warp == kDwConsumerWarps) {
        if (elect_one()) {
            tma_prefetch_desc(&p.tm);
            int it = 0;
            for (int sp = sp0; sp < per_chunk; sp += sp_step, ++it) {
                const int s = it % kDwStages;
                const uint32_t ph = (it / kDwStages) & 1;
                mbar_wait(&empty[s], ph ^ 1);
                const int img = sp / per_img;
                int r = sp - img * per_img;
                const int ty = r / p.tiles_x, tx = r - ty * p.tiles_x;
                item[s] = make_int4(img, ty, tx, chunk);  // ordered before the consumers' reads by the mbarrier
                uint8_t* dst = smem + s * stage_bytes;
                mbar_arrive_expect_tx(&full[s], stage_bytes);
                tma_load_4d(dst, &p.tm, &full[s], chunk * kDwCh, tx * kDwTW - 1, ty * kDwTH - 1, img);
                if (GATE != 0)
                    tma_load_4d(dst + kDwStreamBytes, &p.tm, &full[s], p.hp_off + chunk * kDwCh, tx * kDwTW - 1,
                                ty * kDwTH - 1, img);
            }
        }
        return;
    }

    // ---- consumers: warp = output columns (2*warp, 2*warp + 1), lane = channel pair ----
    const int col0 = warp * 2;
    const uint32_t lane_off = smem_u32(smem) + (col0 * kDwCh + lane * 2) * 2;
    const long long row_step = (long long)p.W * p.ldy;
    const int c0 = chunk * kDwCh + lane * 2;
    const bool cvalid = c0 < p.Cout;
    f32x2 wr[NS][9];
#pragma unroll
    for (int st = 0; st < NS; ++st)
#pragma unroll
        for (int t = 0; t < 9; ++t) {
            float2 w2 = make_float2(0.f, 0.f);
            if (cvalid) w2 = __ldg(reinterpret_cast<const float2*>(p.wt + (long long)t * p.wt_ld + c0 + st * p.hp_off));
            wr[st][t] = pack2(w2.x, w2.y);
        }
    int it = 0;
#ifdef IRB_DW_TRACE  // timing experiment: clock() trace of consumer warp 0 of CTA 0 (gate kernel, Cout == IRB_DW_TRACE)
    const bool trace = GATE != 0 && blockIdx.x == 0 && threadIdx.x == 0 && p.Cout == IRB_DW_TRACE;
    long long tsv[120];
    int tsi[120];
    int nts = 0;
#define IRB_TS(i) do { if (trace && nts < 120) { tsv[nts] = clock64(); tsi[nts++] = (i); } } while (0)
#else
#define IRB_TS(i) do { } while (0)
#endif
    for (int sp = sp0; sp < per_chunk; sp += sp_step, ++it) {
        const int s = it % kDwStages;
        const uint32_t ph = (it / kDwStages) & 1;
        IRB_TS(0);
        mbar_wait(&full[s], ph);
        IRB_TS(1);
        const int4 id = item[s];
        const int img = id.x, ty = id.y, tx = id.z;
        const uint32_t tile = lane_off + s * stage_bytes;
        const int gx = tx * kDwTW + col0;
        const int nrows = min(kDwTH, p.H - ty * kDwTH);
        const bool ok0 = cvalid && gx < p.W, ok1 = cvalid && gx + 1 < p.W;
        bf16* yp0 = p.y + (((long long)img * p.H + ty * kDwTH) * p.W + gx) * p.ldy + c0;
        bf16* yp1 = yp0 + p.ldy;
        const bool fast = ok1 && nrows == kDwTH;  // lane-varying only through cvalid (ragged last chunk)
        f32x2 acc[3][2][NS];  // [output row mod 3][pixel][stream]
#pragma unroll
        for (int hr = 0; hr < kDwHH; ++hr) {
            f32x2 in[NS][4];
#pragma unroll
            for (int st = 0; st < NS; ++st)
#pragma unroll
                for (int d = 0; d < 4; ++d)
                    in[st][d] = bf16x2_to_f32x2_alu(lds_u32(tile + st * kDwStreamBytes + (hr * kDwHW + d) * kDwCh * 2));
            // halo row hr is tap row dr of output row hr - dr
#pragma unroll
            for (int dr = 0; dr < 3; ++dr) {
                const int orow = hr - dr;
                if (orow < 0 || orow >= kDwTH) continue;
                const int sl = orow % 3;
#pragma unroll
                for (int st = 0; st < NS; ++st)
#pragma unroll
                    for (int px = 0; px < 2; ++px) {
                        if (dr == 0)
                            acc[sl][px][st] = mul2(in[st][px], wr[st][0]);
                        else
                            ffma2(acc[sl][px][st], in[st][px], wr[st][dr * 3]);
                        ffma2(acc[sl][px][st], in[st][px + 1], wr[st][dr * 3 + 1]);
                        ffma2(acc[sl][px][st], in[st][px + 2], wr[st][dr * 3 + 2]);
                    }
            }
            const int done = hr - 2;  // this output row is complete
            if (done >= 0) {
                const int sl = done % 3;
                uint32_t pk[2];
#pragma unroll
                for (int px = 0; px < 2; ++px) {
                    f32x2 o = acc[sl][px][0];
                    if (GATE == 1) o = mul2(gelu2_fast(o), acc[sl][px][NS - 1]);
                    if (GATE == 2) o = mul2(gelu2_tanh(o), acc[sl][px][NS - 1]);
                    float o0, o1;
                    unpack2(o, o0, o1);
                    pk[px] = pack_bf16x2(o0, o1);
                }
                if (fast) {
                    *reinterpret_cast<uint32_t*>(yp0) = pk[0];
                    *reinterpret_cast<uint32_t*>(yp1) = pk[1];
                } else if (done < nrows) {
                    if (ok0) *reinterpret_cast<uint32_t*>(yp0) = pk[0];
                    if (ok1) *reinterpret_cast<uint32_t*>(yp1) = pk[1];
                }
                yp0 += row_step;
                yp1 += row_step;
            }
        }
        IRB_TS(2);
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
        IRB_TS(3);
    }
#ifdef IRB_DW_TRACE
    if (trace) {
        static __device__ int printed;
        if (atomicAdd(&printed, 1) == 0)
            for (int i = 1; i < nts; ++i) printf("TS %d %d %lld\n", i, tsi[i], tsv[i] - tsv[i - 1]);
    }
#endif
}

static int dw_tma_launch(bool gate, const bf16* x, long long ldx, int Ctot, bf16* y, long long ldy, int Nimg, int H,
                         int W, int Cout, int hp_off, const float* wt, int wt_ld, cudaStream_t stream) {
    IRB_REQUIRE(ldx % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0, "dw3x3: x must be 16-byte aligned rows");
    IRB_REQUIRE(Cout % 2 == 0 && ldy % 2 == 0 && (reinterpret_cast<uintptr_t>(y) & 3) == 0, "dw3x3: y alignment");
    DwParams p{};
    uint64_t dims[4] = {uint64_t(Ctot), uint64_t(W), uint64_t(H), uint64_t(Nimg)};
    uint64_t str[3] = {uint64_t(ldx) * 2, uint64_t(ldx) * 2 * W, uint64_t(ldx) * 2 * W * H};
    const int kDwTH = gate ? DwGeom<1>::TH : DwGeom<0>::TH;
    const int kDwStreamBytes = gate ? DwGeom<1>::kStream : DwGeom<0>::kStream;
    uint32_t box[4] = {kDwCh, kDwHW, uint32_t(kDwTH + 2), 1};
    IRB_TRY(make_tmap_bf16_sw(&p.tm, x, 4, dims, str, box, 0));
    p.y = y;
    p.ldy = ldy;
    p.H = H;
    p.W = W;
    p.Cout = Cout;
    p.hp_off = hp_off;
    p.wt = wt;
    p.wt_ld = wt_ld;
    p.tiles_x = ceil_div(W, kDwTW);
    p.tiles_y = ceil_div(H, kDwTH);
    p.chunks = ceil_div(Cout, kDwCh);
    const long long total = (long long)p.tiles_x * p.tiles_y * p.chunks * Nimg;
    IRB_REQUIRE(total < (1ll << 31), "dw3x3: too many work items");
    p.total = int(total);
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    IRB_REQUIRE(p.chunks <= sms, "dw3x3: too many channel chunks");
    const int spatial = p.total / p.chunks;
    const int groups = spatial < sms / p.chunks ? spatial : sms / p.chunks;  // CTAs per chunk
    const int grid = groups * p.chunks;
    const int nst = gate ? DwGeom<1>::kStages : DwGeom<0>::kStages;
    const size_t smem = size_t(nst) * (gate ? 2 : 1) * kDwStreamBytes + 128 + nst * 32;
    IRB_SMEM_OPTIN((dw3x3_tma_kernel<2>), 227 * 1024);
    IRB_SMEM_OPTIN((dw3x3_tma_kernel<1>), 227 * 1024);
    IRB_SMEM_OPTIN((dw3x3_tma_kernel<0>), 227 * 1024);
    static int gelu_tanh = -1;
    if (gelu_tanh < 0) {
        // measured equal to the sigmoid form here (6.07 vs 6.06 ms on cfg2: this kernel is not SFU-bound), so the
        // default stays the more accurate sigmoid form
        gelu_tanh = IRB_DIAG_ENV("IRB_DW_GELU_TANH", 0);
    }
    if (gate && gelu_tanh)
        IRB_CHECK_CUDA(launch_pdl(dw3x3_tma_kernel<2>, dim3(grid), dim3(kDwThreads), smem, stream, p));
    else if (gate)
        IRB_CHECK_CUDA(launch_pdl(dw3x3_tma_kernel<1>, dim3(grid), dim3(kDwThreads), smem, stream, p));
    else
        IRB_CHECK_CUDA(launch_pdl(dw3x3_tma_kernel<0>, dim3(grid), dim3(kDwThreads), smem, stream, p));
    const double px = double(Nimg) * H * W, cin = gate ? 2.0 * Cout : double(Cout);
    prof_mark(gate ? "dw3x3_gate_tma" : "dw3x3_tma", Cout, 2, 18.0 * px * cin, px * (cin + Cout) * 2 + 36.0 * cin, stream);
    return 0;
}

int dwconv3x3_bf16_tma(const bf16* x, long long ldx, bf16* y, long long ldy, int Nimg, int H, int W, int C,
                       const float* wt, cudaStream_t stream) {
    return dw_tma_launch(false, x, ldx, C, y, ldy, Nimg, H, W, C, 0, wt, C, stream);
}

int dwconv3x3_gate_bf16_tma(const bf16* x, long long ldx, bf16* y, long long ldy, int Nimg, int H, int W, int hp,
                            const float* wt, cudaStream_t stream) {
    return dw_tma_launch(true, x, ldx, 2 * hp, y, ldy, Nimg, H, W, hp, hp, wt, 2 * hp, stream);
}

}  // namespace irb
