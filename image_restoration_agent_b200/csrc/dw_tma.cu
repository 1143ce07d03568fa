// Depthwise 3x3 (zero pad 1) on channels-last bf16 activations, fed by TMA, for sm_100a.
//   plain : y[p, c]  = dw(x)[p, c]                                  (MDTA qkv_dwconv, restormer_arch.py:106,116)
//   gate  : y[p, c]  = gelu(dw(x)[p, c]) * dw(x)[p, hp + c]          (GDFN, restormer_arch.py:88-93)
//
// Work item = (image, 32x8 pixel tile, 64-channel chunk). A persistent CTA walks its items with a 2-stage
// shared-memory ring: one producer thread issues a 4-D TMA box {64 ch, 34, 10, 1} per stream (the halo'd tile;
// rows / columns outside the image are ZERO-FILLED by TMA, which is exactly the convolution's zero padding), 256
// consumer threads compute from shared memory. Consumer thread = (4-channel group, column); it walks down the 8
// rows of two columns with a 3-row register window, so every input element is read from shared memory 3 times
// (8-byte conflict-free loads) and from HBM once (+ halo). Arithmetic is fp32 with packed FFMA2 (fma.rn.f32x2).
#include "common.cuh"
#include "ptx.cuh"
#include "kernels.h"

namespace irb {

static constexpr int kDwTW = 32, kDwTH = 8;                    // output tile (pixels)
static constexpr int kDwHW = kDwTW + 2, kDwHH = kDwTH + 2;     // halo'd tile
static constexpr int kDwCh = 64;                               // channels per chunk = one 128-byte row
static constexpr int kDwStreamBytes = kDwHH * kDwHW * kDwCh * 2;  // 43 520
static constexpr int kDwStages = 2;

struct alignas(64) DwParams {
    CUtensorMap tm;  // {Ctot, W, H, N}, box {64, 34, 10, 1}, no swizzle
    bf16* y;
    long long ldy;
    int H, W, Cout, hp_off;
    const float* wt;  // [9][wt_ld] fp32 tap-major
    int wt_ld;
    int tiles_x, tiles_y, chunks, total;
};

typedef unsigned long long f32x2;  // two packed fp32 (lo = even channel, hi = odd channel)

__device__ __forceinline__ f32x2 pack2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpack2(f32x2 v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 bf16x2_to_f32x2(uint32_t u) {
    return pack2(__uint_as_float(u << 16), __uint_as_float(u & 0xFFFF0000u));
}
__device__ __forceinline__ void ffma2(f32x2& acc, f32x2 a, f32x2 b) {
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(a), "l"(b));
}

template <bool GATE>
__global__ void __launch_bounds__(288, 1) dw3x3_tma_kernel(const __grid_constant__ DwParams p) {
    constexpr int NS = GATE ? 2 : 1;
    constexpr int stage_bytes = NS * kDwStreamBytes;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 127) & ~uintptr_t(127));
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + kDwStages * stage_bytes);
    uint64_t* empty = full + kDwStages;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int s = 0; s < kDwStages; ++s) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], 8);
        }
        fence_mbar_init();
    }
    __syncthreads();
    const int per_img = p.tiles_x * p.tiles_y * p.chunks;

    if (warp == 8) {
        if (lane == 0) {
            tma_prefetch_desc(&p.tm);
            int it = 0;
            for (int w = blockIdx.x; w < p.total; w += gridDim.x, ++it) {
                const int s = it % kDwStages;
                const uint32_t ph = (it / kDwStages) & 1;
                mbar_wait(&empty[s], ph ^ 1);
                const int img = w / per_img;
                int r = w % per_img;
                const int chunk = r % p.chunks;
                r /= p.chunks;
                const int tx = r % p.tiles_x, ty = r / p.tiles_x;
                uint8_t* dst = smem + s * stage_bytes;
                mbar_arrive_expect_tx(&full[s], stage_bytes);
                tma_load_4d(dst, &p.tm, &full[s], chunk * kDwCh, tx * kDwTW - 1, ty * kDwTH - 1, img);
                if (GATE)
                    tma_load_4d(dst + kDwStreamBytes, &p.tm, &full[s], p.hp_off + chunk * kDwCh, tx * kDwTW - 1,
                                ty * kDwTH - 1, img);
            }
        }
        return;
    }

    // ---- consumers: thread = (4-channel group g, column cl and cl + 16) ----
    const int g = tid & 15, cl = tid >> 4;
    int it = 0;
    for (int w = blockIdx.x; w < p.total; w += gridDim.x, ++it) {
        const int s = it % kDwStages;
        const uint32_t ph = (it / kDwStages) & 1;
        const int img = w / per_img;
        int r = w % per_img;
        const int chunk = r % p.chunks;
        r /= p.chunks;
        const int tx = r % p.tiles_x, ty = r / p.tiles_x;
        const int c0 = chunk * kDwCh + g * 4;
        const bool cvalid = c0 < p.Cout;
        f32x2 wr[NS][9][2];
#pragma unroll
        for (int st = 0; st < NS; ++st)
#pragma unroll
            for (int t = 0; t < 9; ++t) {
                float4 w4 = make_float4(0.f, 0.f, 0.f, 0.f);
                if (cvalid) w4 = __ldg(reinterpret_cast<const float4*>(p.wt + (long long)t * p.wt_ld + c0 + st * p.hp_off));
                wr[st][t][0] = pack2(w4.x, w4.y);
                wr[st][t][1] = pack2(w4.z, w4.w);
            }
        mbar_wait(&full[s], ph);
        const uint8_t* tile = smem + s * stage_bytes;
#pragma unroll 1
        for (int half = 0; half < 2; ++half) {
            const int col = cl + 16 * half;           // output column inside the tile
            const int gx = tx * kDwTW + col;
            f32x2 win[NS][3][3][2];                   // [stream][row slot][dx][channel pair]
            auto load_row = [&](int hr, int slot) {   // hr: halo row 0..9
#pragma unroll
                for (int st = 0; st < NS; ++st)
#pragma unroll
                    for (int d = 0; d < 3; ++d) {
                        const uint2 u = *reinterpret_cast<const uint2*>(
                            tile + st * kDwStreamBytes + ((hr * kDwHW + col + d) * kDwCh + g * 4) * 2);
                        win[st][slot][d][0] = bf16x2_to_f32x2(u.x);
                        win[st][slot][d][1] = bf16x2_to_f32x2(u.y);
                    }
            };
            load_row(0, 0);
            load_row(1, 1);
#pragma unroll
            for (int ry = 0; ry < kDwTH; ++ry) {
                load_row(ry + 2, 2);
                f32x2 acc[NS][2];
#pragma unroll
                for (int st = 0; st < NS; ++st) {
                    acc[st][0] = 0ull;
                    acc[st][1] = 0ull;
#pragma unroll
                    for (int rr = 0; rr < 3; ++rr)
#pragma unroll
                        for (int d = 0; d < 3; ++d) {
                            ffma2(acc[st][0], win[st][rr][d][0], wr[st][rr * 3 + d][0]);
                            ffma2(acc[st][1], win[st][rr][d][1], wr[st][rr * 3 + d][1]);
                        }
                }
                float o[4];
                unpack2(acc[0][0], o[0], o[1]);
                unpack2(acc[0][1], o[2], o[3]);
                if (GATE) {
                    float q[4];
                    unpack2(acc[NS - 1][0], q[0], q[1]);
                    unpack2(acc[NS - 1][1], q[2], q[3]);
#pragma unroll
                    for (int i = 0; i < 4; ++i) o[i] = gelu_as(o[i]) * q[i];
                }
                const int gy = ty * kDwTH + ry;
                if (cvalid && gx < p.W && gy < p.H) {
                    uint2 pk;
                    pk.x = pack_bf16x2(o[0], o[1]);
                    pk.y = pack_bf16x2(o[2], o[3]);
                    *reinterpret_cast<uint2*>(p.y + (((long long)img * p.H + gy) * p.W + gx) * p.ldy + c0) = pk;
                }
#pragma unroll
                for (int st = 0; st < NS; ++st)
#pragma unroll
                    for (int d = 0; d < 3; ++d)
#pragma unroll
                        for (int k = 0; k < 2; ++k) {
                            win[st][0][d][k] = win[st][1][d][k];
                            win[st][1][d][k] = win[st][2][d][k];
                        }
            }
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
    }
}

static int dw_tma_launch(bool gate, const bf16* x, long long ldx, int Ctot, bf16* y, long long ldy, int Nimg, int H,
                         int W, int Cout, int hp_off, const float* wt, int wt_ld, cudaStream_t stream) {
    IRB_REQUIRE(ldx % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0, "dw3x3: x must be 16-byte aligned rows");
    IRB_REQUIRE(Cout % 4 == 0 && ldy % 4 == 0 && (reinterpret_cast<uintptr_t>(y) & 7) == 0, "dw3x3: y alignment");
    DwParams p{};
    uint64_t dims[4] = {uint64_t(Ctot), uint64_t(W), uint64_t(H), uint64_t(Nimg)};
    uint64_t str[3] = {uint64_t(ldx) * 2, uint64_t(ldx) * 2 * W, uint64_t(ldx) * 2 * W * H};
    uint32_t box[4] = {kDwCh, kDwHW, kDwHH, 1};
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
    const int grid = int(total < sms ? total : sms);
    const size_t smem = size_t(kDwStages) * (gate ? 2 : 1) * kDwStreamBytes + 128 + 64;
    static bool attr_set = false;
    if (!attr_set) {
        IRB_CHECK_CUDA(cudaFuncSetAttribute(dw3x3_tma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        IRB_CHECK_CUDA(cudaFuncSetAttribute(dw3x3_tma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_set = true;
    }
    if (gate)
        dw3x3_tma_kernel<true><<<grid, 288, smem, stream>>>(p);
    else
        dw3x3_tma_kernel<false><<<grid, 288, smem, stream>>>(p);
    IRB_CHECK_CUDA(cudaGetLastError());
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
