// Restormer GDFN (restormer_arch.py:88-93) as ONE kernel for sm_100a, bf16 storage:
//
//   x_out[p, :] = x_mid[p, :] + Wout . ( gelu(dw3x3(h1))[p, :] * dw3x3(h2)[p, :] ),   [h1 | h2] = Win' . LN(x_mid)
//
// i.e. LayerNorm (folded, statistics supplied per pixel) -> project_in (1x1, C -> 2*hp) -> depthwise 3x3 (zero pad 1)
// -> GELU gate -> project_out (1x1, hp -> C) -> residual add, plus the LayerNorm statistics of the OUTPUT rows for the
// next block. The 2*hp-channel hidden tensor (512 channels at C = 96: 2 KB per pixel written and read again by the
// unfused path) never leaves the SM: it is recomputed with a one-pixel halo per 32x4 pixel tile.
//
// Per tile (persistent CTAs, one per SM):
//   TMA      : x_mid halo tile {C, 34, 6} -> A1 (128B-swizzled K-major UMMA operand, 204 rows, 2 M-blocks of 128)
//   per hidden chunk j (32 channels of h1 + the matching 32 of h2; Win rows streamed by TMA):
//     MMA1   : D1[j % 3] (TMEM, 2 x [128 x 64] fp32)  = A1 . Win_j^T                         (tcgen05.mma, warp 21)
//     convert: D1 -> registers -> LN fold (rstd * acc + (-mu * rstd) * colsum + t), zero outside the image (= the
//              convolution's zero padding) -> bf16 -> shared halo tile [204][h1 32 | h2 32]  (warps 16..19)
//     dw+gate: 16 warps; warp = 2 output columns, lanes 0..15 convolve h1, lanes 16..31 h2 (channel pair per lane),
//              input-stationary rows with packed fma.rn.f32x2; the two half-warps swap one pixel each, so every lane
//              evaluates one gelu(h1) * h2 and writes it into A2 (128B-swizzled K-major UMMA operand)
//     MMA2   : D2 (TMEM [128 x C]) += A2_j . Wout[:, chunk j]^T                              (tcgen05.mma, warp 22)
//   epilogue : D2 -> registers, + x_mid, shifted-sum LayerNorm statistics, bf16 stores       (warps 16..19)
// D1 is triple-buffered and the hidden tile / A2 double-buffered, so TMA, both MMA streams, the converter and the
// depthwise warps all run concurrently; the FP32 pipe of the depthwise + GELU work is the bound.
#include "common.cuh"
#include "ptx.cuh"
#include "kernels.h"
#include <stdlib.h>

namespace irb {

static constexpr int kFfTW = 32, kFfTH = 4;
static constexpr int kFfHW = kFfTW + 2, kFfHH = kFfTH + 2;
static constexpr int kFfRows = kFfHW * kFfHH;   // 204 halo pixels = rows of GEMM 1
static constexpr int kFfA1Slab = 256 * 128;     // one 64-channel K slab of the halo tile (2 M-blocks)
static constexpr int kFfWinSlab = 64 * 128;     // 64 hidden rows x 64 K
static constexpr int kFfA2Bytes = 128 * 128;    // 128 pixels x 64 gated channels (two chunks)
static constexpr int kFfHidBytes = 26624;       // 204 x 128 B, rounded up to 1 KB
static constexpr int kFfDwWarps = 16, kFfCvWarps = 8;  // converter: one warpgroup per M-block of GEMM 1
static constexpr int kFfCvActive = 7;  // converter warps that own halo rows (rows 224.. of M-block 1 do not exist)
static constexpr int kFfThreads = (kFfDwWarps + kFfCvWarps + 3) * 32;  // + TMA, MMA1, MMA2 warps = 864
static constexpr int kFfD2Col = 384;            // TMEM: D1 buffers at 0 / 128 / 256 (2 x 64 columns each), D2 at 384

struct alignas(64) FfnParams {
    CUtensorMap tmX;     // x_mid {C, W, H, N}, box {64, 34, 6, 1}
    CUtensorMap tmWin;   // Win' {C, 2*hp}, box {64, 32}
    CUtensorMap tmWout;  // Wout {hp, C}, box {64, C}
    const bf16* xmid;
    long long ldm;
    bf16* xout;
    long long ldx;
    const float2* stats_in;
    float2* stats_out;
    float stats_eps;
    const float* cs;   // [2*hp] column sums of Win' (WithBias LayerNorm) or nullptr (BiasFree)
    const float* t;    // [2*hp] Win . ln_bias or nullptr
    const float* dww;  // [9][2*hp] depthwise taps
    int H, W, C, hp, KS, NCH, wout_bytes;
    int tiles_x, tiles_y, total;
};

struct FfnBars {
    uint64_t a1_full, a1_empty, win_full[2], win_empty[2], wout_full[3], wout_empty[3], d1_full[3], d1_empty[3],
        hid_full[2], hid_empty[2], a2_full[4], a2_empty[2], d2_full, d2_empty;
    uint32_t tmem_slot;
};

__global__ void __launch_bounds__(kFfThreads, 1) gdfn_fused_kernel(const __grid_constant__ FfnParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sA1 = smem;
    uint8_t* sWin = sA1 + p.KS * kFfA1Slab;
    uint8_t* sA2 = sWin + 2 * p.KS * kFfWinSlab;
    uint8_t* sWout = sA2 + 2 * kFfA2Bytes;
    uint8_t* sHid = sWout + 3 * p.wout_bytes;
    float* colvec = reinterpret_cast<float*>(sHid + 2 * kFfHidBytes);  // [cs | t], 2*hp floats each
    FfnBars* bars = reinterpret_cast<FfnBars*>(colvec + 4 * p.hp);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int per_img = p.tiles_x * p.tiles_y;
    const int NCH = p.NCH, KS = p.KS;

    if (threadIdx.x == 0) {
        mbar_init(&bars->a1_full, 1);
        mbar_init(&bars->a1_empty, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&bars->win_full[i], 1);
            mbar_init(&bars->win_empty[i], 1);
            mbar_init(&bars->hid_full[i], kFfCvActive);
            mbar_init(&bars->hid_empty[i], kFfDwWarps);
            mbar_init(&bars->a2_empty[i], 1);
        }
        for (int i = 0; i < 3; ++i) {
            mbar_init(&bars->wout_full[i], 1);
            mbar_init(&bars->wout_empty[i], 1);
            mbar_init(&bars->d1_full[i], 1);
            mbar_init(&bars->d1_empty[i], kFfCvActive);
        }
        for (int i = 0; i < 4; ++i) mbar_init(&bars->a2_full[i], kFfDwWarps);
        mbar_init(&bars->d2_full, 1);
        mbar_init(&bars->d2_empty, 4);
        fence_mbar_init();
    }
    if (warp == kFfDwWarps + kFfCvWarps) {
        tmem_alloc(&bars->tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_slot;

    if (warp < kFfDwWarps) {
        // ================= depthwise 3x3 + GELU gate =================
        const int col0 = warp * 2, stream = lane >> 4, cp = lane & 15;
        const int in_chunk = stream * 4 + (cp >> 2);
        const uint32_t sub = (cp & 3) * 4;
        const uint8_t* hid_base = sHid + col0 * 128 + sub;  // plain shared-memory pointers: LDS / STS with immediates
        // output word of this lane inside the A2 row of pixel (col0 + stream): 16-byte chunk (cp >> 2) of the chunk's
        // 64-byte half, XOR-swizzled by the row index (rows of the tile are 32 apart, so only col0 + stream matters)
        const int opx = col0 + stream;
        const uint32_t out_off = opx * 128 + ((((cp >> 2) ^ (opx & 7)) << 4) | sub);
        int g = 0;
        for (int tile = blockIdx.x; tile < p.total; tile += gridDim.x) {
            for (int j = 0; j < NCH; ++j, ++g) {
                const int hb = g & 1;
                const int ch = stream * p.hp + j * 32 + cp * 2;
                f32x2 wr[9];
#pragma unroll
                for (int t = 0; t < 9; ++t) {
                    const float2 w2 = __ldg(reinterpret_cast<const float2*>(p.dww + (long long)t * 2 * p.hp + ch));
                    wr[t] = pack2(w2.x, w2.y);
                }
                // halo pixel (hr, col0 + d) is row hr * 34 + col0 + d of the hidden tile; its swizzle key is
                // (2 * hr + d + col0) & 7, so 8 per-lane bases cover all 24 loads and the rest is an immediate offset
                const uint8_t* hk[8];
#pragma unroll
                for (int k = 0; k < 8; ++k)
                    hk[k] = hid_base + hb * kFfHidBytes + ((in_chunk ^ ((col0 + k) & 7)) << 4);
                mbar_wait(&bars->hid_full[hb], (g >> 1) & 1);
                if (!(j & 1)) mbar_wait(&bars->a2_empty[(g >> 1) & 1], ((g >> 2) & 1) ^ 1);
                uint8_t* a2 = sA2 + ((g >> 1) & 1) * kFfA2Bytes + (out_off ^ ((g & 1) << 6));
                f32x2 acc[3][2];
#pragma unroll
                for (int hr = 0; hr < kFfHH; ++hr) {
                    f32x2 in[4];
#pragma unroll
                    for (int d = 0; d < 4; ++d)
                        in[d] = bf16x2_to_f32x2_alu(*reinterpret_cast<const uint32_t*>(hk[(2 * hr + d) & 7] + (hr * kFfHW + d) * 128));
#pragma unroll
                    for (int dr = 0; dr < 3; ++dr) {
                        const int orow = hr - dr;
                        if (orow < 0 || orow >= kFfTH) continue;
                        const int sl = orow % 3;
#pragma unroll
                        for (int px = 0; px < 2; ++px) {
                            if (dr == 0)
                                acc[sl][px] = mul2(in[px], wr[0]);
                            else
                                ffma2(acc[sl][px], in[px], wr[dr * 3]);
                            ffma2(acc[sl][px], in[px + 1], wr[dr * 3 + 1]);
                            ffma2(acc[sl][px], in[px + 2], wr[dr * 3 + 2]);
                        }
                    }
                    const int done = hr - 2;
                    if (done >= 0) {
                        const int sl = done % 3;
                        // lanes 0..15 hold dw(h1) of pixels (col0, col0 + 1), lanes 16..31 dw(h2) of the same pixels:
                        // lane < 16 finishes pixel col0, its partner pixel col0 + 1
                        const f32x2 keep = stream ? acc[sl][1] : acc[sl][0];
                        const f32x2 send = stream ? acc[sl][0] : acc[sl][1];
                        const f32x2 recv = __shfl_xor_sync(0xffffffffu, send, 16);
                        const f32x2 h1 = stream ? recv : keep, h2 = stream ? keep : recv;
                        *reinterpret_cast<uint32_t*>(a2 + done * kFfTW * 128) = f32x2_to_bf16x2(mul2(gelu2_fast(h1), h2));
                    }
                }
                fence_proxy_async();  // A2 is read by the tensor core (async proxy)
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(&bars->a2_full[g & 3]);
                    mbar_arrive(&bars->hid_empty[hb]);
                }
            }
        }
    } else if (warp < kFfDwWarps + kFfCvWarps) {
        // ================= converter (TMEM D1 -> hidden halo tile) + epilogue (D2 -> x_out) =================
        // warpgroup mb converts M-block mb (halo rows mb * 128 ..); warpgroup 1 has only 76 rows and also runs the
        // output epilogue
        const int q = warp & 3, mb = (warp - kFfDwWarps) >> 2;
        const int ct = threadIdx.x - kFfDwWarps * 32;
        float* s_cs = colvec;
        float* s_t = colvec + 2 * p.hp;
        for (int i = ct; i < 2 * p.hp; i += kFfCvWarps * 32) {
            s_cs[i] = p.cs ? __ldg(p.cs + i) : 0.f;
            s_t[i] = p.t ? __ldg(p.t + i) : 0.f;
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
        const uint32_t tlane = tmem_base + (uint32_t(q * 32) << 16);
        const int r = mb * 128 + q * 32 + lane;      // halo row of this thread
        const bool warp_rows = mb * 128 + q * 32 < kFfRows;  // warp-uniform: this warp has rows to convert
        const int hy = r / kFfHW, hx = r - hy * kFfHW;

        // 16 output channels of this thread's pixel: + residual, shifted-sum statistics, bf16 store
        auto epi16 = [&](const uint32_t* acc, const uint4& r0, const uint4& r1, f32x2 nshift2, f32x2& s1, f32x2& s2,
                         bf16* dst, bool valid) {
            const uint32_t rw[8] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
            uint32_t pk[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                f32x2 y = pack2(__uint_as_float(acc[2 * i]), __uint_as_float(acc[2 * i + 1]));
                y = add2(y, bf16x2_to_f32x2(rw[i]));
                const f32x2 d = add2(y, nshift2);
                s1 = add2(s1, d);
                s2 = fma2(d, d, s2);
                pk[i] = f32x2_to_bf16x2(y);
            }
            if (valid) {
                *reinterpret_cast<uint4*>(dst) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                *reinterpret_cast<uint4*>(dst + 8) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
            }
        };
        // The residual row (x_mid, an L2 hit ~700 cycles away) is fetched 32 channels at a time, one block ahead:
        // the first block is requested before waiting for the accumulator, the next one before the current is used.
        auto epilogue = [&](int etile, int eti) {
            const int eimg = etile / per_img;
            const int er = etile - eimg * per_img;
            const int ety = er / p.tiles_x, etx = er - ety * p.tiles_x;
            const int gy = ety * kFfTH + q, gx = etx * kFfTW + lane;
            const bool valid = gy < p.H && gx < p.W;
            const long long pix = valid ? ((long long)eimg * p.H + gy) * p.W + gx : 0;
            const bf16* rrow = p.xmid + pix * p.ldm;
            bf16* orow = p.xout + pix * p.ldx;
            uint4 rb[3][2];  // 16-channel groups of the residual row, two groups ahead of their use
            rb[0][0] = __ldg(reinterpret_cast<const uint4*>(rrow));
            rb[0][1] = __ldg(reinterpret_cast<const uint4*>(rrow + 8));
            if (16 < p.C) {
                rb[1][0] = __ldg(reinterpret_cast<const uint4*>(rrow + 16));
                rb[1][1] = __ldg(reinterpret_cast<const uint4*>(rrow + 24));
            }
            mbar_wait(&bars->d2_full, eti & 1);
            tc_fence_after();
            f32x2 s1 = 0ull, s2 = 0ull;
            const float shift = __uint_as_float(rb[0][0].x << 16);
            const f32x2 nshift2 = splat2(-shift);
#pragma unroll
            for (int gi = 0; gi < 6; ++gi) {
                const int cb = gi * 16;
                if (cb < p.C) {
                    if (cb + 32 < p.C) {
                        rb[(gi + 2) % 3][0] = __ldg(reinterpret_cast<const uint4*>(rrow + cb + 32));
                        rb[(gi + 2) % 3][1] = __ldg(reinterpret_cast<const uint4*>(rrow + cb + 40));
                    }
                    uint32_t acc[16];
                    tmem_ld16_nowait(tlane + kFfD2Col + cb, acc);
                    tmem_ld_wait();
                    epi16(acc, rb[gi % 3][0], rb[gi % 3][1], nshift2, s1, s2, orow + cb, valid);
                }
            }
            if (valid && p.stats_out) {
                float a0, a1, q0, q1;
                unpack2(s1, a0, a1);
                unpack2(s2, q0, q1);
                const float invn = 1.0f / float(p.C);
                const float m = (a0 + a1) * invn;
                const float var = fmaxf((q0 + q1) * invn - m * m, 0.f);
                p.stats_out[pix] = make_float2(shift + m, rsqrtf(var + p.stats_eps));
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars->d2_empty);
        };

        // (mu, rstd) of this thread's halo pixel; fetched one tile ahead so that the L2 latency is off the critical path
        auto tile_stats = [&](int tile, bool& in) -> float2 {
            const int img = tile / per_img;
            const int rr = tile - img * per_img;
            const int ty = rr / p.tiles_x, tx = rr - ty * p.tiles_x;
            const int gy = ty * kFfTH - 1 + hy, gx = tx * kFfTW - 1 + hx;
            in = r < kFfRows && gy >= 0 && gy < p.H && gx >= 0 && gx < p.W;
            return in ? __ldg(p.stats_in + ((long long)img * p.H + gy) * p.W + gx) : make_float2(0.f, 0.f);
        };
        int g = 0, ti = 0, prev_tile = -1;
        bool in_next = false;
        float2 st_next = make_float2(0.f, 0.f);
        if (blockIdx.x < p.total) st_next = tile_stats(blockIdx.x, in_next);
        for (int tile = blockIdx.x; tile < p.total; tile += gridDim.x, ++ti) {
            const bool inimg = in_next;
            const float fa = inimg ? st_next.y : 0.f;
            const float fb = (inimg && p.cs) ? -st_next.x * st_next.y : 0.f;
            if (tile + gridDim.x < p.total) st_next = tile_stats(tile + gridDim.x, in_next);
            const f32x2 a2 = splat2(fa), b2 = splat2(fb);
            for (int j = 0; j < NCH; ++j, ++g) {
                const int d = g % 3, hb = g & 1;
                if (warp_rows) {
                    mbar_wait(&bars->d1_full[d], (g / 3) & 1);
                    tc_fence_after();
                    mbar_wait(&bars->hid_empty[hb], ((g >> 1) & 1) ^ 1);
                    uint8_t* hrow = sHid + hb * kFfHidBytes + r * 128;
#pragma unroll
                    for (int half = 0; half < 2; ++half) {
                        uint32_t acc[32];
                        tmem_ld32_nowait(tlane + d * 128 + mb * 64 + half * 32, acc);
                        tmem_ld_wait();
                        const float* cs = s_cs + half * p.hp + j * 32;
                        const float* tt = s_t + half * p.hp + j * 32;
                        if (r < kFfRows) {
                            if (inimg) {
#pragma unroll
                                for (int i = 0; i < 4; ++i) {  // 16-byte chunk i of this half: 8 channels
                                    uint32_t pk[4];
#pragma unroll
                                    for (int h4 = 0; h4 < 2; ++h4) {
                                        const int c = i * 8 + h4 * 4;
                                        const ulonglong2 cc = *reinterpret_cast<const ulonglong2*>(cs + c);
                                        const ulonglong2 tv = *reinterpret_cast<const ulonglong2*>(tt + c);
                                        const f32x2 x0 = pack2(__uint_as_float(acc[c]), __uint_as_float(acc[c + 1]));
                                        const f32x2 x1 = pack2(__uint_as_float(acc[c + 2]), __uint_as_float(acc[c + 3]));
                                        pk[h4 * 2] = f32x2_to_bf16x2(fma2(x0, a2, fma2(b2, cc.x, tv.x)));
                                        pk[h4 * 2 + 1] = f32x2_to_bf16x2(fma2(x1, a2, fma2(b2, cc.y, tv.y)));
                                    }
                                    *reinterpret_cast<uint4*>(hrow + (((half * 4 + i) ^ (r & 7)) << 4)) =
                                        make_uint4(pk[0], pk[1], pk[2], pk[3]);
                                }
                            } else {
#pragma unroll
                                for (int i = 0; i < 4; ++i)
                                    *reinterpret_cast<uint4*>(hrow + (((half * 4 + i) ^ (r & 7)) << 4)) =
                                        make_uint4(0u, 0u, 0u, 0u);
                            }
                        }
                    }
                    tc_fence_before();
                }
                __syncwarp();
                if (warp_rows && lane == 0) {
                    mbar_arrive(&bars->d1_empty[d]);
                    mbar_arrive(&bars->hid_full[hb]);
                }
                if (mb == 1 && j == 1 && ti > 0) epilogue(prev_tile, ti - 1);
            }
            prev_tile = tile;
        }
        if (mb == 1 && ti > 0) epilogue(prev_tile, ti - 1);
    } else if (warp == kFfDwWarps + kFfCvWarps) {
        // ================= TMA producer =================
        if (elect_one()) {
            tma_prefetch_desc(&p.tmX);
            tma_prefetch_desc(&p.tmWin);
            tma_prefetch_desc(&p.tmWout);
            int g = 0, pr = 0, ti = 0;
            for (int tile = blockIdx.x; tile < p.total; tile += gridDim.x, ++ti) {
                const int img = tile / per_img;
                const int rr = tile - img * per_img;
                const int ty = rr / p.tiles_x, tx = rr - ty * p.tiles_x;
                mbar_wait(&bars->a1_empty, (ti & 1) ^ 1);
                mbar_arrive_expect_tx(&bars->a1_full, KS * kFfRows * 128);
                for (int ks = 0; ks < KS; ++ks)
                    tma_load_4d(sA1 + ks * kFfA1Slab, &p.tmX, &bars->a1_full, ks * 64, tx * kFfTW - 1, ty * kFfTH - 1, img);
                for (int j = 0; j < NCH; ++j, ++g) {
                    const int b = g & 1;
                    mbar_wait(&bars->win_empty[b], ((g >> 1) & 1) ^ 1);
                    mbar_arrive_expect_tx(&bars->win_full[b], KS * kFfWinSlab);
                    for (int ks = 0; ks < KS; ++ks) {
                        uint8_t* dst = sWin + (b * KS + ks) * kFfWinSlab;
                        tma_load_2d(dst, &p.tmWin, &bars->win_full[b], ks * 64, j * 32);
                        tma_load_2d(dst + kFfWinSlab / 2, &p.tmWin, &bars->win_full[b], ks * 64, p.hp + j * 32);
                    }
                    if (!(j & 1)) {
                        const int wb = pr % 3;
                        mbar_wait(&bars->wout_empty[wb], ((pr / 3) & 1) ^ 1);
                        mbar_arrive_expect_tx(&bars->wout_full[wb], p.C * 128);
                        tma_load_2d(sWout + wb * p.wout_bytes, &p.tmWout, &bars->wout_full[wb], (j >> 1) * 64, 0);
                        ++pr;
                    }
                }
            }
        }
    } else if (warp == kFfDwWarps + kFfCvWarps + 1) {
        // ================= MMA 1: project_in on the halo tile =================
        if (elect_one()) {
            const uint32_t idesc = make_idesc_bf16(128, 64, 0, 0);
            const uint32_t a1 = smem_u32(sA1), win = smem_u32(sWin);
            int g = 0, ti = 0;
            for (int tile = blockIdx.x; tile < p.total; tile += gridDim.x, ++ti) {
                mbar_wait(&bars->a1_full, ti & 1);
                for (int j = 0; j < NCH; ++j, ++g) {
                    const int b = g & 1, d = g % 3;
                    mbar_wait(&bars->win_full[b], (g >> 1) & 1);
                    mbar_wait(&bars->d1_empty[d], ((g / 3) & 1) ^ 1);
                    tc_fence_after();
                    for (int mb = 0; mb < 2; ++mb)
                        for (int ks = 0; ks < KS; ++ks)
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                const uint64_t ad = make_sdesc_sw128(a1 + ks * kFfA1Slab + mb * 16384 + k * 32, 16, 1024);
                                const uint64_t bd = make_sdesc_sw128(win + (b * KS + ks) * kFfWinSlab + k * 32, 16, 1024);
                                umma_bf16(tmem_base + d * 128 + mb * 64, ad, bd, idesc, (ks | k) != 0 ? 1u : 0u);
                            }
                    umma_commit(&bars->win_empty[b]);
                    umma_commit(&bars->d1_full[d]);
                    if (j == NCH - 1) umma_commit(&bars->a1_empty);
                }
            }
        }
    } else {
        // ================= MMA 2: project_out of the gated tile =================
        if (elect_one()) {
            const uint32_t idesc = make_idesc_bf16(128, p.C, 0, 0);
            const uint32_t a2 = smem_u32(sA2), wout = smem_u32(sWout);
            int g = 0, pr = 0, ti = 0;
            for (int tile = blockIdx.x; tile < p.total; tile += gridDim.x, ++ti) {
                for (int j = 0; j < NCH; ++j, ++g) {
                    const int wb = pr % 3, ab = (g >> 1) & 1;
                    if (j == 0) mbar_wait(&bars->d2_empty, (ti & 1) ^ 1);
                    if (!(j & 1)) mbar_wait(&bars->wout_full[wb], (pr / 3) & 1);
                    mbar_wait(&bars->a2_full[g & 3], (g >> 2) & 1);
                    tc_fence_after();
#pragma unroll
                    for (int k2 = 0; k2 < 2; ++k2) {
                        const int k = 2 * (g & 1) + k2;
                        const uint64_t ad = make_sdesc_sw128(a2 + ab * kFfA2Bytes + k * 32, 16, 1024);
                        const uint64_t bd = make_sdesc_sw128(wout + wb * p.wout_bytes + k * 32, 16, 1024);
                        umma_bf16(tmem_base + kFfD2Col, ad, bd, idesc, (j | k2) != 0 ? 1u : 0u);
                    }
                    if (j & 1) {
                        umma_commit(&bars->a2_empty[ab]);
                        umma_commit(&bars->wout_empty[wb]);
                        ++pr;
                    }
                    if (j == NCH - 1) umma_commit(&bars->d2_full);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kFfDwWarps + kFfCvWarps) tmem_dealloc(tmem_base, 512);
}

bool gdfn_fused_shape_ok(int C, int hp) {
    return C >= 16 && C <= 96 && C % 16 == 0 && hp >= 64 && hp <= 256 && hp % 64 == 0;
}

bool gdfn_fused_supported(int C, int hp) {
    static int enabled = -1;
    if (enabled < 0) {
        enabled = IRB_DIAG_ENV("IRB_FUSED_FFN", 0);  // opt-in: at parity with the three-kernel path, see DESIGN.md
    }
    return enabled && gdfn_fused_shape_ok(C, hp);
}

int gdfn_fused_bf16(const bf16* xmid, long long ldm, const float2* stats_in, bf16* xout, long long ldx,
                    float2* stats_out, float stats_eps, int Nimg, int H, int W, int C, int hp, const bf16* win,
                    const float* cs, const float* t, const float* dww, const bf16* wout, cudaStream_t stream) {
    IRB_REQUIRE(C >= 16 && C <= 96 && C % 16 == 0 && hp >= 64 && hp <= 256 && hp % 64 == 0,
                "gdfn_fused: needs C in {16..96} (multiple of 16) and hp in {64,128,192,256}");
    IRB_REQUIRE(ldm % 8 == 0 && ldx % 8 == 0 && (reinterpret_cast<uintptr_t>(xmid) & 15) == 0 &&
                    (reinterpret_cast<uintptr_t>(xout) & 15) == 0,
                "gdfn_fused: rows must be 16-byte aligned");
    IRB_REQUIRE((reinterpret_cast<uintptr_t>(win) & 15) == 0 && (reinterpret_cast<uintptr_t>(wout) & 15) == 0,
                "gdfn_fused: weights must be 16-byte aligned");
    IRB_REQUIRE(xmid != xout && stats_in != stats_out, "gdfn_fused: in-place operation is not supported (halo reads)");
    IRB_REQUIRE(stats_in != nullptr, "gdfn_fused: LayerNorm statistics required");
    FfnParams p{};
    {
        uint64_t dims[4] = {uint64_t(C), uint64_t(W), uint64_t(H), uint64_t(Nimg)};
        uint64_t str[3] = {uint64_t(ldm) * 2, uint64_t(ldm) * 2 * W, uint64_t(ldm) * 2 * W * H};
        uint32_t box[4] = {64, kFfHW, kFfHH, 1};
        IRB_TRY(make_tmap_bf16(&p.tmX, xmid, 4, dims, str, box));
    }
    {
        uint64_t dims[2] = {uint64_t(C), uint64_t(2 * hp)};
        uint64_t str[1] = {uint64_t(C) * 2};
        uint32_t box[2] = {64, 32};
        IRB_TRY(make_tmap_bf16(&p.tmWin, win, 2, dims, str, box));
    }
    {
        uint64_t dims[2] = {uint64_t(hp), uint64_t(C)};
        uint64_t str[1] = {uint64_t(hp) * 2};
        uint32_t box[2] = {64, uint32_t(C)};
        IRB_TRY(make_tmap_bf16(&p.tmWout, wout, 2, dims, str, box));
    }
    p.xmid = xmid;
    p.ldm = ldm;
    p.xout = xout;
    p.ldx = ldx;
    p.stats_in = stats_in;
    p.stats_out = stats_out;
    p.stats_eps = stats_eps;
    p.cs = cs;
    p.t = t;
    p.dww = dww;
    p.H = H;
    p.W = W;
    p.C = C;
    p.hp = hp;
    p.KS = ceil_div(C, 64);
    p.NCH = hp / 32;
    p.wout_bytes = (C * 128 + 1023) / 1024 * 1024;
    p.tiles_x = ceil_div(W, kFfTW);
    p.tiles_y = ceil_div(H, kFfTH);
    const long long total = (long long)p.tiles_x * p.tiles_y * Nimg;
    IRB_REQUIRE(total < (1ll << 31) / 16, "gdfn_fused: too many tiles");
    p.total = int(total);
    const size_t smem = 1024 + size_t(p.KS) * kFfA1Slab + size_t(2) * p.KS * kFfWinSlab + 2 * kFfA2Bytes +
                        3 * size_t(p.wout_bytes) + 2 * kFfHidBytes + size_t(4) * hp * 4 + sizeof(FfnBars) + 64;
    IRB_REQUIRE(smem <= 227 * 1024, "gdfn_fused: shared memory budget exceeded");
    IRB_SMEM_OPTIN((gdfn_fused_kernel), 227 * 1024);
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = p.total < sms ? p.total : sms;
    gdfn_fused_kernel<<<grid, kFfThreads, smem, stream>>>(p);
    IRB_CHECK_CUDA(cudaGetLastError());
    const double px = double(Nimg) * H * W;
    prof_mark("gdfn_fused", C, hp, px * (2.0 * C * 2 * hp + 18.0 * 2 * hp + 2.0 * hp * C),
              px * (2.0 * C * 2 + 16.0) + 2.0 * hp * C * 2 + double(C) * hp * 2 + 72.0 * hp, stream);
    return 0;
}

}  // namespace irb
