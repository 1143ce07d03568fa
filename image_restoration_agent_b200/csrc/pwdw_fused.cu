// Restormer's two "1x1 convolution -> depthwise 3x3" pairs as ONE kernel each, bf16 storage, sm_100a:
//
//   plain : qkv[p, n]   = dw3x3( W1' . LN(x) )[p, n]                       MDTA qkv -> qkv_dwconv  (restormer_arch.py:115-116)
//   gate  : gated[p, c] = gelu(dw3x3(h)[p, c]) * dw3x3(h)[p, hp + c],  h = W1' . LN(x)   GDFN project_in -> dwconv -> gate (:88-93)
//
// The 3C- / 2hp-channel tensor between the 1x1 and the depthwise convolution (`qkv_pre`: 576 B per pixel, `hidden`:
// 1 KB per pixel at C = 96) never goes to HBM: it is recomputed per 32x8 pixel tile on a one-pixel halo (34x10 = 340
// pixels, 1.33x the tile: the tensor pipe idles at ~13 % in the unfused model, so the recompute is free) and handed to
// the depthwise warps through shared memory.
//
// Work item = (image, 32x8 tile); persistent CTAs, one per SM, 18 warps:
//   warp 16 (TMA)      : x halo tile {C, 34, 10} -> A (K-major, 128B swizzle, 3 M-blocks of 128 rows; rows / columns
//                        outside the image are zero-filled) once per item; per 64-column chunk j of W1' -> B ring (2)
//   warp 17 (MMA)      : D[j & 1] (TMEM, 3 x [128 x 64] fp32) = A . B_j^T                         tcgen05.mma
//   warps 0-15, per chunk, two phases (conversion runs one chunk ahead of the stencil; hand-over by mbarriers):
//     convert          : warp w drains TMEM lane quarter w & 3, columns 16 (w >> 2) .. +15 of the three M-blocks:
//                        D -> registers -> folded LayerNorm (rstd * acc + (-mu rstd) * colsum + t), ZERO outside the
//                        image (the depthwise convolution's zero padding applies to the 1x1's OUTPUT) -> bf16 ->
//                        hidden halo tile [340 px][64 ch] in shared memory (2 buffers, 16-byte chunks XOR-swizzled by
//                        the row so that these row-per-thread stores and the pixel-per-warp loads are conflict-free)
//     stencil          : the depthwise code of dw_tma.cu on that tile: warp = 2 output columns, lane = channel pair,
//                        input-stationary rows, packed fma.rn.f32x2; plain: bf16 stores; gate: lanes 0-15 hold the
//                        gelu half, lanes 16-31 the linear half of the SAME 32 gated channels, the half-warps swap one
//                        pixel each and every lane stores one gelu(h1) * h2 pair
//   (a first version with 4 dedicated converter warps left the 16 stencil warps waiting on them 46 % of the time,
//   profiles/r02_s02_ncu_pwdw_v1.txt; here every warp does both, the TMEM double buffer keeps the MMA a chunk ahead).
#include "common.cuh"
#include "ptx.cuh"
#include "kernels.h"

namespace irb {

static constexpr int kPwTW = 32, kPwTH = 8;
static constexpr int kPwHW = kPwTW + 2, kPwHH = kPwTH + 2;
static constexpr int kPwRows = kPwHW * kPwHH;        // 340 halo pixels = rows of the 1x1 GEMM
static constexpr int kPwASlab = 384 * 128;           // one 64-channel K slab of the halo tile (3 M-blocks)
static constexpr int kPwBSlab = 64 * 128;            // 64 output columns x 64 K
static constexpr int kPwHid = kPwRows * 128;         // 43 520 B: hidden halo tile, 64 channels
static constexpr int kPwDwWarps = 16, kPwCvWarps = 0;
static constexpr int kPwThreads = (kPwDwWarps + kPwCvWarps + 2) * 32;  // 576
static constexpr int kPwAccStride = 192;             // TMEM columns per accumulator buffer (3 M-blocks x 64)

struct alignas(64) PwDwParams {
    CUtensorMap tmX;  // x {C, W, H, N}, box {64, 34, 10, 1}, 128B swizzle
    CUtensorMap tmW;  // W1' {C, rows}, box {64, 64} (plain) / {64, 32} (gate), 128B swizzle
    const float2* stats;  // [N*H*W] (mu, rstd) of the rows of x
    const float* cs;      // [rows] column sums of the folded weights (WithBias LayerNorm) or nullptr
    const float* tb;      // [rows] W1 . ln_bias or nullptr
    const float* dww;     // [rows / 2][9][2] depthwise taps, fp32: the 9 taps of a channel pair are contiguous
    bf16* y;
    long long ldy;
    int H, W, KS, klast, nchunks, nout, nrows, hp;
    int tiles_x, tiles_y, total;
};

struct PwBars {
    uint64_t a_full, a_empty, b_full[2], b_empty[2], d_full[2], d_empty[2], h_full[2], h_empty[2];
    uint32_t tmem_slot;
};

template <int GATE>  // 0: plain depthwise; 1: GELU gate (sigmoid-form erf GELU of dw_tma.cu)
__global__ void __launch_bounds__(kPwThreads, 1) pwdw_fused_kernel(const __grid_constant__ PwDwParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* sA = smem;
    uint8_t* sB = sA + p.KS * kPwASlab;
    uint8_t* sH = sB + 2 * p.KS * kPwBSlab;
    PwBars* bars = reinterpret_cast<PwBars*>(sH + 2 * kPwHid);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int per_img = p.tiles_x * p.tiles_y;
    const int KS = p.KS, NCH = p.nchunks;

    if (threadIdx.x == 0) {
        mbar_init(&bars->a_full, 1);
        mbar_init(&bars->a_empty, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(&bars->b_full[i], 1);
            mbar_init(&bars->b_empty[i], 1);
            mbar_init(&bars->d_full[i], 1);
            mbar_init(&bars->d_empty[i], kPwDwWarps);
            mbar_init(&bars->h_full[i], kPwDwWarps);
            mbar_init(&bars->h_empty[i], kPwDwWarps);
        }
        fence_mbar_init();
    }
    if (warp == kPwDwWarps + kPwCvWarps + 1) {
        tmem_alloc(&bars->tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_slot;
    griddep_wait();  // PDL: everything above overlapped the previous kernel's tail
    griddep_launch();

    if (warp == kPwDwWarps + kPwCvWarps) {
        // ================= TMA producer =================
        if (elect_one()) {
            tma_prefetch_desc(&p.tmX);
            tma_prefetch_desc(&p.tmW);
            int it = 0, cj = 0;
            for (int item = blockIdx.x; item < p.total; item += gridDim.x, ++it) {
                const int img = item / per_img;
                const int r = item - img * per_img;
                const int ty = r / p.tiles_x, tx = r - ty * p.tiles_x;
                mbar_wait(&bars->a_empty, (it & 1) ^ 1);
                mbar_arrive_expect_tx(&bars->a_full, uint32_t(KS) * kPwHid);
                for (int ks = 0; ks < KS; ++ks)
                    tma_load_4d(sA + ks * kPwASlab, &p.tmX, &bars->a_full, ks * 64, tx * kPwTW - 1, ty * kPwTH - 1, img);
                for (int j = 0; j < NCH; ++j, ++cj) {
                    const int s = cj & 1;
                    mbar_wait(&bars->b_empty[s], ((cj >> 1) & 1) ^ 1);
                    mbar_arrive_expect_tx(&bars->b_full[s], uint32_t(KS) * kPwBSlab);
                    uint8_t* dst = sB + s * KS * kPwBSlab;
                    for (int ks = 0; ks < KS; ++ks) {
                        if (GATE == 0) {
                            tma_load_2d(dst + ks * kPwBSlab, &p.tmW, &bars->b_full[s], ks * 64, j * 64);
                        } else {  // columns 0..31 = h1 rows j*32.., columns 32..63 = the matching h2 rows
                            tma_load_2d(dst + ks * kPwBSlab, &p.tmW, &bars->b_full[s], ks * 64, j * 32);
                            tma_load_2d(dst + ks * kPwBSlab + 32 * 128, &p.tmW, &bars->b_full[s], ks * 64, p.hp + j * 32);
                        }
                    }
                }
            }
        }
    } else if (warp == kPwDwWarps + kPwCvWarps + 1) {
        // ================= MMA issuer =================
        if (elect_one()) {
            const uint32_t idesc = make_idesc_bf16(128, 64, 0, 0);
            const uint32_t a_base = smem_u32(sA), b_base = smem_u32(sB);
            int it = 0, cj = 0;
            for (int item = blockIdx.x; item < p.total; item += gridDim.x, ++it) {
                mbar_wait(&bars->a_full, it & 1);
                tc_fence_after();
                for (int j = 0; j < NCH; ++j, ++cj) {
                    const int s = cj & 1;
                    const uint32_t ph = (cj >> 1) & 1;
                    mbar_wait(&bars->b_full[s], ph);
                    mbar_wait(&bars->d_empty[s], ph ^ 1);
                    tc_fence_after();
#pragma unroll
                    for (int mb = 0; mb < 3; ++mb) {
                        const uint32_t tacc = tmem_base + s * kPwAccStride + mb * 64;
                        for (int ks = 0; ks < KS; ++ks) {
                            const int nk = ks == KS - 1 ? p.klast : 4;
                            const uint64_t ad = make_sdesc_sw128(a_base + ks * kPwASlab + mb * 16384, 16, 1024);
                            const uint64_t bd = make_sdesc_sw128(b_base + (s * KS + ks) * kPwBSlab, 16, 1024);
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                if (k < nk) umma_bf16(tacc, ad + uint64_t(k * 2), bd + uint64_t(k * 2), idesc, (ks | k) != 0 ? 1u : 0u);
                        }
                    }
                    umma_commit(&bars->b_empty[s]);
                    umma_commit(&bars->d_full[s]);
                }
                umma_commit(&bars->a_empty);  // every MMA of this item has read the halo tile
            }
        }
    } else {
        // ================= warps 0..15: convert chunk t (TMEM -> hidden tile), then the stencil of chunk t - 1 =================
        // A software pipeline over this CTA's flat chunk sequence t = 0, 1, ...: conversion runs ONE chunk ahead of the
        // stencil, and the two hidden buffers are handed over with mbarriers (h_full: all 16 warps converted their part;
        // h_empty: all 16 warps finished the stencil), so warps drift up to a chunk apart: while some wait for TMEM loads
        // others issue stencil math (a CTA-wide barrier per chunk left the SM at IPC 1.3).
        const int q = warp & 3;    // TMEM lane quarter this warp may read
        const int cg = warp >> 2;  // its 16-column group of the chunk
        const int col0 = warp * 2; // its two output columns in the stencil phase
        const uint32_t hid0 = smem_u32(sH);
        const long long row_step = (long long)p.W * p.ldy;
        const int my_items = blockIdx.x < p.total ? (p.total - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
        const int T = my_items * NCH;
        // conversion stage state
        int itemC = blockIdx.x, jC = 0;
        float ar[3], br[3];
        uint32_t okm[3];
        // stencil stage state
        int itemS = blockIdx.x, jS = 0;
        int imgS = 0, tyS = 0, txS = 0;
        for (int t = 0; t <= T; ++t) {
            if (t < T) {
                // ---------------- convert chunk t: columns cg*16 .. +15, rows q*32 + lane of every M-block ----------------
                const int s = t & 1;
                const uint32_t ph = (t >> 1) & 1;
                if (jC == 0) {  // new item: LayerNorm statistics of this thread's three halo rows; mask = 0 outside the image
                    const int img = itemC / per_img;
                    const int rr = itemC - img * per_img;
                    const int ty = rr / p.tiles_x, tx = rr - ty * p.tiles_x;
#pragma unroll
                    for (int mb = 0; mb < 3; ++mb) {
                        const int r = mb * 128 + q * 32 + lane;
                        const int hy = r / kPwHW, hx = r - hy * kPwHW;
                        const int y = ty * kPwTH - 1 + hy, x = tx * kPwTW - 1 + hx;
                        const bool ok = r < kPwRows && y >= 0 && y < p.H && x >= 0 && x < p.W;
                        float2 st = make_float2(0.f, 0.f);
                        if (ok) st = __ldg(p.stats + ((long long)img * p.H + y) * p.W + x);
                        ar[mb] = st.y;
                        br[mb] = -st.x * st.y;
                        okm[mb] = ok ? 0xFFFFFFFFu : 0u;
                    }
                }
                const int j = jC;
                int off;      // index of the chunk's column cg*16 in the colsum / t vectors
                int nv = 16;  // how many of the 16 columns exist (plain: the last chunk may be ragged)
                if (GATE == 0) {
                    off = j * 64 + cg * 16;
                    nv = min(16, max(0, p.nrows - off));
                } else {
                    off = (cg < 2 ? j * 32 + cg * 16 : p.hp + j * 32 + (cg - 2) * 16);
                }
                f32x2 csv[8], tbv[8];
                if (p.cs != nullptr) {
                    const float4* cp = reinterpret_cast<const float4*>(p.cs + off);
                    const float4* tp = reinterpret_cast<const float4*>(p.tb + off);
#pragma unroll
                    for (int g4 = 0; g4 < 4; ++g4) {
                        float4 c4 = make_float4(0.f, 0.f, 0.f, 0.f), t4 = c4;
                        if (g4 * 4 < nv) {
                            c4 = __ldg(cp + g4);
                            t4 = __ldg(tp + g4);
                        }
                        csv[2 * g4] = pack2(c4.x, c4.y);
                        csv[2 * g4 + 1] = pack2(c4.z, c4.w);
                        tbv[2 * g4] = pack2(t4.x, t4.y);
                        tbv[2 * g4 + 1] = pack2(t4.z, t4.w);
                    }
                }
                mbar_wait(&bars->h_empty[s], ph ^ 1);  // every warp finished the stencil of chunk t - 2 on this buffer
                mbar_wait(&bars->d_full[s], ph);
                tc_fence_after();
                const uint32_t hid = hid0 + s * kPwHid;
                const uint32_t tq = tmem_base + s * kPwAccStride + cg * 16 + (uint32_t(q * 32) << 16);
                uint32_t acc[2][16];  // ping-pong: the load of M-block mb + 1 flies while mb is converted
                tmem_ld16_nowait(tq, acc[0]);
#pragma unroll
                for (int mb = 0; mb < 3; ++mb) {
                    if (mb * 128 + q * 32 >= kPwRows) continue;  // warp-uniform: quarter 3 of the last M-block is empty
                    tmem_ld_wait();
                    if (mb + 1 < 3 && (mb + 1) * 128 + q * 32 < kPwRows) tmem_ld16_nowait(tq + (mb + 1) * 64, acc[(mb + 1) & 1]);
                    const int r = mb * 128 + q * 32 + lane;
                    const f32x2 a2 = splat2(ar[mb]), b2 = splat2(br[mb]);
                    uint32_t pk[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        f32x2 x = pack2(__uint_as_float(acc[mb & 1][2 * i]), __uint_as_float(acc[mb & 1][2 * i + 1]));
                        if (p.cs != nullptr)   // WithBias: rstd * acc + ((-mu rstd) * colsum + t)
                            x = fma2(x, a2, fma2(b2, csv[i], tbv[i]));
                        else                   // BiasFree: rstd * acc
                            x = mul2(x, a2);
                        pk[i] = f32x2_to_bf16x2(x) & okm[mb];  // zero padding of the 1x1's OUTPUT
                    }
                    if (r < kPwRows) {
                        const uint32_t row = hid + r * 128;
                        const uint32_t rx = uint32_t(r & 7);
                        sts_v4(row + ((uint32_t(cg * 2) ^ rx) << 4), pk[0], pk[1], pk[2], pk[3]);
                        sts_v4(row + ((uint32_t(cg * 2 + 1) ^ rx) << 4), pk[4], pk[5], pk[6], pk[7]);
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) {
                    mbar_arrive(&bars->d_empty[s]);  // this warp's share of the accumulator is drained
                    mbar_arrive(&bars->h_full[s]);   // and its part of the hidden tile is written
                }
                if (++jC == NCH) {
                    jC = 0;
                    itemC += gridDim.x;
                }
            }
            if (t == 0) continue;
            // ---------------- stencil of chunk t - 1 ----------------
            const int s = (t - 1) & 1;
            const uint32_t ph = ((t - 1) >> 1) & 1;
            if (jS == 0) {
                imgS = itemS / per_img;
                const int rr = itemS - imgS * per_img;
                tyS = rr / p.tiles_x;
                txS = rr - tyS * p.tiles_x;
            }
            const int j = jS;
            const int gx = txS * kPwTW + col0;
            const int nrows = min(kPwTH, p.H - tyS * kPwTH);
            int c0, oc;
            bool cvalid;
            if (GATE == 0) {
                c0 = j * 64 + lane * 2;
                oc = c0;
                cvalid = c0 < p.nout;
            } else {
                oc = j * 32 + (lane & 15) * 2;
                c0 = (lane < 16 ? 0 : p.hp) + oc;
                cvalid = oc < p.nout;
            }
            f32x2 wr[9];  // this lane's channel pair: 9 taps x 2 channels, contiguous in the packed weights
            {
                const float2* wp = reinterpret_cast<const float2*>(p.dww) + (cvalid ? (c0 >> 1) * 9 : 0);
#pragma unroll
                for (int tt = 0; tt < 9; ++tt) {
                    const float2 w2 = __ldg(wp + tt);
                    wr[tt] = cvalid ? pack2(w2.x, w2.y) : pack2(0.f, 0.f);
                }
            }
            // stencil loads: the 16-byte chunk (lane >> 2) of halo row r sits at ((lane >> 2) ^ (r & 7)) << 4; r = col0 + K
            // with a compile-time K, and (col0 + K) & 7 == (col0 + (K & 7)) & 7: eight per-lane bases cover every row
            uint32_t a[8];
#pragma unroll
            for (int k = 0; k < 8; ++k)
                a[k] = hid0 + s * kPwHid + col0 * 128 + ((uint32_t(lane >> 2) ^ uint32_t((col0 + k) & 7)) << 4) + (lane & 3) * 4;
            bf16* yp0;
            bool ok0, ok1;
            if (GATE == 0) {
                yp0 = p.y + (((long long)imgS * p.H + tyS * kPwTH) * p.W + gx) * p.ldy + oc;
                ok0 = cvalid && gx < p.W;
                ok1 = cvalid && gx + 1 < p.W;
            } else {  // this lane stores pixel gx + (lane >> 4)
                yp0 = p.y + (((long long)imgS * p.H + tyS * kPwTH) * p.W + gx + (lane >> 4)) * p.ldy + oc;
                ok0 = ok1 = cvalid && gx + (lane >> 4) < p.W;
            }
            mbar_wait(&bars->h_full[s], ph);
            f32x2 acc[3][2];  // [output row mod 3][pixel]
#pragma unroll
            for (int hr = 0; hr < kPwHH; ++hr) {
                f32x2 in[4];
#pragma unroll
                for (int d = 0; d < 4; ++d) {
                    const int K = hr * kPwHW + d;
                    in[d] = bf16x2_to_f32x2_alu(lds_u32(a[K & 7] + K * 128));
                }
#pragma unroll
                for (int dr = 0; dr < 3; ++dr) {  // halo row hr is tap row dr of output row hr - dr
                    const int orow = hr - dr;
                    if (orow < 0 || orow >= kPwTH) continue;
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
                const int done = hr - 2;  // this output row is complete
                if (done >= 0) {
                    const int sl = done % 3;
                    if (GATE == 0) {
                        float o0, o1;
                        unpack2(acc[sl][0], o0, o1);
                        const uint32_t pk0 = pack_bf16x2(o0, o1);
                        unpack2(acc[sl][1], o0, o1);
                        const uint32_t pk1 = pack_bf16x2(o0, o1);
                        if (done < nrows) {
                            if (ok0) *reinterpret_cast<uint32_t*>(yp0) = pk0;
                            if (ok1) *reinterpret_cast<uint32_t*>(yp0 + p.ldy) = pk1;
                        }
                    } else {
                        // lanes 0-15 hold dw(h1), lanes 16-31 dw(h2) of the same channel pair: the lower half gates
                        // pixel 0, the upper half pixel 1
                        const f32x2 send = lane < 16 ? acc[sl][1] : acc[sl][0];
                        const f32x2 recv = __shfl_xor_sync(0xffffffffu, send, 16);
                        const f32x2 h1 = lane < 16 ? acc[sl][0] : recv;
                        const f32x2 h2 = lane < 16 ? recv : acc[sl][1];
                        const f32x2 g = mul2(gelu2_fast(h1), h2);
                        float o0, o1;
                        unpack2(g, o0, o1);
                        if (done < nrows && ok0) *reinterpret_cast<uint32_t*>(yp0) = pack_bf16x2(o0, o1);
                    }
                    yp0 += row_step;
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&bars->h_empty[s]);
            if (++jS == NCH) {
                jS = 0;
                itemS += gridDim.x;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kPwDwWarps + kPwCvWarps + 1) tmem_dealloc(tmem_base, 512);
}

bool pwdw_fused_shape_ok(int C, int nrows, bool gate, int hp) {
    if (C < 16 || C > 128 || C % 16 != 0) return false;           // K = C in one (C <= 64) or two 64-channel slabs
    if (gate) return hp >= 32 && hp % 32 == 0 && nrows == 2 * hp;
    return nrows % 8 == 0 && nrows >= 8;
}

// plain: y[n,H,W,:nrows] (row stride ldy) = dw3x3(W1' . LN(x)); gate: y[n,H,W,:hp] = gelu(dw(h)[:hp]) * dw(h)[hp:]
int pwdw_fused_bf16(bool gate, const bf16* x, long long ldx, const float2* stats, const bf16* w1, int nrows, int hp,
                    const float* cs, const float* tb, const float* dww, int dw_ld, bf16* y, long long ldy, int Nimg,
                    int H, int W, int C, cudaStream_t stream) {
    IRB_REQUIRE(pwdw_fused_shape_ok(C, nrows, gate, hp), "pwdw_fused: unsupported shape");
    IRB_REQUIRE(ldx % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0, "pwdw_fused: x must have 16-byte aligned rows");
    IRB_REQUIRE(ldy % 2 == 0 && (reinterpret_cast<uintptr_t>(y) & 3) == 0 && stats != nullptr, "pwdw_fused: y / stats");
    IRB_REQUIRE((cs == nullptr) == (tb == nullptr), "pwdw_fused: colsum and t come together");
    PwDwParams p{};
    {
        uint64_t dims[4] = {uint64_t(C), uint64_t(W), uint64_t(H), uint64_t(Nimg)};
        uint64_t str[3] = {uint64_t(ldx) * 2, uint64_t(ldx) * 2 * W, uint64_t(ldx) * 2 * W * H};
        uint32_t box[4] = {64, kPwHW, kPwHH, 1};
        IRB_TRY(make_tmap_bf16(&p.tmX, x, 4, dims, str, box));
    }
    {
        uint64_t dims[2] = {uint64_t(C), uint64_t(nrows)};
        uint64_t str[1] = {uint64_t(C) * 2};
        uint32_t box[2] = {64, gate ? 32u : 64u};
        IRB_TRY(make_tmap_bf16(&p.tmW, w1, 2, dims, str, box));
    }
    p.stats = stats;
    p.cs = cs;
    p.tb = tb;
    p.dww = dww;
    (void)dw_ld;
    p.y = y;
    p.ldy = ldy;
    p.H = H;
    p.W = W;
    p.KS = ceil_div(C, 64);
    p.klast = (C - (p.KS - 1) * 64) / 16;
    p.nrows = nrows;
    p.hp = hp;
    p.nout = gate ? hp : nrows;
    p.nchunks = gate ? hp / 32 : ceil_div(nrows, 64);
    p.tiles_x = ceil_div(W, kPwTW);
    p.tiles_y = ceil_div(H, kPwTH);
    const long long total = (long long)p.tiles_x * p.tiles_y * Nimg;
    IRB_REQUIRE(total < (1ll << 31), "pwdw_fused: too many work items");
    p.total = int(total);
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = int(total < sms ? total : sms);
    const size_t smem = 1024 + size_t(p.KS) * kPwASlab + size_t(2) * p.KS * kPwBSlab + 2 * kPwHid + sizeof(PwBars) + 64;
    IRB_REQUIRE(smem <= 227 * 1024, "pwdw_fused: shared memory budget exceeded");
    IRB_SMEM_OPTIN((pwdw_fused_kernel<0>), 227 * 1024);
    IRB_SMEM_OPTIN((pwdw_fused_kernel<1>), 227 * 1024);
    if (gate)
        IRB_CHECK_CUDA(launch_pdl(pwdw_fused_kernel<1>, dim3(grid), dim3(kPwThreads), smem, stream, p));
    else
        IRB_CHECK_CUDA(launch_pdl(pwdw_fused_kernel<0>, dim3(grid), dim3(kPwThreads), smem, stream, p));
    const double px = double(Nimg) * H * W;
    // algorithmic work: x read once, the result written once, weights once; the 1x1 and the stencil flops
    prof_mark(gate ? "pwdw_gate_fused" : "pwdw_fused", C, gate ? hp : nrows, 2.0 * px * C * nrows + 18.0 * px * nrows,
              px * (C + (gate ? hp : nrows)) * 2 + double(nrows) * (C * 2 + 36 + 8) + px * 8, stream);
    return 0;
}

}  // namespace irb
