// tcgen05 pixel-GEMM / implicit-GEMM 3x3 convolution for sm_100a (bf16 in, fp32 accumulate in TMEM).
//
//   out[row, n] = epilogue( sum_{tap, k} A[row shifted by tap, k] * Wt[n, tap*kpad + k] )
//
// mode 0 (1x1 conv / Linear): A is a [batch, M, K] channels-last activation (row stride lda), loaded by a
//   3-D TMA map {K, M, batch} with a {64, 128, 1} box; K beyond the tensor is zero-filled by TMA.
// mode 1 (3x3 conv, zero pad 1): A is a [N, H, W, C] channels-last activation, loaded by a 4-D TMA map
//   {C, W, H, N} with a {64, BW, BH, 1} box (BW*BH = 128 pixels); the 9 taps are 9 shifted boxes and the
//   out-of-bounds halo is zero-filled by TMA, which is exactly the convolution's zero padding.
//   Narrow outputs (BN <= 64, `dyr`): a 16 x 8 pixel tile, and per (dx, K slab) ONE A box {64, 16, 10} whose rows
//   dy * 16 .. dy * 16 + 127 are the A operand of tap (dy, dx) (a 2 KB offset: plain descriptors), next to the three
//   weight slabs of that dx. Other narrow plain rings put `kg` K slabs into one stage (fewer barrier handshakes).
// Weights Wt are [Nrows, Ktot] bf16, K contiguous (the layout nn.Conv2d / nn.Linear already use), loaded by a
//   2-D TMA map with a {64, BN} box. Both operands land in shared memory in the 128B-swizzled K-major layout
//   that the UMMA descriptors below describe.
//
// Persistent CTAs walk 128 x BN output tiles. Warp 0 lane 0: TMA producer; warp 1 lane 0: tcgen05.mma issuer;
// warps 2..5 / 6..9: two epilogue sets, one per TMEM accumulator (TMEM -> registers -> fused LN-fold / bias / act /
// residual -> swizzled smem -> TMA store). A ring of `stages` shared-memory slots with full/empty mbarriers
// decouples TMA from the tensor core; tfull/tempty mbarriers hand accumulators between MMA and epilogue.
#include "common.cuh"
#include <cstdio>
#include "ptx.cuh"
#include "kernels.h"
#include <stdlib.h>

namespace irb {

static constexpr int kTileM = 128;
static constexpr int kTileK = 64;                 // bf16 elements per 128-byte swizzle row
static constexpr int kABytes = kTileM * 128;      // 16 KB

struct alignas(64) TcGemmParams {
    CUtensorMap tmA;
    CUtensorMap tmB;
    CUtensorMap tmR;  // bf16 residual (mode 0, RES 2): same geometry and box as tmC
    CUtensorMap tmC;  // output (bf16): mode 0 {N, M, batch} box {64, 32, 1}; mode 1 {N, W, H, Nimg} box {64, bw, 32/bw, 1}
    int tma_store;    // 1: epilogue stages 32-row x 64-column blocks in shared memory and stores them with TMA
    int mode;
    int M;        // rows per batch (mode 0)
    int N;        // valid output columns
    int BN;       // columns per tile (multiple of 16, <= 256)
    int ktiles;   // 64-wide K chunks per tap
    int klast;    // 16-wide K steps of the last chunk of a tap that hold data (1..4); the rest is TMA zero fill
    int taps;     // 1 or 9
    int kpad;     // K distance between taps inside Wt
    int H, W, BW, BH, tiles_x;
    int b_rows_per_batch;
    int stages;
    int dyr;          // mode 1, narrow tiles: one A halo box {64, BW, BH + 2} per (dx, K slab) serves the three dy taps
    int kg;           // K slabs (A | B pairs) per ring stage: 1, or more for narrow tiles (see launch_tc)
    int tmem_cols;    // 2 accumulators of acc_stride columns
    int acc_stride;
    int tiles_m, tiles_n, total;  // work items: total = batch * tiles_m * tiles_n, n fastest
    int step_nt, step_units;      // order 0: divmod(grid, tiles_n)
    int m_outer;                  // 1: a CTA walks all n-tiles of its m-tile back to back
    int b_resident;               // 1 (mode 0): the whole weight matrix stays in shared memory, the ring carries A only
    int bres_bytes;
    int a_hold;                   // mode 0, weights not resident: 1 / 2 buffers that keep an m-tile's A slabs while the
                                  // CTA walks its n-tiles (the ring then carries B only); 0 = A travels with B
    int blk_cols;                 // columns per epilogue block: 64, or 48 (PixelShuffle with 48 | N/4, specialised kernels only)
    int res_tma;                  // 1: tmR is valid and the residual is TMA-staged (RES 2)
    int stg_per_warp;             // 4 KB staging blocks per epilogue warp: 2, or 3 when the residual is TMA-staged
    int dbg;                      // IRB_GEMM_DBG: 1 skip output stores, 2 skip LN-stats loads, 4 skip residual loads (diagnosis only)
    Epi epi;
};

static constexpr int kActGeluTanh = 3;    // internal: ACT_GELU evaluated in its tanh form (bf16 outputs, see ptx.cuh)
static constexpr int kGemmThreads = 320;  // warp 0: TMA, warp 1: MMA, warps 2..9: epilogue

// Lean epilogue math for 16 columns of one row: the per-column vectors come from shared memory (staged once
// per tile), the per-row LayerNorm fold is two FMAs per element:
//   y = act(acc * a + (b * colsum[col] + bias[col])) + res         a = rstd (or 1), b = -mu * rstd (or 0)
template <int ACT>
__device__ __forceinline__ void epi_math16(const Epi& e, long long row, int col0, int res_cols, float* v,
                                           const float* s_bias, const float* s_cs, float a, float b) {
#pragma unroll
    for (int j = 0; j < 16; j += 4) {
        const float4 bi = *reinterpret_cast<const float4*>(s_bias + j);
        const float4 cs = *reinterpret_cast<const float4*>(s_cs + j);
        v[j + 0] = fmaf(v[j + 0], a, fmaf(b, cs.x, bi.x));
        v[j + 1] = fmaf(v[j + 1], a, fmaf(b, cs.y, bi.y));
        v[j + 2] = fmaf(v[j + 2], a, fmaf(b, cs.z, bi.z));
        v[j + 3] = fmaf(v[j + 3], a, fmaf(b, cs.w, bi.w));
    }
    if (ACT == ACT_GELU) {
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = gelu_as(v[j]);  // bf16 output: 1.5e-7 abs error is invisible
    } else if (ACT == ACT_LRELU) {
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = v[j] > 0.f ? v[j] : v[j] * e.slope;
    }
    if (res_cols > 0 && res_cols < 16) {  // ragged last group: element-wise
        const long long rbase = row * e.ldr + col0;
#pragma unroll
        for (int j = 0; j < 16; ++j)
            if (j < res_cols) v[j] += ld_as_float(e.res, rbase + j, e.res_dt);
    } else if (res_cols >= 16) {
        const long long rbase = row * e.ldr + col0;
        if (e.res_dt == DT_BF16) {
            const uint4* rp = reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(e.res) + rbase);
#pragma unroll
            for (int q = 0; q < 2; ++q) {
                const uint4 u = __ldg(rp + q);
                float x0, x1;
                unpack_bf16x2(u.x, x0, x1); v[q * 8 + 0] += x0; v[q * 8 + 1] += x1;
                unpack_bf16x2(u.y, x0, x1); v[q * 8 + 2] += x0; v[q * 8 + 3] += x1;
                unpack_bf16x2(u.z, x0, x1); v[q * 8 + 4] += x0; v[q * 8 + 5] += x1;
                unpack_bf16x2(u.w, x0, x1); v[q * 8 + 6] += x0; v[q * 8 + 7] += x1;
            }
        } else {
            const float4* rp = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(e.res) + rbase);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float4 u = __ldg(rp + q);
                v[q * 4 + 0] += u.x; v[q * 4 + 1] += u.y; v[q * 4 + 2] += u.z; v[q * 4 + 3] += u.w;
            }
        }
    }
}
template <int ACT>
__device__ __forceinline__ void epi_fast16(const Epi& e, long long row, int col0, float* v, const float* s_bias,
                                           const float* s_cs, float a, float b) {
    epi_math16<ACT>(e, row, col0, e.res != nullptr ? 16 : 0, v, s_bias, s_cs, a, b);
    const long long obase = row * e.ldo + col0;
    if (e.out_dt == DT_BF16) {
        uint4* op = reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(e.out) + obase);
#pragma unroll
        for (int q = 0; q < 2; ++q) {
            uint4 u;
            u.x = pack_bf16x2(v[q * 8 + 0], v[q * 8 + 1]);
            u.y = pack_bf16x2(v[q * 8 + 2], v[q * 8 + 3]);
            u.z = pack_bf16x2(v[q * 8 + 4], v[q * 8 + 5]);
            u.w = pack_bf16x2(v[q * 8 + 6], v[q * 8 + 7]);
            op[q] = u;
        }
    } else {
        float4* op = reinterpret_cast<float4*>(reinterpret_cast<float*>(e.out) + obase);
#pragma unroll
        for (int q = 0; q < 4; ++q) op[q] = make_float4(v[q * 4], v[q * 4 + 1], v[q * 4 + 2], v[q * 4 + 3]);
    }
}

// Vectorised epilogue math for one full 16-column group held as raw TMEM words: packed FFMA2 / FADD2, column
// vectors read from shared memory as 16-byte loads, bf16 residual pre-fetched by the caller (r0: columns 0..7,
// r1: 8..15). LN 0: y = acc; 1: y = acc * a + bias[col]; 2: y = acc * a + (b * colsum[col] + bias[col]).
// Everything that selects a variant is a template parameter: the unrolled body has no branches.
template <int ACT, int LN, bool RES, bool STATS>
__device__ __forceinline__ void epi_group16(const uint32_t* acc, const float* sb, const float* scs, f32x2 a2, f32x2 b2,
                                            float slope, const uint4& r0, const uint4& r1, uint32_t* pk, f32x2& s1,
                                            f32x2& s2, f32x2 nshift2) {
#pragma unroll
    for (int j4 = 0; j4 < 4; ++j4) {
        f32x2 x0 = pack2(__uint_as_float(acc[4 * j4 + 0]), __uint_as_float(acc[4 * j4 + 1]));
        f32x2 x1 = pack2(__uint_as_float(acc[4 * j4 + 2]), __uint_as_float(acc[4 * j4 + 3]));
        if (LN != 0) {
            const ulonglong2 bb = *reinterpret_cast<const ulonglong2*>(sb + j4 * 4);
            f32x2 bi0 = bb.x, bi1 = bb.y;
            if (LN == 2) {
                const ulonglong2 cc = *reinterpret_cast<const ulonglong2*>(scs + j4 * 4);
                bi0 = fma2(b2, cc.x, bi0);
                bi1 = fma2(b2, cc.y, bi1);
            }
            x0 = fma2(x0, a2, bi0);
            x1 = fma2(x1, a2, bi1);
        }
        if (ACT == ACT_GELU) {
            x0 = gelu2(x0);
            x1 = gelu2(x1);
        } else if (ACT == kActGeluTanh) {
            x0 = gelu2_tanh(x0);
            x1 = gelu2_tanh(x1);
        } else if (ACT == ACT_LRELU) {
            float t0, t1, t2, t3;
            unpack2(x0, t0, t1);
            unpack2(x1, t2, t3);
            x0 = pack2(t0 > 0.f ? t0 : t0 * slope, t1 > 0.f ? t1 : t1 * slope);
            x1 = pack2(t2 > 0.f ? t2 : t2 * slope, t3 > 0.f ? t3 : t3 * slope);
        }
        if (RES) {
            const uint32_t rw0 = j4 == 0 ? r0.x : (j4 == 1 ? r0.z : (j4 == 2 ? r1.x : r1.z));
            const uint32_t rw1 = j4 == 0 ? r0.y : (j4 == 1 ? r0.w : (j4 == 2 ? r1.y : r1.w));
            x0 = add2(x0, bf16x2_to_f32x2(rw0));
            x1 = add2(x1, bf16x2_to_f32x2(rw1));
        }
        if (STATS) {  // shifted sums: sum (y - shift), sum (y - shift)^2
            const f32x2 d0 = add2(x0, nshift2), d1 = add2(x1, nshift2);
            s1 = add2(s1, add2(d0, d1));
            s2 = fma2(d0, d0, s2);
            s2 = fma2(d1, d1, s2);
        }
        pk[2 * j4] = f32x2_to_bf16x2(x0);
        pk[2 * j4 + 1] = f32x2_to_bf16x2(x1);
    }
}

// A group whose 16 columns are not all < N (ragged N such as SwinIR's 180): element-wise, columns >= N give 0.
template <int ACT, int LN, bool RES, bool STATS>
__device__ __forceinline__ void epi_group16_ragged(const uint32_t* acc, int nvalid, const float* sb, const float* scs,
                                                   float fa, float fb, float slope, const bf16* res, uint32_t* pk,
                                                   f32x2& s1, f32x2& s2, float shift) {
    float v[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        float y = __uint_as_float(acc[j]);
        if (LN == 1) y = fmaf(y, fa, sb[j]);
        if (LN == 2) y = fmaf(y, fa, fmaf(fb, scs[j], sb[j]));
        if (ACT == ACT_GELU) y = gelu_as(y);
        if (ACT == kActGeluTanh) {
            const float hy = 0.5f * y;
            y = fmaf(hy, tanh_approx(y * fmaf(y * y, 0.7978845608f * 0.044715f, 0.7978845608f)), hy);
        }
        if (ACT == ACT_LRELU) y = y > 0.f ? y : y * slope;
        if (j < nvalid) {
            if (RES) y += __bfloat162float(res[j]);
            if (STATS) {
                const float d = y - shift;
                s1 = add2(s1, pack2(d, 0.f));
                s2 = fma2(pack2(d, 0.f), pack2(d, 0.f), s2);
            }
        } else {
            y = 0.f;
        }
        v[j] = y;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) pk[j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
}

// One 64-column block (ng <= 4 groups, nvalid columns < N) of one row: math + the swizzled 16-byte stores into the
// staging block.
template <int ACT, int LN, bool RES, bool STATS>
__device__ __forceinline__ void epi_block(const uint32_t (*acc)[16], const uint4* rr, int ng, int nvalid,
                                          const float* sb, const float* scs, float fa, float fb, float slope,
                                          const bf16* res_row, uint8_t* stg_row, int lane, f32x2& s1, f32x2& s2,
                                          float shift) {
    const f32x2 a2 = splat2(fa), b2 = splat2(fb), nshift2 = splat2(-shift);
#pragma unroll
    for (int gi = 0; gi < 4; ++gi) {
        if (gi < ng) {
            uint32_t pk[8];
            if (gi * 16 + 16 <= nvalid || !RES)
                epi_group16<ACT, LN, RES, STATS>(acc[gi], sb + gi * 16, scs + gi * 16, a2, b2, slope, rr[2 * gi],
                                                 rr[2 * gi + 1], pk, s1, s2, nshift2);
            else
                epi_group16_ragged<ACT, LN, RES, STATS>(acc[gi], nvalid - gi * 16, sb + gi * 16, scs + gi * 16, fa, fb,
                                                        slope, res_row + gi * 16, pk, s1, s2, shift);
#pragma unroll
            for (int c = 0; c < 2; ++c)
                *reinterpret_cast<uint4*>(stg_row + (((2 * gi + c) ^ (lane & 7)) << 4)) =
                    make_uint4(pk[c * 4], pk[c * 4 + 1], pk[c * 4 + 2], pk[c * 4 + 3]);
        }
    }
}

// bf16 residual of one 64-column block of a row: whole 16-column groups only (ragged groups are read element-wise)
__device__ __forceinline__ void load_res_block(const bf16* rb, int ng, int nvalid, uint4 (&dst)[8]) {
#pragma unroll
    for (int gi = 0; gi < 4; ++gi) {
        if (gi < ng && gi * 16 + 16 <= nvalid) {
            dst[2 * gi] = __ldg(reinterpret_cast<const uint4*>(rb + gi * 16));
            dst[2 * gi + 1] = __ldg(reinterpret_cast<const uint4*>(rb + gi * 16 + 8));
        }
    }
}

// Work item w = (z * tiles_m + mt) * tiles_n + nt. A CTA takes m-tiles blockIdx.x, + gridDim.x, ... and walks ALL
// n-tiles of an m-tile back to back, so the A tile it re-reads for the next n-tile is still in L2.
__device__ __forceinline__ int first_item(const TcGemmParams& p) {
    return p.m_outer ? blockIdx.x * p.tiles_n : blockIdx.x;
}
__device__ __forceinline__ int next_item(int w, const TcGemmParams& p) {
    if (!p.m_outer) return w + gridDim.x;
    return ((w + 1) % p.tiles_n) ? w + 1 : w + 1 + (gridDim.x - 1) * p.tiles_n;
}

// The same walk with the coordinates carried along: the single-thread producer / MMA loops and the epilogue warps paid
// 8-12 integer divisions by run-time values per tile (each a ~30-instruction dependent chain: ~1 us per tile with
// nothing else to do, measured with IRB_GEMM_DBG=224), now only additions and compares.
struct TileIter {
    int w, nt, mt, z;
};
__device__ __forceinline__ void tile_set(const TcGemmParams& p, TileIter& t, int w) {
    t.w = w;
    t.nt = w % p.tiles_n;
    const int q = w / p.tiles_n;
    t.z = q / p.tiles_m;
    t.mt = q - t.z * p.tiles_m;
}
__device__ __forceinline__ void tile_first(const TcGemmParams& p, TileIter& t) { tile_set(p, t, first_item(p)); }
__device__ __forceinline__ void tile_next(const TcGemmParams& p, TileIter& t) {
    if (!p.m_outer) {  // w += grid: (step_units, step_nt) = divmod(grid, tiles_n), computed on the host
        t.w += gridDim.x;
        t.nt += p.step_nt;
        int du = p.step_units;
        if (t.nt >= p.tiles_n) {
            t.nt -= p.tiles_n;
            ++du;
        }
        t.mt += du;
        if (t.mt >= p.tiles_m) {
            if (t.mt - p.tiles_m < p.tiles_m) {
                t.mt -= p.tiles_m;
                ++t.z;
            } else {
                const int q = t.mt / p.tiles_m;
                t.mt -= q * p.tiles_m;
                t.z += q;
            }
        }
        return;
    }
    if (++t.nt < p.tiles_n) {
        ++t.w;
        return;
    }
    t.nt = 0;
    t.w += 1 + (gridDim.x - 1) * p.tiles_n;
    t.mt += gridDim.x;
    if (t.mt >= p.tiles_m) {
        if (t.mt - p.tiles_m < p.tiles_m) {
            t.mt -= p.tiles_m;
            ++t.z;
        } else {
            const int q = t.mt / p.tiles_m;
            t.mt -= q * p.tiles_m;
            t.z += q;
        }
    }
}


// ----------------------------------------------------------------------------------------------
// Specialised epilogue (TMA-store path, no residual / bf16 residual). Every switch that selects the arithmetic is a
// template parameter and the dispatch happens once per kernel, so the per-block code is one straight run of
// LDTM / (LDS) / FFMA2 / F2FP / STS instead of a switch over inlined variants inside the tile loop.
//   RES 0: no residual; 1: bf16 residual read by each thread from global memory (16-byte loads, mode 1);
//   RES 2: bf16 residual rows TMA-loaded INTO the warp's staging block two blocks ahead (mode 0): the thread reads
//          its 16-byte chunks from the swizzled block, adds, writes the result back in place, and the same block
//          goes out as the TMA store, so no residual latency sits in front of the math.
// ----------------------------------------------------------------------------------------------
// IRB_GEMM_DBG & 256: poll with mbarrier.test_wait instead of the suspending try_wait (handshake-latency diagnosis)
#define GEMM_WAIT(bar, parity)                          \
    do {                                                \
        if (p.dbg & 256) mbar_wait_spin((bar), (parity)); \
        else mbar_wait((bar), (parity));                \
    } while (0)

struct EpiCtx {
    uint32_t tmem_base;
    uint64_t* tfull_bar;
    uint64_t* tempty_bar;
    uint64_t* res_bar;   // [8 warps][3] residual-block barriers (RES 2)
    const float* s_bias_all;
    const float* s_cs_all;
    uint8_t* staging;    // 8 warps x kStgPerWarp blocks of 4 KB
    int stg_per_warp;    // 2, or 3 with RES 2
};

// rows of tile item w that this thread (row r of the tile) owns
__device__ __forceinline__ void epi_row_of(const TcGemmParams& p, const TileIter& ti, int r, long long& grow, bool& valid) {
    const int mt = ti.mt;
    const int z = ti.z;
    if (p.mode == 0) {
        const int m = mt * kTileM + r;
        valid = m < p.M;
        grow = (long long)z * p.M + m;
    } else {
        const int py = (mt / p.tiles_x) * p.BH + r / p.BW;
        const int px = (mt % p.tiles_x) * p.BW + r % p.BW;
        valid = (py < p.H) && (px < p.W);
        grow = ((long long)z * p.H + py) * p.W + px;
    }
}

__device__ __forceinline__ float2 epi_load_stats(const TcGemmParams& p, const TileIter& ti, int r) {
    const int w = ti.w;
    float2 st = make_float2(0.f, 1.f);
    if (w < p.total) {
        long long grow;
        bool valid;
        epi_row_of(p, ti, r, grow, valid);
        if (valid) st = __ldg(p.epi.ln_stats + grow);
    }
    return st;
}

// NG full 16-column groups of one row. sx = shared address of the thread's staging row | ((lane & 7) << 4): chunk j of
// the row lives at sx ^ (j << 4) (128-byte swizzle).
template <int ACT, int LN, int RES, bool STATS, int NG>
__device__ __forceinline__ void epi_block_full(uint32_t taddr_cb, const float* sb, const float* scs, f32x2 a2, f32x2 b2,
                                               float slope, const bf16* rrow_cb, uint32_t sx, bool first_blk,
                                               f32x2& s1, f32x2& s2, float& shift, bool wait_read, int lane,
                                               uint32_t (&accb)[2][16], bool preloaded, uint32_t next_taddr) {
    // TMEM loads run one 16-column group ahead of the math (two register buffers): group gi + 1 is in flight while
    // group gi is computed, and the block's last group fetches group 0 of the NEXT block of the tile (next_taddr != 0;
    // NG even, so that it lands in buffer 0), which then enters with `preloaded`. Only the first load of a tile is exposed.
    if (!preloaded) tmem_ld16_nowait(taddr_cb, accb[0]);
    uint4 rr[RES == 1 ? 2 * NG : 1];
    if (RES == 1) {
#pragma unroll
        for (int c = 0; c < 2 * NG; ++c) rr[c] = __ldg(reinterpret_cast<const uint4*>(rrow_cb + c * 8));
    }
    if (wait_read) {  // the store issued two blocks ago (same staging block) must have finished READING it
        bulk_wait_read1();  // every lane: only the elected issuer has bulk groups, the others fall through
        __syncwarp();
    }
    if (STATS && first_blk)  // x_old[row, 0]
        shift = __uint_as_float((RES == 2 ? lds_u32(sx) : rr[0].x) << 16);
    const f32x2 nshift2 = splat2(-shift);
#pragma unroll
    for (int gi = 0; gi < NG; ++gi) {
        tmem_ld_wait();
        if (gi + 1 < NG)
            tmem_ld16_nowait(taddr_cb + (gi + 1) * 16, accb[(gi + 1) & 1]);
        else if ((NG & 1) == 0 && next_taddr != 0)
            tmem_ld16_nowait(next_taddr, accb[0]);
        uint32_t pk[8];
        uint4 r0 = make_uint4(0, 0, 0, 0), r1 = r0;
        if (RES == 1) {
            r0 = rr[2 * gi];
            r1 = rr[2 * gi + 1];
        } else if (RES == 2) {  // staged by TMA: short latency, read group by group
            r0 = lds_v4(sx ^ ((2 * gi) << 4));
            r1 = lds_v4(sx ^ ((2 * gi + 1) << 4));
        }
        epi_group16<ACT, LN, RES != 0, STATS>(accb[gi & 1], sb + gi * 16, scs + gi * 16, a2, b2, slope, r0, r1, pk, s1,
                                              s2, nshift2);
        sts_v4(sx ^ ((2 * gi) << 4), pk[0], pk[1], pk[2], pk[3]);
        sts_v4(sx ^ ((2 * gi + 1) << 4), pk[4], pk[5], pk[6], pk[7]);
    }
}

template <int ACT, int LN, int RES, bool STATS>
__device__ __forceinline__ void epilogue_fast(const TcGemmParams& p, const EpiCtx& c, int warp, int lane) {
    const int q = warp & 3;
    const int set = (warp - 2) >> 2;
    const int r = q * 32 + lane;
    const int bcols = p.blk_cols;
    const int nblk_max = (p.BN + bcols - 1) / bcols;
    const uint32_t stg_warp = smem_u32(c.staging) + (warp - 2) * c.stg_per_warp * 4096;
    uint8_t* stg_warp_p = c.staging + (warp - 2) * c.stg_per_warp * 4096;
    const uint32_t row_off = lane * 128 + ((lane & 7) << 4);
    uint64_t* rbar = c.res_bar + (warp - 2) * 3;
    const float slope = p.epi.slope;
    // The lane that issues this warp's TMA loads / stores. elect.sync (deterministic for the full mask) instead of
    // lane == 0 lets ptxas issue UTMASTG / UTMALDG / the commit without an election loop around each; the bulk-group
    // waits are executed by all lanes because only the issuer's are non-trivial.
    const bool leader = elect_one() != 0;

    // first tile of this warp set
    TileIter ti;  // this set's current tile; tn = its next tile (two steps of the CTA's walk ahead)
    tile_first(p, ti);
    if (set == 1) tile_next(p, ti);
    TileIter tn = ti;
    tile_next(p, tn);
    tile_next(p, tn);
    int t = set;
    // software pipeline: LayerNorm statistics of the NEXT tile's row are fetched while this tile is processed
    float2 st_next = make_float2(0.f, 1.f);
    if (LN != 0 && p.epi.ln_stats) st_next = epi_load_stats(p, ti, r);

    // RES 2: residual prefetch iterator (tp, pblk), two blocks ahead of the consumer
    TileIter tp = ti;
    int pblk = 0;
    uint32_t nissued = 0;
    auto issue_res = [&]() {
        if (tp.w >= p.total) return;
        const int nt2 = tp.nt, mt2 = tp.mt, z2 = tp.z;
        const int col = nt2 * p.BN + pblk * bcols;
        const uint32_t b = nissued % 3;
        if (leader) {
            mbar_arrive_expect_tx(&rbar[b], 4096);
            tma_load_3d(stg_warp_p + b * 4096, &p.tmR, &rbar[b], col, mt2 * kTileM + q * 32, z2);
        }
        ++nissued;
        ++pblk;
        if (pblk >= nblk_max || nt2 * p.BN + pblk * bcols >= p.N) {
            pblk = 0;
            tile_next(p, tp);
            tile_next(p, tp);
        }
    };
    if (RES == 2 && !(p.dbg & 32)) {
        issue_res();
        issue_res();
    }

    uint32_t nb = 0;  // blocks consumed by this warp
#ifdef IRB_GEMM_TRACE  // timing experiment: clock() trace of the first epilogue warp of CTA 0, launches with N == IRB_GEMM_TRACE
    const bool trace = blockIdx.x == 0 && warp == 2 && lane == 0 && p.N == IRB_GEMM_TRACE;
    long long tsv[160];
    int tsi[160];
    int nts = 0;
#define IRB_TS(i) do { if (trace && nts < 160) { tsv[nts] = clock64(); tsi[nts++] = (i); } } while (0)
#else
#define IRB_TS(i) do { } while (0)
#endif
    for (; ti.w < p.total; t += 2) {
        IRB_TS(0);
        const int nt = ti.nt;
        const int mt = ti.mt;
        const int z = ti.z;
        const int n0 = nt * p.BN;
        const float* s_bias = c.s_bias_all + n0;
        const float* s_cs = c.s_cs_all + n0;
        long long grow;
        bool valid;
        epi_row_of(p, ti, r, grow, valid);
        const TileIter tcur = ti;  // (unused fields fold away) advance now: `continue` below must not skip it
        ti = tn;
        tile_next(p, tn);
        tile_next(p, tn);
        (void)tcur;
        if (RES == 1 && p.mode == 0) {  // pull the residual rows of the NEXT tile of this set into L2
            if (ti.w < p.total) {
                const int m2 = ti.mt * kTileM + r;
                if (m2 < p.M) {
                    const bf16* pr = reinterpret_cast<const bf16*>(p.epi.res) +
                                     ((long long)ti.z * p.M + m2) * p.epi.ldr + ti.nt * p.BN;
                    const int nbytes = min(p.BN, p.N - ti.nt * p.BN) * 2;
                    for (int o = 0; o < nbytes; o += 128) prefetch_l2(reinterpret_cast<const uint8_t*>(pr) + o);
                    prefetch_l2(reinterpret_cast<const uint8_t*>(pr) + nbytes - 2);
                }
            }
        }
        float fa = 1.f, fb = 0.f;
        if (LN != 0 && p.epi.ln_stats) {
            fa = st_next.y;
            fb = (LN == 2) ? -st_next.x * st_next.y : 0.f;
            st_next = epi_load_stats(p, ti, r);  // ti is already the next tile of this set
        }
        const f32x2 a2 = splat2(fa), b2 = splat2(fb);
        f32x2 st_s1 = 0ull, st_s2 = 0ull;
        float st_shift = 0.f;
        const bf16* rrow = reinterpret_cast<const bf16*>(p.epi.res) + (valid ? grow : 0) * p.epi.ldr + n0;
        IRB_TS(1);
        GEMM_WAIT(&c.tfull_bar[set], (t >> 1) & 1);
        IRB_TS(2);
        tc_fence_after();
        const uint32_t taddr = c.tmem_base + set * p.acc_stride + (uint32_t(q * 32) << 16);
        if (p.dbg & 32) {  // diagnosis: no epilogue work at all (results are wrong on purpose)
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&c.tempty_bar[set]);
            continue;
        }
        const int ncols = min(p.BN, p.N - n0);  // valid columns of this tile
        uint32_t accb[2][16];  // TMEM load double buffer (16 columns each)
        bool pre = false;      // accb[0] already holds an in-flight load of this block's first group
        for (int blk = 0; blk < nblk_max; ++blk) {
            const int cb = blk * bcols;
            if (cb >= ncols) break;  // warp-uniform
            const int nvalid = min(bcols, ncols - cb);
            const uint32_t bsel = (RES == 2) ? nb % 3 : nb & 1;
            const uint32_t stg = stg_warp + bsel * 4096;
            const uint32_t sx = stg + row_off;
            if (RES == 2) GEMM_WAIT(&rbar[bsel], (nb / 3) & 1);
            // does the next block of this tile take a pipelined path too (so that its first TMEM load can be issued
            // from this block)?
            const int nv_next = min(bcols, ncols - (cb + bcols));
            const bool full_next = nv_next == 64 || nv_next == 32 || (nv_next == 48 && RES == 0);
            const uint32_t next_t = full_next ? taddr + cb + bcols : 0u;
            if (nvalid == 64 && RES == 1) {  // register-held residual: two halves keep the live set under the cap
                epi_block_full<ACT, LN, RES, STATS, 2>(taddr + cb, s_bias + cb, s_cs + cb, a2, b2, slope, rrow + cb, sx,
                                                       blk == 0, st_s1, st_s2, st_shift, true, lane, accb, pre,
                                                       taddr + cb + 32);
                epi_block_full<ACT, LN, RES, STATS, 2>(taddr + cb + 32, s_bias + cb + 32, s_cs + cb + 32, a2, b2, slope,
                                                       rrow + cb + 32, sx ^ 64u, false, st_s1, st_s2, st_shift, false, lane,
                                                       accb, true, next_t);
                pre = full_next;
            } else if (nvalid == 64) {
                epi_block_full<ACT, LN, RES, STATS, 4>(taddr + cb, s_bias + cb, s_cs + cb, a2, b2, slope, rrow + cb, sx,
                                                       blk == 0, st_s1, st_s2, st_shift, RES != 2, lane, accb, pre, next_t);
                pre = full_next;
            } else if (nvalid == 32) {
                epi_block_full<ACT, LN, RES, STATS, 2>(taddr + cb, s_bias + cb, s_cs + cb, a2, b2, slope, rrow + cb, sx,
                                                       blk == 0, st_s1, st_s2, st_shift, RES != 2, lane, accb, pre, next_t);
                pre = full_next;
            } else if (nvalid == 48 && RES == 0) {
                epi_block_full<ACT, LN, RES, STATS, 3>(taddr + cb, s_bias + cb, s_cs + cb, a2, b2, slope, rrow + cb, sx,
                                                       blk == 0, st_s1, st_s2, st_shift, true, lane, accb, pre, 0u);
                pre = false;
            } else {
                // ragged block (N % 32 != 0): runtime group count, element-wise tail group
                pre = false;
                const int ng = min(4, (nvalid + 15) >> 4);
                uint32_t acc[4][16];
#pragma unroll
                for (int gi = 0; gi < 4; ++gi)
                    if (gi < ng) tmem_ld16_nowait(taddr + cb + gi * 16, acc[gi]);
                uint4 rr[8];
                if (RES == 1) load_res_block(rrow + cb, ng, nvalid, rr);
                if (RES == 2) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) rr[k] = lds_v4(sx ^ (k << 4));
                }
                if (STATS && blk == 0) st_shift = __uint_as_float(rr[0].x << 16);
                if (RES != 2) {
                    bulk_wait_read1();
                    __syncwarp();
                }
                tmem_ld_wait();
                const bf16* rrag = rrow + cb;
                bf16 rtmp[16];
                if (RES == 2) {  // the ragged group's residual as an array (columns >= nvalid are zero-filled by TMA)
                    const int gl = nvalid >> 4;
                    const uint4 u0 = gl == 0 ? rr[0] : (gl == 1 ? rr[2] : (gl == 2 ? rr[4] : rr[6]));
                    const uint4 u1 = gl == 0 ? rr[1] : (gl == 1 ? rr[3] : (gl == 2 ? rr[5] : rr[7]));
                    *reinterpret_cast<uint4*>(rtmp) = u0;
                    *reinterpret_cast<uint4*>(rtmp + 8) = u1;
                    rrag = rtmp - (nvalid >> 4) * 16;
                }
                epi_block<ACT, LN, RES != 0, STATS>(acc, rr, ng, nvalid, s_bias + cb, s_cs + cb, fa, fb, slope, rrag,
                                                    stg_warp_p + bsel * 4096 + lane * 128, lane, st_s1, st_s2, st_shift);
            }
            IRB_TS(3);
            const bool last = cb + bcols >= ncols;
            if (last) {  // the accumulator is in registers / staged: hand it back to the MMA warp
                tc_fence_before();
            }
            fence_proxy_async();
            __syncwarp();
            if (leader) {
                if (last) mbar_arrive(&c.tempty_bar[set]);
                if (!(p.dbg & 1)) {
                    uint8_t* sp = stg_warp_p + bsel * 4096;
                    if (p.mode == 0) {
                        tma_store_3d(&p.tmC, sp, n0 + cb, mt * kTileM + q * 32, z);
                    } else {
                        const int r0 = q * 32;
                        const int sx0 = (mt % p.tiles_x) * p.BW + r0 % p.BW, sy0 = (mt / p.tiles_x) * p.BH + r0 / p.BW;
                        if (p.epi.ps == 3) {
                            const int cout = p.N >> 2, ncol = n0 + cb;
                            const int sub = ncol / cout;
                            tma_store_5d(&p.tmC, sp, ncol - sub * cout, sub & 1, sx0, 2 * sy0 + (sub >> 1), z);
                        } else {
                            tma_store_4d(&p.tmC, sp, n0 + cb, sx0, sy0, z);
                        }
                    }
                }
                bulk_commit();
            }
            IRB_TS(4);
            ++nb;
            if (RES == 2) {
                // block nb+1 goes into the staging block of block nb-2: its store must have finished reading
                bulk_wait_read1();
                __syncwarp();
                issue_res();
            }
        }
        if (STATS && valid) {
            float a0, a1, q0, q1;
            unpack2(st_s1, a0, a1);
            unpack2(st_s2, q0, q1);
            const float invn = 1.0f / float(p.N);
            const float m = (a0 + a1) * invn;
            const float var = fmaxf((q0 + q1) * invn - m * m, 0.f);
            p.epi.stats_out[grow] = make_float2(st_shift + m, rsqrtf(var + p.epi.stats_eps));
        }
        IRB_TS(5);
    }
#ifdef IRB_GEMM_TRACE
    if (trace) {
        static __device__ int printed;
        if (atomicAdd(&printed, 1) == 0)
            for (int i = 1; i < nts; ++i) printf("TS %d %d %lld\n", i, tsi[i], tsv[i] - tsv[i - 1]);
    }
#endif
    bulk_wait0();  // all bulk stores of this warp have landed before the CTA exits
}

template <int VAR>
__global__ void __launch_bounds__(kGemmThreads, 1) tc_gemm_kernel(const __grid_constant__ TcGemmParams p) {
    extern __shared__ uint8_t smem_raw[];
    // pointer arithmetic on the __shared__ symbol (not an integer round trip) keeps the address space known: LDS/STS
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    // b_resident: [weights: tiles_n x ktiles slabs of BN x 128 B][ring of A tiles]; a_hold: [a_hold x ktiles A slabs][ring
    // of B slabs]; otherwise ring of (A, B) stages
    const int slab_bytes = kABytes + p.BN * 128;   // one (A, B) slab pair of the plain ring
    const int halo_bytes = (p.BH + 2) * p.BW * 128;   // dyr: A box of BH + 2 pixel rows
    const int stage_bytes = p.dyr ? halo_bytes + 3 * p.BN * 128
                                  : (p.b_resident ? kABytes : (p.a_hold ? p.BN * 128 : p.kg * slab_bytes));
    uint8_t* ring = smem + p.bres_bytes;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(ring + p.stages * stage_bytes);
    uint64_t* empty_bar = full_bar + p.stages;
    uint64_t* tfull_bar = empty_bar + p.stages;   // [2] accumulator ready (MMA -> epilogue)
    uint64_t* tempty_bar = tfull_bar + 2;         // [2] accumulator drained (epilogue -> MMA)
    uint64_t* bres_full = tempty_bar + 2;         // resident weights loaded (TMA -> MMA)
    uint64_t* bres_empty = bres_full + 1;         // all MMAs that read the resident weights are done (MMA -> TMA)
    uint64_t* res_bar = bres_empty + 1;           // [8][3] residual blocks landed (TMA -> epilogue warp)
    uint64_t* ah_full = res_bar + 24;             // [2] held A tile loaded (TMA -> MMA)
    uint64_t* ah_empty = ah_full + 2;             // [2] all n-tiles of the held A tile are done (MMA -> TMA)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(ah_empty + 2);
    const int ncolv = p.tiles_n * p.BN;                        // padded column count
    float* colvec = reinterpret_cast<float*>(tmem_slot + 4);  // [bias | colsum][ncolv], 16-byte aligned, staged once
    // 8 epilogue warps x stg_per_warp x 4 KB staging blocks for TMA stores ([32 rows][64 bf16], 128B-swizzled), 1024-aligned;
    // double-buffered so that a warp never waits for the store it has just issued (wait_group.read 1)
    uint8_t* stg_lo = reinterpret_cast<uint8_t*>(colvec + 2 * ncolv);
    uint8_t* staging = stg_lo + ((1024u - (smem_u32(stg_lo) & 1023u)) & 1023u);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0) {
        if (lane == 0) {
            tma_prefetch_desc(&p.tmA);
            tma_prefetch_desc(&p.tmB);
            if (p.tma_store) tma_prefetch_desc(&p.tmC);
            for (int s = 0; s < p.stages; ++s) {
                mbar_init(&full_bar[s], 1);
                mbar_init(&empty_bar[s], 1);
            }
            for (int a = 0; a < 2; ++a) {
                mbar_init(&tfull_bar[a], 1);
                mbar_init(&tempty_bar[a], 4);
            }
            mbar_init(bres_full, 1);
            mbar_init(bres_empty, 1);
            for (int i = 0; i < 24; ++i) mbar_init(&res_bar[i], 1);
            for (int i = 0; i < 2; ++i) {
                mbar_init(&ah_full[i], 1);
                mbar_init(&ah_empty[i], 1);
            }
            if (VAR >= 0 && ((VAR >> 4) & 3) == 2) tma_prefetch_desc(&p.tmR);
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc(tmem_slot, p.tmem_cols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int total_it = p.dyr ? 3 * p.ktiles : p.taps * p.ktiles / p.kg;   // ring stages per tile
    // PDL: everything above overlapped the previous kernel's tail; its results are visible from here on
    griddep_wait();
    griddep_launch();

    if (warp == 0) {
        // ===== TMA producer: the whole warp walks the tiles, the elected lane issues (ptxas then knows the issue
        // is single-threaded and drops its per-instruction election loops) =====
        const bool leader = elect_one() != 0;
        int nb = 0, cur_z = -1, nhold = 0;
        int s = 0;         // ring slot and its phase, carried instead of it % stages, (it / stages) & 1
        uint32_t ph = 0;
        TileIter ti;
        for (tile_first(p, ti); ti.w < p.total; tile_next(p, ti)) {
            const int nt = ti.nt;
            const int mt = ti.mt;
            const int z = ti.z;
            const int n0 = nt * p.BN;
            if (p.b_resident && (cur_z < 0 || (p.b_rows_per_batch != 0 && z != cur_z))) {
                // (re)load the whole weight matrix; a new batch index only after the MMAs of the old one are done
                if (nb > 0) mbar_wait(bres_empty, (nb - 1) & 1);
                if (leader) {
                    mbar_arrive_expect_tx(bres_full, p.bres_bytes);
                    for (int t = 0; t < p.tiles_n; ++t)
                        for (int kc = 0; kc < p.ktiles; ++kc)
                            tma_load_2d(smem + (t * p.ktiles + kc) * p.BN * 128, &p.tmB, bres_full, kc * kTileK,
                                        z * p.b_rows_per_batch + t * p.BN);
                }
                __syncwarp();
                cur_z = z;
                ++nb;
            }
            if (p.a_hold && nt == 0) {
                // first n-tile of an m-tile (a CTA walks them in order): its A slabs go into a hold buffer once; the
                // other n-tiles re-use them instead of pulling them through the L2 fabric again
                const int hb = nhold % p.a_hold;
                GEMM_WAIT(&ah_empty[hb], ((nhold / p.a_hold) & 1) ^ 1);
                if (leader) {
                    mbar_arrive_expect_tx(&ah_full[hb], p.ktiles * kABytes);
                    for (int kc = 0; kc < p.ktiles; ++kc)
                        tma_load_3d(smem + (hb * p.ktiles + kc) * kABytes, &p.tmA, &ah_full[hb], kc * kTileK, mt * kTileM, z);
                }
                __syncwarp();
                ++nhold;
            }
            int x0 = 0, y0 = 0;
            if (p.mode == 1) {
                x0 = (mt % p.tiles_x) * p.BW;
                y0 = (mt / p.tiles_x) * p.BH;
            }
            if (p.dyr) {
                // stage (dx, kc): the A box covers pixel rows y0 - 1 .. y0 + BH of the dx-shifted columns; tap (dy, dx)
                // reads it from row dy * BW on (a 1024-byte multiple: plain descriptors), with its own weight slab
                for (int dx = 0; dx < 3; ++dx) {
                    for (int kc = 0; kc < p.ktiles; ++kc) {
                        GEMM_WAIT(&empty_bar[s], ph ^ 1);
                        if (leader) {
                            uint8_t* sa = ring + s * stage_bytes;
                            mbar_arrive_expect_tx(&full_bar[s], stage_bytes);
                            tma_load_4d(sa, &p.tmA, &full_bar[s], kc * kTileK, x0 + dx - 1, y0 - 1, z);
                            for (int dy = 0; dy < 3; ++dy)
                                tma_load_2d(sa + halo_bytes + dy * p.BN * 128, &p.tmB, &full_bar[s],
                                            (dy * 3 + dx) * p.kpad + kc * kTileK, n0);
                        }
                        __syncwarp();
                        if (++s == p.stages) {
                            s = 0;
                            ph ^= 1;
                        }
                    }
                }
                continue;
            }
            for (int tap = 0; tap < p.taps; ++tap) {
                const int dy = (p.taps == 9) ? tap / 3 - 1 : 0;
                const int dx = (p.taps == 9) ? tap % 3 - 1 : 0;
                for (int kc0 = 0; kc0 < p.ktiles; kc0 += p.kg) {
                    GEMM_WAIT(&empty_bar[s], ph ^ 1);
                    if (leader) {
                        if (p.dbg & 128) {  // diagnosis: no operand loads
                            mbar_arrive(&full_bar[s]);
                        } else {
                            mbar_arrive_expect_tx(&full_bar[s], stage_bytes);
                            for (int j = 0; j < p.kg; ++j) {  // kg > 1 only in the plain (A, B) ring
                                const int kc = kc0 + j;
                                uint8_t* sa = ring + s * stage_bytes + j * slab_bytes;
                                uint8_t* sb = sa + kABytes;
                                if (p.a_hold) {  // the stage is a B slab only
                                    tma_load_2d(sa, &p.tmB, &full_bar[s], kc * kTileK, z * p.b_rows_per_batch + n0);
                                } else {
                                    if (p.mode == 0)
                                        tma_load_3d(sa, &p.tmA, &full_bar[s], kc * kTileK, mt * kTileM, z);
                                    else
                                        tma_load_4d(sa, &p.tmA, &full_bar[s], kc * kTileK, x0 + dx, y0 + dy, z);
                                    if (!p.b_resident)
                                        tma_load_2d(sb, &p.tmB, &full_bar[s], tap * p.kpad + kc * kTileK, z * p.b_rows_per_batch + n0);
                                }
                            }
                        }
                    }
                    __syncwarp();
                    if (++s == p.stages) {
                        s = 0;
                        ph ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: the whole warp walks the tiles and polls the barriers (so the loop state and the
        // descriptors stay warp-uniform), lane 0 alone issues the tcgen05 instructions =====
        const bool leader = elect_one() != 0;  // elect.sync is deterministic for the full mask: always the same lane
        const uint32_t idesc = make_idesc_bf16(kTileM, p.BN, 0, 0);
#ifdef IRB_GEMM_TRACE  // timing experiment: clock() trace of the MMA issuer of CTA 0 (lines "TM")
        const bool mtrace = blockIdx.x == 0 && lane == 0 && p.N == IRB_GEMM_TRACE;
        long long mtv[200];
        int mti[200];
        int nmt = 0;
#define IRB_TM(i) do { if (mtrace && nmt < 200) { mtv[nmt] = clock64(); mti[nmt++] = (i); } } while (0)
#else
#define IRB_TM(i) do { } while (0)
#endif
        int t = 0, nb = 0, cur_z = -1, nhold = 0, cur_hb = 0;
        int s = 0;
        uint32_t ph = 0;
        // loop invariants of the issue loops, read from the constant bank once
        const int g_kg = p.kg, g_ktiles = p.ktiles, g_klast = p.klast, g_nst = p.stages;
        const bool g_hold = p.a_hold != 0, g_bres = p.b_resident != 0, g_nomma = (p.dbg & 64) != 0;
        const uint32_t g_bn128 = uint32_t(p.BN) * 128, ring_u32 = smem_u32(ring), hold_base = smem_u32(smem);
        TileIter ti;
        for (tile_first(p, ti); ti.w < p.total; tile_next(p, ti), ++t) {
            const int a = t & 1;
            if (p.a_hold && ti.nt == 0) {
                cur_hb = nhold % p.a_hold;
                GEMM_WAIT(&ah_full[cur_hb], (nhold / p.a_hold) & 1);
                ++nhold;
            }
            if (p.b_resident) {
                const int z = ti.z;
                if (cur_z < 0 || (p.b_rows_per_batch != 0 && z != cur_z)) {
                    if (nb > 0 && leader) umma_commit(bres_empty);  // every MMA issued so far has read the old weights
                    __syncwarp();
                    mbar_wait(bres_full, nb & 1);
                    cur_z = z;
                    ++nb;
                }
            }
            IRB_TM(10);
            GEMM_WAIT(&tempty_bar[a], ((t >> 1) & 1) ^ 1);
            IRB_TM(11);
            tc_fence_after();
            const uint32_t tacc = tmem_base + a * p.acc_stride;
            const uint32_t bres_tile = smem_u32(smem) + ti.nt * p.ktiles * p.BN * 128;
            int kc = 0;  // chunk within the tap
            if (p.dyr) {
                // lean issue loop: everything but the slot address is loop-invariant (the generic loop below spends
                // ~310 cycles per slab on constant-bank loads and register -> uniform moves, profiles/r02_s75)
                const uint32_t dstep = uint32_t(p.BW * 128) >> 4, bstep = uint32_t(p.BN * 128) >> 4;
                for (int i = 0; i < total_it; ++i) {
                    const int ksteps = (kc == g_ktiles - 1) ? g_klast : kTileK / 16;
                    if (++kc == g_ktiles) kc = 0;
                    IRB_TM(12);
                    GEMM_WAIT(&full_bar[s], ph);
                    IRB_TM(13);
                    tc_fence_after();
                    const uint32_t st_addr = ring_u32 + s * stage_bytes;
                    const uint64_t adesc0 = make_sdesc_sw128(st_addr, 16, 1024);
                    const uint64_t bdesc0 = make_sdesc_sw128(st_addr + halo_bytes, 16, 1024);
                    if (leader) {
#pragma unroll
                        for (int dy = 0; dy < 3; ++dy) {
#pragma unroll
                            for (int k = 0; k < kTileK / 16; ++k) {
                                if (k < ksteps)
                                    umma_bf16(tacc, adesc0 + uint64_t(dy * dstep + k * 2), bdesc0 + uint64_t(dy * bstep + k * 2),
                                              idesc, (i | dy | k) != 0 ? 1u : 0u);
                            }
                        }
                        umma_commit(&empty_bar[s]);
                    }
                    IRB_TM(15);
                    __syncwarp();
                    if (++s == g_nst) {
                        s = 0;
                        ph ^= 1;
                    }
                }
            } else {
                // same idea for the generic ring: the loop-invariant parameters live in registers (hoisted above the tile
                // loop), the operand addresses are one select per mode
                const uint32_t hold_a = hold_base + uint32_t(cur_hb * g_ktiles) * kABytes;
                for (int i = 0; i < total_it; ++i) {
                    IRB_TM(12);
                    GEMM_WAIT(&full_bar[s], ph);
                    IRB_TM(13);
                    tc_fence_after();
                    const uint32_t st_addr = ring_u32 + s * stage_bytes;
                    for (int j = 0; j < g_kg; ++j) {  // slabs of this stage (kg > 1 only in the plain (A, B) ring)
                        const int ksteps = (kc == g_ktiles - 1) ? g_klast : kTileK / 16;
                        if (++kc == g_ktiles) kc = 0;
                        const int gi = i * g_kg + j;
                        const uint32_t sl_addr = st_addr + j * slab_bytes;
                        const uint32_t a_addr = g_hold ? hold_a + gi * kABytes : sl_addr;
                        const uint32_t b_addr = g_bres ? bres_tile + gi * g_bn128 : (g_hold ? sl_addr : sl_addr + kABytes);
                        // one descriptor per operand and slab; the 32-byte K steps only move the 16-byte-unit address field
                        const uint64_t adesc0 = make_sdesc_sw128(a_addr, 16, 1024);
                        const uint64_t bdesc0 = make_sdesc_sw128(b_addr, 16, 1024);
                        if (leader) {
#pragma unroll
                            for (int k = 0; k < kTileK / 16; ++k) {
                                if (k < ksteps && !g_nomma)  // K steps past the data are zeros on both sides: skipped
                                    umma_bf16(tacc, adesc0 + uint64_t(k * 2), bdesc0 + uint64_t(k * 2), idesc, (gi | k) != 0 ? 1u : 0u);
                            }
                        }
                        IRB_TM(14);
                    }
                    if (leader) umma_commit(&empty_bar[s]);  // frees the smem slot once these MMAs have read it
                    IRB_TM(15);
                    __syncwarp();
                    IRB_TM(16);
                    if (++s == g_nst) {
                        s = 0;
                        ph ^= 1;
                    }
                }
            }
            IRB_TM(17);
            if (leader) {
                umma_commit(&tfull_bar[a]);  // accumulator complete
                if (p.a_hold && ti.nt == p.tiles_n - 1) umma_commit(&ah_empty[cur_hb]);  // the held A tile is done
            }
            IRB_TM(18);
            __syncwarp();
            IRB_TM(19);
        }
#ifdef IRB_GEMM_TRACE
        if (mtrace) {
            static __device__ int mprinted;
            if (atomicAdd(&mprinted, 1) == 0)
                for (int i = 1; i < nmt; ++i) printf("TM %d %d %lld\n", i, mti[i], mtv[i] - mtv[i - 1]);
        }
#endif
    } else if (warp >= 2) {
        // ===== epilogue: TMEM -> registers -> global. Two sets of 4 warps: set s drains accumulator s, i.e. the
        // tiles t of this CTA with (t & 1) == s, so two tiles are in the epilogue at once and the sets never wait
        // for each other. Warp (warp & 3) owns that TMEM lane quarter (rows q*32 .. q*32+31 of the tile). =====
        const int q = warp & 3;
        const int set = (warp - 2) >> 2;
        const int r = q * 32 + lane;
        const int ncol16 = p.BN >> 4;  // 16-column groups per tile
        const int et = threadIdx.x - 64;  // 0..255
        float* s_bias_all = colvec;
        float* s_cs_all = colvec + ncolv;
        for (int i = et; i < ncolv; i += 256) {
            s_bias_all[i] = (p.epi.bias && i < p.N) ? __ldg(p.epi.bias + i) : 0.f;
            s_cs_all[i] = (p.epi.ln_colsum && i < p.N) ? __ldg(p.epi.ln_colsum + i) : 0.f;
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if constexpr (VAR >= 0) {
            EpiCtx c{tmem_base, tfull_bar, tempty_bar, res_bar, s_bias_all, s_cs_all, staging, p.stg_per_warp};
            epilogue_fast<VAR & 3, (VAR >> 2) & 3, (VAR >> 4) & 3, ((VAR >> 6) & 1) != 0>(p, c, warp, lane);
        } else {
        // fast path: aligned vector access for 16-column groups, no pixel shuffle / post scale
        const bool vec_ok = p.epi.ps == 0 && p.epi.post_scale == 1.0f && p.epi.post_shift == nullptr &&
                            (p.epi.out_dt == DT_BF16 ? (p.epi.ldo % 8 == 0) : (p.epi.ldo % 4 == 0)) &&
                            (reinterpret_cast<uintptr_t>(p.epi.out) % 16 == 0) &&
                            (!p.epi.res || ((p.epi.res_dt == DT_BF16 ? (p.epi.ldr % 8 == 0) : (p.epi.ldr % 4 == 0)) &&
                                            reinterpret_cast<uintptr_t>(p.epi.res) % 16 == 0));
        const int lnmode = (p.dbg & 16) ? 0 : (p.epi.ln_colsum ? 2 : ((p.epi.ln_stats || p.epi.bias) ? 1 : 0));
        const bool res_bf16 = p.epi.res != nullptr && p.epi.res_dt == DT_BF16;
        const bool do_stats = p.epi.stats_out != nullptr;
        int t = 0;
        uint32_t sblk = 0;  // staging-block parity of this warp
        for (int w = first_item(p); w < p.total; w = next_item(w, p), ++t) {
            if ((t & 1) != set) continue;
            const int nt = w % p.tiles_n;
            const int mt = (w / p.tiles_n) % p.tiles_m;
            const int z = w / (p.tiles_n * p.tiles_m);
            const int n0 = nt * p.BN;
            const float* s_bias = s_bias_all + n0;
            const float* s_cs = s_cs_all + n0;
            long long grow;
            bool valid;
            int py = 0, px = 0;
            if (p.mode == 0) {
                const int m = mt * kTileM + r;
                valid = m < p.M;
                grow = (long long)z * p.M + m;
            } else {
                py = (mt / p.tiles_x) * p.BH + r / p.BW;
                px = (mt % p.tiles_x) * p.BW + r % p.BW;
                valid = (py < p.H) && (px < p.W);
                grow = ((long long)z * p.H + py) * p.W + px;
            }
            // The bf16 residual rows come from DRAM (the stream was written by the previous kernel) and their latency
            // sits in front of every block's math: pull the rows of the tile this warp set handles NEXT into L2 now.
            if (res_bf16 && p.mode == 0 && !(p.dbg & 8)) {
                const int w2 = next_item(next_item(w, p), p);
                if (w2 < p.total) {
                    const int m2 = ((w2 / p.tiles_n) % p.tiles_m) * kTileM + r;
                    if (m2 < p.M) {
                        const bf16* pr = reinterpret_cast<const bf16*>(p.epi.res) +
                                         ((long long)(w2 / (p.tiles_n * p.tiles_m)) * p.M + m2) * p.epi.ldr + (w2 % p.tiles_n) * p.BN;
                        const int nb = min(p.BN, p.N - (w2 % p.tiles_n) * p.BN) * 2;
                        for (int o = 0; o < nb; o += 128) prefetch_l2(reinterpret_cast<const uint8_t*>(pr) + o);
                        prefetch_l2(reinterpret_cast<const uint8_t*>(pr) + nb - 2);
                    }
                }
            }
            // the row's LayerNorm statistics do not depend on the accumulator: fetch them before waiting for it
            float fa = 1.f, fb = 0.f;
            if (p.epi.ln_stats && valid && !(p.dbg & 2)) {
                const float2 st = __ldg(p.epi.ln_stats + grow);
                fa = st.y;
                fb = p.epi.ln_colsum ? -st.x * st.y : 0.f;
            }
            f32x2 st_s1 = 0ull, st_s2 = 0ull;  // shifted sums for Epi::stats_out
            float st_shift = 0.f;
            mbar_wait(&tfull_bar[set], (t >> 1) & 1);
            tc_fence_after();
            const uint32_t taddr = tmem_base + set * p.acc_stride + (uint32_t(q * 32) << 16);
            if (p.tma_store) {
                // ---- 64-column blocks: TMEM -> registers (math) -> swizzled smem block -> TMA store ----
                const int nblk = (p.BN + 63) >> 6;
                uint8_t* stg_base = staging + (warp - 2) * p.stg_per_warp * 4096;
                const bool has_res = p.epi.res != nullptr && valid;
                const bool has_res_any = p.epi.res != nullptr;  // warp-uniform (rows past M read row 0, are clipped)
                for (int blk = 0; blk < nblk; ++blk) {
                    const int cb = blk * 64;
                    if (n0 + cb >= p.N) break;  // warp-uniform
                    const int ng = min(4, (p.BN - cb) >> 4);
                    uint8_t* stg = stg_base + (sblk & 1) * 4096;
                    ++sblk;
                    // fast path: no residual or a bf16 residual
                    if (res_bf16 || p.epi.res == nullptr) {
                        // all TMEM loads and residual loads of the block in flight before the first wait
                        uint32_t acc[4][16];
#pragma unroll
                        for (int gi = 0; gi < 4; ++gi)
                            if (gi < ng) tmem_ld16_nowait(taddr + cb + gi * 16, acc[gi]);
                        const int nvalid = min(64, p.N - (n0 + cb));
                        uint4 rr[8];
                        const bf16* rrow = reinterpret_cast<const bf16*>(p.epi.res) + (valid ? grow : 0) * p.epi.ldr + n0 + cb;
                        if (has_res_any && !(p.dbg & 4)) load_res_block(rrow, ng, nvalid, rr);
                        if (do_stats && blk == 0) st_shift = __uint_as_float(rr[0].x << 16);  // x_old[row, 0]
                        // the store issued two blocks ago (same staging buffer) must have finished READING it
                        if (lane == 0) bulk_wait_read1();
                        __syncwarp();
                        tmem_ld_wait();
                        const float* sb = s_bias + cb;
                        const float* sc = s_cs + cb;
                        uint8_t* srow = stg + lane * 128;
#define IRB_EPI(A, L, R, S) \
    epi_block<A, L, R, S>(acc, rr, ng, nvalid, sb, sc, fa, fb, p.epi.slope, rrow, srow, lane, st_s1, st_s2, st_shift)
#define IRB_EPI_LR(A)                                                  \
    switch (lnmode * 2 + (has_res_any ? 1 : 0)) {                      \
        case 0: IRB_EPI(A, 0, false, false); break;                    \
        case 1: IRB_EPI(A, 0, true, false); break;                     \
        case 2: IRB_EPI(A, 1, false, false); break;                    \
        case 3: IRB_EPI(A, 1, true, false); break;                     \
        case 4: IRB_EPI(A, 2, false, false); break;                    \
        default: IRB_EPI(A, 2, true, false); break;                    \
    }
                        if (do_stats) {  // ACT_NONE + residual only (checked on the host)
                            if (lnmode == 0) {
                                IRB_EPI(ACT_NONE, 0, true, true);
                            } else {
                                IRB_EPI(ACT_NONE, 1, true, true);
                            }
                        } else if (p.epi.act == ACT_GELU) {
                            IRB_EPI_LR(ACT_GELU)
                        } else if (p.epi.act == ACT_LRELU) {
                            IRB_EPI_LR(ACT_LRELU)
                        } else {
                            IRB_EPI_LR(ACT_NONE)
                        }
#undef IRB_EPI_LR
#undef IRB_EPI
                    } else {
                        if (lane == 0) bulk_wait_read1();
                        __syncwarp();
#pragma unroll
                        for (int gi = 0; gi < 4; ++gi) {
                            if (gi < ng) {
                                float v[16];
                                tmem_ld16(taddr + cb + gi * 16, v);
                                const int col0 = n0 + cb + gi * 16;
                                const int lr = has_res ? min(16, p.N - col0) : 0;
                                if (p.epi.act == ACT_GELU)
                                    epi_math16<ACT_GELU>(p.epi, grow, col0, lr, v, s_bias + cb + gi * 16, s_cs + cb + gi * 16, fa, fb);
                                else if (p.epi.act == ACT_LRELU)
                                    epi_math16<ACT_LRELU>(p.epi, grow, col0, lr, v, s_bias + cb + gi * 16, s_cs + cb + gi * 16, fa, fb);
                                else
                                    epi_math16<ACT_NONE>(p.epi, grow, col0, lr, v, s_bias + cb + gi * 16, s_cs + cb + gi * 16, fa, fb);
#pragma unroll
                                for (int c = 0; c < 2; ++c)
                                    *reinterpret_cast<uint4*>(stg + lane * 128 + (((2 * gi + c) ^ (lane & 7)) << 4)) =
                                        make_uint4(pack_bf16x2(v[c * 8 + 0], v[c * 8 + 1]), pack_bf16x2(v[c * 8 + 2], v[c * 8 + 3]),
                                                   pack_bf16x2(v[c * 8 + 4], v[c * 8 + 5]), pack_bf16x2(v[c * 8 + 6], v[c * 8 + 7]));
                            }
                        }
                    }
                    fence_proxy_async();
                    __syncwarp();
                    if (lane == 0) {
                        if (p.mode == 0) {
                            if (!(p.dbg & 1)) tma_store_3d(&p.tmC, stg, n0 + cb, mt * kTileM + q * 32, z);
                        } else {
                            const int r0 = q * 32;
                            const int sx = (mt % p.tiles_x) * p.BW + r0 % p.BW, sy = (mt / p.tiles_x) * p.BH + r0 / p.BW;
                            if (p.epi.ps == 3) {  // {c, j, x, 2y+i, image}: 32 LR pixels of one row -> every other HR pixel
                                const int cout = p.N >> 2, ncol = n0 + cb;
                                const int sub = ncol / cout;
                                tma_store_5d(&p.tmC, stg, ncol - sub * cout, sub & 1, sx, 2 * sy + (sub >> 1), z);
                            } else {
                                tma_store_4d(&p.tmC, stg, n0 + cb, sx, sy, z);
                            }
                        }
                        bulk_commit();
                    }
                }
            } else {
                for (int g = 0; g < ncol16; ++g) {
                    float v[16];
                    tmem_ld16(taddr + g * 16, v);
                    const int col0 = n0 + g * 16;
                    if (!valid || col0 >= p.N) continue;
                    if (vec_ok && col0 + 16 <= p.N) {
                        if (p.epi.act == ACT_GELU)
                            epi_fast16<ACT_GELU>(p.epi, grow, col0, v, s_bias + g * 16, s_cs + g * 16, fa, fb);
                        else if (p.epi.act == ACT_LRELU)
                            epi_fast16<ACT_LRELU>(p.epi, grow, col0, v, s_bias + g * 16, s_cs + g * 16, fa, fb);
                        else
                            epi_fast16<ACT_NONE>(p.epi, grow, col0, v, s_bias + g * 16, s_cs + g * 16, fa, fb);
                    } else {
                        epi_apply_store<16>(p.epi, grow, col0, p.N, v, z, py, px);
                    }
                }
            }
            if (do_stats && valid) {
                float a0, a1, q0, q1;
                unpack2(st_s1, a0, a1);
                unpack2(st_s2, q0, q1);
                const float invn = 1.0f / float(p.N);
                const float m = (a0 + a1) * invn;
                const float var = fmaxf((q0 + q1) * invn - m * m, 0.f);
                p.epi.stats_out[grow] = make_float2(st_shift + m, rsqrtf(var + p.epi.stats_eps));
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty_bar[set]);
        }
        if (p.tma_store && lane == 0) bulk_wait0();  // all bulk stores of this warp have landed before the CTA exits
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, p.tmem_cols);
}

// ----------------------------------------------------------------------------------------------
// host side
// ----------------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
    static PFN_encodeTiled fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<PFN_encodeTiled>(p);
    }
    return fn;
}

// dims/strides innermost first; strides[i] is the byte stride of dim i+1.
int make_tmap_bf16(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                   const uint32_t* box) {
    return make_tmap_bf16_sw(m, base, rank, dims, strides_bytes, box, 128);
}

// swizzle_bytes: 128 (UMMA operand tiles) or 0 (plain row-major box, read by ordinary shared-memory loads)
int make_tmap_bf16_sw(CUtensorMap* m, const void* base, int rank, const uint64_t* dims, const uint64_t* strides_bytes,
                      const uint32_t* box, int swizzle_bytes) {
    PFN_encodeTiled enc = get_encode();
    if (!enc) {
        set_error("cuTensorMapEncodeTiled entry point not available (driver too old?)");
        return 2;
    }
    cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    cuuint64_t d[5], s[5];
    cuuint32_t b[5];
    for (int i = 0; i < rank; ++i) {
        d[i] = dims[i];
        b[i] = box[i];
    }
    for (int i = 0; i + 1 < rank; ++i) s[i] = strides_bytes[i];
    // L2 promotion widens every TMA read to whole 64/128/256-byte L2 requests. Measured on cfg2 (profiles/
    // r01_s25_tma_promo.txt): 128 B and none are equal and ~1 % faster than 256 B, which drags neighbouring columns
    // out of DRAM when a box covers a slice of each row (v of a qkv buffer: 192 of 576 bytes, 1.67x over-read in ncu).
    // IRB_TMA_PROMO=0..3 forces none/64/128/256 (diagnosis).
    static int promo_env = -2;
    if (promo_env == -2) {
        promo_env = IRB_DIAG_ENV("IRB_TMA_PROMO", -1);
    }
    const int promo = promo_env >= 0 ? promo_env : 2;
    const CUtensorMapL2promotion pr = promo == 0 ? CU_TENSOR_MAP_L2_PROMOTION_NONE
                                      : promo == 1 ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B
                                      : promo == 2 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B
                                                   : CU_TENSOR_MAP_L2_PROMOTION_L2_256B;
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), d, s, b, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE,
                     swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                     pr,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        char buf[256];
        snprintf(buf, sizeof buf,
                 "cuTensorMapEncodeTiled failed (%d): rank=%d base=%p dims=[%llu,%llu,%llu,%llu] stride0=%llu box=[%u,%u,%u,%u]",
                 int(r), rank, base, (unsigned long long)d[0], (unsigned long long)(rank > 1 ? d[1] : 0),
                 (unsigned long long)(rank > 2 ? d[2] : 0), (unsigned long long)(rank > 3 ? d[3] : 0),
                 (unsigned long long)s[0], b[0], rank > 1 ? b[1] : 0, rank > 2 ? b[2] : 0, rank > 3 ? b[3] : 0);
        set_error(buf);
        return 2;
    }
    return 0;
}

static int pick_bn(int N) {
    // one tile when N fits; otherwise a multiple of 64 (the TMA-store block width) with the least padded work
    const int n16 = (N + 15) / 16 * 16;
    if (n16 <= 256) return n16;
    static int force = -1;  // IRB_GEMM_BN: tile width for N > 256 (diagnosis)
    if (force < 0) {
        force = IRB_DIAG_ENV("IRB_GEMM_BN", 0);
    }
    if (force >= 16 && force <= 256 && force % 16 == 0) return force;
    int best = 256, best_pad = ceil_div(N, 256) * 256;
    for (int bn : {192, 128}) {
        const int pad = ceil_div(N, bn) * bn;
        if (pad < best_pad) {
            best = bn;
            best_pad = pad;
        }
    }
    return best;
}

static bool can_tma_store(const Epi& e) {
    return e.out_dt == DT_BF16 && e.ps == 0 && e.post_scale == 1.0f && e.post_shift == nullptr && e.ldo % 8 == 0 &&
           (reinterpret_cast<uintptr_t>(e.out) & 15) == 0 &&
           (!e.res || ((e.res_dt == DT_BF16 ? e.ldr % 8 == 0 : e.ldr % 4 == 0) &&
                       (reinterpret_cast<uintptr_t>(e.res) & 15) == 0));
}

bool tc_can_fuse_stats(int N, const Epi& e) {
    return N <= 256 && can_tma_store(e) && e.res != nullptr && e.res_dt == DT_BF16 && e.act == ACT_NONE &&
           e.ln_stats == nullptr;
}

static int tmem_cols_for(int bn) {
    int c = 32;
    while (c < bn) c <<= 1;
    return c;
}

static int sm_count() {
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (sms <= 0) sms = 148;
    }
    return sms;
}

typedef void (*GemmKernelFn)(const TcGemmParams);

static int gemm_spec_on() {  // IRB_GEMM_SPEC=0: generic kernel only (diagnosis)
    static int on = -1;
    if (on < 0) {
        on = IRB_DIAG_ENV("IRB_GEMM_SPEC", 1);
    }
    return on;
}

// Specialised epilogue variants that exist (VAR = act | ln << 2 | res << 4 | stats << 6); anything else runs the
// generic kernel (VAR -1), which also covers fp32 outputs / residuals, PixelShuffle(1) / PixelUnshuffle stores.
#define IRB_GEMM_VARIANTS(X)                                                                                    \
    X(0, 0, 0, 0) X(0, 1, 0, 0) X(0, 2, 0, 0) X(0, 0, 1, 0) X(0, 1, 1, 0) X(0, 2, 1, 0) X(0, 0, 2, 0) X(0, 1, 2, 0) \
    X(0, 2, 2, 0) X(1, 0, 0, 0) X(1, 1, 0, 0) X(1, 2, 0, 0) X(2, 0, 0, 0) X(2, 1, 0, 0) X(0, 0, 1, 1) X(0, 1, 1, 1) \
    X(0, 0, 2, 1) X(0, 1, 2, 1) X(3, 0, 0, 0) X(3, 1, 0, 0) X(3, 2, 0, 0)

static GemmKernelFn pick_gemm_kernel(int& var, int* slot) {
    int i = 1;
#define X(A, L, R, S)                                             \
    if (var == ((A) | ((L) << 2) | ((R) << 4) | ((S) << 6))) {    \
        *slot = i;                                                \
        return tc_gemm_kernel<(A) | ((L) << 2) | ((R) << 4) | ((S) << 6)>; \
    }                                                             \
    ++i;
    IRB_GEMM_VARIANTS(X)
#undef X
    var = -1;
    *slot = 0;
    return tc_gemm_kernel<-1>;
}

// does the (act, ln, TMA-staged residual, stats) combination of this epilogue have a specialised kernel?
static bool gemm_variant_exists(const Epi& e) {
    if (!gemm_spec_on()) return false;
    const int ln = e.ln_colsum ? 2 : ((e.ln_stats || e.bias) ? 1 : 0);
    int var = e.act | (ln << 2) | (2 << 4) | (e.stats_out ? 64 : 0), slot = 0;
    pick_gemm_kernel(var, &slot);
    return var >= 0;
}

// shared-memory bytes of everything but the operand buffers
static size_t gemm_fixed_smem(int tiles_n, int BN, int stg_per_warp) {
    return 1024 /*align*/ + 512 /*barriers*/ + size_t(2) * tiles_n * BN * 4 /*column vectors*/ + 1024 +
           8 * size_t(stg_per_warp) * 4096 /*TMA-store staging*/;
}
static bool gemm_m_outer(long long m_units, int tiles_n) {
    static int order = -2;
    if (order == -2) {
        order = IRB_DIAG_ENV("IRB_GEMM_ORDER", -1);
    }
    return order >= 0 ? order != 0 : !(tiles_n > 1 && m_units < 8ll * sm_count());
}
// A-hold plan of a 1x1 GEMM whose weights do not fit in shared memory (SwinIR's qkv / fc layers: K = 180 / 240, N = 576 /
// 768): with (A, B) stages every (m, n) tile pulls its A slabs AND its B slabs through the L2 fabric, 885 MB per launch of
// the [180 -> 576] GEMM against 94 MB of unique A (ncu l1tex__m_xbar2l1tex_read_bytes, profiles/r02_s52), at 7.6 TB/s
// the binding resource; holding the m-tile's A for all its n-tiles removes the A re-reads. Returns the number of hold
// buffers (0 = none).
static int plan_a_hold(int M, int K, int N, int batch, int stg_per_warp, int* BN) {
    static int on = -1;
    if (on < 0) on = IRB_DIAG_ENV("IRB_GEMM_AHOLD", 1);
    if (!on) return 0;
    const int ktiles = ceil_div(K, kTileK);
    const int bn0 = *BN;
    if (ceil_div(N, bn0) < 2) return 0;
    if (!gemm_m_outer((long long)batch * ceil_div(M, kTileM), ceil_div(N, bn0))) return 0;
    {   // weights resident? then the ring carries A only anyway
        const size_t bres = size_t(ceil_div(N, bn0)) * ktiles * bn0 * 128;
        if (gemm_fixed_smem(ceil_div(N, bn0), bn0, stg_per_warp) + bres + 3 * size_t(kABytes) <= 227 * 1024) return 0;
    }
    // Measured (same-box A/B, profiles/r02_s54 / r02_s55): the [180 -> 576] GEMM of SwinIR-M / HAT with ONE hold buffer and
    // the planned BN = 192: 4.60 -> 4.25 ms per 36 launches (cfg3 tile: all GEMMs 14.35 -> 13.45 ms). Narrowing BN to make
    // room loses more than the hold gains (BN = 128, two buffers: 5.07 ms; SwinIR-L's [240 -> 768] at BN = 128: cfg4 GEMMs
    // 36.1 -> 40.1 ms), and with K <= 128 (Restormer: two K slabs, little A to re-read) the hold is a small loss
    // (cfg2 GEMMs 17.9 -> 18.3 ms). So: the planned width only, three or more K slabs.
    if (ktiles < 3 || bn0 % 64 != 0) return 0;
    for (int nh : {2, 1}) {
        const size_t need = gemm_fixed_smem(ceil_div(N, bn0), bn0, stg_per_warp) + size_t(nh) * ktiles * kABytes + 3 * size_t(bn0) * 128;
        if (need <= 227 * 1024) return nh;
    }
    return 0;
}

static int launch_tc(TcGemmParams& p, int batch, cudaStream_t stream) {
    const int stage_bytes = kABytes + p.BN * 128;
    p.acc_stride = (p.BN + 31) / 32 * 32;
    p.tmem_cols = tmem_cols_for(2 * p.acc_stride);
    p.tiles_n = ceil_div(p.N, p.BN);
    p.stg_per_warp = p.res_tma ? 3 : 2;
    if (p.blk_cols == 0) p.blk_cols = 64;
    const size_t fixed = gemm_fixed_smem(p.tiles_n, p.BN, p.stg_per_warp);
    IRB_REQUIRE(fixed + 2 * size_t(stage_bytes) <= 227 * 1024, "tc_gemm: N too large for the shared-memory column vectors");
    int stages = int((227 * 1024 - fixed) / stage_bytes);
    if (stages > 8) stages = 8;
    // 1x1 GEMMs whose whole weight matrix fits next to >= 3 A stages keep it resident: otherwise every CTA re-reads
    // the same N x K weights from L2 for each of its tiles (3x the A traffic at K = 96, all SMs on the same lines)
    static int bres_on = -1;
    if (bres_on < 0) {
        bres_on = IRB_DIAG_ENV("IRB_GEMM_BRES", 1);
    }
    static int dbg = -1;
    if (dbg < 0) {
        dbg = IRB_DIAG_ENV("IRB_GEMM_DBG", 0);
    }
    p.dbg = dbg;
    p.b_resident = 0;
    p.bres_bytes = 0;
    size_t ring_stage = stage_bytes;
    if (bres_on && p.mode == 0) {
        const size_t bres = size_t(p.tiles_n) * p.ktiles * p.BN * 128;
        if (fixed + bres + 3 * size_t(kABytes) <= 227 * 1024) {
            p.b_resident = 1;
            p.bres_bytes = int(bres);
            ring_stage = kABytes;
            stages = int((227 * 1024 - fixed - bres) / kABytes);
            if (stages > 8) stages = 8;
        }
    }
    const long long m_units = (long long)batch * p.tiles_m;
    p.m_outer = gemm_m_outer(m_units, p.tiles_n) ? 1 : 0;
    if (p.a_hold) {  // planned by tc_gemm (plan_a_hold): re-checked against the final layout
        const size_t hold = size_t(p.a_hold) * p.ktiles * kABytes, bst = size_t(p.BN) * 128;
        if (p.mode == 0 && !p.b_resident && p.m_outer && p.tiles_n >= 2 && fixed + hold + 3 * bst <= 227 * 1024) {
            p.bres_bytes = int(hold);
            ring_stage = bst;
            stages = int((227 * 1024 - fixed - hold) / bst);
            if (stages > 8) stages = 8;
        } else {
            p.a_hold = 0;
        }
    }
    // Narrow tiles: one ring handshake of the MMA issuer (full wait, descriptors, commit, empty wait on the producer side)
    // costs 150 - 200 cycles, a 64-column slab is only 4 x 32 = 128 cycles of tensor-core time: HAT's [180 -> 60] CAB
    // convolution ran at the SAME speed without its operand loads and without its MMAs (IRB_GEMM_DBG 128 / 64,
    // profiles/r02_s72): 27 handshakes per tile bound it. Such tiles put kg K slabs into one stage (the smallest divisor
    // of ktiles that gives a stage >= 384 tensor-core cycles and still leaves two stages).
    if (p.dyr) {  // planned by tc_conv3x3 (narrow tiles only): A halo box + three weight slabs per stage
        ring_stage = size_t(p.BH + 2) * p.BW * 128 + 3 * size_t(p.BN) * 128;
        stages = int((227 * 1024 - fixed) / ring_stage);
        IRB_REQUIRE(stages >= 2, "tc_conv3x3: dy-reuse stage does not fit");
        if (stages > 8) stages = 8;
    }
    p.kg = 1;
    static int kg_on = -1;
    if (kg_on < 0) kg_on = IRB_DIAG_ENV("IRB_GEMM_KG", 1);
    if (kg_on && !p.dyr && !p.b_resident && !p.a_hold && p.ktiles > 1 && p.BN <= (kg_on >= 2 ? 128 : 64)) {
        for (int kg = 2; kg <= p.ktiles; ++kg) {
            if (p.ktiles % kg) continue;
            const int st = int((227 * 1024 - fixed) / (size_t(kg) * stage_bytes));
            if (st < 2) break;
            p.kg = kg;
            stages = st > 8 ? 8 : st;
            ring_stage = size_t(kg) * stage_bytes;
            if (2 * p.BN * kg >= 384) break;
        }
    }
    p.stages = stages;
    const long long total = (long long)batch * p.tiles_m * p.tiles_n;
    IRB_REQUIRE(total < (1ll << 31), "tc_gemm: too many tiles");
    p.total = int(total);
    const size_t smem = size_t(stages) * ring_stage + p.bres_bytes + fixed;
    int var = -1;
    const int spec_on = gemm_spec_on();
    const bool res_bf16 = p.epi.res != nullptr && p.epi.res_dt == DT_BF16;
    if (spec_on && p.tma_store && (p.epi.res == nullptr || res_bf16)) {
        const int ln = p.epi.ln_colsum ? 2 : ((p.epi.ln_stats || p.epi.bias) ? 1 : 0);
        const int res = p.epi.res == nullptr ? 0 : (p.res_tma ? 2 : 1);
        static int gelu_tanh = -1;
        if (gelu_tanh < 0) {
            gelu_tanh = IRB_DIAG_ENV("IRB_GELU_TANH", 1);
        }
        const int act = (p.epi.act == ACT_GELU && gelu_tanh && p.epi.out_dt == DT_BF16) ? kActGeluTanh : p.epi.act;
        var = act | (ln << 2) | (res << 4) | (p.epi.stats_out ? 64 : 0);
    }
    int slot = 0;
    GemmKernelFn fn = pick_gemm_kernel(var, &slot);
    IRB_REQUIRE(var >= 0 || (!p.res_tma && p.blk_cols == 64), "tc_gemm: internal: res_tma / 48-column blocks need a specialised variant");
    IRB_SMEM_OPTIN((fn), 227 * 1024);
    // Order 1 (a CTA walks all n-tiles of an m-tile) re-reads A from L2 back to back, but hands out whole m-tiles: with
    // few m-tiles per CTA (the deep Restormer levels: 192 m-tiles on 148 SMs) a third of the SMs idle in the second
    // wave. Then the (m, n) items are dealt round-robin instead (order 0, gemm_m_outer()). IRB_GEMM_ORDER=0/1 forces one.
    const long long units = p.m_outer ? m_units : total;
    const int grid = int(units < sm_count() ? units : sm_count());
    p.step_nt = grid % p.tiles_n;
    p.step_units = grid / p.tiles_n;
    IRB_CHECK_CUDA(launch_pdl(fn, dim3(grid), dim3(kGemmThreads), smem, stream, p));
    return 0;
}

// C[b, m, :] = epi(A[b, m, :K] . Wt[b*w_rows_per_batch + n, :K])
int tc_gemm(const bf16* A, long long lda, int M, int K, int batch, long long a_batch_stride, const bf16* Wt, int ldw,
            int N, int w_rows_total, int w_rows_per_batch, const Epi& epi, cudaStream_t stream) {
    IRB_REQUIRE((lda % 8) == 0 && (ldw % 8) == 0, "tc_gemm: row strides must be multiples of 8 bf16 (16 B)");
    IRB_REQUIRE((reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(Wt) & 15) == 0,
                "tc_gemm: operands must be 16-byte aligned");
    IRB_REQUIRE(M > 0 && K > 0 && N > 0 && batch > 0, "tc_gemm: empty problem");
    TcGemmParams p{};
    {
        uint64_t dims[3] = {uint64_t(K), uint64_t(M), uint64_t(batch)};
        uint64_t str[2] = {uint64_t(lda) * 2, uint64_t(a_batch_stride ? a_batch_stride : (long long)M * lda) * 2};
        uint32_t box[3] = {kTileK, kTileM, 1};
        IRB_TRY(make_tmap_bf16(&p.tmA, A, 3, dims, str, box));
    }
    p.mode = 0;
    p.M = M;
    p.N = N;
    p.BN = pick_bn(N);
    p.a_hold = plan_a_hold(M, K, N, batch, epi.res != nullptr ? 3 : 2, &p.BN);
    p.ktiles = ceil_div(K, kTileK);
    p.klast = ceil_div(K - (p.ktiles - 1) * kTileK, 16);
    p.taps = 1;
    p.kpad = 0;
    p.b_rows_per_batch = w_rows_per_batch;
    {
        uint64_t dims[2] = {uint64_t(K), uint64_t(w_rows_total)};
        uint64_t str[1] = {uint64_t(ldw) * 2};
        uint32_t box[2] = {kTileK, uint32_t(p.BN)};
        IRB_TRY(make_tmap_bf16(&p.tmB, Wt, 2, dims, str, box));
    }
    p.epi = epi;
    IRB_REQUIRE(!epi.stats_out || tc_can_fuse_stats(N, epi), "tc_gemm: stats_out needs N <= 256, bf16 residual, TMA-store epilogue");
    p.tma_store = can_tma_store(epi) ? 1 : 0;
    if (p.tma_store) {
        uint64_t dims[3] = {uint64_t(N), uint64_t(M), uint64_t(batch)};
        uint64_t str[2] = {uint64_t(epi.ldo) * 2, uint64_t(epi.ldo) * 2 * M};
        uint32_t box[3] = {64, 32, 1};
        IRB_TRY(make_tmap_bf16(&p.tmC, epi.out, 3, dims, str, box));
        static int restma_on = -1;
        if (restma_on < 0) {
            restma_on = IRB_DIAG_ENV("IRB_GEMM_RESTMA", 1);
        }
        if (restma_on && epi.res != nullptr && epi.res_dt == DT_BF16 && gemm_variant_exists(epi)) {
            uint64_t rstr[2] = {uint64_t(epi.ldr) * 2, uint64_t(epi.ldr) * 2 * M};
            IRB_TRY(make_tmap_bf16(&p.tmR, epi.res, 3, dims, rstr, box));
            p.res_tma = 1;
        }
    }
    p.tiles_m = ceil_div(M, kTileM);
    IRB_TRY(launch_tc(p, batch, stream));
    {
        const double rows = double(M) * batch;
        const double wrows = w_rows_per_batch ? double(N) * batch : double(N);
        prof_mark("tc_gemm", K, N, 2.0 * rows * K * N, rows * K * 2 + wrows * K * 2 + epi_bytes(epi, rows, N), stream);
    }
    return 0;
}

// 3x3 convolution, stride 1, zero pad 1, channels-last. Wt is [Nrows, 9*kpad] with K index = tap*kpad + c.
int tc_conv3x3(const bf16* A, long long lda, int Nimg, int H, int W, int C, const bf16* Wt, int kpad, int N,
               int w_rows_total, const Epi& epi, cudaStream_t stream) {
    IRB_REQUIRE((lda % 8) == 0 && (kpad % 64) == 0, "tc_conv3x3: lda %8, kpad %64");
    IRB_REQUIRE((reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(Wt) & 15) == 0,
                "tc_conv3x3: operands must be 16-byte aligned");
    IRB_REQUIRE(kpad >= C, "tc_conv3x3: kpad < C");
    TcGemmParams p{};
    // pixel box of 128 pixels: the shape that covers the image with the fewest boxes (a 544-wide HAT tile takes five
    // 128 x 1 boxes per row, 15 % of them padding, but exactly 17 x 136 boxes of 32 x 4), the widest one on a tie
    int BW = 128;
    while (BW > 8 && BW > W) BW >>= 1;
    if (BW > W) BW = 8;
    static int shape_on = -1;
    if (shape_on < 0) shape_on = IRB_DIAG_ENV("IRB_CONV_SHAPE", 1);
    if (shape_on) {
        long long best = (long long)ceil_div(W, BW) * ceil_div(H, 128 / BW);
        const int bw_min = epi.ps == 3 ? 32 : 8;   // the PixelShuffle TMA store wants >= 32 pixels of one row
        for (int bw = BW >> 1; bw >= bw_min; bw >>= 1) {
            const long long n = (long long)ceil_div(W, bw) * ceil_div(H, 128 / bw);
            if (n < best) {
                best = n;
                BW = bw;
            }
        }
    }
    // Narrow outputs (BN <= 64: HAT's / SwinIR's C -> C / 3 compress convolutions): a 16 x 8 pixel box, and ONE A box of
    // 16 x 10 pixels per (dx, K slab) that serves the three dy taps (p.dyr): 396 KB instead of 648 KB through the L2
    // fabric per tile at K = 180, which at 10 TB/s was one of the two bounds of that convolution (profiles/r02_s72-s76)
    static int dyr_on = -1;
    if (dyr_on < 0) dyr_on = IRB_DIAG_ENV("IRB_CONV_DYR", 1);
    p.dyr = (dyr_on && pick_bn(N) <= 64 && W >= 16 && H >= 8 && epi.ps != 3) ? 1 : 0;   // not with the 5-D PixelShuffle TMA store (needs >= 32 pixels of a row)
    if (p.dyr) BW = 16;
    const int BH = 128 / BW;
    {
        uint64_t dims[4] = {uint64_t(C), uint64_t(W), uint64_t(H), uint64_t(Nimg)};
        uint64_t str[3] = {uint64_t(lda) * 2, uint64_t(lda) * 2 * W, uint64_t(lda) * 2 * W * H};
        uint32_t box[4] = {kTileK, uint32_t(BW), uint32_t(p.dyr ? BH + 2 : BH), 1};
        IRB_TRY(make_tmap_bf16(&p.tmA, A, 4, dims, str, box));
    }
    p.mode = 1;
    p.N = N;
    p.BN = pick_bn(N);
    p.ktiles = kpad / kTileK;
    {
        const int rem = C - (p.ktiles - 1) * kTileK;  // channels in the last chunk of a tap
        p.klast = (rem >= 1 && rem <= kTileK) ? ceil_div(rem, 16) : kTileK / 16;
    }
    p.taps = 9;
    p.kpad = kpad;
    p.H = H;
    p.W = W;
    p.BW = BW;
    p.BH = BH;
    p.tiles_x = ceil_div(W, BW);
    p.b_rows_per_batch = 0;
    {
        uint64_t dims[2] = {uint64_t(9 * kpad), uint64_t(w_rows_total)};
        uint64_t str[1] = {uint64_t(9 * kpad) * 2};
        uint32_t box[2] = {kTileK, uint32_t(p.BN)};
        IRB_TRY(make_tmap_bf16(&p.tmB, Wt, 2, dims, str, box));
    }
    p.epi = epi;
    IRB_REQUIRE(!epi.stats_out || tc_can_fuse_stats(N, epi), "tc_conv3x3: stats_out needs N <= 256, bf16 residual, TMA-store epilogue");
    p.tma_store = can_tma_store(epi) ? 1 : 0;
    if (epi.ps == 3) {
        IRB_REQUIRE(N % 4 == 0, "tc_conv3x3: PixelShuffle needs N % 4 == 0");
        Epi e0 = epi;
        e0.ps = 0;
        // 64-column blocks must not straddle sub-pixels: N/4 a multiple of 64, or of 48 with 48-column blocks (the
        // store box stays 64 wide; the map's channel extent N/4 clips columns 48..63 when N/4 == 48, and for
        // N/4 == 96 the second block of a sub-pixel starts at channel 48 and is clipped at 96)
        const int cout = N / 4;
        const int bc = cout % 64 == 0 ? 64 : ((cout % 48 == 0 && p.BN % 48 == 0 && gemm_spec_on() && epi.act == ACT_NONE) ? 48 : 0);
        if (can_tma_store(e0) && BW >= 32 && bc != 0 && epi.res == nullptr) {
            p.tma_store = 1;
            p.blk_cols = bc;
            uint64_t dims[5] = {uint64_t(cout), 2, uint64_t(W), uint64_t(2 * H), uint64_t(Nimg)};
            uint64_t str[4] = {uint64_t(epi.ldo) * 2, uint64_t(epi.ldo) * 4, uint64_t(epi.ldo) * 2 * 2 * W,
                               uint64_t(epi.ldo) * 2 * 2 * W * 2 * H};
            uint32_t box[5] = {64, 1, 32, 1, 1};
            IRB_TRY(make_tmap_bf16(&p.tmC, epi.out, 5, dims, str, box));
        }
    } else if (p.tma_store) {
        const int bw = BW < 32 ? BW : 32;
        uint64_t dims[4] = {uint64_t(N), uint64_t(W), uint64_t(H), uint64_t(Nimg)};
        uint64_t str[3] = {uint64_t(epi.ldo) * 2, uint64_t(epi.ldo) * 2 * W, uint64_t(epi.ldo) * 2 * W * H};
        uint32_t box[4] = {64, uint32_t(bw), uint32_t(32 / bw), 1};
        IRB_TRY(make_tmap_bf16(&p.tmC, epi.out, 4, dims, str, box));
    }
    p.tiles_m = p.tiles_x * ceil_div(H, BH);
    IRB_TRY(launch_tc(p, Nimg, stream));
    {
        const double rows = double(H) * W * Nimg;
        prof_mark("tc_conv3x3", C, N, 18.0 * rows * C * N, rows * C * 2 + 9.0 * N * C * 2 + epi_bytes(epi, rows, N), stream);
    }
    return 0;
}

}  // namespace irb
