// SwinIR's MLP half-block as ONE kernel, bf16 storage, sm_100a (network_swinir.py:24-30 Mlp.forward with the block's
// norm2 and residual, :277):
//
//     x[p, :] <- x[p, :] + W2 . gelu(W1' . LN(x[p, :]) + b1') + b2        in place, plus (mu, rstd) of the new row
//
// The `hidden` tensor (2C = 360 channels, 720 B per token written and read back by the two-GEMM path: 46 % of the two
// GEMMs' HBM bytes) never leaves the SM. Unlike Restormer's depthwise pairs (pwdw_fused.cu) there is no CUDA-core
// stencil between the two contractions, only the GELU, so the fusion removes bytes without loading the issue slots.
//
// Work item = 128 tokens (one UMMA M); persistent CTAs, one per SM, 19 warps:
//   warp 16 (TMA)  : x tile [128][C] -> A (K-major, 128B swizzle, 2 buffers); per 64-column group g of the hidden
//                    layer: W1' rows [64g, 64g+64) -> ring (2), W2 columns [64g, 64g+64) -> ring (2). Weights stream
//                    from L2 for every tile (276 KB per 128 tokens at C = 180; they do not fit next to the tiles).
//   warp 17 (MMA)  : D1[g & 1] (TMEM [128 x 64]) = A . W1'_g^T ; D2[tile & 1] (TMEM [128 x Npad]) += A2_g . W2_g^T with
//                    A2_g read FROM TENSOR MEMORY (tcgen05.mma with a TMEM A operand); a polling issue loop over two
//                    cursors, the first GEMM at most one group ahead of the second (A2 sits in D1's columns)
//   warps 0-15     : convert: D1 -> folded LayerNorm (rstd * acc + (-mu rstd) * colsum + b1') -> tanh-form GELU -> bf16
//                    pairs written back IN PLACE over the first 8 of the thread's 16 accumulator columns (tcgen05.st):
//                    the GELU output never touches shared memory. Warp w: TMEM lane quarter w & 3, columns
//                    16 (w >> 2) .. +15 = one K = 16 step of the second GEMM.
//                    epilogue: D2 -> + b2 + residual (read from the A tile that is still in shared memory: the residual
//                    IS the GEMM input) -> LayerNorm statistics of the new row -> bf16 written IN PLACE into the A tile
//   warp 18 (store): TMA-stores the A tile back to x (same tensor map as the load), then frees the A buffer
// Measured per tile with the clock trace (-DIRB_MLP_DBG=5): six groups of ~1 450 cycles (D1 wait 240, TMEM load 240,
// GELU + TMEM store 650, the rest barriers) and an epilogue of ~4 800 (D2 wait 1 100, 98 KB of TMEM reads + residual +
// statistics 3 000, barriers 600).
#include "common.cuh"
#include "ptx.cuh"
#include "kernels.h"

#if defined(IRB_MLP_DBG) && IRB_MLP_DBG == 5  // timing experiment: clock() trace of warp 0 / lane 0 of CTA 0, printed at exit
#include <cstdio>
#define IRB_TS(i) do { if (trace && nts < 160) { tsv[nts] = clock64(); tsi[nts++] = (i); } } while (0)
#else
#define IRB_TS(i) do { } while (0)
#endif

namespace irb {

static constexpr int kMfWarps = 16;                      // conversion / epilogue warps: 4 TMEM lane quarters x 4 column groups
static constexpr int kMfCG = kMfWarps / 4;               // column groups
static constexpr int kMfThreads = (kMfWarps + 3) * 32;  // 608
static constexpr int kMfATile = 128 * 128;               // one 64-channel K chunk of a 128-token tile (16 KB)
static constexpr int kMfW1Slab = 64 * 128;               // 64 hidden rows x 64 K (8 KB)
static constexpr int kMfAux = 16 * 1024;                 // per-column vectors + LayerNorm partial sums

struct alignas(64) MlpParams {
    CUtensorMap tmX;   // x {C, P, 1}, box {64, 128, 1}: load and in-place store
    CUtensorMap tmW1;  // W1' {C, hid, 1}, box {64, 64, 1}
    CUtensorMap tmW2;  // W2 {hid, C, 1}, box {64, NB, 1}
    const float2* stats_in;  // [P] (mu, rstd) of the rows of x (norm2)
    float2* stats_out;       // [P] (mu, rstd) of the new rows, or nullptr
    const float* cs1;        // [hid] colsum of the folded W1'
    const float* b1;         // [hid] fc1 bias + W1 . ln_bias
    const float* b2;         // [C]
    int P, C, hid, KS, klast, NG, NB;  // KS = K chunks of 64 in C; NG = hidden groups of 64; NB = C rounded up to 16
    int tiles;
    float eps;
};

struct MfBars {
    uint64_t a_full[2], a_empty[2], w1_full[2], w1_empty[2], w2_full[2], w2_empty[2], d1_full[2], d1_empty[2],
        a2_full[2], a2_empty[2], d2_full[2], d2_empty[2], out_ready[2];
    uint32_t tmem_slot;
};

__global__ void __launch_bounds__(kMfThreads, 1) mlp_fused_kernel(const __grid_constant__ MlpParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int KS = p.KS, NG = p.NG;
    uint8_t* sA = smem;                                   // 2 x KS x 16 KB
    uint8_t* sW1 = sA + 2 * KS * kMfATile;                // 2 x KS x 8 KB
    uint8_t* sAux = sW1 + 2 * KS * kMfW1Slab;             // 16 KB
    uint8_t* sW2 = sAux + kMfAux;                         // 2 x NB x 128 B
    const int w2_bytes = p.NB * 128;
    uint8_t* sRed = sAux;  // [kMfCG - 1][128] float2 partial LayerNorm sums (3 KB)
    // The per-column vectors (colsum | b1' padded to NG * 64, then b2 padded to NB) live in shared memory. Read through
    // __ldg they cost 1 000 - 1 500 cycles per group and 3 500 - 5 000 per epilogue (clock trace of a conversion warp,
    // DESIGN.md section 5): with 231 KB of shared memory carved out almost no L1 is left, every such load went to L2, and
    // its latency sat on the conversion chain. The room is what the GELU output used to occupy: it now stays in tensor memory.
    float* sVec = reinterpret_cast<float*>(sAux + 4096);
    float* sVecB1 = sVec + NG * 64;
    float* sVecB2 = sVec + 2 * NG * 64;
    for (int i = threadIdx.x; i < NG * 64; i += kMfThreads) {
        sVec[i] = i < p.hid ? __ldg(p.cs1 + i) : 0.f;
        sVecB1[i] = i < p.hid ? __ldg(p.b1 + i) : 0.f;
    }
    for (int i = threadIdx.x; i < p.NB; i += kMfThreads) sVecB2[i] = i < p.C ? __ldg(p.b2 + i) : 0.f;
    MfBars* bars = reinterpret_cast<MfBars*>(sW2 + 2 * w2_bytes);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(&bars->a_full[i], 1);
            mbar_init(&bars->a_empty[i], 1);
            mbar_init(&bars->w1_full[i], 1);
            mbar_init(&bars->w1_empty[i], 1);
            mbar_init(&bars->w2_full[i], 1);
            mbar_init(&bars->w2_empty[i], 1);
            mbar_init(&bars->d1_full[i], 1);
            mbar_init(&bars->d1_empty[i], kMfWarps);
            mbar_init(&bars->a2_full[i], kMfWarps);
            mbar_init(&bars->a2_empty[i], 1);
            mbar_init(&bars->d2_full[i], 1);
            mbar_init(&bars->d2_empty[i], kMfWarps);
            mbar_init(&bars->out_ready[i], kMfWarps);
        }
        fence_mbar_init();
    }
    if (warp == kMfWarps + 1) {
        tmem_alloc(&bars->tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = bars->tmem_slot;
    const uint32_t d2_col0 = 128;  // D1 buffers at columns 0 / 64, D2 buffers at 128 / 128 + 192 (NB <= 192)
    griddep_wait();
    griddep_launch();

    if (warp == kMfWarps) {
        // ================= TMA loader =================
        if (elect_one()) {
            tma_prefetch_desc(&p.tmX);
            tma_prefetch_desc(&p.tmW1);
            tma_prefetch_desc(&p.tmW2);
            int it = 0, G = 0;
            for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
                const int ab = it & 1;
                mbar_wait(&bars->a_empty[ab], ((it >> 1) & 1) ^ 1);
                mbar_arrive_expect_tx(&bars->a_full[ab], uint32_t(KS) * kMfATile);
                for (int ks = 0; ks < KS; ++ks)
                    tma_load_3d(sA + (ab * KS + ks) * kMfATile, &p.tmX, &bars->a_full[ab], ks * 64, tile * 128, 0);
                for (int g = 0; g < NG; ++g, ++G) {
                    const int s = G & 1;
                    const uint32_t ph = ((G >> 1) & 1) ^ 1;
                    mbar_wait(&bars->w1_empty[s], ph);
                    mbar_arrive_expect_tx(&bars->w1_full[s], uint32_t(KS) * kMfW1Slab);
                    for (int ks = 0; ks < KS; ++ks)
                        tma_load_3d(sW1 + (s * KS + ks) * kMfW1Slab, &p.tmW1, &bars->w1_full[s], ks * 64, g * 64, 0);
                    mbar_wait(&bars->w2_empty[s], ph);
                    mbar_arrive_expect_tx(&bars->w2_full[s], uint32_t(w2_bytes));
                    tma_load_3d(sW2 + s * w2_bytes, &p.tmW2, &bars->w2_full[s], g * 64, 0, 0);
                }
            }
        }
    } else if (warp == kMfWarps + 1) {
        // ================= MMA issuer =================
        // One thread, a polling issue loop over two independent cursors: G1 = next first-GEMM group, G2 = next second-GEMM
        // group (G2 < G1). With blocking waits in program order (MMA1(G), then MMA2(G - 1)) the thread sat in MMA2(G - 1)'s
        // wait for the conversion of G - 1 and could not issue MMA1(G + 1), whose accumulator buffer had been free since
        // that conversion STARTED: the conversion warps then waited for D1 (ncu r02_s38: 10 % of their samples on d1_full,
        // 30 % of the issuer's on a2_full). Descriptors are built once; inside the loop a descriptor is a 32-bit add.
        if (elect_one()) {
            const uint32_t idesc1 = make_idesc_bf16(128, 64, 0, 0);
            const uint32_t idesc2 = make_idesc_bf16(128, p.NB, 0, 0);
            const uint64_t a_desc0 = make_sdesc_sw128(smem_u32(sA), 16, 1024);
            const uint64_t w1_desc0 = make_sdesc_sw128(smem_u32(sW1), 16, 1024);
            const uint64_t w2_desc0 = make_sdesc_sw128(smem_u32(sW2), 16, 1024);
            const uint32_t w2_step = uint32_t(w2_bytes) >> 4;
            int my_tiles = 0;
            for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x) ++my_tiles;
            const int total = my_tiles * NG;
            int G1 = 0, it1 = 0, g1 = 0;   // first GEMM: global group counter, tile iteration, group inside the tile
            int G2 = 0, it2 = 0, g2 = 0;   // second GEMM
            bool a_ok = false, w1_ok = false, d1_ok = false, a2_ok = false, w2_ok = false, d2_ok = false;
            uint32_t spins = 0;
            while (G2 < total) {
                bool progress = false;
                if (G2 < G1) {  // ---- D2 += A2[G2 & 1] . W2[G2 & 1]^T  (K = 64) ----
                    const int s = G2 & 1, ab = it2 & 1;
                    const uint32_t ph = (G2 >> 1) & 1;
                    if (!a2_ok && mbar_test_wait(&bars->a2_full[s], ph)) a2_ok = true;
                    if (!w2_ok && mbar_test_wait(&bars->w2_full[s], ph)) w2_ok = true;
                    if (!d2_ok && (g2 != 0 || mbar_test_wait(&bars->d2_empty[ab], ((it2 >> 1) & 1) ^ 1))) d2_ok = true;
                    if (a2_ok && w2_ok && d2_ok) {
                        tc_fence_after();
                        const uint32_t d2 = tmem_base + d2_col0 + ab * 192;
                        const uint64_t bd = w2_desc0 + uint64_t(uint32_t(s) * w2_step);
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma_bf16_ts(d2, tmem_base + s * 64 + k * 16, bd + uint64_t(k * 2), idesc2, (g2 | k) != 0 ? 1u : 0u);
                        umma_commit(&bars->a2_empty[s]);
                        umma_commit(&bars->w2_empty[s]);
                        if (g2 == NG - 1) umma_commit(&bars->d2_full[ab]);
                        a2_ok = w2_ok = d2_ok = false;
                        ++G2;
                        if (++g2 == NG) {
                            g2 = 0;
                            ++it2;
                        }
                        progress = true;
                    }
                }
                if (G1 < total && G1 - G2 <= 1) {
                    const int s = G1 & 1, ab = it1 & 1;
                    const uint32_t ph = (G1 >> 1) & 1;
                    if (!a_ok && (g1 != 0 || mbar_test_wait(&bars->a_full[ab], (it1 >> 1) & 1))) a_ok = true;
                    if (!w1_ok && mbar_test_wait(&bars->w1_full[s], ph)) w1_ok = true;
                    if (!d1_ok && mbar_test_wait(&bars->d1_empty[s], ph ^ 1)) d1_ok = true;
                    if (a_ok && w1_ok && d1_ok) {
                        tc_fence_after();
                        const uint32_t d1 = tmem_base + s * 64;
                        uint64_t ad = a_desc0 + uint64_t(uint32_t(ab * KS) * (kMfATile >> 4));
                        uint64_t bd = w1_desc0 + uint64_t(uint32_t(s * KS) * (kMfW1Slab >> 4));
                        for (int ks = 0; ks < KS; ++ks) {
                            const int nk = ks == KS - 1 ? p.klast : 4;
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                if (k < nk) umma_bf16(d1, ad + uint64_t(k * 2), bd + uint64_t(k * 2), idesc1, (ks | k) != 0 ? 1u : 0u);
                            ad += kMfATile >> 4;
                            bd += kMfW1Slab >> 4;
                        }
                        umma_commit(&bars->w1_empty[s]);
                        umma_commit(&bars->d1_full[s]);
                        w1_ok = d1_ok = false;
                        ++G1;
                        if (++g1 == NG) {
                            g1 = 0;
                            ++it1;
                            a_ok = false;
                        }
                        progress = true;
                    }
                }
                if (progress) {
                    spins = 0;
                } else {
                    __nanosleep(20);
                    if (++spins > (1u << 24)) __trap();  // protocol bug: fail the launch instead of hanging
                }
            }
        }
    } else if (warp == kMfWarps + 2) {
        // ================= store warp: the finished tile goes back to x, then the A buffer is free =================
        if (elect_one()) {
            int it = 0;
            for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
                const int ab = it & 1;
                mbar_wait(&bars->out_ready[ab], (it >> 1) & 1);
                for (int ks = 0; ks < KS; ++ks)
                    tma_store_3d(&p.tmX, sA + (ab * KS + ks) * kMfATile, ks * 64, tile * 128, 0);
                bulk_commit();
                bulk_wait_read0();  // shared memory has been read: the loader may refill this buffer
                mbar_arrive(&bars->a_empty[ab]);
            }
            bulk_wait0();  // all stores have landed before the CTA exits
        }
    } else {
        // ================= warps 0..7: GELU conversion of every hidden group, then the tile's epilogue =================
        const int q = warp & 3;    // TMEM lane quarter
        const int ch = warp >> 2;  // column group: conversion columns CW ch .. +CW-1 of a hidden group; epilogue columns NB/kMfCG ch ..
        constexpr int CW = 64 / kMfCG;         // 16 conversion columns per thread and group
        const int r = q * 32 + lane;           // row of the tile
        const uint32_t rx = uint32_t(r & 7);   // swizzle term of this row
        const uint32_t lane_t = uint32_t(q * 32) << 16;
        float2* red = reinterpret_cast<float2*>(sRed);
#if defined(IRB_MLP_DBG) && IRB_MLP_DBG == 5
        const bool trace = blockIdx.x == 0 && threadIdx.x == 0;
        long long tsv[160];
        int tsi[160];
        int nts = 0;
#endif
        int it = 0, G = 0;
        // (mu, rstd) of this thread's row, fetched one tile ahead
        float2 st_next = make_float2(0.f, 1.f);
        if ((long long)blockIdx.x * 128 + r < p.P) st_next = __ldg(p.stats_in + (long long)blockIdx.x * 128 + r);
        for (int tile = blockIdx.x; tile < p.tiles; tile += gridDim.x, ++it) {
            const int ab = it & 1;
            const long long grow = (long long)tile * 128 + r;
            const bool valid = grow < p.P;
            const float2 st = st_next;
            {
                const long long gnext = grow + (long long)gridDim.x * 128;
                st_next = make_float2(0.f, 1.f);
                if (tile + (int)gridDim.x < p.tiles && gnext < p.P) st_next = __ldg(p.stats_in + gnext);
            }
            const f32x2 a2v = splat2(st.y), b2v = splat2(-st.x * st.y);
            for (int g = 0; g < NG; ++g, ++G) {
                const int s = G & 1;
                const uint32_t ph = (G >> 1) & 1;
                const int col0 = g * 64 + ch * CW;  // first hidden column of this thread's CW
                // column vectors first: they do not depend on the accumulator, so their latency hides behind the wait
                f32x2 tbv[CW / 2];
#pragma unroll
                for (int i = 0; i < CW / 4; ++i) {
                    const int col = col0 + i * 4;
                    const float4 c4 = *reinterpret_cast<const float4*>(sVec + col);
                    const float4 t4 = *reinterpret_cast<const float4*>(sVecB1 + col);
                    tbv[2 * i] = fma2(b2v, pack2(c4.x, c4.y), pack2(t4.x, t4.y));      // (-mu rstd) * colsum + b1'
                    tbv[2 * i + 1] = fma2(b2v, pack2(c4.z, c4.w), pack2(t4.z, t4.w));
                }
                IRB_TS(0);
                mbar_wait(&bars->d1_full[s], ph);
                IRB_TS(1);
                tc_fence_after();
                uint32_t acc[CW];
                tmem_ld16_nowait(tmem_base + s * 64 + ch * CW + lane_t, acc);
                mbar_wait(&bars->a2_empty[s], ph ^ 1);  // the second GEMM of group G - 2 has read this A2 buffer
                IRB_TS(2);
                tmem_ld_wait();
                IRB_TS(3);
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&bars->d1_empty[s]);  // the accumulator is in registers: D1[s] is free
                uint32_t pk[CW / 2];
#pragma unroll
                for (int h = 0; h < CW / 2; ++h) {
                    f32x2 x = pack2(__uint_as_float(acc[h * 2]), __uint_as_float(acc[h * 2 + 1]));
                    x = fma2(x, a2v, tbv[h]);
                    pk[h] = f32x2_to_bf16x2(gelu2_tanh(x));
                }
                // in place: the packed bf16 pairs of this thread's 16 columns overwrite the first 8 of those columns
                tmem_st8(tmem_base + s * 64 + ch * CW + lane_t, pk);
                tmem_st_wait();
                tc_fence_before();
                IRB_TS(4);
                __syncwarp();
                if (lane == 0) mbar_arrive(&bars->a2_full[s]);
                IRB_TS(5);
            }
            IRB_TS(6);
            // ---- epilogue: columns [ch * NB/kMfCG, (ch + 1) * NB/kMfCG) of this row ----
            mbar_wait(&bars->d2_full[ab], (it >> 1) & 1);
            IRB_TS(7);
            tc_fence_after();
            const int half = p.NB / kMfCG;         // columns per thread: a multiple of 8 (NB % 32 == 0)
            const int ecol0 = ch * half;
            const uint32_t d2 = tmem_base + d2_col0 + ab * 192 + lane_t;
            const uint32_t a_tile = smem_u32(sA) + ab * KS * kMfATile;
            // shift of the LayerNorm sums: the OLD row's mean (both column halves of a row must use the same value, and
            // x[row, 0] itself may already have been overwritten in place by the other half's thread)
            const float shift = st.x;
            const f32x2 nshift = splat2(-shift);
            f32x2 s1 = pack2(0.f, 0.f), s2 = s1;
            for (int c32 = 0; c32 < half; c32 += 32) {
                // up to 32 columns per TMEM load (half = 96 at C = 180: three loads per row instead of twelve)
                uint32_t a32[32];
                const int ncol = min(32, half - c32);
                if (ncol == 32) {
                    tmem_ld32_nowait(d2 + ecol0 + c32, a32);
                } else {
                    for (int c8 = 0; c8 < ncol; c8 += 8)
                        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                                     : "=r"(a32[c8 + 0]), "=r"(a32[c8 + 1]), "=r"(a32[c8 + 2]), "=r"(a32[c8 + 3]),
                                       "=r"(a32[c8 + 4]), "=r"(a32[c8 + 5]), "=r"(a32[c8 + 6]), "=r"(a32[c8 + 7])
                                     : "r"(d2 + ecol0 + c32 + c8)
                                     : "memory");
                }
                tmem_ld_wait();
#pragma unroll
                for (int c8 = 0; c8 < 4; ++c8) {
                    if (c8 * 8 >= ncol) break;
                    const int col = ecol0 + c32 + c8 * 8;
                    float v[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(a32[c8 * 8 + i]);
                    // residual chunk: K chunk col / 64 of the A tile, 16-byte chunk (col % 64) / 8 of the row
                    const uint32_t addr = a_tile + (col >> 6) * kMfATile + r * 128 + ((uint32_t((col & 63) >> 3) ^ rx) << 4);
                    const uint4 res = lds_v4(addr);
                    const float4 bb0 = *reinterpret_cast<const float4*>(sVecB2 + col);
                    const float4 bb1 = *reinterpret_cast<const float4*>(sVecB2 + col + 4);
                    const f32x2 y0 = add2(add2(pack2(v[0], v[1]), pack2(bb0.x, bb0.y)), bf16x2_to_f32x2(res.x));
                    const f32x2 y1 = add2(add2(pack2(v[2], v[3]), pack2(bb0.z, bb0.w)), bf16x2_to_f32x2(res.y));
                    const f32x2 y2 = add2(add2(pack2(v[4], v[5]), pack2(bb1.x, bb1.y)), bf16x2_to_f32x2(res.z));
                    const f32x2 y3 = add2(add2(pack2(v[6], v[7]), pack2(bb1.z, bb1.w)), bf16x2_to_f32x2(res.w));
                    if (col + 8 <= p.C) {
                        const f32x2 d0 = add2(y0, nshift), d1 = add2(y1, nshift), d2v = add2(y2, nshift), d3 = add2(y3, nshift);
                        s1 = add2(s1, add2(add2(d0, d1), add2(d2v, d3)));
                        s2 = fma2(d0, d0, s2);
                        s2 = fma2(d1, d1, s2);
                        s2 = fma2(d2v, d2v, s2);
                        s2 = fma2(d3, d3, s2);
                    } else if (col < p.C) {  // the ragged last chunk (C = 180: columns 176..179)
                        float yy[8];
                        unpack2(y0, yy[0], yy[1]);
                        unpack2(y1, yy[2], yy[3]);
                        unpack2(y2, yy[4], yy[5]);
                        unpack2(y3, yy[6], yy[7]);
                        float t1 = 0.f, t2 = 0.f;
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            if (col + i < p.C) {
                                const float d = yy[i] - shift;
                                t1 += d;
                                t2 = fmaf(d, d, t2);
                            }
                        s1 = add2(s1, pack2(t1, 0.f));
                        s2 = add2(s2, pack2(t2, 0.f));
                    }
                    sts_v4(addr, f32x2_to_bf16x2(y0), f32x2_to_bf16x2(y1), f32x2_to_bf16x2(y2), f32x2_to_bf16x2(y3));
                }
            }
            IRB_TS(8);
            tc_fence_before();
            fence_proxy_async();  // the TMA store (async proxy) reads the tile
            IRB_TS(9);
            // LayerNorm statistics of the new row: the two column halves meet through shared memory
            float a0, a1, q0, q1;
            unpack2(s1, a0, a1);
            unpack2(s2, q0, q1);
            if (ch > 0) red[(ch - 1) * 128 + r] = make_float2(a0 + a1, q0 + q1);
            asm volatile("bar.sync 1, %0;" ::"n"(kMfWarps * 32) : "memory");
            if (ch == 0 && valid && p.stats_out != nullptr) {
                float t1 = a0 + a1, t2 = q0 + q1;
#pragma unroll
                for (int k = 0; k < kMfCG - 1; ++k) {
                    const float2 o = red[k * 128 + r];
                    t1 += o.x;
                    t2 += o.y;
                }
                const float invn = 1.0f / float(p.C);
                const float m = t1 * invn;
                const float var = fmaxf(t2 * invn - m * m, 0.f);
                p.stats_out[grow] = make_float2(shift + m, rsqrtf(var + p.eps));
            }
            asm volatile("bar.sync 1, %0;" ::"n"(kMfWarps * 32) : "memory");  // red[] (= A2) is free again; the tile is written
            __syncwarp();
            if (lane == 0) {
                mbar_arrive(&bars->d2_empty[ab]);
                mbar_arrive(&bars->out_ready[ab]);
            }
            IRB_TS(10);
        }
#if defined(IRB_MLP_DBG) && IRB_MLP_DBG == 5
        if (trace) {
            static __device__ int printed;
            if (atomicAdd(&printed, 1) == 0)
                for (int i = 1; i < nts; ++i) printf("TS %d %d %lld\n", i, tsi[i], tsv[i] - tsv[i - 1]);
        }
#endif
    }
    tc_fence_before();
    __syncthreads();
    if (warp == kMfWarps + 1) tmem_dealloc(tmem_base, 512);
}

bool mlp_fused_shape_ok(int C, int hid) {
    // K = C in <= 3 chunks of 64 and two D2 buffers of <= 192 TMEM columns next to the two D1 buffers
    return C >= 32 && C <= 192 && C % 4 == 0 && ((C + 15) / 16 * 16) % (8 * kMfCG) == 0 && hid % 4 == 0 && hid >= 64 &&
           hid <= 1024;
}

// x [P][ldx] bf16 (in place); w1 bf16 [hid][ldw1] = fc1 with norm2's gain folded in; w2 bf16 [C][ldw2]
int mlp_fused_bf16(bf16* x, long long ldx, long long P, int C, int hid, const float2* stats_in, const bf16* w1,
                   int ldw1, const float* cs1, const float* b1, const bf16* w2, int ldw2, const float* b2,
                   float2* stats_out, float eps, cudaStream_t stream) {
    IRB_REQUIRE(mlp_fused_shape_ok(C, hid), "mlp_fused: unsupported shape");
    IRB_REQUIRE(ldx % 8 == 0 && ldw1 % 8 == 0 && ldw2 % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0,
                "mlp_fused: rows must be 16-byte aligned");
    IRB_REQUIRE(stats_in && cs1 && b1 && b2 && P > 0 && P < (1ll << 31), "mlp_fused: missing operand");
    MlpParams p{};
    p.P = int(P);
    p.C = C;
    p.hid = hid;
    p.KS = ceil_div(C, 64);
    p.klast = ceil_div(C - (p.KS - 1) * 64, 16);
    p.NG = ceil_div(hid, 64);
    p.NB = (C + 15) / 16 * 16;
    p.tiles = int(ceil_div_ll(P, 128));
    p.stats_in = stats_in;
    p.stats_out = stats_out;
    p.cs1 = cs1;
    p.b1 = b1;
    p.b2 = b2;
    p.eps = eps;
    {
        uint64_t dims[3] = {uint64_t(C), uint64_t(P), 1};
        uint64_t str[2] = {uint64_t(ldx) * 2, uint64_t(ldx) * 2 * uint64_t(P)};
        uint32_t box[3] = {64, 128, 1};
        IRB_TRY(make_tmap_bf16(&p.tmX, x, 3, dims, str, box));
    }
    {
        uint64_t dims[3] = {uint64_t(C), uint64_t(hid), 1};
        uint64_t str[2] = {uint64_t(ldw1) * 2, uint64_t(ldw1) * 2 * uint64_t(hid)};
        uint32_t box[3] = {64, 64, 1};
        IRB_TRY(make_tmap_bf16(&p.tmW1, w1, 3, dims, str, box));
    }
    {
        uint64_t dims[3] = {uint64_t(hid), uint64_t(C), 1};
        uint64_t str[2] = {uint64_t(ldw2) * 2, uint64_t(ldw2) * 2 * uint64_t(C)};
        uint32_t box[3] = {64, uint32_t(p.NB), 1};
        IRB_TRY(make_tmap_bf16(&p.tmW2, w2, 3, dims, str, box));
    }
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = p.tiles < sms ? p.tiles : sms;
    const size_t smem = 1024 + size_t(2) * p.KS * kMfATile + size_t(2) * p.KS * kMfW1Slab + kMfAux +
                        size_t(2) * p.NB * 128 + sizeof(MfBars) + 64;
    IRB_REQUIRE(smem <= 227 * 1024, "mlp_fused: shared memory budget exceeded");
    IRB_SMEM_OPTIN((mlp_fused_kernel), 227 * 1024);
    IRB_CHECK_CUDA(launch_pdl(mlp_fused_kernel, dim3(grid), dim3(kMfThreads), smem, stream, p));
    // algorithmic work: x read once and written once, weights once; both contractions
    prof_mark("mlp_fused", C, hid, 4.0 * double(P) * C * hid, double(P) * C * 2 * 2 + 2.0 * C * hid * 2 + double(P) * 16,
              stream);
    return 0;
}

}  // namespace irb
