// Restormer MDTA (restormer_arch.py:111-132): channel-transposed attention.
//   G[b,h,i,j]  = sum_p q[b,p,hc+i] * k[b,p,hc+j]      (c x c Gram over K = H*W pixels)
//   nq[i], nk[j] = sum_p q^2, sum_p k^2                 (the F.normalize denominators, :121-122)
//   attn        = softmax_j( G / (max(|q_i|,1e-12) max(|k_j|,1e-12)) * temperature[h] )
//   Weff[b]     = Wp . blockdiag_h(attn[b,h])           (attn @ v followed by project_out is ONE per-image GEMM)
//
// The pixel axis is split into `nchunks` chunks per image; every (chunk, head, image) CTA writes its partial
// Gram AND the partial squared norms (the diagonals of q^T q and k^T k, computed by two more MMAs on the tiles that
// are already in shared memory) to its own slot; the softmax kernel sums the slots in a fixed order. No atomics:
// the result is bit-reproducible run to run.
//
// The Gram runs on tcgen05 with BOTH operands MN-major: q and k are channels-last, so the contraction index
// (pixels) is the slow axis of the TMA box and the channel axis is the 128-byte swizzled row.
#include "common.cuh"
#include "ptx.cuh"
#include "kernels.h"
#include <stdlib.h>

namespace irb {

static constexpr int kGramKP = 64;             // pixels per pipeline stage
static constexpr int kGramBox = kGramKP * 128; // bytes of one [64 ch x KP px] box
// A stage holds nbox 64-channel boxes of q and of k: 2 + 2 for heads wider than 64 channels, 1 + 1 otherwise (the UMMA
// still runs at M = 128: its rows 64..127 alias rows 0..63 through a zero leading-dimension offset and are ignored),
// which halves the shared-memory fill traffic of the 48-channel heads and lets 2-3 CTAs share an SM.

// Pixel chunks per image: about 2048 pixels each, at most 32 slots. A function of P ONLY, so that the summation
// order (hence every output bit) does not depend on how many tiles are batched into one launch.
int mdta_nchunks(int P, int heads, int Nimg) {
    (void)heads;
    (void)Nimg;
    long long want = P / 2048;
    if (want > 32) want = 32;
    if (want < 1) want = 1;
    const int chunk_px = int((((P + want - 1) / want) + kGramKP - 1) / kGramKP * kGramKP);
    return (P + chunk_px - 1) / chunk_px;
}
static int chunk_px_for(int P, int nchunks) {
    return ((P + nchunks - 1) / nchunks + kGramKP - 1) / kGramKP * kGramKP;
}

struct alignas(64) GramParams {
    CUtensorMap tm;   // {channels(3C), P, Nimg}, box {64, KP, 1}
    int C, c, heads, P;
    int chunk_px;     // pixels per CTA
    int ncols;        // UMMA N (c rounded up to 16)
    int acc_stride;   // TMEM columns between the three accumulators (ncols rounded up to 32)
    int stages;
    int nbox;         // 64-channel boxes per operand and stage: 1 (c <= 64) or 2
    int tmem_cols;
    float* Gpart;     // [Nimg][heads][nchunks][c][c]
    float* ssq_part;  // [Nimg][heads][nchunks][2][c]
};

__global__ void __launch_bounds__(192) mdta_gram_tc_kernel(const __grid_constant__ GramParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int stage_bytes = 2 * p.nbox * kGramBox;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + p.stages * stage_bytes);
    uint64_t* empty_bar = full_bar + p.stages;
    uint64_t* accum_bar = empty_bar + p.stages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int head = blockIdx.y, img = blockIdx.z;
    const int px0 = blockIdx.x * p.chunk_px;
    const int px1 = min(px0 + p.chunk_px, p.P);
    const int total_it = (px1 - px0 + kGramKP - 1) / kGramKP;

    if (warp == 0) {
        if (lane == 0) {
            tma_prefetch_desc(&p.tm);
            for (int s = 0; s < p.stages; ++s) {
                mbar_init(&full_bar[s], 1);
                mbar_init(&empty_bar[s], 1);
            }
            mbar_init(accum_bar, 1);
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
    const int qc = head * p.c;         // first q channel of this head
    const int kc = p.C + head * p.c;   // first k channel of this head

    // elect.sync leaders: ptxas then issues TMA / tcgen05 directly instead of through per-instruction election loops
    if (warp == 0 && elect_one()) {
        for (int it = 0; it < total_it; ++it) {
            const int s = it % p.stages;
            const uint32_t ph = (it / p.stages) & 1;
            mbar_wait(&empty_bar[s], ph ^ 1);
            uint8_t* sa = smem + s * stage_bytes;
            mbar_arrive_expect_tx(&full_bar[s], stage_bytes);
            const int pp = px0 + it * kGramKP;  // pixels >= P are zero-filled by TMA: they add nothing
            tma_load_3d(sa, &p.tm, &full_bar[s], qc, pp, img);
            if (p.nbox == 2) tma_load_3d(sa + kGramBox, &p.tm, &full_bar[s], qc + 64, pp, img);
            tma_load_3d(sa + p.nbox * kGramBox, &p.tm, &full_bar[s], kc, pp, img);
            if (p.nbox == 2) tma_load_3d(sa + 3 * kGramBox, &p.tm, &full_bar[s], kc + 64, pp, img);
        }
    } else if (warp == 1 && elect_one()) {
        const uint32_t idesc = make_idesc_bf16(128, p.ncols, 1, 1);  // both operands MN-major
        for (int it = 0; it < total_it; ++it) {
            const int s = it % p.stages;
            const uint32_t ph = (it / p.stages) & 1;
            mbar_wait(&full_bar[s], ph);
            tc_fence_after();
            const uint32_t q_addr = smem_u32(smem + s * stage_bytes);
            const uint32_t k_addr = q_addr + p.nbox * kGramBox;
            const uint32_t lbo = p.nbox == 2 ? kGramBox : 0;
#pragma unroll
            for (int k = 0; k < kGramKP / 16; ++k) {
                // 16 pixels = two 8-pixel groups (SBO = 1024 B apart); next 64 channels are one box (LBO) away
                const uint64_t qd = make_sdesc_sw128(q_addr + k * 2048, lbo, 1024);
                const uint64_t kd = make_sdesc_sw128(k_addr + k * 2048, lbo, 1024);
                const uint32_t acc = (it | k) != 0 ? 1u : 0u;
                umma_bf16(tmem_base, qd, kd, idesc, acc);                // q^T k
                umma_bf16(tmem_base + p.acc_stride, qd, qd, idesc, acc);      // q^T q (diagonal = |q_i|^2)
                umma_bf16(tmem_base + 2 * p.acc_stride, kd, kd, idesc, acc);  // k^T k (diagonal = |k_j|^2)
            }
            umma_commit(&empty_bar[s]);
        }
        umma_commit(accum_bar);
    } else if (warp >= 2) {
        mbar_wait(accum_bar, 0);
        tc_fence_after();
        const int q = warp & 3;
        const int i = q * 32 + lane;  // channel within the head (row of all three accumulators)
        const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16);
        const long long slot = ((long long)img * p.heads + head) * gridDim.x + blockIdx.x;
        float* g = p.Gpart + (slot * p.c + i) * p.c;
        for (int cb = 0; cb < p.ncols; cb += 16) {
            float v[16];
            tmem_ld16(taddr + cb, v);
            if (i < p.c) {
#pragma unroll
                for (int j = 0; j < 16; j += 4)
                    if (cb + j < p.c)  // c % 8 == 0
                        *reinterpret_cast<float4*>(g + cb + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
            }
        }
        if (q * 32 < p.ncols) {  // warp-uniform: rows of this warp that can be < c
            float d[2];
#pragma unroll
            for (int a = 0; a < 2; ++a) {
                float lo[16], hi[16];
                tmem_ld16(taddr + (a + 1) * p.acc_stride + q * 32, lo);
                tmem_ld16(taddr + (a + 1) * p.acc_stride + q * 32 + 16, hi);
                float r = 0.f;
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    r = (lane == j) ? lo[j] : r;
                    r = (lane == j + 16) ? hi[j] : r;
                }
                d[a] = r;
            }
            if (i < p.c) {
                p.ssq_part[(slot * 2 + 0) * p.c + i] = d[0];
                p.ssq_part[(slot * 2 + 1) * p.c + i] = d[1];
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, p.tmem_cols);
}

int mdta_gram_tc(const bf16* qkv, long long ld, int Nimg, int P, int C, int heads, int nchunks, float* Gpart,
                 float* ssq_part, cudaStream_t stream) {
    const int c = C / heads;
    IRB_REQUIRE(C % heads == 0 && c <= 128 && c % 8 == 0, "mdta_gram_tc: channels per head must be <=128 and %8");
    IRB_REQUIRE(ld % 8 == 0 && (reinterpret_cast<uintptr_t>(qkv) & 15) == 0, "mdta_gram_tc: alignment");
    IRB_REQUIRE(nchunks >= 1 && nchunks <= 64, "mdta_gram_tc: 1..64 chunks");
    GramParams p{};
    uint64_t dims[3] = {uint64_t(3 * C), uint64_t(P), uint64_t(Nimg)};
    uint64_t str[2] = {uint64_t(ld) * 2, uint64_t(ld) * 2 * P};
    uint32_t box[3] = {64, kGramKP, 1};
    IRB_TRY(make_tmap_bf16(&p.tm, qkv, 3, dims, str, box));
    p.C = C;
    p.c = c;
    p.heads = heads;
    p.P = P;
    p.ncols = (c + 15) / 16 * 16;
    p.acc_stride = (p.ncols + 31) / 32 * 32;
    p.tmem_cols = 32;
    while (p.tmem_cols < 3 * p.acc_stride + 32) p.tmem_cols <<= 1;  // +32: the diagonal loads may run past the last column
    IRB_REQUIRE(p.tmem_cols <= 512, "mdta_gram_tc: head too wide for TMEM");
    p.chunk_px = chunk_px_for(P, nchunks);
    IRB_REQUIRE((P + p.chunk_px - 1) / p.chunk_px == nchunks, "mdta_gram_tc: nchunks must come from mdta_nchunks()");
    static int nbox1 = -1;
    if (nbox1 < 0) {
        nbox1 = IRB_DIAG_ENV("IRB_GRAM_NBOX1", 1);
    }
    p.nbox = (c <= 64 && nbox1) ? 1 : 2;
    p.stages = 4;
    p.Gpart = Gpart;
    p.ssq_part = ssq_part;
    const size_t smem = size_t(p.stages) * 2 * p.nbox * kGramBox + 1024 + 256;
    IRB_SMEM_OPTIN((mdta_gram_tc_kernel), 200 * 1024);
    dim3 grid(nchunks, heads, Nimg);
    mdta_gram_tc_kernel<<<grid, 192, smem, stream>>>(p);
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("mdta_gram_tc", C, heads, 2.0 * Nimg * P * C * c,
              double(Nimg) * P * 2 * C * 2 + double(Nimg) * nchunks * (C * c + 2 * C) * 4, stream);
    return 0;
}

// ---- fp32 FMA Gram (fp32 precision mode) -----------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) mdta_gram_simt_kernel(const T* __restrict__ qkv, long long ld, int P, int C,
                                                             int c, int heads, int chunk_px, float* __restrict__ Gpart,
                                                             float* __restrict__ ssq_part) {
    constexpr int TP = 16;
    extern __shared__ float sm[];  // qs[TP][c], ks[TP][c]
    float* qs = sm;
    float* ks = sm + TP * c;
    const int head = blockIdx.y, img = blockIdx.z;
    const int px0 = blockIdx.x * chunk_px, px1 = min(px0 + chunk_px, P);
    const T* base = qkv + (long long)img * P * ld;
    const int nout = c * c;
    constexpr int MAXO = 36;  // 96*96/256
    float acc[MAXO];
#pragma unroll
    for (int m = 0; m < MAXO; ++m) acc[m] = 0.f;
    float sq = 0.f, sk = 0.f;  // thread t < c: |q_t|^2 and |k_t|^2
    for (int pp = px0; pp < px1; pp += TP) {
        for (int e = threadIdx.x; e < TP * c; e += 256) {
            const int r = e / c, ch = e % c;
            const int px = pp + r;
            float a = 0.f, b = 0.f;
            if (px < px1) {
                a = ld_as_float(base, (long long)px * ld + head * c + ch, sizeof(T) == 4 ? DT_F32 : DT_BF16);
                b = ld_as_float(base, (long long)px * ld + C + head * c + ch, sizeof(T) == 4 ? DT_F32 : DT_BF16);
            }
            qs[e] = a;
            ks[e] = b;
        }
        __syncthreads();
#pragma unroll
        for (int m = 0; m < MAXO; ++m) {
            const int o = threadIdx.x + 256 * m;
            if (o < nout) {
                const int i = o / c, j = o % c;
                float a = acc[m];
#pragma unroll
                for (int r = 0; r < TP; ++r) a = fmaf(qs[r * c + i], ks[r * c + j], a);
                acc[m] = a;
            }
        }
        if (threadIdx.x < c) {
#pragma unroll
            for (int r = 0; r < TP; ++r) {
                sq = fmaf(qs[r * c + threadIdx.x], qs[r * c + threadIdx.x], sq);
                sk = fmaf(ks[r * c + threadIdx.x], ks[r * c + threadIdx.x], sk);
            }
        }
        __syncthreads();
    }
    const long long slot = ((long long)img * heads + head) * gridDim.x + blockIdx.x;
    float* g = Gpart + slot * nout;
#pragma unroll
    for (int m = 0; m < MAXO; ++m) {
        const int o = threadIdx.x + 256 * m;
        if (o < nout) g[o] = acc[m];
    }
    if (threadIdx.x < c) {
        ssq_part[(slot * 2 + 0) * c + threadIdx.x] = sq;
        ssq_part[(slot * 2 + 1) * c + threadIdx.x] = sk;
    }
}

int mdta_gram_simt(const void* qkv, int dt, long long ld, int Nimg, int P, int C, int heads, int nchunks,
                   float* Gpart, float* ssq_part, cudaStream_t stream) {
    const int c = C / heads;
    IRB_REQUIRE(C % heads == 0 && c * c <= 36 * 256 && c <= 256, "mdta_gram_simt: channels per head must be <= 96");
    IRB_REQUIRE(nchunks >= 1 && nchunks <= 64, "mdta_gram_simt: 1..64 chunks");
    const int chunk = chunk_px_for(P, nchunks);
    IRB_REQUIRE((P + chunk - 1) / chunk == nchunks, "mdta_gram_simt: nchunks must come from mdta_nchunks()");
    dim3 grid(nchunks, heads, Nimg);
    const size_t smem = size_t(2) * 16 * c * sizeof(float);
    if (dt == DT_F32)
        mdta_gram_simt_kernel<float><<<grid, 256, smem, stream>>>((const float*)qkv, ld, P, C, c, heads, chunk, Gpart, ssq_part);
    else
        mdta_gram_simt_kernel<bf16><<<grid, 256, smem, stream>>>((const bf16*)qkv, ld, P, C, c, heads, chunk, Gpart, ssq_part);
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("mdta_gram_simt", C, heads, 2.0 * Nimg * P * C * c,
              double(Nimg) * P * 2 * C * (dt == DT_F32 ? 4 : 2) + double(Nimg) * nchunks * (C * c + 2 * C) * 4, stream);
    return 0;
}

// ---- fixed-order reduction of the partials + softmax, then the fold into project_out ------------
// Kernel A: CTA = 8 attention rows of one (image, head); warp = one row. Sums the per-chunk Gram / sum-of-squares
// slots in slot order (bit-reproducible), applies F.normalize's denominators and the temperature, softmax over
// the row, writes attn [n][heads][c][c] fp32.
__global__ void __launch_bounds__(256) mdta_softmax_kernel(const float* __restrict__ Gpart,
                                                           const float* __restrict__ ssq_part, int nchunks,
                                                           const float* __restrict__ temperature, int c, int heads,
                                                           float* __restrict__ attn_out) {
    __shared__ float nrm_k[128];
    const int head = blockIdx.y, img = blockIdx.z;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long slot0 = ((long long)img * heads + head) * nchunks;
    const float* g = Gpart + slot0 * c * c;
    const float* ss = ssq_part + slot0 * 2 * c;
    // F.normalize(dim=-1) divides by max(||.||_2, 1e-12) (restormer_arch.py:121-122)
    for (int v = threadIdx.x; v < c; v += 256) {
        float a = 0.f;
        for (int k = 0; k < nchunks; ++k) a += ss[(long long)k * 2 * c + c + v];
        nrm_k[v] = fmaxf(sqrtf(a), 1e-12f);
    }
    __syncthreads();
    const int i = blockIdx.x * 8 + warp;
    if (i >= c) return;
    float nq = 0.f;
    for (int k = 0; k < nchunks; ++k) nq += ss[(long long)k * 2 * c + i];
    nq = fmaxf(sqrtf(nq), 1e-12f);
    const float temp = __ldg(temperature + head);
    float sv[4];
    float mx = -INFINITY;
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
        const int j = lane + jj * 32;
        sv[jj] = -INFINITY;
        if (j < c) {
            float a = 0.f;
            const float* gp = g + (long long)i * c + j;
#pragma unroll 8
            for (int k = 0; k < nchunks; ++k) a += gp[(long long)k * c * c];
            sv[jj] = a / (nq * nrm_k[j]) * temp;
            mx = fmaxf(mx, sv[jj]);
        }
    }
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.f;
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
        const int j = lane + jj * 32;
        sv[jj] = j < c ? expf(sv[jj] - mx) : 0.f;
        sum += sv[jj];
    }
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float inv = 1.0f / sum;
    float* arow = attn_out + (((long long)img * heads + head) * c + i) * c;
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
        const int j = lane + jj * 32;
        if (j < c) arow[j] = sv[jj] * inv;
    }
}

// Kernel B: Weff[img][o][head*c + j] = sum_i Wp[o][head*c + i] * attn[img][head][i][j]; CTA = 64 rows o of one
// (head, image): the c x c attention block is staged once per 64 rows (16 rows per CTA made the staging the whole
// cost: 62 us for 340 MFLOP at level 4), each thread owns 4 rows x 1 column so that one attn load feeds 4 FMAs.
// RPT rows per thread (16 * RPT rows per CTA): 4 when that still gives about two CTAs per SM, else 2 or 1.
template <int RPT>
__global__ void __launch_bounds__(256) mdta_weff_kernel(const float* __restrict__ attn, const float* __restrict__ Wp,
                                                        int C, int heads, void* __restrict__ Weff, int w_dt) {
    extern __shared__ __align__(16) float sm[];  // attn [c][cs] then Wp rows [16 * RPT][c + 1]
    const int c = C / heads;
    const int head = blockIdx.y, img = blockIdx.z;
    constexpr int kWeffRows = 16 * RPT;
    const int o0 = blockIdx.x * kWeffRows;
    float* s_attn = sm;
    const int wld = c + 1;  // odd-ish stride: the 4 rows of a thread and the row groups of a warp spread over banks
    const int cs = (c + 3) & ~3;  // attn row stride in shared memory (16-byte rows; the pad columns are zero)
    float* s_w2 = sm + c * cs;
    const float* a = attn + ((long long)img * heads + head) * c * c;
    for (int e = threadIdx.x; e < c * cs; e += 256) {
        const int i = e / cs, j = e - i * cs;
        s_attn[e] = j < c ? a[i * c + j] : 0.f;
    }
    for (int e = threadIdx.x; e < kWeffRows * c; e += 256) {
        const int r = e / c, i = e - r * c;
        s_w2[r * wld + i] = (o0 + r < C) ? __ldg(Wp + (long long)(o0 + r) * C + head * c + i) : 0.f;
    }
    __syncthreads();
    // thread = RPT rows x 4 columns: per i one 16-byte attn load and RPT weight loads feed 4 * RPT FMAs (the first version's
    // RPT rows x 1 column issued 5 shared-memory loads per 4 FMAs and was bound by them: 45 us for 0.64 GFLOP at level 4)
    const int c4 = cs >> 2;
    for (int e = threadIdx.x; e < 16 * c4; e += 256) {
        const int rq = e / c4, j = (e - rq * c4) * 4;
        float acc[RPT][4];
#pragma unroll
        for (int k = 0; k < RPT; ++k) acc[k][0] = acc[k][1] = acc[k][2] = acc[k][3] = 0.f;
        const float* w0 = s_w2 + rq * wld;
#pragma unroll 4
        for (int i = 0; i < c; ++i) {
            const float4 av = *reinterpret_cast<const float4*>(s_attn + i * cs + j);
#pragma unroll
            for (int k = 0; k < RPT; ++k) {
                const float w = w0[k * 16 * wld + i];
                acc[k][0] = fmaf(w, av.x, acc[k][0]);
                acc[k][1] = fmaf(w, av.y, acc[k][1]);
                acc[k][2] = fmaf(w, av.z, acc[k][2]);
                acc[k][3] = fmaf(w, av.w, acc[k][3]);
            }
        }
        const long long ob = ((long long)img * C + o0 + rq) * C + head * c + j;
#pragma unroll
        for (int k = 0; k < RPT; ++k) {
            if (o0 + rq + 16 * k < C) {
#pragma unroll
                for (int jj = 0; jj < 4; ++jj)
                    if (j + jj < c) st_from_float(Weff, ob + (long long)(16 * k) * C + jj, w_dt, acc[k][jj]);
            }
        }
    }
}

int mdta_softmax_weff(const float* Gpart, const float* ssq_part, int nchunks, const float* temperature,
                      const float* Wp, int Nimg, int C, int heads, void* Weff, int w_dt, float* attn,
                      cudaStream_t stream) {
    const int c = C / heads;
    IRB_REQUIRE(C % heads == 0 && c <= 104, "mdta_softmax_weff: channels per head too large");
    IRB_REQUIRE(nchunks >= 1, "mdta_softmax_weff: nchunks");
    IRB_REQUIRE(attn != nullptr, "mdta_softmax_weff: the attn buffer [n][heads][c][c] is required");
    {
        dim3 grid(ceil_div(c, 8), heads, Nimg);
        mdta_softmax_kernel<<<grid, 256, 0, stream>>>(Gpart, ssq_part, nchunks, temperature, c, heads, attn);
        IRB_CHECK_CUDA(cudaGetLastError());
        prof_mark("mdta_softmax", C, heads, 0.0, double(Nimg) * (nchunks * (C * c + 2.0 * C) + C * c) * 4.0, stream);
    }
    {
        const long long per = (long long)heads * Nimg;
        const int rpt = ceil_div(C, 64) * per >= 296 ? 4 : (ceil_div(C, 32) * per >= 296 ? 2 : 1);
        const size_t smem = (size_t(c) * ((c + 3) & ~3) + size_t(16 * rpt) * (c + 1)) * sizeof(float);
        IRB_SMEM_OPTIN((mdta_weff_kernel<1>), 100 * 1024);
        IRB_SMEM_OPTIN((mdta_weff_kernel<2>), 100 * 1024);
        IRB_SMEM_OPTIN((mdta_weff_kernel<4>), 100 * 1024);
        dim3 grid(ceil_div(C, 16 * rpt), heads, Nimg);
        if (rpt == 4)
            mdta_weff_kernel<4><<<grid, 256, smem, stream>>>(attn, Wp, C, heads, Weff, w_dt);
        else if (rpt == 2)
            mdta_weff_kernel<2><<<grid, 256, smem, stream>>>(attn, Wp, C, heads, Weff, w_dt);
        else
            mdta_weff_kernel<1><<<grid, 256, smem, stream>>>(attn, Wp, C, heads, Weff, w_dt);
        IRB_CHECK_CUDA(cudaGetLastError());
        prof_mark("mdta_weff", C, heads, 2.0 * Nimg * C * C * c,
                  double(Nimg) * (C * c * 4.0 + C * C * (w_dt == DT_BF16 ? 2 : 4)) + C * C * 4.0, stream);
    }
    return 0;
}

}  // namespace irb
