// Restormer MDTA (restormer_arch.py:111-132): channel-transposed attention.
//   G[b,h,i,j] = sum_p q[b,p,hc+i] * k[b,p,hc+j]      (c x c Gram over K = H*W pixels, split-K over CTAs)
//   attn       = softmax_j( G / (|q_i| |k_j|) * temperature[h] )
//   Weff[b]    = Wp . blockdiag_h(attn[b,h])           (attn @ v followed by project_out is ONE per-image GEMM)
// The Gram runs on tcgen05 with BOTH operands MN-major: q and k are channels-last, so the contraction index
// (pixels) is the slow axis of the TMA box and the channel axis is the 128-byte swizzled row.
#include "common.cuh"
#include "ptx.cuh"
#include "kernels.h"

namespace irb {

static constexpr int kGramKP = 64;             // pixels per pipeline stage
static constexpr int kGramBox = kGramKP * 128; // bytes of one [64 ch x KP px] box

struct alignas(64) GramParams {
    CUtensorMap tm;   // {channels(3C), P, Nimg}, box {64, KP, 1}
    int C, c, heads, P;
    int chunk_px;     // pixels per CTA
    int nB;           // 64-channel boxes of k per stage (1 for c<=64, 2 for c<=128)
    int ncols;        // UMMA N (c rounded up to 16)
    int stages;
    int tmem_cols;
    float* G;
};

__global__ void __launch_bounds__(192) mdta_gram_tc_kernel(const __grid_constant__ GramParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int stage_bytes = (2 + p.nB) * kGramBox;
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

    if (warp == 0 && lane == 0) {
        for (int it = 0; it < total_it; ++it) {
            const int s = it % p.stages;
            const uint32_t ph = (it / p.stages) & 1;
            mbar_wait(&empty_bar[s], ph ^ 1);
            uint8_t* sa = smem + s * stage_bytes;
            mbar_arrive_expect_tx(&full_bar[s], stage_bytes);
            const int pp = px0 + it * kGramKP;
            tma_load_3d(sa, &p.tm, &full_bar[s], qc, pp, img);
            tma_load_3d(sa + kGramBox, &p.tm, &full_bar[s], qc + 64, pp, img);
            for (int j = 0; j < p.nB; ++j)
                tma_load_3d(sa + (2 + j) * kGramBox, &p.tm, &full_bar[s], kc + 64 * j, pp, img);
        }
    } else if (warp == 1 && lane == 0) {
        const uint32_t idesc = make_idesc_bf16(128, p.ncols, 1, 1);  // both operands MN-major
        for (int it = 0; it < total_it; ++it) {
            const int s = it % p.stages;
            const uint32_t ph = (it / p.stages) & 1;
            mbar_wait(&full_bar[s], ph);
            tc_fence_after();
            const uint32_t a_addr = smem_u32(smem + s * stage_bytes);
            const uint32_t b_addr = a_addr + 2 * kGramBox;
#pragma unroll
            for (int k = 0; k < kGramKP / 16; ++k) {
                // 16 pixels = two 8-pixel groups (SBO = 1024 B apart); next 64 channels are one box (LBO) away
                const uint64_t adesc = make_sdesc_sw128(a_addr + k * 2048, kGramBox, 1024);
                const uint64_t bdesc = make_sdesc_sw128(b_addr + k * 2048, kGramBox, 1024);
                umma_bf16(tmem_base, adesc, bdesc, idesc, (it | k) != 0 ? 1u : 0u);
            }
            umma_commit(&empty_bar[s]);
        }
        umma_commit(accum_bar);
    } else if (warp >= 2) {
        mbar_wait(accum_bar, 0);
        tc_fence_after();
        const int q = warp & 3;
        const int i = q * 32 + lane;  // q channel within the head
        const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16);
        float* g = p.G + (((long long)img * p.heads + head) * p.c + i) * p.c;
        for (int cb = 0; cb < p.ncols; cb += 16) {
            float v[16];
            tmem_ld16(taddr + cb, v);
            if (i < p.c) {
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    if (cb + j < p.c) atomicAdd(g + cb + j, v[j]);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, p.tmem_cols);
}

int mdta_gram_tc(const bf16* qkv, long long ld, int Nimg, int P, int C, int heads, float* G, cudaStream_t stream) {
    const int c = C / heads;
    IRB_REQUIRE(C % heads == 0 && c <= 128 && c % 8 == 0, "mdta_gram_tc: channels per head must be <=128 and %8");
    IRB_REQUIRE(ld % 8 == 0 && (reinterpret_cast<uintptr_t>(qkv) & 15) == 0, "mdta_gram_tc: alignment");
    GramParams p{};
    uint64_t dims[3] = {uint64_t(3 * C), uint64_t(P), uint64_t(Nimg)};
    uint64_t str[2] = {uint64_t(ld) * 2, uint64_t(ld) * 2 * P};
    uint32_t box[3] = {64, kGramKP, 1};
    IRB_TRY(make_tmap_bf16(&p.tm, qkv, 3, dims, str, box));
    p.C = C;
    p.c = c;
    p.heads = heads;
    p.P = P;
    p.nB = c <= 64 ? 1 : 2;
    p.ncols = (c + 15) / 16 * 16;
    p.tmem_cols = 32;
    while (p.tmem_cols < p.ncols) p.tmem_cols <<= 1;
    // split-K: aim for >= 4 waves of CTAs but at least 512 pixels per CTA
    int chunk = 2048;
    while (chunk > 512 && (long long)ceil_div(P, chunk) * heads * Nimg < 4 * 148) chunk >>= 1;
    p.chunk_px = chunk;
    const int stage_bytes = (2 + p.nB) * kGramBox;
    p.stages = 4;
    p.G = G;
    const size_t smem = size_t(p.stages) * stage_bytes + 1024 + 256;
    static bool attr_set = false;
    if (!attr_set) {
        IRB_CHECK_CUDA(cudaFuncSetAttribute(mdta_gram_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_set = true;
    }
    dim3 grid(ceil_div(P, chunk), heads, Nimg);
    mdta_gram_tc_kernel<<<grid, 192, smem, stream>>>(p);
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("mdta_gram_tc", C, heads, 2.0 * Nimg * P * C * c, double(Nimg) * P * 2 * C * 2 + double(Nimg) * C * c * 4, stream);
    return 0;
}

// ---- fp32 FMA Gram (fp32 precision mode) -----------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) mdta_gram_simt_kernel(const T* __restrict__ qkv, long long ld, int P, int C,
                                                             int c, int heads, int chunk_px, float* __restrict__ G) {
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
        __syncthreads();
    }
    float* g = G + ((long long)img * heads + head) * nout;
#pragma unroll
    for (int m = 0; m < MAXO; ++m) {
        const int o = threadIdx.x + 256 * m;
        if (o < nout) atomicAdd(g + o, acc[m]);
    }
}

int mdta_gram_simt(const void* qkv, int dt, long long ld, int Nimg, int P, int C, int heads, float* G,
                   cudaStream_t stream) {
    const int c = C / heads;
    IRB_REQUIRE(C % heads == 0 && c * c <= 36 * 256, "mdta_gram_simt: channels per head must be <= 96");
    int chunk = 1024;
    dim3 grid(ceil_div(P, chunk), heads, Nimg);
    const size_t smem = size_t(2) * 16 * c * sizeof(float);
    if (dt == DT_F32)
        mdta_gram_simt_kernel<float><<<grid, 256, smem, stream>>>((const float*)qkv, ld, P, C, c, heads, chunk, G);
    else
        mdta_gram_simt_kernel<bf16><<<grid, 256, smem, stream>>>((const bf16*)qkv, ld, P, C, c, heads, chunk, G);
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("mdta_gram_simt", C, heads, 2.0 * Nimg * P * C * c, double(Nimg) * P * 2 * C * (dt == DT_F32 ? 4 : 2) + double(Nimg) * C * c * 4, stream);
    return 0;
}

// ---- softmax + fold into project_out ----------------------------------------------------------
__global__ void __launch_bounds__(256) mdta_softmax_weff_kernel(const float* __restrict__ G,
                                                                const float* __restrict__ sumsq,
                                                                const float* __restrict__ temperature,
                                                                const float* __restrict__ Wp, int C, int heads,
                                                                void* __restrict__ Weff, int w_dt,
                                                                float* __restrict__ attn_out) {
    extern __shared__ float attn[];  // [c][c+1]
    const int c = C / heads;
    const int head = blockIdx.x, img = blockIdx.y;
    const int ldA = c + 1;
    const float* g = G + ((long long)img * heads + head) * c * c;
    const float* ssq = sumsq + (long long)img * 2 * C + head * c;
    const float* ssk = ssq + C;
    const float temp = __ldg(temperature + head);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // one warp per row i: F.normalize(dim=-1) divides by max(||.||_2, 1e-12) (restormer_arch.py:121-122)
    for (int i = warp; i < c; i += 8) {
        const float nq = fmaxf(sqrtf(ssq[i]), 1e-12f);
        float mx = -INFINITY;
        for (int j = lane; j < c; j += 32) {
            const float nk = fmaxf(sqrtf(ssk[j]), 1e-12f);
            const float s = g[i * c + j] / (nq * nk) * temp;
            attn[i * ldA + j] = s;
            mx = fmaxf(mx, s);
        }
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        float sum = 0.f;
        for (int j = lane; j < c; j += 32) {
            const float e = expf(attn[i * ldA + j] - mx);
            attn[i * ldA + j] = e;
            sum += e;
        }
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        const float inv = 1.0f / sum;
        for (int j = lane; j < c; j += 32) {
            const float a = attn[i * ldA + j] * inv;
            attn[i * ldA + j] = a;
            if (attn_out) attn_out[(((long long)img * heads + head) * c + i) * c + j] = a;
        }
    }
    __syncthreads();
    // Weff[img][o][head*c + j] = sum_i Wp[o][head*c + i] * attn[i][j]
    for (int e = threadIdx.x; e < C * c; e += 256) {
        const int o = e / c, j = e % c;
        const float* wrow = Wp + (long long)o * C + head * c;
        float acc = 0.f;
        for (int i = 0; i < c; ++i) acc = fmaf(__ldg(wrow + i), attn[i * ldA + j], acc);
        st_from_float(Weff, ((long long)img * C + o) * C + head * c + j, w_dt, acc);
    }
}

int mdta_softmax_weff(const float* G, const float* sumsq, const float* temperature, const float* Wp, int Nimg, int C,
                      int heads, void* Weff, int w_dt, float* attn_out, cudaStream_t stream) {
    const int c = C / heads;
    IRB_REQUIRE(C % heads == 0 && c <= 104, "mdta_softmax_weff: channels per head too large");
    const size_t smem = size_t(c) * (c + 1) * sizeof(float);
    dim3 grid(heads, Nimg);
    mdta_softmax_weff_kernel<<<grid, 256, smem, stream>>>(G, sumsq, temperature, Wp, C, heads, Weff, w_dt, attn_out);
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("mdta_softmax_weff", C, heads, 2.0 * Nimg * C * C * c, double(Nimg) * (C * c * 4.0 + C * C * (w_dt == DT_BF16 ? 2 : 4)) + C * C * 4.0, stream);
    return 0;
}

}  // namespace irb
