// Memory-bound kernels on channels-last activations: LayerNorm over channels, depthwise 3x3 (+ the GDFN
// GELU gate, + the MDTA sum-of-squares side accumulator), nearest-neighbour 2x upsample.
// All are vectorised 4 channels per lane (8 B for bf16, 16 B for fp32), coalesced along the channel axis,
// and judged against the HBM roofline (bytes per element in DESIGN.md).
#include "common.cuh"
#include "kernels.h"

namespace irb {

template <typename T>
struct Vec4;
template <>
struct Vec4<float> {
    static __device__ __forceinline__ void load(const float* p, float* v) {
        const float4 t = *reinterpret_cast<const float4*>(p);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
    static __device__ __forceinline__ void store(float* p, const float* v) {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    }
};
template <>
struct Vec4<bf16> {
    static __device__ __forceinline__ void load(const bf16* p, float* v) {
        const uint2 t = *reinterpret_cast<const uint2*>(p);
        unpack_bf16x2(t.x, v[0], v[1]);
        unpack_bf16x2(t.y, v[2], v[3]);
    }
    static __device__ __forceinline__ void store(bf16* p, const float* v) {
        uint2 t;
        t.x = pack_bf16x2(v[0], v[1]);
        t.y = pack_bf16x2(v[2], v[3]);
        *reinterpret_cast<uint2*>(p) = t;
    }
};

// ----------------------------------------------------------------------------------------------
// LayerNorm (restormer_arch.py:25-57 WithBias / BiasFree; network_swinir.py nn.LayerNorm eps 1e-5)
// LPP lanes cooperate on one pixel; each lane owns chunks lane, lane+LPP, ... of 4 channels.
// ----------------------------------------------------------------------------------------------
template <typename TI, typename TO, int LPP>
__global__ void __launch_bounds__(256) layernorm_kernel(const TI* __restrict__ x, long long ldx, TO* __restrict__ y,
                                                        long long ldy, long long rows, int C,
                                                        const float* __restrict__ w, const float* __restrict__ bias,
                                                        int mode, float eps) {
    constexpr int MAXCH = 4;
    const int lane = threadIdx.x % LPP;
    const long long row = ((long long)blockIdx.x * blockDim.x + threadIdx.x) / LPP;
    const bool active = row < rows;
    const int nch = C >> 2;
    const bool vec_wb = ((reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(bias)) & 15) == 0;
    float v[MAXCH][4];
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < MAXCH; ++j) {
        const int ch = lane + j * LPP;
        if (active && ch < nch) {
            Vec4<TI>::load(x + row * ldx + ch * 4, v[j]);
            s += v[j][0] + v[j][1] + v[j][2] + v[j][3];
        } else {
            v[j][0] = v[j][1] = v[j][2] = v[j][3] = 0.f;
        }
    }
#pragma unroll
    for (int o = LPP / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o, LPP);
    const float mu = s / float(C);
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < MAXCH; ++j) {
        const int ch = lane + j * LPP;
        if (ch < nch) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float d = v[j][i] - mu;
                q += d * d;
            }
        }
    }
#pragma unroll
    for (int o = LPP / 2; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o, LPP);
    const float rstd = 1.0f / sqrtf(q / float(C) + eps);
    if (!active) return;
#pragma unroll
    for (int j = 0; j < MAXCH; ++j) {
        const int ch = lane + j * LPP;
        if (ch < nch) {
            float o4[4], g4[4], b4[4] = {0.f, 0.f, 0.f, 0.f};
            if (vec_wb) {  // one 16-byte load per vector and chunk instead of four (eight) scalar ones
                const float4 g = __ldg(reinterpret_cast<const float4*>(w) + ch);
                g4[0] = g.x, g4[1] = g.y, g4[2] = g.z, g4[3] = g.w;
                if (mode == 0 && bias) {
                    const float4 b = __ldg(reinterpret_cast<const float4*>(bias) + ch);
                    b4[0] = b.x, b4[1] = b.y, b4[2] = b.z, b4[3] = b.w;
                }
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    g4[i] = __ldg(w + ch * 4 + i);
                    if (mode == 0 && bias) b4[i] = __ldg(bias + ch * 4 + i);
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                if (mode == 0)
                    o4[i] = (v[j][i] - mu) * rstd * g4[i] + b4[i];
                else
                    o4[i] = v[j][i] * rstd * g4[i];
            }
            Vec4<TO>::store(y + row * ldy + ch * 4, o4);
        }
    }
}

template <typename TI, typename TO>
static int layernorm_launch(const void* x, long long ldx, void* y, long long ldy, long long rows, int C, const float* w,
                            const float* b, int mode, float eps, cudaStream_t stream) {
    const int nch = C / 4;
    // lanes per pixel: the group size in {8, 16, 32} (each lane owns up to 4 chunks of 4 channels) that leaves the fewest
    // idle lane slots: C = 180 -> 45 chunks: 16 lanes x 3 passes (48 slots) instead of 32 x 2 (64 slots, 30 % idle)
    int lpp = 32, best = 1 << 30;
    for (int cand = 8; cand <= 32; cand <<= 1) {
        if (nch > 4 * cand) continue;
        const int slots = (nch + cand - 1) / cand * cand;
        if (slots < best) {
            best = slots;
            lpp = cand;
        }
    }
    const long long threads = rows * lpp;
    const int blocks = int(ceil_div_ll(threads, 256));
    const TI* xi = reinterpret_cast<const TI*>(x);
    TO* yo = reinterpret_cast<TO*>(y);
    if (lpp == 8)
        layernorm_kernel<TI, TO, 8><<<blocks, 256, 0, stream>>>(xi, ldx, yo, ldy, rows, C, w, b, mode, eps);
    else if (lpp == 16)
        layernorm_kernel<TI, TO, 16><<<blocks, 256, 0, stream>>>(xi, ldx, yo, ldy, rows, C, w, b, mode, eps);
    else
        layernorm_kernel<TI, TO, 32><<<blocks, 256, 0, stream>>>(xi, ldx, yo, ldy, rows, C, w, b, mode, eps);
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("layernorm", C, int(sizeof(TI)), 8.0 * rows * C, double(rows) * C * (sizeof(TI) + sizeof(TO)), stream);
    return 0;
}

int layernorm(const void* x, int x_dt, long long ldx, void* y, int y_dt, long long ldy, long long rows, int C,
              const float* w, const float* b, int mode, float eps, cudaStream_t stream) {
    IRB_REQUIRE(C % 4 == 0 && C <= 512 && C > 0, "layernorm: C must be a multiple of 4 and <= 512");
    IRB_REQUIRE(ldx % 4 == 0 && ldy % 4 == 0, "layernorm: row strides must be multiples of 4");
    if (rows <= 0) return 0;
    if (x_dt == DT_F32 && y_dt == DT_F32) return layernorm_launch<float, float>(x, ldx, y, ldy, rows, C, w, b, mode, eps, stream);
    if (x_dt == DT_BF16 && y_dt == DT_BF16) return layernorm_launch<bf16, bf16>(x, ldx, y, ldy, rows, C, w, b, mode, eps, stream);
    if (x_dt == DT_F32 && y_dt == DT_BF16) return layernorm_launch<float, bf16>(x, ldx, y, ldy, rows, C, w, b, mode, eps, stream);
    set_error("layernorm: unsupported dtype combination");
    return 1;
}

// ----------------------------------------------------------------------------------------------
// Row statistics for the LayerNorm that is folded into the following GEMM (Epi::ln_stats): per pixel
// (mu, 1/sqrt(var + eps)) with the population variance about mu, two-pass in registers like layernorm_kernel.
// Reads x once (C*e bytes per row), writes 8 bytes per row.
// ----------------------------------------------------------------------------------------------
// 4 lanes cooperate on one row; every lane owns 16-byte chunks lane, lane+4, ... (8 bf16 / 4 fp32 channels), so a
// warp has 8 rows and up to 8 x C x e bytes in flight. NCH = chunks per lane (compile-time upper bound).
template <typename TI, int NCH>
__global__ void __launch_bounds__(256) row_stats_kernel(const TI* __restrict__ x, long long ldx, float2* __restrict__ st,
                                                        long long rows, int C, float eps) {
    constexpr int EPC = 16 / sizeof(TI);  // elements per 16-byte chunk
    const int lane = threadIdx.x & 3;
    const long long row = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 2;
    const bool active = row < rows;
    float v[NCH][EPC];
    float s = 0.f;
    const TI* xr = x + (active ? row : 0) * ldx;
#pragma unroll
    for (int j = 0; j < NCH; ++j) {
        const int c0 = (lane + j * 4) * EPC;
        if (active && c0 < C) {
            const uint4 u = __ldg(reinterpret_cast<const uint4*>(xr + c0));
            if (sizeof(TI) == 2) {
                unpack_bf16x2(u.x, v[j][0], v[j][1]);
                unpack_bf16x2(u.y, v[j][2], v[j][3]);
                unpack_bf16x2(u.z, v[j][4 % EPC], v[j][5 % EPC]);
                unpack_bf16x2(u.w, v[j][6 % EPC], v[j][7 % EPC]);
            } else {
                v[j][0] = __uint_as_float(u.x); v[j][1] = __uint_as_float(u.y);
                v[j][2] = __uint_as_float(u.z); v[j][3] = __uint_as_float(u.w);
            }
#pragma unroll
            for (int i = 0; i < EPC; ++i) {
                if (c0 + i >= C) v[j][i] = 0.f;  // ragged last chunk (C % EPC != 0): padding columns are not data
                s += v[j][i];
            }
        } else {
#pragma unroll
            for (int i = 0; i < EPC; ++i) v[j][i] = 0.f;
        }
    }
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    const float mu = s / float(C);
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < NCH; ++j) {
        const int c0 = (lane + j * 4) * EPC;
#pragma unroll
        for (int i = 0; i < EPC; ++i) {
            const float d = (c0 + i < C) ? v[j][i] - mu : 0.f;
            q = fmaf(d, d, q);
        }
    }
    q += __shfl_xor_sync(0xffffffffu, q, 1);
    q += __shfl_xor_sync(0xffffffffu, q, 2);
    if (active && lane == 0) st[row] = make_float2(mu, 1.0f / sqrtf(q / float(C) + eps));
}

int row_stats(const void* x, int x_dt, long long ldx, float2* stats, long long rows, int C, float eps,
              cudaStream_t stream) {
    const int epc = x_dt == DT_BF16 ? 8 : 4;
    IRB_REQUIRE(C % 4 == 0 && C <= 512 && C > 0 && ldx % epc == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0,
                "row_stats: C must be a multiple of 4 and <= 512, rows 16-byte aligned");
    IRB_REQUIRE(ldx >= ceil_div(C, epc) * epc, "row_stats: row stride must cover the last 16-byte chunk");
    if (rows <= 0) return 0;
    const int nch = ceil_div(ceil_div(C, epc), 4);  // chunks per lane
    const int blocks = int(ceil_div_ll(rows * 4, 256));
#define IRB_RS(T, L) row_stats_kernel<T, L><<<blocks, 256, 0, stream>>>((const T*)x, ldx, stats, rows, C, eps)
    if (x_dt == DT_BF16) {
        if (nch <= 2) IRB_RS(bf16, 2); else if (nch <= 3) IRB_RS(bf16, 3); else if (nch <= 6) IRB_RS(bf16, 6);
        else if (nch <= 8) IRB_RS(bf16, 8); else if (nch <= 12) IRB_RS(bf16, 12); else IRB_RS(bf16, 16);
    } else {
        if (nch <= 3) IRB_RS(float, 3); else if (nch <= 6) IRB_RS(float, 6); else if (nch <= 12) IRB_RS(float, 12);
        else if (nch <= 16) IRB_RS(float, 16); else if (nch <= 24) IRB_RS(float, 24); else IRB_RS(float, 32);
    }
#undef IRB_RS
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("row_stats", C, x_dt == DT_BF16 ? 2 : 4, 4.0 * rows * C, double(rows) * (C * (x_dt == DT_BF16 ? 2 : 4) + 8), stream);
    return 0;
}

// ----------------------------------------------------------------------------------------------
// Depthwise 3x3, zero pad 1 (restormer_arch.py:82,106 qkv_dwconv / ffn.dwconv, bias=False).
// Thread = (x, 4-channel group); it walks RY rows with a 3x3 register window, so every input element is
// loaded 3x from L1 and once from HBM. GATE: the GDFN gate gelu(dw(x1)) * dw(x2) (restormer_arch.py:90-91).
// ----------------------------------------------------------------------------------------------
template <typename T, bool GATE>
__global__ void __launch_bounds__(256) dwconv3x3_kernel(const T* __restrict__ x, long long ldx, T* __restrict__ y,
                                                        long long ldy, int H, int W, int C /*groups*4 = C out*/,
                                                        int hp_off /*GATE: channel offset of x2*/,
                                                        const float* __restrict__ wt, int wt_ld, int RY,
                                                        float* __restrict__ sumsq, int sumsq_channels) {
    extern __shared__ float ssum[];  // [sumsq_channels] when sumsq != nullptr
    const int CG = C >> 2;
    const long long lin = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = lin < (long long)W * CG;
    const int px = active ? int(lin / CG) : 0;
    const int cg = active ? int(lin % CG) : 0;
    const int c0 = cg * 4;
    const int img = blockIdx.z;
    const int y0 = blockIdx.y * RY;
    const int y1 = min(y0 + RY, H);
    if (sumsq) {
        for (int i = threadIdx.x; i < sumsq_channels; i += blockDim.x) ssum[i] = 0.f;
        __syncthreads();
    }
    if (active) {
        constexpr int NS = GATE ? 2 : 1;
        float wreg[NS][9][4];
#pragma unroll
        for (int s = 0; s < NS; ++s)
#pragma unroll
            for (int t = 0; t < 9; ++t) {
                const float4 w4 = __ldg(reinterpret_cast<const float4*>(wt + (long long)t * wt_ld + c0 + s * hp_off));
                wreg[s][t][0] = w4.x; wreg[s][t][1] = w4.y; wreg[s][t][2] = w4.z; wreg[s][t][3] = w4.w;
            }
        const T* xb = x + (long long)img * H * W * ldx;
        T* yb = y + (long long)img * H * W * ldy;
        float win[NS][3][3][4];  // [stream][row slot][dx][ch]
        auto load_row = [&](int yy, int slot) {
#pragma unroll
            for (int s = 0; s < NS; ++s)
#pragma unroll
                for (int d = 0; d < 3; ++d) {
                    const int xx = px + d - 1;
                    if (yy >= 0 && yy < H && xx >= 0 && xx < W)
                        Vec4<T>::load(xb + ((long long)yy * W + xx) * ldx + c0 + s * hp_off, win[s][slot][d]);
                    else
                        win[s][slot][d][0] = win[s][slot][d][1] = win[s][slot][d][2] = win[s][slot][d][3] = 0.f;
                }
        };
        load_row(y0 - 1, 0);
        load_row(y0, 1);
        float ss[4] = {0.f, 0.f, 0.f, 0.f};
        for (int yy = y0; yy < y1; ++yy) {
            load_row(yy + 1, 2);
            float acc[NS][4];
#pragma unroll
            for (int s = 0; s < NS; ++s) {
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[s][i] = 0.f;
#pragma unroll
                for (int r = 0; r < 3; ++r)
#pragma unroll
                    for (int d = 0; d < 3; ++d)
#pragma unroll
                        for (int i = 0; i < 4; ++i) acc[s][i] = fmaf(win[s][r][d][i], wreg[s][r * 3 + d][i], acc[s][i]);
            }
            float o4[4];
            if (GATE) {
#pragma unroll
                for (int i = 0; i < 4; ++i) o4[i] = gelu_erf(acc[0][i]) * acc[NS - 1][i];
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    o4[i] = acc[0][i];
                    ss[i] = fmaf(acc[0][i], acc[0][i], ss[i]);
                }
            }
            Vec4<T>::store(yb + ((long long)yy * W + px) * ldy + c0, o4);
            // slide the window down one row
#pragma unroll
            for (int s = 0; s < NS; ++s)
#pragma unroll
                for (int d = 0; d < 3; ++d)
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        win[s][0][d][i] = win[s][1][d][i];
                        win[s][1][d][i] = win[s][2][d][i];
                    }
        }
        if (!GATE && sumsq && c0 < sumsq_channels) {
#pragma unroll
            for (int i = 0; i < 4; ++i) atomicAdd(&ssum[c0 + i], ss[i]);
        }
    }
    if (sumsq) {
        __syncthreads();
        for (int i = threadIdx.x; i < sumsq_channels; i += blockDim.x) {
            const float v = ssum[i];
            if (v != 0.f) atomicAdd(sumsq + (long long)img * sumsq_channels + i, v);
        }
    }
}

template <typename T>
static int dwconv_launch(bool gate, const void* x, long long ldx, void* y, long long ldy, int Nimg, int H, int W,
                         int Cout, int hp_off, const float* wt, int wt_ld, float* sumsq, int sumsq_channels,
                         cudaStream_t stream) {
    const int RY = H < 32 ? H : 32;
    const long long lin = (long long)W * (Cout / 4);
    dim3 grid((unsigned)ceil_div_ll(lin, 256), ceil_div(H, RY), Nimg);
    const size_t smem = sumsq ? size_t(sumsq_channels) * sizeof(float) : 0;
    const T* xi = reinterpret_cast<const T*>(x);
    T* yo = reinterpret_cast<T*>(y);
    if (gate)
        dwconv3x3_kernel<T, true><<<grid, 256, 0, stream>>>(xi, ldx, yo, ldy, H, W, Cout, hp_off, wt, wt_ld, RY, nullptr, 0);
    else
        dwconv3x3_kernel<T, false><<<grid, 256, smem, stream>>>(xi, ldx, yo, ldy, H, W, Cout, 0, wt, wt_ld, RY, sumsq, sumsq_channels);
    IRB_CHECK_CUDA(cudaGetLastError());
    {
        const double px = double(Nimg) * H * W, cin = gate ? 2.0 * Cout : double(Cout);
        prof_mark(gate ? "dwconv3x3_gate" : "dwconv3x3", Cout, int(sizeof(T)), 18.0 * px * cin, px * (cin + Cout) * sizeof(T) + 36.0 * cin, stream);
    }
    return 0;
}

int dwconv3x3(const void* x, int dt, long long ldx, void* y, long long ldy, int Nimg, int H, int W, int C,
              const float* wt, float* sumsq, int sumsq_channels, cudaStream_t stream) {
    IRB_REQUIRE(C % 4 == 0 && ldx % 4 == 0 && ldy % 4 == 0, "dwconv3x3: C and strides must be multiples of 4");
    IRB_REQUIRE(sumsq_channels % 4 == 0 && sumsq_channels <= C && sumsq_channels <= 8192, "dwconv3x3: bad sumsq_channels");
    if (dt == DT_F32) return dwconv_launch<float>(false, x, ldx, y, ldy, Nimg, H, W, C, 0, wt, C, sumsq, sumsq_channels, stream);
    if (!sumsq && ldx % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(y) & 7) == 0)
        return dwconv3x3_bf16_tma((const bf16*)x, ldx, (bf16*)y, ldy, Nimg, H, W, C, wt, stream);
    return dwconv_launch<bf16>(false, x, ldx, y, ldy, Nimg, H, W, C, 0, wt, C, sumsq, sumsq_channels, stream);
}

int dwconv3x3_gate(const void* x, int dt, long long ldx, void* y, long long ldy, int Nimg, int H, int W, int hp,
                   const float* wt, cudaStream_t stream) {
    IRB_REQUIRE(hp % 4 == 0 && ldx % 4 == 0 && ldy % 4 == 0, "dwconv3x3_gate: hp and strides must be multiples of 4");
    if (dt == DT_F32) return dwconv_launch<float>(true, x, ldx, y, ldy, Nimg, H, W, hp, hp, wt, 2 * hp, nullptr, 0, stream);
    if (ldx % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(y) & 7) == 0)
        return dwconv3x3_gate_bf16_tma((const bf16*)x, ldx, (bf16*)y, ldy, Nimg, H, W, hp, wt, stream);
    return dwconv_launch<bf16>(true, x, ldx, y, ldy, Nimg, H, W, hp, hp, wt, 2 * hp, nullptr, 0, stream);
}

// ----------------------------------------------------------------------------------------------
// nearest-neighbour x2 (network_swinir.py:828-830, F.interpolate(scale_factor=2, mode='nearest'))
// ----------------------------------------------------------------------------------------------
template <typename T>
__global__ void nearest_up2_kernel(const T* __restrict__ x, long long ldx, T* __restrict__ y, long long ldy, int Nimg,
                                   int H, int W, int CG) {
    const long long total = (long long)Nimg * 2 * H * 2 * W * CG;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int cg = int(i % CG);
        long long p = i / CG;
        const int ox = int(p % (2 * W));
        p /= 2 * W;
        const int oy = int(p % (2 * H));
        const int n = int(p / (2 * H));
        float v[4];
        Vec4<T>::load(x + (((long long)n * H + (oy >> 1)) * W + (ox >> 1)) * ldx + cg * 4, v);
        Vec4<T>::store(y + (((long long)n * 2 * H + oy) * 2 * W + ox) * ldy + cg * 4, v);
    }
}

int nearest_up2(const void* x, int dt, long long ldx, void* y, long long ldy, int Nimg, int H, int W, int C,
                cudaStream_t stream) {
    IRB_REQUIRE(C % 4 == 0 && ldx % 4 == 0 && ldy % 4 == 0, "nearest_up2: C and strides must be multiples of 4");
    const long long total = (long long)Nimg * 4 * H * W * (C / 4);
    const int blocks = int(ceil_div_ll(total, 256) > 148 * 32 ? 148 * 32 : ceil_div_ll(total, 256));
    if (dt == DT_F32)
        nearest_up2_kernel<float><<<blocks, 256, 0, stream>>>((const float*)x, ldx, (float*)y, ldy, Nimg, H, W, C / 4);
    else
        nearest_up2_kernel<bf16><<<blocks, 256, 0, stream>>>((const bf16*)x, ldx, (bf16*)y, ldy, Nimg, H, W, C / 4);
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("nearest_up2", C, dt, 0.0, double(Nimg) * H * W * C * 5.0 * (dt == DT_F32 ? 4 : 2), stream);
    return 0;
}

// ----------------------------------------------------------------------------------------------
// x = hi + lo split of the fp32 network input for the tensor-core first convolution (see pack_conv3x3_split).
// ----------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) split_hi_lo_kernel(const float* __restrict__ x, int cin, bf16* __restrict__ y,
                                                          long long rows) {
    const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= rows) return;
    float hi[4] = {0.f, 0.f, 0.f, 0.f}, lo[4] = {0.f, 0.f, 0.f, 0.f};
    for (int c = 0; c < cin; ++c) {
        const float v = x[r * cin + c];
        hi[c] = __bfloat162float(__float2bfloat16_rn(v));
        lo[c] = v - hi[c];
    }
    float o[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int c = 0; c < cin; ++c) {
        o[c] = hi[c];
        o[cin + c] = lo[c];
    }
    uint4 u;
    u.x = pack_bf16x2(o[0], o[1]);
    u.y = pack_bf16x2(o[2], o[3]);
    u.z = pack_bf16x2(o[4], o[5]);
    u.w = pack_bf16x2(o[6], o[7]);
    *reinterpret_cast<uint4*>(y + r * 8) = u;
}

// fp32 [rows][K] -> bf16 [rows][6K] = h | h | h | m | m | l with x = h + m + l (three bf16 terms carry the 24 mantissa
// bits of x): against weights laid out Wh | Wm | Wl | Wh | Wm | Wh one bf16 GEMM over 6K computes
// h.Wh + h.Wm + h.Wl + m.Wh + m.Wm + l.Wh, i.e. x.W up to terms of relative size 2^-24 (fp32 accumulation in TMEM).
__global__ void split3_kernel(const float* __restrict__ x, long long ldx, bf16* __restrict__ y, long long ldy,
                              long long rows, int K) {
    const int K2 = K >> 1;
    const long long total = rows * K2;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / K2;
        const int c = int(i - r * K2) * 2;
        const float2 v = *reinterpret_cast<const float2*>(x + r * ldx + c);
        const bf16 h0 = __float2bfloat16_rn(v.x), h1 = __float2bfloat16_rn(v.y);
        const float r0 = v.x - __bfloat162float(h0), r1 = v.y - __bfloat162float(h1);   // exact in fp32
        const bf16 m0 = __float2bfloat16_rn(r0), m1 = __float2bfloat16_rn(r1);
        const bf16 l0 = __float2bfloat16_rn(r0 - __bfloat162float(m0)), l1 = __float2bfloat16_rn(r1 - __bfloat162float(m1));
        __nv_bfloat162 h, m, l;
        h.x = h0; h.y = h1; m.x = m0; m.y = m1; l.x = l0; l.y = l1;
        __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(y + r * ldy + c);
        o[0] = h;
        o[K2] = h;
        o[2 * K2] = h;
        o[3 * K2] = m;
        o[4 * K2] = m;
        o[5 * K2] = l;
    }
}

int split3_bf16(const float* x, long long ldx, bf16* y, long long ldy, long long rows, int K, cudaStream_t stream) {
    IRB_REQUIRE(K % 2 == 0 && ldx % 2 == 0 && ldy % 2 == 0 && ldy >= 6ll * K, "split3_bf16: K and the row strides must be even");
    if (rows <= 0) return 0;
    const long long total = rows * (K / 2);
    const long long b = ceil_div_ll(total, 256);
    split3_kernel<<<int(b > 148 * 32 ? 148 * 32 : b), 256, 0, stream>>>(x, ldx, y, ldy, rows, K);
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("split3", K, 0, 0.0, double(rows) * K * (4 + 12), stream);
    return 0;
}

int split_hi_lo_bf16(const float* x, int cin, bf16* y, long long rows, cudaStream_t stream) {
    IRB_REQUIRE(cin >= 1 && cin <= 4, "split_hi_lo_bf16: cin must be <= 4");
    if (rows <= 0) return 0;
    split_hi_lo_kernel<<<int(ceil_div_ll(rows, 256)), 256, 0, stream>>>(x, cin, y, rows);
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("split_hi_lo", cin, 0, 0.0, double(rows) * (cin * 4 + 16), stream);
    return 0;
}

}  // namespace irb
