// SwinIR (shifted-)window multi-head self-attention (network_swinir.py:114-145 WindowAttention.forward,
// :216-237 calculate_mask, :239-279 the roll / window_partition / window_reverse / roll-back around it).
//
// Input : qkv [n, H, W, 3*heads*dstride] channels-last, sections q | k | v, each head-major with `dstride`
//         channels per head of which the first d = C/heads are real (q already multiplied by d^-0.5).
// Output: out [n, H, W, >=C] with channel = head*d + dd, written at the ORIGINAL (un-rolled) pixel positions.
// The cyclic shift is pure addressing: window token (ty,tx) of window (wy,wx) lives at shifted-frame coordinate
// (wy*ws+ty, wx*ws+tx) = original coordinate ((y+shift) mod H, (x+shift) mod W). The -100 mask is derived from the
// shifted-frame region ids, the relative-position bias from the token coordinates: nothing is materialised.
//
// This file holds the fp32 FMA kernel (fp32 precision mode, any window <= 8, any head_dim <= 32).
#include "../../include/irb200.h"
#include "common.cuh"
#include "kernels.h"

namespace irb {

struct WinAttnParams {
    const void* qkv;
    int dt;
    long long ld;
    void* out;
    long long ldo;
    int H, W, C, heads, d, dstride, ws, shift;
    const float* bias_table;  // [(2ws-1)^2][heads]
};

__device__ __forceinline__ int region_id(int v, int n, int ws, int shift) {
    return v < n - ws ? 0 : (v < n - shift ? 1 : 2);
}

__global__ void __launch_bounds__(256) window_attention_simt_kernel(const WinAttnParams p) {
    constexpr int MAXN = 64, DP = 33;
    __shared__ float qs[MAXN][DP], ks[MAXN][DP], vs[MAXN][DP];
    __shared__ float S[MAXN][MAXN + 1];
    __shared__ int tok_row[MAXN];  // global pixel index of each token (original frame)
    __shared__ int tok_rid[MAXN];  // shifted-frame region id
    const int ws = p.ws, N = ws * ws, d = p.d;
    const int nwx = p.W / ws, nwy = p.H / ws;
    const int win = blockIdx.x, head = blockIdx.y, img = blockIdx.z;
    const int wy = win / nwx, wx = win % nwx;
    (void)nwy;
    const int tid = threadIdx.x;
    if (tid < N) {
        const int ty = tid / ws, tx = tid % ws;
        const int sy = wy * ws + ty, sx = wx * ws + tx;  // shifted-frame coordinates
        const int oy = (sy + p.shift) % p.H, ox = (sx + p.shift) % p.W;
        tok_row[tid] = (img * p.H + oy) * p.W + ox;
        tok_rid[tid] = p.shift > 0 ? 3 * region_id(sy, p.H, ws, p.shift) + region_id(sx, p.W, ws, p.shift) : 0;
    }
    __syncthreads();
    const int sec = p.heads * p.dstride;
    for (int e = tid; e < N * 32; e += 256) {
        const int t = e >> 5, dd = e & 31;
        float q = 0.f, k = 0.f, v = 0.f;
        if (dd < d) {
            const long long base = (long long)tok_row[t] * p.ld + head * p.dstride + dd;
            q = ld_as_float(p.qkv, base, p.dt);
            k = ld_as_float(p.qkv, base + sec, p.dt);
            v = ld_as_float(p.qkv, base + 2 * sec, p.dt);
        }
        qs[t][dd] = q;
        ks[t][dd] = k;
        vs[t][dd] = v;
    }
    __syncthreads();
    // scores: thread -> row i = tid/4, columns j = (tid%4) + 4*jj
    const int i = tid >> 2, jl = tid & 3;
    float sc[16];
    float mx = -INFINITY;
    if (i < N) {
        const int iy = i / ws, ix = i % ws;
#pragma unroll
        for (int jj = 0; jj < 16; ++jj) {
            const int j = jl + 4 * jj;
            float a = -INFINITY;
            if (j < N) {
                a = 0.f;
#pragma unroll
                for (int dd = 0; dd < 32; ++dd) a = fmaf(qs[i][dd], ks[j][dd], a);
                const int jy = j / ws, jx = j % ws;
                const int idx = (iy - jy + ws - 1) * (2 * ws - 1) + (ix - jx + ws - 1);  // :93-103
                a += __ldg(p.bias_table + idx * p.heads + head);
                if (tok_rid[i] != tok_rid[j]) a += -100.0f;  // :236
                mx = fmaxf(mx, a);
            }
            sc[jj] = a;
        }
    }
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
    float sum = 0.f;
    if (i < N) {
#pragma unroll
        for (int jj = 0; jj < 16; ++jj) {
            const float e = (jl + 4 * jj < N) ? expf(sc[jj] - mx) : 0.f;
            sc[jj] = e;
            sum += e;
        }
    }
    sum += __shfl_xor_sync(0xffffffffu, sum, 1);
    sum += __shfl_xor_sync(0xffffffffu, sum, 2);
    if (i < N) {
        const float inv = 1.0f / sum;
#pragma unroll
        for (int jj = 0; jj < 16; ++jj) {
            const int j = jl + 4 * jj;
            if (j < N) S[i][j] = sc[jj] * inv;
        }
    }
    __syncthreads();
    // out[i][dd], dd = (tid%4) + 4*e
    if (i < N) {
        float o[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = 0.f;
        for (int j = 0; j < N; ++j) {
            const float pij = S[i][j];
#pragma unroll
            for (int e = 0; e < 8; ++e) o[e] = fmaf(pij, vs[j][jl + 4 * e], o[e]);
        }
        const long long ob = (long long)tok_row[i] * p.ldo + head * d;
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int dd = jl + 4 * e;
            if (dd < d) st_from_float(p.out, ob + dd, p.dt, o[e]);
        }
    }
}

int window_attention_simt(const void* qkv, int dt, long long ld, void* out, long long ldo, int n, int H, int W, int C,
                          int heads, int dstride, int ws, int shift, const float* bias_table, cudaStream_t stream) {
    IRB_REQUIRE(ws >= 1 && ws <= 8, "window_attention: window size must be <= 8");
    IRB_REQUIRE(H % ws == 0 && W % ws == 0, "window_attention: H, W must be multiples of the window");
    IRB_REQUIRE(C % heads == 0 && C / heads <= 32 && dstride >= C / heads, "window_attention: head_dim must be <= 32");
    IRB_REQUIRE(shift >= 0 && shift < ws, "window_attention: 0 <= shift < window");
    WinAttnParams p{qkv, dt, ld, out, ldo, H, W, C, heads, C / heads, dstride, ws, shift, bias_table};
    dim3 grid((H / ws) * (W / ws), heads, n);
    window_attention_simt_kernel<<<grid, 256, 0, stream>>>(p);
    IRB_CHECK_CUDA(cudaGetLastError());
    {
        const double px = double(n) * H * W, e = dt == DT_F32 ? 4 : 2;
        prof_mark("window_attention_simt", C, shift, 4.0 * px * ws * ws * C, px * 4.0 * C * e, stream);
    }
    return 0;
}

// swin_attn_tc.cu
bool window_attention_tc_ok(int dt, long long ld, long long ldo, int H, int W, int C, int heads, int dstride, int ws,
                            int shift);
int window_attention_tc(const bf16* qkv, long long ld, bf16* out, long long ldo, int n, int H, int W, int C, int heads,
                        int shift, const float* bias_table, cudaStream_t stream);

int window_attention(const void* qkv, int dt, long long ld, void* out, long long ldo, int n, int H, int W, int C,
                     int heads, int dstride, int ws, int shift, const float* bias_table, cudaStream_t stream) {
    // bf16, 8x8 windows, head_dim <= 32 padded to 32: tcgen05 kernel; everything else (fp32 mode, window 7): FMA kernel
    if (window_attention_tc_ok(dt, ld, ldo, H, W, C, heads, dstride, ws, shift))
        return window_attention_tc((const bf16*)qkv, ld, (bf16*)out, ldo, n, H, W, C, heads, shift, bias_table, stream);
    return window_attention_simt(qkv, dt, ld, out, ldo, n, H, W, C, heads, dstride, ws, shift, bias_table, stream);
}

}  // namespace irb

extern "C" int irb_window_attention(const void* d_qkv, int dtype, long long ld, void* d_out, long long ldo, int n,
                                    int H, int W, int C, int heads, int dstride, int window, int shift,
                                    const float* d_bias_table, void* stream) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        irb::set_error("no CUDA device: irb200 has no CPU fallback");
        return 2;
    }
    return irb::window_attention(d_qkv, dtype, ld, d_out, ldo, n, H, W, C, heads, dstride, window, shift,
                                 d_bias_table, (cudaStream_t)stream);
}
