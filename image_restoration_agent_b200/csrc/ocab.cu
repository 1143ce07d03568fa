// X-Restormer overlapping cross-attention (xrestormer_arch.py:190-237, OCAB.forward) on a channels-last qkv tensor.
//
//   q: the ws x ws tokens of a window; k, v: the ows x ows tokens of the window enlarged by (ows - ws) / 2 on every side
//   (nn.Unfold with zero padding: positions outside the image have k = v = 0 but still take part in the softmax);
//   logit(q, key (i, j)) = (q * d^-0.5) . k + (q * d^-0.5) . rel_width[j - y + ows - 1] + (q * d^-0.5) . rel_height[i - x + ows - 1]
//   for the query at row x, column y of the window (RelPosEmb, :64-103); softmax over the ows^2 keys; out = P v.
//
// CTA = (window, head), one thread per query token (ws^2 = 64 threads). K and V of the enlarged window are staged in
// shared memory as fp32; every thread first forms its 2 (2 ows - 1) relative-position logits, then runs an online
// softmax over the keys. fp32 arithmetic in both precision modes (bf16 only changes the storage type). This op is
// 2.5 % of the X-Restormer FLOPs; it is written for parity, not yet for the tensor pipe.
#include "common.cuh"
#include "kernels.h"

namespace irb {

template <typename T, int D>
__global__ void __launch_bounds__(64) ocab_kernel(const T* __restrict__ qkv, long long ld, T* __restrict__ out,
                                                  long long ldo, int H, int W, int heads, int ws, int ows,
                                                  const float* __restrict__ rel_h, const float* __restrict__ rel_w,
                                                  float scale) {
    extern __shared__ float sm[];
    const int nk = ows * ows;
    float* sk = sm;                 // [nk][D]
    float* sv = sk + nk * D;        // [nk][D]
    float* srh = sv + nk * D;       // [2 ows - 1][D]
    float* srw = srh + (2 * ows - 1) * D;
    const int nwx = W / ws, nwy = H / ws;
    const int win = blockIdx.x, head = blockIdx.y, img = blockIdx.z;
    const int wy = win / nwx, wx = win - wy * nwx;
    (void)nwy;
    const int inner = heads * D;
    const int pad = (ows - ws) / 2;
    const int tid = threadIdx.x;
    const T* base = qkv + (long long)img * H * W * ld;
    // stage K, V (zero outside the image) and the two relative-position tables
    for (int e = tid; e < nk * D; e += blockDim.x) {
        const int key = e / D, c = e - key * D;
        const int ki = key / ows, kj = key - ki * ows;
        const int y = wy * ws - pad + ki, x = wx * ws - pad + kj;
        float kv = 0.f, vv = 0.f;
        if (y >= 0 && y < H && x >= 0 && x < W) {
            const long long px = ((long long)y * W + x) * ld;
            kv = ld_as_float(base, px + inner + head * D + c, sizeof(T) == 4 ? DT_F32 : DT_BF16);
            vv = ld_as_float(base, px + 2 * inner + head * D + c, sizeof(T) == 4 ? DT_F32 : DT_BF16);
        }
        sk[e] = kv;
        sv[e] = vv;
    }
    for (int e = tid; e < (2 * ows - 1) * D; e += blockDim.x) {
        srh[e] = __ldg(rel_h + e);
        srw[e] = __ldg(rel_w + e);
    }
    __syncthreads();
    if (tid >= ws * ws) return;
    const int qx = tid / ws, qy = tid - qx * ws;  // row, column of the query inside the window
    const long long qpx = (long long)(wy * ws + qx) * W + (wx * ws + qy);
    float q[D];
#pragma unroll
    for (int c = 0; c < D; ++c)
        q[c] = ld_as_float(base, qpx * ld + head * D + c, sizeof(T) == 4 ? DT_F32 : DT_BF16) * scale;
    // relative-position logits of this query: lw[j] for key column j, lh[i] for key row i (MAXO >= ows)
    constexpr int MAXO = 16;
    float lw[MAXO], lh[MAXO];
#pragma unroll
    for (int j = 0; j < MAXO; ++j) {
        lw[j] = 0.f;
        lh[j] = 0.f;
        if (j < ows) {
            const float* rw = srw + (j - qy + ows - 1) * D;
            const float* rh = srh + (j - qx + ows - 1) * D;
            float a = 0.f, b = 0.f;
#pragma unroll
            for (int c = 0; c < D; ++c) {
                a = fmaf(q[c], rw[c], a);
                b = fmaf(q[c], rh[c], b);
            }
            lw[j] = a;
            lh[j] = b;
        }
    }
    float m = -INFINITY, l = 0.f, o[D];
#pragma unroll
    for (int c = 0; c < D; ++c) o[c] = 0.f;
#pragma unroll 1
    for (int ki = 0; ki < ows; ++ki) {
        float lhi = 0.f;
#pragma unroll
        for (int j = 0; j < MAXO; ++j) lhi = (j == ki) ? lh[j] : lhi;
#pragma unroll
        for (int kj = 0; kj < MAXO; ++kj) {
            if (kj >= ows) break;
            const float* kr = sk + (ki * ows + kj) * D;
            float s = lhi + lw[kj];
#pragma unroll
            for (int c = 0; c < D; ++c) s = fmaf(q[c], kr[c], s);
            const float mn = fmaxf(m, s);
            const float corr = __expf(m - mn), pj = __expf(s - mn);
            l = l * corr + pj;
            const float* vr = sv + (ki * ows + kj) * D;
#pragma unroll
            for (int c = 0; c < D; ++c) o[c] = fmaf(pj, vr[c], o[c] * corr);
            m = mn;
        }
    }
    const float inv = 1.0f / l;
    T* op = out + ((long long)img * H * W + qpx) * ldo + head * D;
#pragma unroll
    for (int c = 0; c < D; ++c) st_from_float(op, c, sizeof(T) == 4 ? DT_F32 : DT_BF16, o[c] * inv);
}

// qkv [n, H, W, 3 * heads * 16] (q | k | v, head-major) -> out [n, H, W, heads * 16]
int ocab_attention(const void* qkv, int dt, long long ld, void* out, long long ldo, int Nimg, int H, int W, int heads,
                   int dim_head, int ws, int ows, const float* rel_h, const float* rel_w, cudaStream_t stream) {
    IRB_REQUIRE(dim_head == 16, "ocab: spatial_dim_head must be 16 (every X-Restormer config of the agent)");
    IRB_REQUIRE(ws >= 1 && ws * ws <= 64 && ows >= ws && ows <= 16 && (ows - ws) % 2 == 0, "ocab: window sizes");
    IRB_REQUIRE(H % ws == 0 && W % ws == 0, "ocab: H and W must be multiples of the window size");
    IRB_REQUIRE(heads >= 1 && heads <= 65535 && Nimg <= 65535, "ocab: grid limits");
    const float scale = 1.0f / sqrtf(float(dim_head));
    const size_t smem = (size_t(2) * ows * ows * 16 + size_t(2) * (2 * ows - 1) * 16) * sizeof(float);
    dim3 grid((H / ws) * (W / ws), heads, Nimg);
    if (dt == DT_F32)
        ocab_kernel<float, 16><<<grid, 64, smem, stream>>>((const float*)qkv, ld, (float*)out, ldo, H, W, heads, ws, ows,
                                                           rel_h, rel_w, scale);
    else
        ocab_kernel<bf16, 16><<<grid, 64, smem, stream>>>((const bf16*)qkv, ld, (bf16*)out, ldo, H, W, heads, ws, ows,
                                                          rel_h, rel_w, scale);
    IRB_CHECK_CUDA(cudaGetLastError());
    const double px = double(Nimg) * H * W, e = dt == DT_F32 ? 4 : 2;
    const double nk = double(ows) * ows;
    prof_mark("ocab", heads * 16, ows, px * heads * (4.0 * nk * 16 + 4.0 * (2 * ows - 1) * 16),
              px * heads * 16 * e * (1 + 1) + px * heads * 16 * e * 2 * (nk / (ws * ws)), stream);
    return 0;
}

}  // namespace irb
