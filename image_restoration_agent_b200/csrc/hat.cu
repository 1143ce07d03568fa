// HAT kernels (AgenticIR/executor/super_resolution/tools/HAT/hat/archs/hat_arch.py) that SwinIR does not have:
//
//   hat_attention  : window attention with up to 16 x 16 = 256 queries per window, for both
//                      (a) the (shifted-)window self-attention of HAB (:165-197 WindowAttention.forward, :266-303 roll /
//                          partition / reverse, :921-941 calculate_mask), window 16, and
//                      (b) the overlapping cross-attention of OCAB (:392-432): queries from the ws x ws window, keys /
//                          values from the ows x ows window around it (nn.Unfold, stride ws, zero padding AFTER the qkv
//                          Linear: a padded key is an exact zero vector whose logit is its position bias alone).
//                    fp32 arithmetic in both precision modes; streaming (online) softmax over 64-key chunks, so neither
//                    the 256 x 576 logits nor the unfolded keys are ever materialised.
//   chan_pool / ca_scale / add_scaled : the channel attention of CAB (:40-58): per-image channel means of the second
//                    convolution, the squeeze / excite 1x1 pair + sigmoid (times conv_scale), and
//                    out = x + cab * scale[img, c] (the conv branch of HAB.forward :306).
#include "../../include/irb200.h"
#include "common.cuh"
#include "kernels.h"
#include <type_traits>
#include <cstring>

namespace irb {

struct HatAttnParams {
    const void* qkv;  // [n*H*W][ld]: q | k | v sections of heads*dstride channels (q already scaled)
    int dt;
    long long ld;
    void* out;        // [n*H*W][ldo], channel head*d + dd
    long long ldo;
    int H, W, heads, d, dstride, ws, ows, pad, shift;
    const float* table;  // [tab_n][heads]
    int tab_w, tab_off, tab_sign, tab_n;
};

static constexpr int kHatKC = 64;   // keys per shared-memory chunk
static constexpr int kHatDP = 32;   // head_dim padded
static constexpr int kHatSub = 16;  // keys per online-softmax step

__device__ __forceinline__ int hat_region(int v, int n, int ws, int shift) {
    return v < n - ws ? 0 : (v < n - shift ? 1 : 2);
}

// CTA = (window, head, image); thread = one query token of the window (blockDim = ws*ws rounded up to a warp).
__global__ void __launch_bounds__(256, 2) hat_attention_kernel(const HatAttnParams p) {
    extern __shared__ float hat_smem[];
    __shared__ __align__(16) float ks[kHatKC][kHatDP];
    __shared__ __align__(16) float vs[kHatKC][kHatDP];
    __shared__ int key_yx[kHatKC];   // (ky << 8) | kx inside the key window, -1 = past the last key
    __shared__ int key_rid[kHatKC];  // shifted-frame region id (self-attention with shift), else 0
    float* tab = hat_smem;           // this head's column of the bias table
    const int ws = p.ws, ows = p.ows, nq = ws * ws, nk = ows * ows, d = p.d;
    const int nwx = p.W / ws;
    const int win = blockIdx.x, head = blockIdx.y, img = blockIdx.z;
    const int wy = win / nwx, wx = win - wy * nwx;
    const int tid = threadIdx.x;
    for (int i = tid; i < p.tab_n; i += blockDim.x) tab[i] = __ldg(p.table + (long long)i * p.heads + head);
    const int sec = p.heads * p.dstride;
    // this thread's query
    const bool qvalid = tid < nq;
    const int qy = qvalid ? tid / ws : 0, qx = qvalid ? tid - qy * ws : 0;
    long long qpix;
    int qrid = 0;
    {
        const int sy = wy * ws + qy, sx = wx * ws + qx;  // shifted-frame coordinates (shift = 0: the frame itself)
        int oy = sy + p.shift, ox = sx + p.shift;
        if (oy >= p.H) oy -= p.H;
        if (ox >= p.W) ox -= p.W;
        qpix = ((long long)img * p.H + oy) * p.W + ox;
        if (p.shift > 0) qrid = 3 * hat_region(sy, p.H, ws, p.shift) + hat_region(sx, p.W, ws, p.shift);
    }
    float q[kHatDP], o[kHatDP];
#pragma unroll
    for (int dd = 0; dd < kHatDP; ++dd) {
        q[dd] = (qvalid && dd < d) ? ld_as_float(p.qkv, qpix * p.ld + head * p.dstride + dd, p.dt) : 0.f;
        o[dd] = 0.f;
    }
    float m = -INFINITY, l = 0.f;
    constexpr float L2E = 1.4426950408889634f;
    for (int k0 = 0; k0 < nk; k0 += kHatKC) {
        __syncthreads();  // the previous chunk is consumed (and, first time round, tab[] is complete)
        for (int e = tid; e < kHatKC * kHatDP; e += blockDim.x) {
            const int kk = e >> 5, dd = e & 31;
            const int j = k0 + kk;
            float kv = 0.f, vv = 0.f;
            int yx = -1, rid = 0;
            if (j < nk) {
                const int ky = j / ows, kx = j - ky * ows;
                yx = (ky << 8) | kx;
                const int sy = wy * ws - p.pad + ky, sx = wx * ws - p.pad + kx;
                if (sy >= 0 && sy < p.H && sx >= 0 && sx < p.W) {  // always true for self-attention (pad = 0)
                    int oy = sy + p.shift, ox = sx + p.shift;
                    if (oy >= p.H) oy -= p.H;
                    if (ox >= p.W) ox -= p.W;
                    if (dd < d) {
                        const long long base = (((long long)img * p.H + oy) * p.W + ox) * p.ld + head * p.dstride + dd;
                        kv = ld_as_float(p.qkv, base + sec, p.dt);
                        vv = ld_as_float(p.qkv, base + 2 * sec, p.dt);
                    }
                    if (p.shift > 0) rid = 3 * hat_region(sy, p.H, ws, p.shift) + hat_region(sx, p.W, ws, p.shift);
                }
            }
            ks[kk][dd] = kv;
            vs[kk][dd] = vv;
            if (dd == 0) {
                key_yx[kk] = yx;
                key_rid[kk] = rid;
            }
        }
        __syncthreads();
#pragma unroll 1
        for (int s0 = 0; s0 < kHatKC; s0 += kHatSub) {
            if (k0 + s0 >= nk) break;
            float s[kHatSub];
            float mloc = -INFINITY;
#pragma unroll
            for (int jj = 0; jj < kHatSub; ++jj) {
                const int kk = s0 + jj;
                const int yx = key_yx[kk];
                float a = 0.f;
                const float4* kr = reinterpret_cast<const float4*>(ks[kk]);
#pragma unroll
                for (int c4 = 0; c4 < kHatDP / 4; ++c4) {
                    const float4 kv = kr[c4];
                    a = fmaf(q[4 * c4 + 0], kv.x, a);
                    a = fmaf(q[4 * c4 + 1], kv.y, a);
                    a = fmaf(q[4 * c4 + 2], kv.z, a);
                    a = fmaf(q[4 * c4 + 3], kv.w, a);
                }
                const int ky = yx >> 8, kx = yx & 255;
                int idx = (p.tab_sign * (ky - qy) + p.tab_off) * p.tab_w + p.tab_sign * (kx - qx) + p.tab_off;
                if (idx < 0) idx += p.tab_n;  // hat_arch.py:911-915: the OCA index goes negative and wraps (PyTorch indexing)
                idx = yx < 0 ? 0 : idx;
                a += tab[idx];
                if (key_rid[kk] != qrid) a += -100.0f;  // :939
                a = yx < 0 ? -INFINITY : a;
                s[jj] = a;
                mloc = fmaxf(mloc, a);
            }
            const float mn = fmaxf(m, mloc);
            const float corr = exp2f((m - mn) * L2E);
            m = mn;
            l *= corr;
#pragma unroll
            for (int dd = 0; dd < kHatDP; ++dd) o[dd] *= corr;
#pragma unroll
            for (int jj = 0; jj < kHatSub; ++jj) {
                const float pj = exp2f((s[jj] - mn) * L2E);
                l += pj;
                const float4* vr = reinterpret_cast<const float4*>(vs[s0 + jj]);
#pragma unroll
                for (int c4 = 0; c4 < kHatDP / 4; ++c4) {
                    const float4 vv = vr[c4];
                    o[4 * c4 + 0] = fmaf(pj, vv.x, o[4 * c4 + 0]);
                    o[4 * c4 + 1] = fmaf(pj, vv.y, o[4 * c4 + 1]);
                    o[4 * c4 + 2] = fmaf(pj, vv.z, o[4 * c4 + 2]);
                    o[4 * c4 + 3] = fmaf(pj, vv.w, o[4 * c4 + 3]);
                }
            }
        }
    }
    if (qvalid) {
        const float inv = 1.0f / l;
        const long long ob = qpix * p.ldo + head * d;
#pragma unroll
        for (int dd = 0; dd < kHatDP; ++dd)
            if (dd < d) st_from_float(p.out, ob + dd, p.dt, o[dd] * inv);
    }
}

// hat_attn_tc.cu
bool hat_attention_tc_ok(int dt, long long ld, long long ldo, int H, int W, int heads, int d, int dstride, int ws, int ows,
                         int shift);
int hat_attention_tc(const bf16* qkv, long long ld, bf16* out, long long ldo, int n, int H, int W, int heads, int d,
                     int ows, int shift, const float* table, cudaStream_t stream);

// mode 0: self-attention (ows = ws; bias index (q - k + ws - 1), shift mask); mode 1: overlapping cross-attention
int hat_attention(const void* qkv, int dt, long long ld, void* out, long long ldo, int n, int H, int W, int heads, int d,
                  int dstride, int ws, int ows, int shift, const float* table, cudaStream_t stream) {
    IRB_REQUIRE(ws >= 1 && ws <= 16, "hat_attention: window size must be <= 16");
    IRB_REQUIRE(ows >= ws && ows <= 64 && (ows - ws) % 2 == 0, "hat_attention: bad overlapping window");
    IRB_REQUIRE(H % ws == 0 && W % ws == 0, "hat_attention: H, W must be multiples of the window");
    IRB_REQUIRE(d >= 1 && d <= kHatDP && dstride >= d, "hat_attention: head_dim must be <= 32");
    IRB_REQUIRE(shift >= 0 && shift < ws && (shift == 0 || ows == ws), "hat_attention: bad shift");
    // bf16, window 16 (overlapping window 24), head_dim <= 32 in a stride-32 layout: tcgen05 kernel; everything else
    // (fp32 mode, other windows): the FMA kernel below
    if (IRB_DIAG_ENV("IRB_HAT_ATTN_TC", 1) && hat_attention_tc_ok(dt, ld, ldo, H, W, heads, d, dstride, ws, ows, shift))
        return hat_attention_tc((const bf16*)qkv, ld, (bf16*)out, ldo, n, H, W, heads, d, ows, shift, table, stream);
    HatAttnParams p{};
    p.qkv = qkv;
    p.dt = dt;
    p.ld = ld;
    p.out = out;
    p.ldo = ldo;
    p.H = H;
    p.W = W;
    p.heads = heads;
    p.d = d;
    p.dstride = dstride;
    p.ws = ws;
    p.ows = ows;
    p.pad = (ows - ws) / 2;
    p.shift = shift;
    p.table = table;
    if (ows == ws) {  // calculate_rpi_sa :882-894: (qy - ky + ws - 1) * (2 ws - 1) + (qx - kx + ws - 1)
        p.tab_w = 2 * ws - 1;
        p.tab_off = ws - 1;
        p.tab_sign = -1;
    } else {          // calculate_rpi_oca :896-919: (ky - qy + ws - ows + 1) * (ws + ows - 1) + (kx - qx + ws - ows + 1)
        p.tab_w = ws + ows - 1;
        p.tab_off = ws - ows + 1;
        p.tab_sign = 1;
    }
    p.tab_n = p.tab_w * p.tab_w;
    const int threads = (ws * ws + 31) / 32 * 32;
    const size_t smem = size_t(p.tab_n) * sizeof(float);
    IRB_REQUIRE(smem <= 24 * 1024, "hat_attention: bias table too large for shared memory");
    dim3 grid((H / ws) * (W / ws), heads, n);
    hat_attention_kernel<<<grid, threads, smem, stream>>>(p);
    IRB_CHECK_CUDA(cudaGetLastError());
    {
        const double px = double(n) * H * W, e = dt == DT_F32 ? 4 : 2;
        prof_mark(ows == ws ? "hat_attention_sa" : "hat_attention_oca", heads * d, shift, 4.0 * px * ows * ows * heads * d,
                  px * 4.0 * heads * d * e, stream);
    }
    return 0;
}

// ---- channel attention --------------------------------------------------------------------------------------------
// sums[img][slab][c] = sum over the slab's pixels of x[img, p, c] (one slot per slab, no atomics: the result does not
// depend on scheduling). Thread = (pixel phase, channel): the block's 1024 threads form 1024 / C phases that stride
// through the slab's pixels with 4 independent accumulators each and meet in shared memory.
static constexpr int kPoolThreads = 1024;
__global__ void __launch_bounds__(kPoolThreads) chan_pool_kernel(const void* __restrict__ x, int dt, long long ldx,
                                                                 long long hw, int C, int slabs,
                                                                 float* __restrict__ sums) {
    extern __shared__ float pool_sm[];  // [phases][C]
    const int img = blockIdx.y, slab = blockIdx.x;
    const int phases = kPoolThreads / C > 0 ? kPoolThreads / C : 1;
    const long long per = (hw + slabs - 1) / slabs;
    const long long p0 = slab * per, p1 = p0 + per < hw ? p0 + per : hw;
    const long long base = (long long)img * hw;
    for (int t = threadIdx.x; t < phases * C; t += kPoolThreads) {
        const int ph = t / C, c = t - ph * C;
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
        long long px = p0 + ph;
        for (; px + 3 * phases < p1; px += 4 * phases) {
            a0 += ld_as_float(x, (base + px) * ldx + c, dt);
            a1 += ld_as_float(x, (base + px + phases) * ldx + c, dt);
            a2 += ld_as_float(x, (base + px + 2 * phases) * ldx + c, dt);
            a3 += ld_as_float(x, (base + px + 3 * phases) * ldx + c, dt);
        }
        for (; px < p1; px += phases) a0 += ld_as_float(x, (base + px) * ldx + c, dt);
        pool_sm[t] = (a0 + a1) + (a2 + a3);
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += kPoolThreads) {
        float acc = 0.f;
        for (int ph = 0; ph < phases; ++ph) acc += pool_sm[ph * C + c];
        sums[((long long)img * slabs + slab) * C + c] = acc;
    }
}

// The same with 4 channels per thread (8-byte bf16 / 16-byte fp32 loads): C % 4 == 0, rows 8 / 16-byte aligned.
// 512-thread blocks at <= 64 registers so that two are resident per SM (the first version's 1024-thread block at 63
// registers was alone on its SM, 64 KB of loads in flight, 4 waves of blocks: 0.34 of the DRAM peak, ncu
// profiles/r02_s82); every pass is a batch of U raw loads (bf16: 16 x 8 bytes, 128 KB in flight per SM), the last one
// predicated instead of a one-load-at-a-time remainder loop.
static constexpr int kPool4Threads = 512;
template <typename T>
__global__ void __launch_bounds__(kPool4Threads, 2) chan_pool4_kernel(const T* __restrict__ x, long long ldx, long long hw,
                                                                      int C, int slabs, float* __restrict__ sums) {
    extern __shared__ float pool_sm[];  // [phases][C]
    const int img = blockIdx.y, slab = blockIdx.x;
    const int C4 = C >> 2;
    const int phases = kPool4Threads / C4;
    const long long per = (hw + slabs - 1) / slabs;
    const long long p0 = slab * per, p1 = p0 + per < hw ? p0 + per : hw;
    const int t = threadIdx.x;
    if (t < phases * C4) {
        const int ph = t / C4, c = (t - ph * C4) * 4;
        const T* xb = x + (long long)img * hw * ldx + c;
        constexpr int U = sizeof(T) == 4 ? 8 : 16;  // independent loads in flight per thread
        typedef typename std::conditional<sizeof(T) == 4, float4, uint2>::type Raw;
        float4 a[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) a[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        const long long step = (long long)phases * ldx;
        for (long long px = p0 + ph; px < p1; px += U * phases) {
            Raw v[U];
            const T* q = xb + px * ldx;
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (px + u * phases < p1)
                    v[u] = *reinterpret_cast<const Raw*>(q + u * step);
                else
                    memset(&v[u], 0, sizeof(Raw));   // +0.0 in both formats
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                float4 f;
                if constexpr (sizeof(T) == 4) {
                    f = v[u];
                } else {
                    unpack_bf16x2(v[u].x, f.x, f.y);
                    unpack_bf16x2(v[u].y, f.z, f.w);
                }
                a[u & 3].x += f.x;
                a[u & 3].y += f.y;
                a[u & 3].z += f.z;
                a[u & 3].w += f.w;
            }
        }
        float* o = pool_sm + ph * C + c;
        o[0] = (a[0].x + a[1].x) + (a[2].x + a[3].x);
        o[1] = (a[0].y + a[1].y) + (a[2].y + a[3].y);
        o[2] = (a[0].z + a[1].z) + (a[2].z + a[3].z);
        o[3] = (a[0].w + a[1].w) + (a[2].w + a[3].w);
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += kPool4Threads) {
        float acc = 0.f;
        for (int ph = 0; ph < phases; ++ph) acc += pool_sm[ph * C + c];
        sums[((long long)img * slabs + slab) * C + c] = acc;
    }
}

// scale[img][c] = out_scale * sigmoid(W2 . relu(W1 . mean + b1) + b2)   (ChannelAttention.forward, hat_arch.py:47-58)
// slab groups of the first phase: one group per (C / 4)-thread slice of the block when C % 4 == 0, per C threads otherwise
__host__ __device__ inline int ca_groups(int threads, int C) {
    const int per = (C % 4 == 0) ? C / 4 : C;
    return threads / per > 0 ? threads / per : 1;
}
__global__ void ca_scale_kernel(const float* __restrict__ sums, int slabs, float inv_hw, int C, int Cs, const float* __restrict__ w1,
                                const float* __restrict__ b1, const float* __restrict__ w2, const float* __restrict__ b2,
                                float out_scale, float* __restrict__ scale) {
    extern __shared__ float ca_sm[];
    float* mean = ca_sm;             // [C]
    float* hid = ca_sm + C;          // [Cs]
    float* part = ca_sm + C + Cs;    // [groups][C]
    const int img = blockIdx.x;
    const int groups = ca_groups(blockDim.x, C);
    if (C % 4 == 0) {
        // thread = (slab group, 4 channels): 16-byte loads, blockDim / (C / 4) groups (22 at C = 180: a few hundred slabs
        // are 4 batches of 4 independent loads per thread, not 30 rounds of scalar ones)
        const int C4 = C >> 2;
        const float4* s4 = reinterpret_cast<const float4*>(sums + (long long)img * slabs * C);
        for (int t = threadIdx.x; t < groups * C4; t += blockDim.x) {
            const int gq = t / C4, c4 = t - gq * C4;
            float4 a[4];   // fixed association: the result does not depend on scheduling
#pragma unroll
            for (int u = 0; u < 4; ++u) a[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int sl = gq; sl < slabs; sl += 4 * groups) {
                float4 v[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    v[u] = sl + u * groups < slabs ? s4[(long long)(sl + u * groups) * C4 + c4] : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    a[u].x += v[u].x;
                    a[u].y += v[u].y;
                    a[u].z += v[u].z;
                    a[u].w += v[u].w;
                }
            }
            float* o = part + gq * C + c4 * 4;
            o[0] = (a[0].x + a[1].x) + (a[2].x + a[3].x);
            o[1] = (a[0].y + a[1].y) + (a[2].y + a[3].y);
            o[2] = (a[0].z + a[1].z) + (a[2].z + a[3].z);
            o[3] = (a[0].w + a[1].w) + (a[2].w + a[3].w);
        }
    } else {
        for (int t = threadIdx.x; t < groups * C; t += blockDim.x) {
            const int gq = t / C, c = t - gq * C;
            float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
            int sl = gq;
            for (; sl + 3 * groups < slabs; sl += 4 * groups) {
                a0 += sums[((long long)img * slabs + sl) * C + c];
                a1 += sums[((long long)img * slabs + sl + groups) * C + c];
                a2 += sums[((long long)img * slabs + sl + 2 * groups) * C + c];
                a3 += sums[((long long)img * slabs + sl + 3 * groups) * C + c];
            }
            for (; sl < slabs; sl += groups) a0 += sums[((long long)img * slabs + sl) * C + c];
            part[t] = (a0 + a1) + (a2 + a3);
        }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float a = 0.f;
        for (int gq = 0; gq < groups; ++gq) a += part[gq * C + c];
        mean[c] = a * inv_hw;
    }
    __syncthreads();
    for (int j = threadIdx.x >> 5; j < Cs; j += blockDim.x >> 5) {  // one warp per squeezed channel
        float a = 0.f;
        for (int c = threadIdx.x & 31; c < C; c += 32) a = fmaf(w1[(long long)j * C + c], mean[c], a);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
        if ((threadIdx.x & 31) == 0) hid[j] = fmaxf(a + b1[j], 0.f);
    }
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float a = b2[c];
        for (int j = 0; j < Cs; ++j) a = fmaf(w2[(long long)c * Cs + j], hid[j], a);
        scale[(long long)img * C + c] = out_scale / (1.0f + expf(-a));
    }
}

// out[p, c] = x[p, c] + y[p, c] * scale[img(p), c]; four channels per thread (8-byte bf16 / 16-byte fp32 accesses)
template <typename T>
__global__ void add_scaled_kernel(const T* __restrict__ x, long long ldx, const T* __restrict__ y, long long ldy,
                                  T* __restrict__ out, long long ldo, long long rows, long long hw, int C,
                                  const float* __restrict__ scale) {
    const int c4 = C >> 2;
    const long long total = rows * c4;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const long long r = i / c4;
        const int c = int(i - r * c4) * 4;
        const float4 sc = *reinterpret_cast<const float4*>(scale + (r / hw) * C + c);
        float4 a, b;
        if (sizeof(T) == 4) {
            a = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(x) + r * ldx + c);
            b = *reinterpret_cast<const float4*>(reinterpret_cast<const float*>(y) + r * ldy + c);
        } else {
            const uint2 ua = *reinterpret_cast<const uint2*>(reinterpret_cast<const bf16*>(x) + r * ldx + c);
            const uint2 ub = *reinterpret_cast<const uint2*>(reinterpret_cast<const bf16*>(y) + r * ldy + c);
            unpack_bf16x2(ua.x, a.x, a.y);
            unpack_bf16x2(ua.y, a.z, a.w);
            unpack_bf16x2(ub.x, b.x, b.y);
            unpack_bf16x2(ub.y, b.z, b.w);
        }
        const float4 o = make_float4(fmaf(b.x, sc.x, a.x), fmaf(b.y, sc.y, a.y), fmaf(b.z, sc.z, a.z), fmaf(b.w, sc.w, a.w));
        if (sizeof(T) == 4) {
            *reinterpret_cast<float4*>(reinterpret_cast<float*>(out) + r * ldo + c) = o;
        } else {
            uint2 uo;
            uo.x = pack_bf16x2(o.x, o.y);
            uo.y = pack_bf16x2(o.z, o.w);
            *reinterpret_cast<uint2*>(reinterpret_cast<bf16*>(out) + r * ldo + c) = uo;
        }
    }
}

int channel_attention_slabs(int n, long long hw) {
    int slabs = int(hw / 256 > 0 ? hw / 256 : 1);
    const int cap = 148 * 4 / (n > 0 ? n : 1);
    if (slabs > cap) slabs = cap > 0 ? cap : 1;
    return slabs;
}

// scale[img][c] for CAB's channel attention of y (the second convolution's output); sums is scratch
// [n][channel_attention_slabs(n, hw)][C]
int channel_attention_scale(const void* y, int dt, long long ldy, int n, long long hw, int C, int Cs, const float* w1,
                            const float* b1, const float* w2, const float* b2, float out_scale, float* sums,
                            float* scale, cudaStream_t stream) {
    IRB_REQUIRE(C >= 1 && C <= 1024 && Cs >= 1 && Cs <= 256, "channel_attention: bad channel counts");
    const int slabs = channel_attention_slabs(n, hw);
    const size_t esz = dt == DT_F32 ? 4 : 2;
    if (C % 4 == 0 && C / 4 <= kPool4Threads && (ldy * esz) % (4 * esz) == 0 &&
        (reinterpret_cast<uintptr_t>(y) & (4 * esz - 1)) == 0) {
        const int phases = kPool4Threads / (C / 4);
        const size_t sm = size_t(phases) * C * sizeof(float);
        if (dt == DT_F32)
            chan_pool4_kernel<float><<<dim3(slabs, n), kPool4Threads, sm, stream>>>((const float*)y, ldy, hw, C, slabs, sums);
        else
            chan_pool4_kernel<bf16><<<dim3(slabs, n), kPool4Threads, sm, stream>>>((const bf16*)y, ldy, hw, C, slabs, sums);
    } else {
        const int phases = kPoolThreads / C > 0 ? kPoolThreads / C : 1;
        chan_pool_kernel<<<dim3(slabs, n), kPoolThreads, size_t(phases) * C * sizeof(float), stream>>>(y, dt, ldy, hw, C,
                                                                                                      slabs, sums);
    }
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("chan_pool", C, 0, 0.0, double(n) * hw * C * (dt == DT_F32 ? 4 : 2), stream);
    {
        const int groups = ca_groups(1024, C);
        ca_scale_kernel<<<n, 1024, size_t(C + Cs + groups * C) * sizeof(float), stream>>>(sums, slabs, 1.0f / float(hw), C,
                                                                                         Cs, w1, b1, w2, b2, out_scale,
                                                                                         scale);
    }
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("ca_scale", C, Cs, 0.0, 0.0, stream);
    return 0;
}

int add_scaled(const void* x, long long ldx, const void* y, long long ldy, void* out, long long ldo, int dt, long long rows,
               long long hw, int C, const float* scale, cudaStream_t stream) {
    IRB_REQUIRE(C % 4 == 0 && ldx % 4 == 0 && ldy % 4 == 0 && ldo % 4 == 0, "add_scaled: channel counts must be multiples of 4");
    const long long total = rows * (C / 4);
    const long long b = ceil_div_ll(total, 256);
    const int blocks = int(b > 148 * 16 ? 148 * 16 : b);
    if (dt == DT_F32)
        add_scaled_kernel<float><<<blocks, 256, 0, stream>>>((const float*)x, ldx, (const float*)y, ldy, (float*)out, ldo,
                                                             rows, hw, C, scale);
    else
        add_scaled_kernel<bf16><<<blocks, 256, 0, stream>>>((const bf16*)x, ldx, (const bf16*)y, ldy, (bf16*)out, ldo, rows,
                                                            hw, C, scale);
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("add_scaled", C, 0, 0.0, 3.0 * rows * C * (dt == DT_F32 ? 4 : 2), stream);
    return 0;
}

}  // namespace irb

extern "C" int irb_hat_attention(const void* d_qkv, int dtype, long long ld, void* d_out, long long ldo, int n, int H,
                                 int W, int heads, int head_dim, int dstride, int window, int overlap_window, int shift,
                                 const float* d_bias_table, void* stream) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        irb::set_error("no CUDA device: irb200 has no CPU fallback");
        return 2;
    }
    return irb::hat_attention(d_qkv, dtype, ld, d_out, ldo, n, H, W, heads, head_dim, dstride, window, overlap_window,
                              shift, d_bias_table, (cudaStream_t)stream);
}
