// SwinIR (shifted-)window attention on tcgen05 for sm_100a (bf16 in, fp32 accumulate in TMEM).
//   network_swinir.py:114-145 WindowAttention.forward, :216-237 calculate_mask, :239-279 roll / partition / reverse.
//
// Unit of work = a PAIR of 8x8 windows (128 tokens = the M of one UMMA); a persistent CTA walks its units and, per
// unit, the heads two at a time (64 channels = one 128-byte swizzled row):
//   loaders (4 warps)  : cp.async gather of the Q | K | V rows of the 128 tokens (the cyclic shift is plain address
//                        arithmetic here, which is why this is not TMA) into the UMMA 128B-swizzled layout, 2 stages
//   MMA (1 thread)     : S_h = Q_h K_h^T  (M 128, N 128, K 32; the two off-diagonal 64x64 blocks are ignored)
//                        O_h = P_h V      (M 128, N 64 = both heads of the pair, K 128 keys; P is block diagonal)
//   softmax (2 x 4 warps, one group per head of the pair, thread = token row):
//                        S row from TMEM -> * log2(e) + rel-pos bias (+ shift mask) -> max / exp2 / sum in registers
//                        -> unnormalised P row (bf16) into shared memory in the UMMA A layout -> after the PV MMA:
//                        O row from TMEM * 1/sum -> bf16 -> per-warp transpose buffer -> coalesced global stores.
// The relative-position bias is looked up from a shared-memory copy of the (2w-1)^2 x heads table with the index
// (yi-yj+7)*15 + (xi-xj+7) folded into immediate offsets; the -100 mask is derived from the token coordinates of the
// windows in the last window row / column of the shifted frame. Nothing is materialised in HBM.
#include <cstdio>
#include "common.cuh"
#include "ptx.cuh"
#include "kernels.h"

// The MMA thread polls its barriers with the non-blocking mbarrier.test_wait: try_wait suspends for a few hundred
// cycles per not-ready barrier, and a polling round over six of them reacted to events ~1 us late.
#ifdef IRB_OLD_POLL  // A/B switch (build.py --variant): the blocking poll of the first version
#define IRB_MMA_POLL mbar_try_wait
#define IRB_MMA_POLL_SLEEP 64
#else
#define IRB_MMA_POLL mbar_test_wait
#define IRB_MMA_POLL_SLEEP 20
#endif

namespace irb {

static constexpr int kWaThreads = 11 * 32;      // warp 0: MMA, 1..8: softmax groups A/B, 9..10: loaders (352 threads: 168 registers each)
static constexpr int kWaTile = 128 * 128;       // one [128 tokens][64 ch] bf16 operand tile (16 KB)
static constexpr int kWaStage = 3 * kWaTile;    // Q | K | V
static constexpr int kWaStages = 2;
static constexpr int kWaPBytes = 2 * kWaTile;   // P of one head: 2 k-tiles of [128 rows][64 keys]
static constexpr float kLog2e = 1.4426950408889634f;
// The (2w-1)^2 = 15 x 15 bias table sits in shared memory MIRRORED in x (column = xj - xi + 7, so the biases of keys
// xj, xj + 1 of one key row are adjacent words in ascending order) and TWICE per head, the second copy shifted by one
// word: a token with odd xi reads its (even, odd) key pairs from copy 0, a token with even xi from copy 1, and either
// way every pair is one aligned LDS.64 with an immediate offset (-yj * 16 + xj) from a per-(token, head) pointer. Row
// stride 16 and copy offset 246 words make the 16 lanes of a half-warp (2 token rows x 8 columns) hit 32 distinct banks
// for every key pair (brute-forced; the kernel is sensitive to shared-memory wavefronts: 32 lookups per token and head).
// Two keys per instruction everywhere else too: packed fma.rn.f32x2 / add.rn.f32x2 (the softmax warps' instruction
// count, not memory, was what bounded the kernel).
static constexpr int kWaTabStride = 16;
static constexpr int kWaTabCopy = 246;
static constexpr int kWaTabSize = 488;           // words per head (both copies)

struct WaParams {
    const bf16* qkv;
    long long ld;
    bf16* out;
    long long ldo;
    int n, H, W, heads, d, Cq, shift;
    int nwx, nwy, total_windows, units;
    const float* bias_table;  // [225][heads]
};

// bar[] layout (uint64 each)
// WB_FULL / WB_EMPTY guard the Q | K halves of a stage (free again as soon as the S MMAs that read them are done),
// WB_VFULL / WB_VEMPTY its V third (free after the P V MMAs): Q and K of step n + 2 are loaded a whole step earlier than
// if the stage were released as one
enum { WB_FULL = 0, WB_EMPTY = 2, WB_SFULL = 4, WB_SEMPTY = 6, WB_PFULL = 8, WB_OFULL = 10, WB_OEMPTY = 12, WB_VFULL = 14,
       WB_VEMPTY = 16, WB_COUNT = 18 };

__device__ __forceinline__ long long wa_token_pixel(const WaParams& p, int unit, int r, bool& valid, int& masky, int& maskx) {
    // row r of the unit -> window, token, original-frame pixel; masky / maskx: 0 none, 1: token in the first
    // (ws - shift) rows/cols of a last-row/col window, 2: in the last `shift` rows/cols
    const int gw = unit * 2 + (r >> 6);
    valid = gw < p.total_windows;
    const int per = p.nwx * p.nwy;
    const int img = gw / per, wi = gw - img * per;
    const int wy = wi / p.nwx, wx = wi - wy * p.nwx;
    const int ty = (r & 63) >> 3, tx = r & 7;
    const int sy = wy * 8 + ty, sx = wx * 8 + tx;
    int oy = sy + p.shift, ox = sx + p.shift;
    if (oy >= p.H) oy -= p.H;
    if (ox >= p.W) ox -= p.W;
    masky = (p.shift > 0 && wy == p.nwy - 1) ? (ty < 8 - p.shift ? 1 : 2) : 0;
    maskx = (p.shift > 0 && wx == p.nwx - 1) ? (tx < 8 - p.shift ? 1 : 2) : 0;
    return valid ? ((long long)img * p.H + oy) * p.W + ox : 0;
}

// D2T: compile-time head_dim / 2 (15: SwinIR-M / -L, 16), or 0 = runtime. With a constant the output loop below is
// fully unrolled and its index arithmetic folds into constants (it was 31 % of the kernel's stall samples).
// The same split in two: the window of a row depends only on (unit, r >> 6), which is uniform per loader warp, so the
// run-time divisions are done once per unit instead of once per row.
struct WaWindow {
    int img, wy, wx;
    bool valid;
};
__device__ __forceinline__ WaWindow wa_window(const WaParams& p, int gw) {
    WaWindow w;
    w.valid = gw < p.total_windows;
    const int per = p.nwx * p.nwy;
    w.img = gw / per;
    const int wi = gw - w.img * per;
    w.wy = wi / p.nwx;
    w.wx = wi - w.wy * p.nwx;
    return w;
}
__device__ __forceinline__ long long wa_window_pixel(const WaParams& p, const WaWindow& w, int r) {
    const int ty = (r & 63) >> 3, tx = r & 7;
    int oy = w.wy * 8 + ty + p.shift, ox = w.wx * 8 + tx + p.shift;
    if (oy >= p.H) oy -= p.H;
    if (ox >= p.W) ox -= p.W;
    return w.valid ? ((long long)w.img * p.H + oy) * p.W + ox : 0;
}

// tm: 4-D tensor map {channel, x, y, image} over the qkv tensor, box {64, 8, 8, 1} (one window of one head pair), 128-byte
// swizzle: a window that does not wrap around the image (all but the last window row / column of a shifted frame) is an
// axis-aligned box and lands in shared memory in the UMMA layout with one TMA load per operand. tm4: the same tensor with
// box {64, 4, 1, 1} = half a token row: a wrapping window (shift 4) is 16 such boxes, none of which straddles the wrap;
// TMA derives the swizzle from the destination address, so each lands at its rows of the 1024-byte-aligned tile.
template <int D2T>
__global__ void __launch_bounds__(kWaThreads, 1) window_attention_tc_kernel(const WaParams p,
                                                                            const __grid_constant__ CUtensorMap tm,
                                                                            const __grid_constant__ CUtensorMap tm4) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* stage_base = smem;                                   // [stages][Q|K|V]
    uint8_t* p_base = smem + kWaStages * kWaStage;                // [2 groups][2 k-tiles][128][128 B]
    uint32_t* tbuf = reinterpret_cast<uint32_t*>(p_base + 2 * kWaPBytes);  // [8 warps][32 rows][<= 17 words] output transpose
    float* s_table = reinterpret_cast<float*>(tbuf + 8 * 32 * 17);  // [heads][15][24] bias * log2(e)
    uint64_t* bar = reinterpret_cast<uint64_t*>(s_table + ((p.heads * kWaTabSize + 3) & ~3));
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + WB_COUNT);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // ---- one-time setup: zero the operand tiles (stale rows must be finite: 0 * NaN would poison valid rows),
    // stage the bias table, init barriers, allocate TMEM ----
    {
        uint4* z = reinterpret_cast<uint4*>(smem);
        const int n16 = (kWaStages * kWaStage + 2 * kWaPBytes) / 16;
        for (int i = threadIdx.x; i < n16; i += kWaThreads) z[i] = make_uint4(0u, 0u, 0u, 0u);
        for (int i = threadIdx.x; i < p.heads * kWaTabSize; i += kWaThreads) {
            const int h = i / kWaTabSize, e2 = i - h * kWaTabSize;
            const int copy = e2 >= kWaTabCopy ? 1 : 0;
            const int w = e2 - copy * (kWaTabCopy + 1);          // copy 1 holds the same words one position later
            float v = 0.f;
            if (w >= 0) {
                const int row = w / kWaTabStride, col = w - row * kWaTabStride;  // row = yi - yj + 7, col = xj - xi + 7
                if (row < 15 && col < 15) v = __ldg(p.bias_table + (row * 15 + (14 - col)) * p.heads + h) * kLog2e;
            }
            s_table[i] = v;
        }
    }
    if (warp == 0) {
        if (lane == 0) {
            for (int s = 0; s < 2; ++s) {
                mbar_init(&bar[WB_FULL + s], 1);
                mbar_init(&bar[WB_EMPTY + s], 1);
                mbar_init(&bar[WB_VFULL + s], 1);
                mbar_init(&bar[WB_VEMPTY + s], 1);
                mbar_init(&bar[WB_SFULL + s], 1);
                mbar_init(&bar[WB_SEMPTY + s], 4);
                mbar_init(&bar[WB_PFULL + s], 4);
                mbar_init(&bar[WB_OFULL + s], 1);
                mbar_init(&bar[WB_OEMPTY + s], 4);
            }
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int npairs = p.heads >> 1;
    // TMEM columns: S of group g at g*128, O of group g at 256 + g*64

    if (warp >= 9) {
        // ===================== loaders: warp 9 streams Q | K, warp 10 streams V =====================
        // One lane per stream issues TMA loads: one box per window and operand, or 16 half-row boxes for a window in the
        // last window row / column of a shifted frame, which wraps around the image (the cyclic shift of :243-247).
        const int lw = warp - 9;             // 0: Q | K, 1: V
        const int nt = lw == 0 ? 2 : 1;      // operand tiles of this warp's stream
        const int t0 = lw == 0 ? 0 : 2;
        uint64_t* full = &bar[lw == 0 ? WB_FULL : WB_VFULL];
        uint64_t* empty = &bar[lw == 0 ? WB_EMPTY : WB_VEMPTY];
        if (lane == 0) {
            int cnt = 0;
            for (int u = blockIdx.x; u < p.units; u += gridDim.x) {
                const WaWindow w0 = wa_window(p, u * 2), w1 = wa_window(p, u * 2 + 1);
                for (int hp = 0; hp < npairs; ++hp, ++cnt) {
                    const int s = cnt & 1;
                    mbar_wait(&empty[s], ((cnt >> 1) & 1) ^ 1);
                    uint8_t* st = stage_base + s * kWaStage;
                    mbar_arrive_expect_tx(&full[s], (w1.valid ? 2 : 1) * nt * 8192);
                    for (int wi = 0; wi < 2; ++wi) {
                        const WaWindow& w = wi ? w1 : w0;
                        if (!w.valid) continue;
                        const bool wrap = p.shift > 0 && (w.wy == p.nwy - 1 || w.wx == p.nwx - 1);
                        for (int t = t0; t < t0 + nt; ++t) {
                            uint8_t* dst = st + t * kWaTile + wi * 8192;
                            const int ch = t * p.Cq + hp * 64;
                            if (!wrap) {
                                tma_load_4d(dst, &tm, &full[s], ch, w.wx * 8 + p.shift, w.wy * 8 + p.shift, w.img);
                            } else {
                                for (int hr = 0; hr < 16; ++hr) {  // token row ty = hr >> 1, tokens tx = 4 * (hr & 1) .. + 3
                                    int oy = w.wy * 8 + (hr >> 1) + p.shift, ox = w.wx * 8 + (hr & 1) * 4 + p.shift;
                                    if (oy >= p.H) oy -= p.H;
                                    if (ox >= p.W) ox -= p.W;
                                    tma_load_4d(dst + hr * 512, &tm4, &full[s], ch, ox, oy, w.img);
                                }
                            }
                        }
                    }
                }
            }
        }
    } else if (warp == 0) {
        // ===================== MMA issuer =====================
        // elect.sync rather than lane == 0: ptxas then knows the region is single-threaded and issues the tcgen05
        // instructions directly instead of wrapping each in an election loop
        if (elect_one()) {
            const uint32_t idesc_s = make_idesc_bf16(128, 128, 0, 0);  // Q, K both K-major
            const uint32_t idesc_o = make_idesc_bf16(128, 64, 0, 1);   // P K-major, V MN-major
            int my_pairs = 0;
            for (int u = blockIdx.x; u < p.units; u += gridDim.x) my_pairs += npairs;
            // Polling issue loop: S(cnt) needs its stage loaded and the group's S accumulator drained; P V(cnt) needs the
            // group's P in shared memory and its O accumulator drained. Whichever is ready goes first: with blocking
            // waits in program order, P V(cnt) sat behind the LOAD of cnt + 1 (which itself waits for P V(cnt - 1)),
            // so every head pair paid a full gather latency.
            int s_next = 0, pv_next = 0;
            bool s_done[2] = {false, false}, pv_done[2] = {false, false};
            bool s_loaded = false, v_loaded = false;
            uint32_t spins = 0;
            while (pv_next < my_pairs) {
                bool progress = false;
                if (s_next < my_pairs) {
                    const int st = s_next & 1;
                    if (!s_loaded && IRB_MMA_POLL(&bar[WB_FULL + st], (s_next >> 1) & 1)) s_loaded = true;
                    if (s_loaded) {
                        const uint32_t q_addr = smem_u32(stage_base + st * kWaStage);
                        const uint32_t k_addr = q_addr + kWaTile;
                        for (int g = 0; g < 2; ++g) {
                            if (s_done[g] || !IRB_MMA_POLL(&bar[WB_SEMPTY + g], (s_next & 1) ^ 1)) continue;
                            tc_fence_after();
#pragma unroll
                            for (int ks = 0; ks < 2; ++ks) {
                                const uint64_t ad = make_sdesc_sw128(q_addr + g * 64 + ks * 32, 16, 1024);
                                const uint64_t bd = make_sdesc_sw128(k_addr + g * 64 + ks * 32, 16, 1024);
                                umma_bf16(tmem_base + g * 128, ad, bd, idesc_s, ks);
                            }
                            umma_commit(&bar[WB_SFULL + g]);
                            s_done[g] = true;
                            progress = true;
                        }
                        if (s_done[0] && s_done[1]) {
                            umma_commit(&bar[WB_EMPTY + st]);  // Q, K of this stage are no longer read
                            s_done[0] = s_done[1] = false;
                            s_loaded = false;
                            ++s_next;
                        }
                    }
                }
                if (pv_next < s_next && !v_loaded && IRB_MMA_POLL(&bar[WB_VFULL + (pv_next & 1)], (pv_next >> 1) & 1))
                    v_loaded = true;
                if (pv_next < s_next && v_loaded) {
                    const int st = pv_next & 1;
                    const uint32_t v_addr = smem_u32(stage_base + st * kWaStage) + 2 * kWaTile;
                    for (int g = 0; g < 2; ++g) {
                        if (pv_done[g] || !IRB_MMA_POLL(&bar[WB_PFULL + g], pv_next & 1) ||
                            !IRB_MMA_POLL(&bar[WB_OEMPTY + g], (pv_next & 1) ^ 1))
                            continue;
                        tc_fence_after();
                        const uint32_t pa = smem_u32(p_base + g * kWaPBytes);
#pragma unroll
                        for (int ks = 0; ks < 8; ++ks) {
                            const uint64_t ad = make_sdesc_sw128(pa + (ks >> 2) * kWaTile + (ks & 3) * 32, 16, 1024);
                            const uint64_t bd = make_sdesc_sw128(v_addr + ks * 2048, kWaTile, 1024);
                            umma_bf16(tmem_base + 256 + g * 64, ad, bd, idesc_o, ks);
                        }
                        umma_commit(&bar[WB_OFULL + g]);
                        pv_done[g] = true;
                        progress = true;
                    }
                    if (pv_done[0] && pv_done[1]) {
                        umma_commit(&bar[WB_VEMPTY + st]);  // V of this stage is no longer read
                        pv_done[0] = pv_done[1] = false;
                        v_loaded = false;
                        ++pv_next;
                    }
                }
                if (progress) {
                    spins = 0;
                } else {
                    // back off: this thread shares its scheduler with softmax and loader warps, and the kernel is
                    // sensitive to issue pressure
                    __nanosleep(IRB_MMA_POLL_SLEEP);
                    if (++spins > (1u << 24)) __trap();  // protocol bug: fail the launch instead of hanging
                }
            }
        }
    } else {
        // ===================== softmax + output groups: warps 1..4 (head 2*hp) and 5..8 (head 2*hp + 1) ==========
        const int g = (warp - 1) >> 2;
        const int q = warp & 3;                  // TMEM lane quarter this warp may read
        const int r = q * 32 + lane;             // token row of the unit
        const int win = r >> 6;
        const int yi = (r & 63) >> 3, xi = r & 7;
        const int d2 = D2T > 0 ? D2T : (p.d >> 1);  // output words (bf16 pairs) per head
        const int tstride = d2 | 1;              // odd stride: conflict-free transpose
        uint32_t* tb = tbuf + (warp - 1) * 32 * 17;
        uint8_t* prow = p_base + g * kWaPBytes + win * kWaTile + r * 128;
        const int sw = r & 7;
        const uint32_t lane_addr = uint32_t(q * 32) << 16;
        uint32_t* out32 = reinterpret_cast<uint32_t*>(p.out);
        const long long ldo2 = p.ldo >> 1;
        // Software pipeline over the CTA's (unit, head pair) sequence: P(cnt) is handed to the tensor core, then the
        // softmax of cnt + 1 runs WHILE the P V product of cnt is in flight, and only then O(cnt) is drained. The
        // S -> softmax -> P -> PV -> O chain of one head pair was fully serial before (ncu: the groups idled on the
        // O barrier and the kernel sat at 0.3 of HBM with 35 % of the issue slots used).
        const float NEG = -100.0f * kLog2e;
        long long orow_s = -1;               // softmax-side unit state (cnt + 1 once the loop runs)
        float m00 = 0.f, m01 = 0.f, m10 = 0.f, m11 = 0.f;
        bool any_mask = false;
        auto unit_info = [&](int u) {
            bool valid;
            int my, mx;
            const long long pix = wa_token_pixel(p, u, r, valid, my, mx);
            orow_s = valid ? pix * ldo2 : -1;
            // additive mask per key quadrant (yj >= 4, xj >= 4), in log2 units (network_swinir.py:216-237: the
            // region ids differ when the y parts or the x parts differ)
            const bool ylo = my == 2, yhi = my == 1, xlo = mx == 2, xhi = mx == 1;  // which key halves are masked
            m00 = (ylo || xlo) ? NEG : 0.f;
            m01 = (ylo || xhi) ? NEG : 0.f;
            m10 = (yhi || xlo) ? NEG : 0.f;
            m11 = (yhi || xhi) ? NEG : 0.f;
            any_mask = __any_sync(0xffffffffu, (my | mx) != 0);
        };
        uint32_t pk[32];
        float inv_s = 0.f;
        // S(cnt) row from TMEM -> unnormalised probabilities (bf16 pairs) in pk, 1 / sum in inv_s
#ifdef IRB_WA_TRACE  // timing experiment: clock() trace of softmax warp 1 / lane 0 of CTA 0, printed at exit
        const bool trace = blockIdx.x == 0 && threadIdx.x == 32;
        long long tsv[150];
        int tsi[150];
        int nts = 0;
#define IRB_TS(i) do { if (trace && nts < 150) { tsv[nts] = clock64(); tsi[nts++] = (i); } } while (0)
#else
#define IRB_TS(i) do { } while (0)
#endif
        auto softmax_phase = [&](int cnt, int h) {
            const int copy = (xi & 1) ^ 1;
            const uint32_t tab = smem_u32(s_table + h * kWaTabSize + copy * (kWaTabCopy + 1) + (yi + 7) * kWaTabStride + (7 - xi));
            // the bias pairs do not depend on S: the first half is in flight while this warp waits for the S barrier, the
            // second half while the first is consumed (left to the compiler, every LDS fed the next FFMA2 through ONE
            // register pair: 32 shared-memory round trips in series, ncu r02_s33)
            f32x2 bq[16];
#pragma unroll
            for (int k = 0; k < 16; ++k) bq[k] = lds_f32x2(tab + 4 * (((k & 3) * 2) - (k >> 2) * kWaTabStride));
            IRB_TS(1);
            mbar_wait(&bar[WB_SFULL + g], cnt & 1);
            IRB_TS(2);
            tc_fence_after();
            uint32_t sr[64];
            tmem_ld32_nowait(tmem_base + lane_addr + g * 128 + win * 64, sr);
            tmem_ld32_nowait(tmem_base + lane_addr + g * 128 + win * 64 + 32, sr + 32);
            tmem_ld_wait();
            IRB_TS(10);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar[WB_SEMPTY + g]);
            f32x2 t2[32];
            const f32x2 l2e = splat2(kLog2e);
#pragma unroll
            for (int k = 0; k < 16; ++k)
                t2[k] = fma2(pack2(__uint_as_float(sr[2 * k]), __uint_as_float(sr[2 * k + 1])), l2e, bq[k]);
#pragma unroll
            for (int k = 16; k < 32; ++k) bq[k - 16] = lds_f32x2(tab + 4 * (((k & 3) * 2) - (k >> 2) * kWaTabStride));
#pragma unroll
            for (int k = 16; k < 32; ++k)
                t2[k] = fma2(pack2(__uint_as_float(sr[2 * k]), __uint_as_float(sr[2 * k + 1])), l2e, bq[k - 16]);
            if (any_mask) {
                const f32x2 q00 = splat2(m00), q01 = splat2(m01), q10 = splat2(m10), q11 = splat2(m11);
#pragma unroll
                for (int k = 0; k < 32; ++k) {
                    const int yj = k >> 2, xj = (k & 3) * 2;
                    t2[k] = add2(t2[k], (yj >= 4) ? ((xj >= 4) ? q11 : q10) : ((xj >= 4) ? q01 : q00));
                }
            }
            IRB_TS(11);
            float mxs[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};  // four independent chains
#pragma unroll
            for (int k = 0; k < 32; ++k) {
                float a, b;
                unpack2(t2[k], a, b);
                mxs[k & 3] = fmaxf(mxs[k & 3], fmaxf(a, b));
            }
            const float mx_ = fmaxf(fmaxf(mxs[0], mxs[1]), fmaxf(mxs[2], mxs[3]));
            IRB_TS(12);
            const f32x2 nmx = splat2(-mx_);
            f32x2 s2a = pack2(0.f, 0.f), s2b = pack2(0.f, 0.f);
#pragma unroll
            for (int k = 0; k < 32; ++k) {
                float a, b;
                unpack2(add2(t2[k], nmx), a, b);
                const float e0 = ex2_approx(a), e1 = ex2_approx(b);
                if (k & 1)
                    s2b = add2(s2b, pack2(e0, e1));
                else
                    s2a = add2(s2a, pack2(e0, e1));
                pk[k] = pack_bf16x2(e0, e1);
            }
            float sum;
            {
                float a, b;
                unpack2(add2(s2a, s2b), a, b);
                sum = a + b;
            }
            inv_s = 1.0f / sum;
        };
        long long o_base = -1;                 // output pointers of the store pattern, per unit (see the O phase)
        int o_off[D2T > 0 ? D2T : 1] = {};
        int u_s = blockIdx.x, hp_s = 0;
        if (u_s < p.units) {
            unit_info(u_s);
            softmax_phase(0, g);
        }
        for (int cnt = 0; u_s < p.units; ++cnt) {
            IRB_TS(3);
            // ---- P(cnt) -> shared memory (UMMA A layout) -> tensor core ----
#pragma unroll
            for (int c = 0; c < 8; ++c)
                *reinterpret_cast<uint4*>(prow + ((c ^ sw) << 4)) =
                    make_uint4(pk[c * 4], pk[c * 4 + 1], pk[c * 4 + 2], pk[c * 4 + 3]);
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar[WB_PFULL + g]);
            IRB_TS(4);
            const long long orow = orow_s;
            const float inv = inv_s;
            const int h = hp_s * 2 + g;
            // ---- softmax of cnt + 1 while P V of cnt runs ----
            if (++hp_s == npairs) {
                hp_s = 0;
                u_s += gridDim.x;
                if (u_s < p.units) unit_info(u_s);
            }
            if (u_s < p.units) softmax_phase(cnt + 1, hp_s * 2 + g);
            // ---- O(cnt) = P V / sum ----
            IRB_TS(5);
            mbar_wait(&bar[WB_OFULL + g], cnt & 1);
            IRB_TS(6);
            tc_fence_after();
            uint32_t orr[32];
            tmem_ld32_nowait(tmem_base + lane_addr + 256 + g * 64 + g * 32, orr);
            tmem_ld_wait();
            IRB_TS(8);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar[WB_OEMPTY + g]);
#pragma unroll
            for (int k = 0; k < 16; ++k)
                if (k < d2)
                    tb[lane * tstride + k] =
                        pack_bf16x2(__uint_as_float(orr[2 * k]) * inv, __uint_as_float(orr[2 * k + 1]) * inv);
            __syncwarp();
            IRB_TS(9);
            // coalesced: consecutive lanes write consecutive words of a token's head slice
            if (D2T > 0) {
                // The store pattern (word wi = lane + 32 * it of the warp's 32 x D2T output words -> token row wi / D2T)
                // needs the row pointer of ANOTHER lane's token. Fetching it with two shuffles per word cost ~100
                // cycles per word in series (clock trace r02_s59: 1 500 of the step's 4 200 cycles): the pointers only
                // change with the unit, so they are gathered once per unit (first head pair) as 32-bit offsets from the
                // warp's first token and kept in registers for its other head pairs.
                if ((h >> 1) == 0) {
                    o_base = __shfl_sync(0xffffffffu, orow, 0);
                    const int rel = int(orow - o_base);
#pragma unroll
                    for (int it = 0; it < D2T; ++it) {
                        const int wi = lane + 32 * it;
                        const int row = wi / D2T;
                        o_off[it] = __shfl_sync(0xffffffffu, rel, row) + (wi - row * D2T);
                    }
                }
                if (o_base >= 0) {  // warp-uniform: the 32 rows of a warp belong to one window
                    uint32_t v[D2T > 0 ? D2T : 1];
#pragma unroll
                    for (int it = 0; it < D2T; ++it) {
                        const int wi = lane + 32 * it;
                        const int row = wi / D2T;
                        v[it] = tb[row * tstride + (wi - row * D2T)];
                    }
                    uint32_t* ob = out32 + o_base + h * D2T;
#pragma unroll
                    for (int it = 0; it < D2T; ++it) ob[o_off[it]] = v[it];
                }
            } else {
                const int nwords = 32 * d2;
                for (int wi = lane; wi < nwords; wi += 32) {
                    const int row = wi / d2, k = wi - row * d2;
                    const uint32_t v = tb[row * tstride + k];
                    const long long ob = __shfl_sync(0xffffffffu, orow, row);
                    if (ob >= 0) out32[ob + h * d2 + k] = v;
                }
            }
            __syncwarp();
            IRB_TS(7);
        }
#ifdef IRB_WA_TRACE
        if (trace) {
            static __device__ int printed;
            if (atomicAdd(&printed, 1) == 0)
                for (int i = 1; i < nts; ++i) printf("TS %d %d %lld\n", i, tsi[i], tsv[i] - tsv[i - 1]);
        }
#endif
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, 512);
}

bool window_attention_tc_ok(int dt, long long ld, long long ldo, int H, int W, int C, int heads, int dstride, int ws,
                            int shift) {
    return dt == DT_BF16 && ws == 8 && dstride == 32 && heads >= 2 && heads % 2 == 0 && heads <= 8 &&
           (C / heads) % 2 == 0 && C / heads <= 32 && (shift == 0 || shift == 4) && H % 8 == 0 && W % 8 == 0 &&
           ld % 8 == 0 && ldo % 2 == 0;
}

int window_attention_tc(const bf16* qkv, long long ld, bf16* out, long long ldo, int n, int H, int W, int C, int heads,
                        int shift, const float* bias_table, cudaStream_t stream) {
    IRB_REQUIRE((reinterpret_cast<uintptr_t>(qkv) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 3) == 0,
                "window_attention_tc: alignment");
    WaParams p{};
    p.qkv = qkv;
    p.ld = ld;
    p.out = out;
    p.ldo = ldo;
    p.n = n;
    p.H = H;
    p.W = W;
    p.heads = heads;
    p.d = C / heads;
    p.Cq = heads * 32;
    p.shift = shift;
    p.nwx = W / 8;
    p.nwy = H / 8;
    const long long tw = (long long)n * p.nwx * p.nwy;
    IRB_REQUIRE(tw < (1ll << 30), "window_attention_tc: too many windows");
    p.total_windows = int(tw);
    p.units = int((tw + 1) / 2);
    p.bias_table = bias_table;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = p.units < sms ? p.units : sms;
    const size_t smem = 1024 + size_t(kWaStages) * kWaStage + 2 * kWaPBytes + 8 * 32 * 17 * 4 +
                        ((heads * kWaTabSize + 3) & ~3) * 4 + WB_COUNT * 8 + 64;
    IRB_SMEM_OPTIN((window_attention_tc_kernel<0>), 227 * 1024);
    IRB_SMEM_OPTIN((window_attention_tc_kernel<15>), 227 * 1024);
    IRB_SMEM_OPTIN((window_attention_tc_kernel<16>), 227 * 1024);
    CUtensorMap tm, tm4;
    {
        const uint64_t dims[4] = {uint64_t(3 * p.Cq), uint64_t(W), uint64_t(H), uint64_t(n)};
        const uint64_t strides[3] = {uint64_t(ld) * 2, uint64_t(W) * ld * 2, uint64_t(H) * W * ld * 2};
        const uint32_t box[4] = {64, 8, 8, 1}, box4[4] = {64, 4, 1, 1};
        IRB_TRY(make_tmap_bf16_sw(&tm, qkv, 4, dims, strides, box, 128));
        IRB_TRY(make_tmap_bf16_sw(&tm4, qkv, 4, dims, strides, box4, 128));
    }
    if (p.d == 30)
        window_attention_tc_kernel<15><<<grid, kWaThreads, smem, stream>>>(p, tm, tm4);
    else if (p.d == 32)
        window_attention_tc_kernel<16><<<grid, kWaThreads, smem, stream>>>(p, tm, tm4);
    else
        window_attention_tc_kernel<0><<<grid, kWaThreads, smem, stream>>>(p, tm, tm4);
    IRB_CHECK_CUDA(cudaGetLastError());
    const double px = double(n) * H * W;
    prof_mark("window_attention_tc", C, shift, 4.0 * px * 64 * C, px * (3.0 * p.Cq + C) * 2, stream);
    return 0;
}

}  // namespace irb
