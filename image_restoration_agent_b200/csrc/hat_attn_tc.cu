// HAT window attention on tcgen05 for sm_100a (bf16 in, fp32 accumulate in TMEM): window 16 (256 queries), for
//   MODE 0: the (shifted-)window self-attention of HAB   (hat_arch.py:165-197, :266-303, mask :921-941), 256 keys
//   MODE 1: the overlapping cross-attention of OCAB       (hat_arch.py:392-432), 24 x 24 = 576 zero-padded keys
//
// A window's 256 queries are two M = 128 blocks; the keys are streamed in chunks of whole key rows (MODE 0: 4 rows of
// 16 = 64 keys, MODE 1: 2 rows of 24 = 48 keys) so that the relative-position bias of key jj of a chunk sits at a
// compile-time offset from a per-(token, chunk) base pointer. One STEP = (window, head pair, M block, key chunk):
//   loaders (3 warps, one per ring stage)
//                     : Q (128 tokens), K and V (the chunk's keys) of the head pair (64 channels = one 128-byte swizzled
//                       row) into the UMMA layout. A window that does not wrap around the image is an axis-aligned box of
//                       the qkv tensor: three TMA loads by one lane ({64 ch, 16, 8} for Q, {64, 16, 4} / {64, 24, 2} for a key
//                       chunk), and the out-of-image keys of the overlapping window arrive as zeros (TMA out-of-bounds
//                       fill = nn.Unfold's padding). The last window row / column of a shifted frame wraps around the
//                       image (~6 % of the windows of a 544 x 544 tile): its tiles are loaded as half-row boxes {64, 8, 1}.
//                       ncu (profiles/r02_s21) showed the all-cp.async loader's issue time on the critical path: the
//                       softmax warps sat on the S barrier for 22 % of the samples; TMA took 0.52 -> 0.34 ms per launch.
//   MMA (1 thread)    : S_h = Q_h K_h^T (M 128, N = chunk, K 32) per head of the pair; O_h = P_h V (M 128, N 64, K = chunk)
//   softmax (2 x 4 warps, one group per head of the pair, thread = query row):
//                       S row from TMEM -> * log2(e) + bias (+ shift mask) -> chunk max / exp2 / sum -> unnormalised P
//                       row (bf16) to shared memory in the UMMA A layout -> after the P V MMA: the chunk's O row from
//                       TMEM is merged into the running (max, sum, o[32]) in registers (streaming softmax: no
//                       rescaling of TMEM accumulators); after the last chunk o / sum goes out through a per-warp
//                       transpose buffer as coalesced stores.
// The softmax of step n + 1 runs while the P V product of step n is in flight (same software pipeline as
// swin_attn_tc.cu). Nothing is materialised in HBM: neither the logits nor the unfolded 24 x 24 key windows.
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

static constexpr int kHtWarps = 12;              // warp 0: MMA, 1..8: softmax groups, 9..11: loaders
static constexpr int kHtThreads = kHtWarps * 32;
static constexpr int kHtStages = 3;
static constexpr int kHtQBytes = 128 * 128;      // [128 tokens][64 ch] bf16
static constexpr int kHtKBytes = 64 * 128;       // [<= 64 keys][64 ch] bf16
static constexpr int kHtStage = kHtQBytes + 2 * kHtKBytes;  // Q | K | V
static constexpr int kHtPBytes = 128 * 128;      // P of one head: [128 rows][<= 64 keys] bf16
static constexpr int kHtTabStride = 48;          // bias rows padded to 48 words: a warp's two token rows hit disjoint banks
static constexpr float kHtLog2e = 1.4426950408889634f;

template <int MODE>
struct HtCfg {
    static constexpr int KW = MODE == 0 ? 16 : 24;    // keys per key row
    static constexpr int KR = MODE == 0 ? 4 : 2;      // key rows per chunk
    static constexpr int KC = KW * KR;                // keys per chunk (64 / 48)
    static constexpr int NC = KW / KR;                // chunks (4 / 12)
    static constexpr int TR = MODE == 0 ? 31 : 39;    // rows (and used columns) of the bias table
};

struct HtParams {
    const bf16* qkv;
    long long ld;
    bf16* out;
    long long ldo;
    int n, H, W, heads, d, Cq, shift;
    int nwx, nwy, windows;
    const float* table;  // [TR*TR][heads]
};

enum { HB_FULL = 0, HB_EMPTY = 3, HB_SFULL = 6, HB_SEMPTY = 8, HB_PFULL = 10, HB_OFULL = 12, HB_OEMPTY = 14, HB_COUNT = 16 };

// the CTA's step sequence: windows u = blockIdx.x, + gridDim.x, ...; per window head pairs, M blocks, key chunks
template <int NC>
struct HtIter {
    int u, hp, m, c;
    __device__ __forceinline__ bool advance(int npairs, int stride) {  // returns true when a new (u, hp, m) job starts
        if (++c < NC) return false;
        c = 0;
        if (++m < 2) return true;
        m = 0;
        if (++hp < npairs) return true;
        hp = 0;
        u += stride;
        return true;
    }
};

// tm_q / tm_kv: 4-D tensor maps {channel, x, y, image} over the qkv tensor with boxes {64, 16, 8, 1} (a Q block: half a
// window) and {64, KW, KR, 1} (a key chunk), 128-byte swizzle: the box lands in shared memory exactly in the UMMA
// layout, and out-of-image key positions of the overlapping window arrive as zeros (nn.Unfold's padding).
template <int MODE, int D2T>
__global__ void __launch_bounds__(kHtThreads, 1) hat_attention_tc_kernel(const HtParams p,
                                                                         const __grid_constant__ CUtensorMap tm_q,
                                                                         const __grid_constant__ CUtensorMap tm_kv,
                                                                         const __grid_constant__ CUtensorMap tm_h) {
    using Cfg = HtCfg<MODE>;
    constexpr int KC = Cfg::KC, NC = Cfg::NC, KW = Cfg::KW, KR = Cfg::KR, TR = Cfg::TR;
    constexpr int kTabSize = TR * kHtTabStride;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    uint8_t* stage_base = smem;                                                // [stages][Q | K | V]
    uint8_t* p_base = smem + kHtStages * kHtStage;                             // [2 groups][128][128 B]
    uint32_t* tbuf = reinterpret_cast<uint32_t*>(p_base + 2 * kHtPBytes);      // [8 warps][32 rows][17 words]
    float* s_table = reinterpret_cast<float*>(tbuf + 8 * 32 * 17);             // [heads][TR][48] bias * log2(e)
    uint64_t* bar = reinterpret_cast<uint64_t*>(s_table + ((p.heads * kTabSize + 3) & ~3));
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + HB_COUNT);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    {
        uint4* z = reinterpret_cast<uint4*>(smem);
        const int n16 = (kHtStages * kHtStage + 2 * kHtPBytes) / 16;
        for (int i = threadIdx.x; i < n16; i += kHtThreads) z[i] = make_uint4(0u, 0u, 0u, 0u);
        for (int i = threadIdx.x; i < p.heads * kTabSize; i += kHtThreads) {
            const int h = i / kTabSize, e2 = i - h * kTabSize;
            const int a = e2 / kHtTabStride, b = e2 - a * kHtTabStride;
            float v = 0.f;
            if (b < TR) {
                int idx;
                if (MODE == 0) {
                    idx = a * TR + b;  // calculate_rpi_sa: (yi - yj + 15) * 31 + (xi - xj + 15); a = dy + 15, b = dx + 15
                } else {
                    // calculate_rpi_oca: (ky - yi - 7) * 39 + (kx - xi - 7), negative values index the table from its
                    // end (hat_arch.py:911-915 + PyTorch indexing); a = ky - yi + 15, b = kx - xi + 15
                    idx = (a - 22) * TR + (b - 22);
                    if (idx < 0) idx += TR * TR;
                }
                v = __ldg(p.table + (long long)idx * p.heads + h) * kHtLog2e;
            }
            s_table[i] = v;
        }
    }
    if (warp == 0) {
        if (lane == 0) {
            for (int s = 0; s < kHtStages; ++s) {
                mbar_init(&bar[HB_FULL + s], 1);
                mbar_init(&bar[HB_EMPTY + s], 1);
            }
            for (int s = 0; s < 2; ++s) {
                mbar_init(&bar[HB_SFULL + s], 1);
                mbar_init(&bar[HB_SEMPTY + s], 4);
                mbar_init(&bar[HB_PFULL + s], 4);
                mbar_init(&bar[HB_OFULL + s], 1);
                mbar_init(&bar[HB_OEMPTY + s], 4);
            }
            fence_mbar_init();
        }
        __syncwarp();
        tmem_alloc(tmem_slot, 256);
        tmem_relinquish();
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int npairs = p.heads >> 1;
    const int per_img = p.nwx * p.nwy;
    // TMEM columns: S of group g at g * 64, O of group g at 128 + g * 64

    if (warp >= 9) {
        // ===================== loaders: warp lw fills ring stage lw (steps lw, lw + 3, ...) =====================
        const int lw = warp - 9;
        HtIter<NC> it{(int)blockIdx.x, 0, 0, 0};
        for (int cnt = 0; it.u < p.windows; ++cnt, it.advance(npairs, gridDim.x)) {
            if (cnt % kHtStages != lw) continue;
            const int img = it.u / per_img, wi = it.u - img * per_img;
            const int wy = wi / p.nwx, wx = wi - wy * p.nwx;
            if (lane == 0) mbar_wait(&bar[HB_EMPTY + lw], ((cnt / kHtStages) & 1) ^ 1);
            __syncwarp();
            // Q: token t = m * 128 + row of the window, at shifted-frame (wy*16 + t/16, wx*16 + t%16). Only the last
            // window row / column of a shifted frame wraps around the image.
            const bool wrap = p.shift > 0 && (wy == p.nwy - 1 || wx == p.nwx - 1);
            if (!wrap) {
                // the 16 x 16 tokens are an axis-aligned box of the image: three TMA loads by one lane
                if (lane == 0) {
                    uint8_t* st = stage_base + lw * kHtStage;
                    mbar_arrive_expect_tx(&bar[HB_FULL + lw], kHtQBytes + 2 * KC * 128);
                    const int x0 = wx * 16 + p.shift, y0 = wy * 16 + p.shift;
                    const int kx0 = MODE == 0 ? x0 : x0 - 4, ky0 = MODE == 0 ? y0 + it.c * KR : y0 - 4 + it.c * KR;
                    tma_load_4d(st, &tm_q, &bar[HB_FULL + lw], it.hp * 64, x0, y0 + it.m * 8, img);
                    tma_load_4d(st + kHtQBytes, &tm_kv, &bar[HB_FULL + lw], p.Cq + it.hp * 64, kx0, ky0, img);
                    tma_load_4d(st + kHtQBytes + kHtKBytes, &tm_kv, &bar[HB_FULL + lw], 2 * p.Cq + it.hp * 64, kx0, ky0, img);
                }
                __syncwarp();
                continue;
            }
            // the last window row / column of a shifted frame (shift 8) wraps around the image: half token rows (8 tokens) never
            // straddle the wrap, so the same tiles are 16 (Q) + 8 + 8 (K, V) half-row boxes; TMA derives the swizzle from
            // the destination address, so every piece lands at its rows of the 1024-byte-aligned tile
            if (lane == 0) {
                uint8_t* st = stage_base + lw * kHtStage;
                mbar_arrive_expect_tx(&bar[HB_FULL + lw], kHtQBytes + 2 * KC * 128);
                for (int hr = 0; hr < 16; ++hr) {  // Q: token row it.m * 8 + (hr >> 1), tokens 8 * (hr & 1) .. + 7
                    int oy = wy * 16 + it.m * 8 + (hr >> 1) + p.shift, ox = wx * 16 + (hr & 1) * 8 + p.shift;
                    if (oy >= p.H) oy -= p.H;
                    if (ox >= p.W) ox -= p.W;
                    tma_load_4d(st + hr * 1024, &tm_h, &bar[HB_FULL + lw], it.hp * 64, ox, oy, img);
                }
                for (int hr = 0; hr < 2 * KR; ++hr) {  // K, V: key row it.c * KR + (hr >> 1)
                    int oy = wy * 16 + it.c * KR + (hr >> 1) + p.shift, ox = wx * 16 + (hr & 1) * 8 + p.shift;
                    if (oy >= p.H) oy -= p.H;
                    if (ox >= p.W) ox -= p.W;
                    tma_load_4d(st + kHtQBytes + hr * 1024, &tm_h, &bar[HB_FULL + lw], p.Cq + it.hp * 64, ox, oy, img);
                    tma_load_4d(st + kHtQBytes + kHtKBytes + hr * 1024, &tm_h, &bar[HB_FULL + lw], 2 * p.Cq + it.hp * 64, ox, oy, img);
                }
            }
            __syncwarp();
        }
    } else if (warp == 0) {
        // ===================== MMA issuer =====================
        if (elect_one()) {
            const uint32_t idesc_s = make_idesc_bf16(128, KC, 0, 0);   // Q, K both K-major
            const uint32_t idesc_o = make_idesc_bf16(128, 64, 0, 1);   // P K-major, V MN-major
            int my_steps = 0;
            for (int u = blockIdx.x; u < p.windows; u += gridDim.x) my_steps += npairs * 2 * NC;
            // polling issue loop (see swin_attn_tc.cu): S(step) needs its stage loaded and the group's S accumulator
            // drained; P V(step) needs the group's P in shared memory and its O accumulator drained
            int s_next = 0, pv_next = 0;
            bool s_done[2] = {false, false}, pv_done[2] = {false, false};
            bool s_loaded = false;
            uint32_t spins = 0;
            while (pv_next < my_steps) {
                bool progress = false;
                if (s_next < my_steps) {
                    const int st = s_next % kHtStages;
                    if (!s_loaded && IRB_MMA_POLL(&bar[HB_FULL + st], (s_next / kHtStages) & 1)) s_loaded = true;
                    if (s_loaded) {
                        const uint32_t q_addr = smem_u32(stage_base + st * kHtStage);
                        const uint32_t k_addr = q_addr + kHtQBytes;
                        for (int g = 0; g < 2; ++g) {
                            if (s_done[g] || !IRB_MMA_POLL(&bar[HB_SEMPTY + g], (s_next & 1) ^ 1)) continue;
                            tc_fence_after();
#pragma unroll
                            for (int ks = 0; ks < 2; ++ks) {
                                const uint64_t ad = make_sdesc_sw128(q_addr + g * 64 + ks * 32, 16, 1024);
                                const uint64_t bd = make_sdesc_sw128(k_addr + g * 64 + ks * 32, 16, 1024);
                                umma_bf16(tmem_base + g * 64, ad, bd, idesc_s, ks);
                            }
                            umma_commit(&bar[HB_SFULL + g]);
                            s_done[g] = true;
                            progress = true;
                        }
                        if (s_done[0] && s_done[1]) {
                            s_done[0] = s_done[1] = false;
                            s_loaded = false;
                            ++s_next;
                        }
                    }
                }
                if (pv_next < s_next) {
                    const int st = pv_next % kHtStages;
                    const uint32_t v_addr = smem_u32(stage_base + st * kHtStage) + kHtQBytes + kHtKBytes;
                    for (int g = 0; g < 2; ++g) {
                        if (pv_done[g] || !IRB_MMA_POLL(&bar[HB_PFULL + g], pv_next & 1) ||
                            !IRB_MMA_POLL(&bar[HB_OEMPTY + g], (pv_next & 1) ^ 1))
                            continue;
                        tc_fence_after();
                        const uint32_t pa = smem_u32(p_base + g * kHtPBytes);
#pragma unroll
                        for (int ks = 0; ks < KC / 16; ++ks) {
                            const uint64_t ad = make_sdesc_sw128(pa + ks * 32, 16, 1024);
                            const uint64_t bd = make_sdesc_sw128(v_addr + ks * 2048, kHtKBytes, 1024);
                            umma_bf16(tmem_base + 128 + g * 64, ad, bd, idesc_o, ks);
                        }
                        umma_commit(&bar[HB_OFULL + g]);
                        pv_done[g] = true;
                        progress = true;
                    }
                    if (pv_done[0] && pv_done[1]) {
                        umma_commit(&bar[HB_EMPTY + st]);  // Q, K, V of this stage are no longer read
                        pv_done[0] = pv_done[1] = false;
                        ++pv_next;
                    }
                }
                if (progress) {
                    spins = 0;
                } else {
                    __nanosleep(IRB_MMA_POLL_SLEEP);
                    if (++spins > (1u << 24)) __trap();  // protocol bug: fail the launch instead of hanging
                }
            }
        }
    } else {
        // ===================== softmax + output groups: warps 1..4 (head 2*hp) and 5..8 (head 2*hp + 1) ==========
        const int g = (warp - 1) >> 2;
        const int q = warp & 3;                  // TMEM lane quarter this warp may read
        const int r = q * 32 + lane;             // query row of the M block
        const int xi = r & 15;
        const int d2 = D2T > 0 ? D2T : (p.d >> 1);  // output words (bf16 pairs) per head
        const int tstride = d2 | 1;              // odd stride: conflict-free transpose
        uint32_t* tb = tbuf + (warp - 1) * 32 * 17;
        uint8_t* prow = p_base + g * kHtPBytes + r * 128;
        const int sw = r & 7;
        const uint32_t lane_addr = uint32_t(q * 32) << 16;
        uint32_t* out32 = reinterpret_cast<uint32_t*>(p.out);
        const long long ldo2 = p.ldo >> 1;
        const float NEG = -100.0f * kHtLog2e;
        // job state (window, head pair, M block) of the softmax side
        long long orow_s = -1;
        int yi = 0;
        bool lastrow = false, lastcol = false;
        auto job_info = [&](const HtIter<NC>& it) {
            const int img = it.u / per_img, wi = it.u - img * per_img;
            const int wy = wi / p.nwx, wx = wi - wy * p.nwx;
            const int t = it.m * 128 + r;
            yi = t >> 4;
            int oy = wy * 16 + yi + p.shift, ox = wx * 16 + xi + p.shift;
            if (oy >= p.H) oy -= p.H;
            if (ox >= p.W) ox -= p.W;
            orow_s = (((long long)img * p.H + oy) * p.W + ox) * ldo2;
            lastrow = p.shift > 0 && wy == p.nwy - 1;
            lastcol = p.shift > 0 && wx == p.nwx - 1;
        };
        uint32_t pk[KC / 2];
        float mloc = 0.f, lloc = 0.f;
        // S(cnt) row from TMEM -> unnormalised chunk probabilities (bf16 pairs) in pk, chunk max in mloc, chunk sum in lloc
#ifdef IRB_HT_TRACE  // timing experiment: clock() trace of softmax warp 1 / lane 0 of CTA 0 (MODE == IRB_HT_TRACE - 1)
        const bool trace = blockIdx.x == 0 && threadIdx.x == 32 && MODE == IRB_HT_TRACE - 1;
        long long tsv[150];
        int tsi[150];
        int nts = 0;
#define IRB_TS(i) do { if (trace && nts < 150) { tsv[nts] = clock64(); tsi[nts++] = (i); } } while (0)
#else
#define IRB_TS(i) do { } while (0)
#endif
        auto softmax_phase = [&](int cnt, const HtIter<NC>& it) {
            const int h = it.hp * 2 + g;
            const float* tab;
            if (MODE == 0)   // (yi - yj + 15) * 48 + (xi - xj + 15), yj = 4 c + jj / 16, xj = jj % 16
                tab = s_table + h * kTabSize + (yi + 15 - it.c * KR) * kHtTabStride + xi + 15;
            else             // (ky - yi + 15) * 48 + (kx - xi + 15), ky = 2 c + jj / 24, kx = jj % 24
                tab = s_table + h * kTabSize + (it.c * KR - yi + 15) * kHtTabStride + 15 - xi;
            IRB_TS(1);
            mbar_wait(&bar[HB_SFULL + g], cnt & 1);
            IRB_TS(2);
            tc_fence_after();
            uint32_t sr[64];
            tmem_ld32_nowait(tmem_base + lane_addr + g * 64, sr);
            tmem_ld32_nowait(tmem_base + lane_addr + g * 64 + 32, sr + 32);  // MODE 1 uses 16 of these 32 columns
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar[HB_SEMPTY + g]);
            float t[KC];
#pragma unroll
            for (int j = 0; j < KC; ++j) {
                const int off = (j / KW) * kHtTabStride + (j % KW);
                t[j] = fmaf(__uint_as_float(sr[j]), kHtLog2e, MODE == 0 ? tab[-off] : tab[off]);
            }
            if (MODE == 0 && (lastrow || lastcol)) {
                // calculate_mask (hat_arch.py:921-941) with shift = 8: in the last window row (column) of the shifted
                // frame the tokens of the first and of the second half belong to different regions; this chunk's
                // keys are rows 4c .. 4c + 3, i.e. in the second half iff c >= 2
                const bool ymis = lastrow && ((it.c >= NC / 2) != (yi >= 8));
                const float mlo = (ymis || (lastcol && xi >= 8)) ? NEG : 0.f;   // keys with xj < 8
                const float mhi = (ymis || (lastcol && xi < 8)) ? NEG : 0.f;    // keys with xj >= 8
#pragma unroll
                for (int j = 0; j < KC; ++j) t[j] += ((j % KW) >= 8) ? mhi : mlo;
            }
            float mx_ = t[0];
#pragma unroll
            for (int j = 1; j < KC; ++j) mx_ = fmaxf(mx_, t[j]);
            float sum = 0.f;
#pragma unroll
            for (int j = 0; j < KC; j += 2) {
                const float e0 = ex2_approx(t[j] - mx_), e1 = ex2_approx(t[j + 1] - mx_);
                sum += e0 + e1;
                pk[j >> 1] = pack_bf16x2(e0, e1);
            }
            mloc = mx_;
            lloc = sum;
        };
        // running streaming-softmax state of the current job, in log2 units
        float M = -INFINITY, L = 0.f;
        float o[32];
#pragma unroll
        for (int k = 0; k < 32; ++k) o[k] = 0.f;
        HtIter<NC> it{(int)blockIdx.x, 0, 0, 0};
        if (it.u < p.windows) {
            job_info(it);
            softmax_phase(0, it);
        }
        for (int cnt = 0; it.u < p.windows; ++cnt) {
            IRB_TS(3);
            // ---- P(cnt) -> shared memory (UMMA A layout) -> tensor core ----
#pragma unroll
            for (int c = 0; c < KC / 8; ++c)
                *reinterpret_cast<uint4*>(prow + ((c ^ sw) << 4)) =
                    make_uint4(pk[c * 4], pk[c * 4 + 1], pk[c * 4 + 2], pk[c * 4 + 3]);
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar[HB_PFULL + g]);
            IRB_TS(4);
            const long long orow = orow_s;
            const float mc = mloc, lc = lloc;
            const int h = it.hp * 2 + g;
            const bool last_chunk = it.c == NC - 1;
            // ---- softmax of step cnt + 1 while P V of cnt runs ----
            const bool new_job = it.advance(npairs, gridDim.x);
            if (it.u < p.windows) {
                if (new_job) job_info(it);
                softmax_phase(cnt + 1, it);
            }
            // ---- O(cnt) = P V: merge the chunk into the running state ----
            IRB_TS(5);
            mbar_wait(&bar[HB_OFULL + g], cnt & 1);
            IRB_TS(6);
            tc_fence_after();
            uint32_t orr[32];
            tmem_ld32_nowait(tmem_base + lane_addr + 128 + g * 64 + g * 32, orr);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&bar[HB_OEMPTY + g]);
            {
                const float Mn = fmaxf(M, mc);
                const float a = ex2_approx(M - Mn), b = ex2_approx(mc - Mn);
                M = Mn;
                L = fmaf(L, a, lc * b);
#pragma unroll
                for (int k = 0; k < 32; ++k) o[k] = fmaf(o[k], a, __uint_as_float(orr[k]) * b);
            }
            IRB_TS(7);
            if (!last_chunk) continue;
            const float inv = 1.0f / L;
#pragma unroll
            for (int k = 0; k < 16; ++k)
                if (k < d2) tb[lane * tstride + k] = pack_bf16x2(o[2 * k] * inv, o[2 * k + 1] * inv);
            M = -INFINITY;
            L = 0.f;
#pragma unroll
            for (int k = 0; k < 32; ++k) o[k] = 0.f;
            __syncwarp();
            // coalesced: consecutive lanes write consecutive words of a token's head slice
            if (D2T > 0) {
                // row pointers of the store pattern as 32-bit offsets from the warp's first token, all gathered before the
                // first store (a 64-bit shuffle + store per word ran in series: ~100 cycles per word in swin_attn_tc.cu)
                const long long o_base = __shfl_sync(0xffffffffu, orow, 0);
                const int rel = int(orow - o_base);
                int o_off[D2T > 0 ? D2T : 1];
                uint32_t v[D2T > 0 ? D2T : 1];
#pragma unroll
                for (int it2 = 0; it2 < D2T; ++it2) {
                    const int wi = lane + 32 * it2;
                    const int row = wi / D2T;
                    o_off[it2] = __shfl_sync(0xffffffffu, rel, row) + (wi - row * D2T);
                    v[it2] = tb[row * tstride + (wi - row * D2T)];
                }
                uint32_t* ob = out32 + o_base + h * D2T;
#pragma unroll
                for (int it2 = 0; it2 < D2T; ++it2) ob[o_off[it2]] = v[it2];
            } else {
                const int nwords = 32 * d2;
                for (int wi = lane; wi < nwords; wi += 32) {
                    const int row = wi / d2, k = wi - row * d2;
                    const uint32_t v = tb[row * tstride + k];
                    const long long ob = __shfl_sync(0xffffffffu, orow, row);
                    out32[ob + h * d2 + k] = v;
                }
            }
            __syncwarp();
            IRB_TS(8);
        }
#ifdef IRB_HT_TRACE
        if (trace) {
            static __device__ int printed;
            if (atomicAdd(&printed, 1) == 0)
                for (int i = 1; i < nts; ++i) printf("TS %d %d %lld\n", i, tsi[i], tsv[i] - tsv[i - 1]);
        }
#endif
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, 256);
}

bool hat_attention_tc_ok(int dt, long long ld, long long ldo, int H, int W, int heads, int d, int dstride, int ws, int ows,
                         int shift) {
    return dt == DT_BF16 && ws == 16 && (ows == 16 || ows == 24) && dstride == 32 && heads >= 2 && heads % 2 == 0 &&
           heads <= 8 && d % 2 == 0 && d <= 32 && (shift == 0 || (shift == 8 && ows == 16)) && H % 16 == 0 &&
           W % 16 == 0 && ld % 8 == 0 && ldo % 2 == 0;
}

template <int MODE, int D2T>
static int hat_attention_tc_launch(const HtParams& p, int grid, cudaStream_t stream) {
    using Cfg = HtCfg<MODE>;
    CUtensorMap tm_q, tm_kv, tm_h;
    {
        const uint64_t dims[4] = {uint64_t(3 * p.Cq), uint64_t(p.W), uint64_t(p.H), uint64_t(p.n)};
        const uint64_t strides[3] = {uint64_t(p.ld) * 2, uint64_t(p.W) * p.ld * 2, uint64_t(p.H) * p.W * p.ld * 2};
        const uint32_t box_q[4] = {64, 16, 8, 1}, box_kv[4] = {64, uint32_t(Cfg::KW), uint32_t(Cfg::KR), 1};
        IRB_TRY(make_tmap_bf16_sw(&tm_q, p.qkv, 4, dims, strides, box_q, 128));
        IRB_TRY(make_tmap_bf16_sw(&tm_kv, p.qkv, 4, dims, strides, box_kv, 128));
        const uint32_t box_h[4] = {64, 8, 1, 1};  // half a token row: the pieces of a window that wraps around the image
        IRB_TRY(make_tmap_bf16_sw(&tm_h, p.qkv, 4, dims, strides, box_h, 128));
    }
    const size_t smem = 1024 + size_t(kHtStages) * kHtStage + 2 * kHtPBytes + 8 * 32 * 17 * 4 +
                        ((p.heads * Cfg::TR * kHtTabStride + 3) & ~3) * 4 + HB_COUNT * 8 + 64;
    IRB_REQUIRE(smem <= 227 * 1024, "hat_attention_tc: shared memory");
    IRB_SMEM_OPTIN((hat_attention_tc_kernel<MODE, D2T>), 227 * 1024);
    hat_attention_tc_kernel<MODE, D2T><<<grid, kHtThreads, smem, stream>>>(p, tm_q, tm_kv, tm_h);
    IRB_CHECK_CUDA(cudaGetLastError());
    return 0;
}

int hat_attention_tc(const bf16* qkv, long long ld, bf16* out, long long ldo, int n, int H, int W, int heads, int d,
                     int ows, int shift, const float* table, cudaStream_t stream) {
    IRB_REQUIRE((reinterpret_cast<uintptr_t>(qkv) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 3) == 0,
                "hat_attention_tc: alignment");
    HtParams p{};
    p.qkv = qkv;
    p.ld = ld;
    p.out = out;
    p.ldo = ldo;
    p.n = n;
    p.H = H;
    p.W = W;
    p.heads = heads;
    p.d = d;
    p.Cq = heads * 32;
    p.shift = shift;
    p.nwx = W / 16;
    p.nwy = H / 16;
    const long long tw = (long long)n * p.nwx * p.nwy;
    IRB_REQUIRE(tw < (1ll << 30), "hat_attention_tc: too many windows");
    p.windows = int(tw);
    p.table = table;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int grid = p.windows < sms ? p.windows : sms;
    if (ows == 16) {
        if (d == 30) IRB_TRY((hat_attention_tc_launch<0, 15>(p, grid, stream)));
        else IRB_TRY((hat_attention_tc_launch<0, 0>(p, grid, stream)));
    } else {
        if (d == 30) IRB_TRY((hat_attention_tc_launch<1, 15>(p, grid, stream)));
        else IRB_TRY((hat_attention_tc_launch<1, 0>(p, grid, stream)));
    }
    const double px = double(n) * H * W;
    prof_mark(ows == 16 ? "hat_attention_tc_sa" : "hat_attention_tc_oca", heads * d, shift, 4.0 * px * ows * ows * heads * d,
              px * (3.0 * p.Cq + heads * d) * 2, stream);
    return 0;
}

}  // namespace irb
