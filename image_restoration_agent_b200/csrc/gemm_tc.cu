// tcgen05 pixel-GEMM / implicit-GEMM 3x3 convolution for sm_100a (bf16 in, fp32 accumulate in TMEM).
//
//   out[row, n] = epilogue( sum_{tap, k} A[row shifted by tap, k] * Wt[n, tap*kpad + k] )
//
// mode 0 (1x1 conv / Linear): A is a [batch, M, K] channels-last activation (row stride lda), loaded by a
//   3-D TMA map {K, M, batch} with a {64, 128, 1} box; K beyond the tensor is zero-filled by TMA.
// mode 1 (3x3 conv, zero pad 1): A is a [N, H, W, C] channels-last activation, loaded by a 4-D TMA map
//   {C, W, H, N} with a {64, BW, BH, 1} box (BW*BH = 128 pixels); the 9 taps are 9 shifted boxes and the
//   out-of-bounds halo is zero-filled by TMA, which is exactly the convolution's zero padding.
// Weights Wt are [Nrows, Ktot] bf16, K contiguous (the layout nn.Conv2d / nn.Linear already use), loaded by a
//   2-D TMA map with a {64, BN} box. Both operands land in shared memory in the 128B-swizzled K-major layout
//   that the UMMA descriptors below describe.
//
// One CTA = one 128 x BN output tile. Warp 0 lane 0: TMA producer; warp 1 lane 0: tcgen05.mma issuer;
// warps 2..5: epilogue (TMEM -> registers -> fused bias/act/residual -> global). A ring of `stages`
// shared-memory slots with full/empty mbarriers decouples TMA from the tensor core.
#include "common.cuh"
#include "ptx.cuh"
#include "kernels.h"

namespace irb {

static constexpr int kTileM = 128;
static constexpr int kTileK = 64;                 // bf16 elements per 128-byte swizzle row
static constexpr int kABytes = kTileM * 128;      // 16 KB

struct alignas(64) TcGemmParams {
    CUtensorMap tmA;
    CUtensorMap tmB;
    int mode;
    int M;        // rows per batch (mode 0)
    int N;        // valid output columns
    int BN;       // columns per CTA (multiple of 16, <= 256)
    int ktiles;   // 64-wide K chunks per tap
    int taps;     // 1 or 9
    int kpad;     // K distance between taps inside Wt
    int H, W, BW, BH, tiles_x;
    int b_rows_per_batch;
    int stages;
    int tmem_cols;
    Epi epi;
};

__global__ void __launch_bounds__(192) tc_gemm_kernel(const __grid_constant__ TcGemmParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int stage_bytes = kABytes + p.BN * 128;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + p.stages * stage_bytes);
    uint64_t* empty_bar = full_bar + p.stages;
    uint64_t* accum_bar = empty_bar + p.stages;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (warp == 0) {
        if (lane == 0) {
            tma_prefetch_desc(&p.tmA);
            tma_prefetch_desc(&p.tmB);
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

    const int n0 = blockIdx.y * p.BN;
    const int total_it = p.taps * p.ktiles;

    // tile origin
    int x0 = 0, y0 = 0;
    if (p.mode == 1) {
        const int ty = blockIdx.x / p.tiles_x, tx = blockIdx.x % p.tiles_x;
        x0 = tx * p.BW;
        y0 = ty * p.BH;
    }

    if (warp == 0 && lane == 0) {
        // ===== TMA producer =====
        int it = 0;
        for (int tap = 0; tap < p.taps; ++tap) {
            const int dy = (p.taps == 9) ? tap / 3 - 1 : 0;
            const int dx = (p.taps == 9) ? tap % 3 - 1 : 0;
            for (int kc = 0; kc < p.ktiles; ++kc, ++it) {
                const int s = it % p.stages;
                const uint32_t ph = (it / p.stages) & 1;
                mbar_wait(&empty_bar[s], ph ^ 1);
                uint8_t* sa = smem + s * stage_bytes;
                uint8_t* sb = sa + kABytes;
                mbar_arrive_expect_tx(&full_bar[s], stage_bytes);
                if (p.mode == 0)
                    tma_load_3d(sa, &p.tmA, &full_bar[s], kc * kTileK, blockIdx.x * kTileM, blockIdx.z);
                else
                    tma_load_4d(sa, &p.tmA, &full_bar[s], kc * kTileK, x0 + dx, y0 + dy, blockIdx.z);
                tma_load_2d(sb, &p.tmB, &full_bar[s], tap * p.kpad + kc * kTileK,
                            blockIdx.z * p.b_rows_per_batch + n0);
            }
        }
    } else if (warp == 1 && lane == 0) {
        // ===== MMA issuer =====
        const uint32_t idesc = make_idesc_bf16(kTileM, p.BN, 0, 0);
        for (int it = 0; it < total_it; ++it) {
            const int s = it % p.stages;
            const uint32_t ph = (it / p.stages) & 1;
            mbar_wait(&full_bar[s], ph);
            tc_fence_after();
            const uint32_t a_addr = smem_u32(smem + s * stage_bytes);
            const uint32_t b_addr = a_addr + kABytes;
#pragma unroll
            for (int k = 0; k < kTileK / 16; ++k) {
                const uint64_t adesc = make_sdesc_sw128(a_addr + k * 32, 16, 1024);
                const uint64_t bdesc = make_sdesc_sw128(b_addr + k * 32, 16, 1024);
                umma_bf16(tmem_base, adesc, bdesc, idesc, (it | k) != 0 ? 1u : 0u);
            }
            umma_commit(&empty_bar[s]);  // frees the smem slot once these MMAs have read it
        }
        umma_commit(accum_bar);  // accumulator complete
    } else if (warp >= 2) {
        // ===== epilogue: TMEM -> registers -> global =====
        mbar_wait(accum_bar, 0);
        tc_fence_after();
        const int q = warp & 3;  // TMEM lane quarter this warp may access
        const int r = q * 32 + lane;
        long long grow;
        bool valid;
        int py = 0, px = 0;
        if (p.mode == 0) {
            const int m = blockIdx.x * kTileM + r;
            valid = m < p.M;
            grow = (long long)blockIdx.z * p.M + m;
        } else {
            py = y0 + r / p.BW;
            px = x0 + r % p.BW;
            valid = (py < p.H) && (px < p.W);
            grow = ((long long)blockIdx.z * p.H + py) * p.W + px;
        }
        const uint32_t taddr = tmem_base + (uint32_t(q * 32) << 16);
        for (int c = 0; c < p.BN; c += 16) {
            float v[16];
            tmem_ld16(taddr + c, v);
            if (valid && n0 + c < p.N) epi_apply_store<16>(p.epi, grow, n0 + c, p.N, v, blockIdx.z, py, px);
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
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(base), d, s, b, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE,
                     swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
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
    // largest multiple of 16 that is <= 256 and splits N (rounded up to 16) evenly
    const int n16 = (N + 15) / 16 * 16;
    if (n16 <= 256) return n16;
    for (int parts = 2; parts <= 64; ++parts) {
        if (n16 % (16 * parts) == 0 && n16 / parts <= 256) return n16 / parts;
    }
    return 128;
}

static int tmem_cols_for(int bn) {
    int c = 32;
    while (c < bn) c <<= 1;
    return c;
}

static int launch_tc(TcGemmParams& p, dim3 grid, cudaStream_t stream) {
    const int stage_bytes = kABytes + p.BN * 128;
    int stages = (100 * 1024) / stage_bytes;
    const int total_it = p.taps * p.ktiles;
    if (stages > 4) stages = 4;
    if (stages > total_it) stages = total_it;
    if (stages < 1) stages = 1;
    p.stages = stages;
    p.tmem_cols = tmem_cols_for(p.BN);
    const size_t smem = size_t(stages) * stage_bytes + 1024 /*align*/ + 256 /*barriers*/;
    static bool attr_set = false;
    if (!attr_set) {
        IRB_CHECK_CUDA(cudaFuncSetAttribute(tc_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_set = true;
    }
    tc_gemm_kernel<<<grid, 192, smem, stream>>>(p);
    IRB_CHECK_CUDA(cudaGetLastError());
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
    p.ktiles = ceil_div(K, kTileK);
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
    dim3 grid(ceil_div(M, kTileM), ceil_div(N, p.BN), batch);
    IRB_TRY(launch_tc(p, grid, stream));
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
    // pixel box: as wide as the row allows, 128 pixels in total
    int BW = 128;
    while (BW > 8 && BW > W) BW >>= 1;
    if (BW > W) BW = 8;
    const int BH = 128 / BW;
    {
        uint64_t dims[4] = {uint64_t(C), uint64_t(W), uint64_t(H), uint64_t(Nimg)};
        uint64_t str[3] = {uint64_t(lda) * 2, uint64_t(lda) * 2 * W, uint64_t(lda) * 2 * W * H};
        uint32_t box[4] = {kTileK, uint32_t(BW), uint32_t(BH), 1};
        IRB_TRY(make_tmap_bf16(&p.tmA, A, 4, dims, str, box));
    }
    p.mode = 1;
    p.N = N;
    p.BN = pick_bn(N);
    p.ktiles = kpad / kTileK;
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
    dim3 grid(p.tiles_x * ceil_div(H, BH), ceil_div(N, p.BN), Nimg);
    IRB_TRY(launch_tc(p, grid, stream));
    {
        const double rows = double(H) * W * Nimg;
        prof_mark("tc_conv3x3", C, N, 18.0 * rows * C * N, rows * C * 2 + 9.0 * N * C * 2 + epi_bytes(epi, rows, N), stream);
    }
    return 0;
}

}  // namespace irb
