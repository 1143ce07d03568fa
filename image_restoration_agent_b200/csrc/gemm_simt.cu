// fp32 FMA pixel-GEMM / 3x3 implicit-GEMM convolution. This is the arithmetic of the fp32 precision mode
// (north_star: fp32 outputs within 1e-4 of the reference) and of the few convolutions whose channel counts are
// too small for a tensor-core tile (3-channel image in / out). Same epilogue as the tcgen05 path.
#include "common.cuh"
#include "kernels.h"

namespace irb {

struct SimtParams {
    const void* A;
    int a_dt;
    long long lda;
    long long a_batch_stride;
    const float* Wt;
    int ldw;
    int w_rows_per_batch;
    int M, N, K, kpad, taps, H, W;
    Epi epi;
};

template <int TAPS>
__global__ void __launch_bounds__(256) simt_gemm_kernel(const SimtParams p) {
    constexpr int BM = 64, BN = 64, BK = 16;
    __shared__ float As[BK][BM + 1];
    __shared__ float Bs[BK][BN + 1];
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN, b = blockIdx.z;
    const char* Ab = reinterpret_cast<const char*>(p.A);
    const long long a_base = (long long)b * p.a_batch_stride;
    const float* Wb = p.Wt + (long long)b * p.w_rows_per_batch * p.ldw;

    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    // each thread loads 4 A and 4 B elements per k-step: element e = tid + 256*i -> (row = e / 16, k = e % 16)
    for (int tap = 0; tap < TAPS; ++tap) {
        const int dy = TAPS == 9 ? tap / 3 - 1 : 0;
        const int dx = TAPS == 9 ? tap % 3 - 1 : 0;
        for (int k0 = 0; k0 < p.K; k0 += BK) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int e = tid + 256 * i;
                const int r = e >> 4, kk = e & 15;
                const int m = m0 + r, k = k0 + kk;
                float v = 0.f;
                if (m < p.M && k < p.K) {
                    long long row = m;
                    bool ok = true;
                    if (TAPS == 9) {
                        const int y = m / p.W + dy, x = m % p.W + dx;
                        ok = (y >= 0) && (y < p.H) && (x >= 0) && (x < p.W);
                        row = (long long)y * p.W + x;
                    }
                    if (ok) v = ld_as_float(Ab, a_base + row * p.lda + k, p.a_dt);
                }
                As[kk][r] = v;
                const int n = n0 + r;
                float w = 0.f;
                if (n < p.N && k < p.K) w = __ldg(Wb + (long long)n * p.ldw + tap * p.kpad + k);
                Bs[kk][r] = w;
            }
            __syncthreads();
#pragma unroll
            for (int kk = 0; kk < BK; ++kk) {
                float a[4], w[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
                for (int j = 0; j < 4; ++j) w[j] = Bs[kk][tx * 4 + j];
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], w[j], acc[i][j]);
            }
            __syncthreads();
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = m0 + ty * 4 + i;
        if (m >= p.M) continue;
        const long long grow = (long long)b * p.M + m;
        int y = 0, x = 0;
        if (TAPS == 9) {
            y = m / p.W;
            x = m % p.W;
        }
        epi_apply_store<4>(p.epi, grow, n0 + tx * 4, p.N, acc[i], b, y, x);
    }
}

int simt_gemm(const void* A, int a_dt, long long lda, int M, int K, int batch, long long a_batch_stride,
              const float* Wt, int ldw, int N, int w_rows_per_batch, const Epi& epi, cudaStream_t stream) {
    IRB_REQUIRE(M > 0 && K > 0 && N > 0 && batch > 0, "simt_gemm: empty problem");
    IRB_REQUIRE(epi.stats_out == nullptr, "simt_gemm: Epi::stats_out is a tcgen05-epilogue feature");
    SimtParams p{};
    p.A = A;
    p.a_dt = a_dt;
    p.lda = lda;
    p.a_batch_stride = a_batch_stride ? a_batch_stride : (long long)M * lda;
    p.Wt = Wt;
    p.ldw = ldw;
    p.w_rows_per_batch = w_rows_per_batch;
    p.M = M;
    p.N = N;
    p.K = K;
    p.kpad = 0;
    p.taps = 1;
    p.epi = epi;
    dim3 grid(ceil_div(M, 64), ceil_div(N, 64), batch);
    simt_gemm_kernel<1><<<grid, 256, 0, stream>>>(p);
    IRB_CHECK_CUDA(cudaGetLastError());
    {
        const double rows = double(M) * batch, ea = a_dt == DT_BF16 ? 2 : 4;
        prof_mark("simt_gemm", K, N, 2.0 * rows * K * N, rows * K * ea + double(N) * K * 4 + epi_bytes(epi, rows, N), stream);
    }
    return 0;
}

int simt_conv3x3(const void* A, int a_dt, long long lda, int Nimg, int H, int W, int C, const float* Wt, int kpad,
                 int N, const Epi& epi, cudaStream_t stream) {
    IRB_REQUIRE(Nimg > 0 && H > 0 && W > 0 && C > 0 && N > 0, "simt_conv3x3: empty problem");
    IRB_REQUIRE(epi.stats_out == nullptr, "simt_conv3x3: Epi::stats_out is a tcgen05-epilogue feature");
    SimtParams p{};
    p.A = A;
    p.a_dt = a_dt;
    p.lda = lda;
    p.a_batch_stride = (long long)H * W * lda;
    p.Wt = Wt;
    p.ldw = 9 * kpad;
    p.w_rows_per_batch = 0;
    p.M = H * W;
    p.N = N;
    p.K = C;
    p.kpad = kpad;
    p.taps = 9;
    p.H = H;
    p.W = W;
    p.epi = epi;
    dim3 grid(ceil_div(H * W, 64), ceil_div(N, 64), Nimg);
    simt_gemm_kernel<9><<<grid, 256, 0, stream>>>(p);
    IRB_CHECK_CUDA(cudaGetLastError());
    {
        const double rows = double(H) * W * Nimg, ea = a_dt == DT_BF16 ? 2 : 4;
        prof_mark("simt_conv3x3", C, N, 18.0 * rows * C * N, rows * C * ea + 9.0 * N * C * 4 + epi_bytes(epi, rows, N), stream);
    }
    return 0;
}

}  // namespace irb
