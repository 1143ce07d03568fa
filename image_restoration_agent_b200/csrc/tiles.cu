// Tiling / IO layer: image -> padded planar float, tile gather, overlap-blend stitch with the
// reference's accumulation order and analytic coverage count, crop, clamp, round-half-even quantisation.
#include "common.cuh"
#include "kernels.h"

namespace irb {

static constexpr int kMaxIdx = 128;
struct IdxLists {
    int ys[kMaxIdx];
    int xs[kMaxIdx];
    int ny, nx;
};

static int fill_idx(IdxLists& L, const int* ys, const int* xs, int ny, int nx) {
    IRB_REQUIRE(ny > 0 && nx > 0 && ny <= kMaxIdx && nx <= kMaxIdx, "tile index lists must hold 1..128 entries");
    for (int i = 0; i < ny; ++i) L.ys[i] = ys[i];
    for (int i = 0; i < nx; ++i) L.xs[i] = xs[i];
    L.ny = ny;
    L.nx = nx;
    return 0;
}

// pad_mode 0: F.pad(..., 'reflect') (demo.py:127, network_swinir.py:787); 1: cat([x, flip(x)]) (inference.py:79-80)
__device__ __forceinline__ int pad_index(int i, int n, int pad_mode) {
    if (i < n) return i;
    return pad_mode == 0 ? 2 * (n - 1) - i : 2 * n - 1 - i;
}

__global__ void image_prepare_kernel(const uint8_t* __restrict__ src, int h, int w, int c, int rev,
                                     float* __restrict__ dst, int hp, int wp, int pad_mode) {
    const long long total = (long long)hp * wp;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int y = int(i / wp), x = int(i % wp);
        const int sy = pad_index(y, h, pad_mode), sx = pad_index(x, w, pad_mode);
        const uint8_t* s = src + ((long long)sy * w + sx) * c;
        for (int ch = 0; ch < c; ++ch) {
            const int sc = rev ? (c - 1 - ch) : ch;
            dst[(long long)ch * hp * wp + i] = float(s[sc]) / 255.0f;
        }
    }
}

int image_prepare(const uint8_t* src, int h, int w, int c, int reverse_channels, float* dst, int hp, int wp,
                  int pad_mode, cudaStream_t stream) {
    IRB_REQUIRE(h > 0 && w > 0 && c > 0 && hp >= h && wp >= w, "image_prepare: bad sizes");
    IRB_REQUIRE(hp - h <= h - (pad_mode == 0 ? 1 : 0) && wp - w <= w - (pad_mode == 0 ? 1 : 0),
                "image_prepare: padding larger than the image");
    const long long total = (long long)hp * wp;
    const int blocks = int(ceil_div_ll(total, 256) > 148 * 16 ? 148 * 16 : ceil_div_ll(total, 256));
    image_prepare_kernel<<<blocks, 256, 0, stream>>>(src, h, w, c, reverse_channels, dst, hp, wp, pad_mode);
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("image_prepare", c, pad_mode, 0.0, 5.0 * total * c, stream);
    return 0;
}

__global__ void image_pad_kernel(const float* __restrict__ src, int c, int h, int w, float* __restrict__ dst, int hp,
                                 int wp, int pad_mode) {
    const long long total = (long long)c * hp * wp;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int x = int(i % wp);
        const int y = int((i / wp) % hp);
        const int ch = int(i / ((long long)wp * hp));
        dst[i] = src[((long long)ch * h + pad_index(y, h, pad_mode)) * w + pad_index(x, w, pad_mode)];
    }
}

int image_pad_f32(const float* src, int c, int h, int w, float* dst, int hp, int wp, int pad_mode,
                  cudaStream_t stream) {
    IRB_REQUIRE(h > 0 && w > 0 && c > 0 && hp >= h && wp >= w, "image_pad_f32: bad sizes");
    IRB_REQUIRE(hp - h <= h - (pad_mode == 0 ? 1 : 0) && wp - w <= w - (pad_mode == 0 ? 1 : 0),
                "image_pad_f32: padding larger than the image");
    const long long total = (long long)c * hp * wp;
    const int blocks = int(ceil_div_ll(total, 256) > 148 * 16 ? 148 * 16 : ceil_div_ll(total, 256));
    image_pad_kernel<<<blocks, 256, 0, stream>>>(src, c, h, w, dst, hp, wp, pad_mode);
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("image_pad", c, pad_mode, 0.0, 8.0 * total, stream);
    return 0;
}

// planar [c][h][w] -> tiles [n = iy*nx+ix][th][tw][c], with the SwinIR mean/range normalisation folded in
__global__ void tiles_gather_kernel(const float* __restrict__ img, int c, int h, int w, IdxLists L, int th, int tw,
                                    float* __restrict__ tiles, float m0, float m1, float m2, float scale) {
    const long long per_tile = (long long)th * tw;
    const long long total = per_tile * L.ny * L.nx;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int n = int(i / per_tile);
        const int r = int(i % per_tile);
        const int ty = r / tw, tx = r % tw;
        const int y = L.ys[n / L.nx] + ty, x = L.xs[n % L.nx] + tx;
        for (int ch = 0; ch < c; ++ch) {
            const float mean = ch == 0 ? m0 : (ch == 1 ? m1 : m2);
            tiles[i * c + ch] = (img[((long long)ch * h + y) * w + x] - mean) * scale;
        }
    }
}

int tiles_gather(const float* img, int c, int h, int w, const int* ys, const int* xs, int ny, int nx, int th, int tw,
                 float* tiles, const float* mean, float scale, cudaStream_t stream) {
    IdxLists L;
    IRB_TRY(fill_idx(L, ys, xs, ny, nx));
    IRB_REQUIRE(c >= 1 && c <= 3, "tiles_gather: 1..3 channels");
    for (int i = 0; i < ny; ++i) IRB_REQUIRE(ys[i] >= 0 && ys[i] + th <= h, "tiles_gather: tile row outside the image");
    for (int i = 0; i < nx; ++i) IRB_REQUIRE(xs[i] >= 0 && xs[i] + tw <= w, "tiles_gather: tile column outside the image");
    const long long total = (long long)th * tw * ny * nx;
    const int blocks = int(ceil_div_ll(total, 256) > 148 * 16 ? 148 * 16 : ceil_div_ll(total, 256));
    tiles_gather_kernel<<<blocks, 256, 0, stream>>>(img, c, h, w, L, th, tw, tiles, mean ? mean[0] : 0.f,
                                                    (mean && c > 1) ? mean[1] : 0.f, (mean && c > 2) ? mean[2] : 0.f,
                                                    scale);
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("tiles_gather", c, ny * nx, 0.0, 8.0 * total * c, stream);
    return 0;
}

// E/W of inference.py:309-320 / demo.py:142-153 as a gather: every output pixel sums the tiles that cover it in
// the reference's loop order (h_idx outer, w_idx inner; fp32, starting from 0), and divides by their count.
__global__ void tiles_stitch_kernel(const float* __restrict__ tiles, int c, int th, int tw, int sf, IdxLists L, int h,
                                    int w, int row0, int rows, int crop_w, float* __restrict__ out_f32,
                                    uint8_t* __restrict__ out_u8, int rev) {
    // the launch covers output rows [row0, row0 + rows) (a rank's strip; the whole crop on one GPU); the outputs are
    // strip-local: out_u8 [rows][crop_w][c], out_f32 [c][rows][crop_w]
    const int TH = th * sf, TW = tw * sf;
    const long long total = (long long)rows * crop_w;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int oy = row0 + int(i / crop_w), ox = int(i % crop_w);
        float e[3] = {0.f, 0.f, 0.f};
        float cnt = 0.f;
        for (int iy = 0; iy < L.ny; ++iy) {
            const int ry = oy - L.ys[iy] * sf;
            if (ry < 0 || ry >= TH) continue;
            for (int ix = 0; ix < L.nx; ++ix) {
                const int rx = ox - L.xs[ix] * sf;
                if (rx < 0 || rx >= TW) continue;
                const float* t = tiles + (((long long)(iy * L.nx + ix) * TH + ry) * TW + rx) * c;
                for (int ch = 0; ch < c; ++ch) e[ch] += t[ch];
                cnt += 1.f;
            }
        }
        for (int ch = 0; ch < c; ++ch) {
            const float v = e[ch] / cnt;
            if (out_f32) out_f32[(long long)ch * total + i] = v;
            if (out_u8) {
                const float cl = fminf(fmaxf(v, 0.f), 1.f);
                const int oc = rev ? (c - 1 - ch) : ch;
                out_u8[i * c + oc] = (uint8_t)rintf(cl * 255.0f);
            }
        }
    }
}

int tiles_stitch_rows(const float* tiles, int c, int th, int tw, int sf, const int* ys, const int* xs, int ny, int nx,
                      int h, int w, int crop_h, int crop_w, int row0, int rows, float* out_f32, uint8_t* out_u8,
                      int u8_reverse_channels, cudaStream_t stream) {
    IdxLists L;
    IRB_TRY(fill_idx(L, ys, xs, ny, nx));
    IRB_REQUIRE(c >= 1 && c <= 3 && sf >= 1, "tiles_stitch: 1..3 channels, sf >= 1");
    IRB_REQUIRE(crop_h <= h * sf && crop_w <= w * sf && crop_h > 0 && crop_w > 0, "tiles_stitch: crop outside the image");
    IRB_REQUIRE(row0 >= 0 && rows >= 0 && row0 + rows <= crop_h, "tiles_stitch: row range outside the crop");
    if (rows == 0) return 0;
    const long long total = (long long)rows * crop_w;
    const int blocks = int(ceil_div_ll(total, 256) > 148 * 16 ? 148 * 16 : ceil_div_ll(total, 256));
    tiles_stitch_kernel<<<blocks, 256, 0, stream>>>(tiles, c, th, tw, sf, L, h, w, row0, rows, crop_w, out_f32, out_u8,
                                                    u8_reverse_channels);
    IRB_CHECK_CUDA(cudaGetLastError());
    // algorithmic bytes: the tile rows that cover the strip, read once, + the strip written once
    const double frac = double(rows) / crop_h;
    prof_mark("tiles_stitch", c, ny * nx, 0.0,
              frac * 4.0 * ny * nx * th * tw * sf * sf * c + double(total) * c * ((out_f32 ? 4 : 0) + (out_u8 ? 1 : 0)),
              stream);
    return 0;
}

// HATModel.tile_process, hat_model.py:92-105: the tile's own area (its output minus the tile_pad border) replaces the
// output rectangle; no blending. tile: channels-last [TH][TW][c]; the rectangle [ph][pw] read at (oy, ox) of the tile is
// written at (Y0, X0) of the outputs out_f32 [c][out_h][out_w] / out_u8 [out_h][out_w][c], clipped to them
// (post_process :107-109 crops the window padding).
__global__ void tile_paste_kernel(const float* __restrict__ tile, int c, int TW, int oy, int ox, int ph, int pw, int Y0,
                                  int X0, int out_h, int out_w, float* __restrict__ out_f32,
                                  uint8_t* __restrict__ out_u8, int rev) {
    const long long total = (long long)ph * pw;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        const int y = int(i / pw), x = int(i % pw);
        const float* t = tile + ((long long)(oy + y) * TW + ox + x) * c;
        const long long o = (long long)(Y0 + y) * out_w + X0 + x;
        for (int ch = 0; ch < c; ++ch) {
            const float v = t[ch];
            if (out_f32) out_f32[(long long)ch * out_h * out_w + o] = v;
            if (out_u8) {
                const float cl = fminf(fmaxf(v, 0.f), 1.f);
                out_u8[o * c + (rev ? (c - 1 - ch) : ch)] = (uint8_t)rintf(cl * 255.0f);
            }
        }
    }
}

int tile_paste(const float* tile, int c, int TH, int TW, int oy, int ox, int ph, int pw, int Y0, int X0, int out_h,
               int out_w, float* out_f32, uint8_t* out_u8, int u8_reverse_channels, cudaStream_t stream) {
    IRB_REQUIRE(c >= 1 && c <= 3, "tile_paste: 1..3 channels");
    IRB_REQUIRE(oy >= 0 && ox >= 0 && oy + ph <= TH && ox + pw <= TW, "tile_paste: rectangle outside the tile");
    if (Y0 + ph > out_h) ph = out_h - Y0;
    if (X0 + pw > out_w) pw = out_w - X0;
    if (ph <= 0 || pw <= 0) return 0;  // the rectangle lies entirely in the cropped padding
    const long long total = (long long)ph * pw;
    const int blocks = int(ceil_div_ll(total, 256) > 148 * 16 ? 148 * 16 : ceil_div_ll(total, 256));
    tile_paste_kernel<<<blocks, 256, 0, stream>>>(tile, c, TW, oy, ox, ph, pw, Y0, X0, out_h, out_w, out_f32, out_u8,
                                                  u8_reverse_channels);
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("tile_paste", c, 0, 0.0, double(total) * c * (4 + (out_f32 ? 4 : 0) + (out_u8 ? 1 : 0)), stream);
    return 0;
}

int tiles_stitch(const float* tiles, int c, int th, int tw, int sf, const int* ys, const int* xs, int ny, int nx, int h,
                 int w, int crop_h, int crop_w, float* out_f32, uint8_t* out_u8, int u8_reverse_channels,
                 cudaStream_t stream) {
    return tiles_stitch_rows(tiles, c, th, tw, sf, ys, xs, ny, nx, h, w, crop_h, crop_w, 0, crop_h, out_f32, out_u8,
                             u8_reverse_channels, stream);
}

}  // namespace irb
