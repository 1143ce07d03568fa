// Weights blob parsing, weight packing, Model plumbing, and the model-level C ABI (irb_create ... irb_restore_image).
#include "model.h"
#include <algorithm>
#include <string.h>

namespace irb {

int Blob::parse(const void* p, size_t nbytes) {
    const unsigned char* b = reinterpret_cast<const unsigned char*>(p);
    if (!p || nbytes < 16 || memcmp(b, "IRBW", 4) != 0) {
        set_error("weights blob: bad magic (expected IRBW)");
        return 3;
    }
    uint32_t ver, n;
    memcpy(&ver, b + 4, 4);
    memcpy(&n, b + 8, 4);
    if (ver != 1) {
        set_error("weights blob: unsupported version");
        return 3;
    }
    const size_t entry = 96 + 4 + 16 + 8 + 8;
    if (16 + size_t(n) * entry > nbytes) {
        set_error("weights blob: truncated table");
        return 3;
    }
    for (uint32_t i = 0; i < n; ++i) {
        const unsigned char* e = b + 16 + size_t(i) * entry;
        char name[97];
        memcpy(name, e, 96);
        name[96] = 0;
        uint32_t ndim, dims[4];
        uint64_t off, numel;
        memcpy(&ndim, e + 96, 4);
        memcpy(dims, e + 100, 16);
        memcpy(&off, e + 116, 8);
        memcpy(&numel, e + 124, 8);
        if (ndim > 4 || off % 4 != 0 || off + numel * 4 > nbytes) {
            set_error(std::string("weights blob: bad entry ") + name);
            return 3;
        }
        BlobTensor t{};
        t.data = reinterpret_cast<const float*>(b + off);
        t.ndim = int(ndim);
        long long prod = 1;
        for (int k = 0; k < 4; ++k) {
            t.dims[k] = k < int(ndim) ? dims[k] : 1;
            prod *= t.dims[k];
        }
        t.numel = (long long)numel;
        if (prod != t.numel) {
            set_error(std::string("weights blob: numel mismatch for ") + name);
            return 3;
        }
        this->t[name] = t;
    }
    return 0;
}

int Blob::get(const std::string& name, BlobTensor& out, long long d0, long long d1, long long d2, long long d3) const {
    auto it = t.find(name);
    if (it == t.end()) {
        set_error("weights blob: missing tensor " + name);
        return 3;
    }
    // unspecified trailing dims must be 1 (dims beyond ndim are stored as 1)
    const long long want[4] = {d0, d1 < 0 ? 1 : d1, d2 < 0 ? 1 : d2, d3 < 0 ? 1 : d3};
    bool ok = true;
    for (int k = 0; k < 4; ++k)
        if (it->second.dims[k] != want[k]) ok = false;
    if (!ok) {
        char buf[256];
        snprintf(buf, sizeof buf, "weights blob: tensor %s has shape [%lld,%lld,%lld,%lld], expected [%lld,%lld,%lld,%lld]",
                 name.c_str(), it->second.dims[0], it->second.dims[1], it->second.dims[2], it->second.dims[3], d0, d1, d2, d3);
        set_error(buf);
        return 3;
    }
    out = it->second;
    return 0;
}

Model::~Model() {
    for (void* p : dev_allocs) cudaFree(p);
}
int Model::dev_alloc(size_t bytes, void** out) {
    void* p = nullptr;
    IRB_CHECK_CUDA(cudaMalloc(&p, bytes < 16 ? 16 : bytes));
    dev_allocs.push_back(p);
    *out = p;
    return 0;
}
int Model::upload_f32(const float* host, size_t n, float** out) {
    void* p;
    IRB_TRY(dev_alloc(n * 4, &p));
    IRB_CHECK_CUDA(cudaMemcpy(p, host, n * 4, cudaMemcpyHostToDevice));
    *out = reinterpret_cast<float*>(p);
    return 0;
}
int Model::upload_act(const float* host, size_t n, void** out) {
    if (dt == DT_F32) {
        float* f;
        IRB_TRY(upload_f32(host, n, &f));
        *out = f;
        return 0;
    }
    std::vector<bf16> tmp(n);
    for (size_t i = 0; i < n; ++i) tmp[i] = __float2bfloat16_rn(host[i]);
    void* p;
    IRB_TRY(dev_alloc(n * 2, &p));
    IRB_CHECK_CUDA(cudaMemcpy(p, tmp.data(), n * 2, cudaMemcpyHostToDevice));
    *out = p;
    return 0;
}

int Model::upload_bf16(const float* host, size_t n, void** out) {
    std::vector<bf16> tmp(n);
    for (size_t i = 0; i < n; ++i) tmp[i] = __float2bfloat16_rn(host[i]);
    void* p;
    IRB_TRY(dev_alloc(n * 2, &p));
    IRB_CHECK_CUDA(cudaMemcpy(p, tmp.data(), n * 2, cudaMemcpyHostToDevice));
    *out = p;
    return 0;
}

std::vector<float> pack_conv3x3(const BlobTensor& w, int& kpad, int& npad) {
    const int N = int(w.dims[0]), C = int(w.dims[1]);
    kpad = (C + 63) / 64 * 64;
    npad = (N + 15) / 16 * 16;
    std::vector<float> o(size_t(npad) * 9 * kpad, 0.f);
    for (int n = 0; n < N; ++n)
        for (int c = 0; c < C; ++c)
            for (int t = 0; t < 9; ++t) o[(size_t(n) * 9 + t) * kpad + c] = w.data[(size_t(n) * C + c) * 9 + t];
    return o;
}
// PixelShuffle(2) convolutions: permute the packed weight rows (and the bias) sub-pixel major, row (i*2+j) * N/4 + c
// <- row c*4 + i*2 + j, so that Epi::ps = 3 applies (common.cuh).
void permute_rows_ps_major(std::vector<float>& w, int N, int row_len, std::vector<float>* bias) {
    const int cout = N / 4;
    std::vector<float> o(w.size(), 0.f);
    for (int c = 0; c < cout; ++c)
        for (int sub = 0; sub < 4; ++sub)
            std::copy(w.begin() + size_t(c * 4 + sub) * row_len, w.begin() + size_t(c * 4 + sub + 1) * row_len,
                      o.begin() + size_t(sub * cout + c) * row_len);
    std::copy(w.begin() + size_t(N) * row_len, w.end(), o.begin() + size_t(N) * row_len);  // padding rows
    w.swap(o);
    if (bias) {
        std::vector<float> b(bias->size());
        for (int c = 0; c < cout; ++c)
            for (int sub = 0; sub < 4; ++sub) b[sub * cout + c] = (*bias)[c * 4 + sub];
        bias->swap(b);
    }
}

// First-layer convolutions (1 or 3 input channels) run on the tensor core with the fp32 input split into two bf16
// terms x = hi + lo (split_hi_lo_bf16, pointwise.cu): channel c holds hi, channel C + c holds lo, and both get the
// same weight, so the input keeps ~16 mantissa bits at no cost (K is padded to 64 anyway).
std::vector<float> pack_conv3x3_split(const BlobTensor& w, int& kpad, int& npad) {
    const int N = int(w.dims[0]), C = int(w.dims[1]);
    kpad = 64;
    npad = (N + 15) / 16 * 16;
    std::vector<float> o(size_t(npad) * 9 * kpad, 0.f);
    for (int n = 0; n < N; ++n)
        for (int c = 0; c < C; ++c)
            for (int t = 0; t < 9; ++t) {
                const float v = w.data[(size_t(n) * C + c) * 9 + t];
                o[(size_t(n) * 9 + t) * kpad + c] = v;
                o[(size_t(n) * 9 + t) * kpad + C + c] = v;
            }
    return o;
}
std::vector<float> pack_linear(const float* w, int N, int K, int& kpad, int& npad) {
    kpad = (K + 7) / 8 * 8;
    npad = (N + 15) / 16 * 16;
    std::vector<float> o(size_t(npad) * kpad, 0.f);
    for (int n = 0; n < N; ++n)
        for (int k = 0; k < K; ++k) o[size_t(n) * kpad + k] = w[size_t(n) * K + k];
    return o;
}
std::vector<float> pack_dw(const BlobTensor& w) {
    const int C = int(w.dims[0]);
    std::vector<float> o(size_t(9) * C);
    for (int c = 0; c < C; ++c)
        for (int t = 0; t < 9; ++t) o[size_t(t) * C + c] = w.data[size_t(c) * 9 + t];
    return o;
}

__global__ void permute_f32_kernel(const float* __restrict__ in, float* __restrict__ out, int n, int c, long long hw,
                                   int to_nhwc) {
    const long long total = (long long)n * c * hw;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (long long)gridDim.x * blockDim.x) {
        // i indexes the OUTPUT
        if (to_nhwc) {
            const int ch = int(i % c);
            const long long p = (i / c) % hw;
            const long long b = i / (c * hw);
            out[i] = in[(b * c + ch) * hw + p];
        } else {
            const long long p = i % hw;
            const int ch = int((i / hw) % c);
            const long long b = i / (hw * c);
            out[i] = in[(b * hw + p) * c + ch];
        }
    }
}
static int permute_launch(const float* in, float* out, int n, int c, int h, int w, int to_nhwc, cudaStream_t s) {
    const long long total = (long long)n * c * h * w;
    const int blocks = int(ceil_div_ll(total, 256) > 148 * 16 ? 148 * 16 : ceil_div_ll(total, 256));
    permute_f32_kernel<<<blocks, 256, 0, s>>>(in, out, n, c, (long long)h * w, to_nhwc);
    IRB_CHECK_CUDA(cudaGetLastError());
    prof_mark("permute_f32", c, to_nhwc, 0.0, 8.0 * total, s);
    return 0;
}
int nchw_to_nhwc_f32(const float* in, float* out, int n, int c, int h, int w, cudaStream_t s) {
    return permute_launch(in, out, n, c, h, w, 1, s);
}
int nhwc_to_nchw_f32(const float* in, float* out, int n, int c, int h, int w, cudaStream_t s) {
    return permute_launch(in, out, n, c, h, w, 0, s);
}

}  // namespace irb

using namespace irb;

// ---------------------------------------------------------------------------------------------------------
static int pick_batch(Model* m, int nt, int th, int tw) {
    if (m->tile_batch > 0) return int(m->tile_batch < nt ? m->tile_batch : nt);
    int nb = nt;
    const size_t cap = size_t(40) << 30;
    while (nb > 1 && m->forward_workspace(nb, th, tw) > cap) nb = (nb + 1) / 2;
    return nb;
}

struct RestorePlan {
    int hp, wp, th, tw, ny, nx, nb, sf, c;
    int ys[128], xs[128];
    // workspace pointers
    float *padded, *tiles_in, *tiles_out;
    void* model_ws;
    size_t model_ws_bytes, total;
};

static int make_plan(Model* m, int img_h, int img_w, int tile, int overlap, void* ws, RestorePlan& P) {
    IRB_REQUIRE(img_h > 0 && img_w > 0, "restore: empty image");
    IRB_TRY(irb_padded_size(m->desc.arch, m->desc.arch == IRB_ARCH_XRESTORMER ? m->tile_multiple() : m->desc.window_size,
                            img_h, img_w, &P.hp, &P.wp));
    P.sf = m->scale;
    P.c = m->cin;
    if (tile <= 0) {
        P.th = P.hp;
        P.tw = P.wp;
        P.ny = P.nx = 1;
        P.ys[0] = P.xs[0] = 0;
    } else {
        int t = tile < P.hp ? tile : P.hp;
        t = t < P.wp ? t : P.wp;
        // inference.py:302 / demo.py:136
        IRB_REQUIRE((m->desc.arch == IRB_ARCH_SWINIR && !m->strict_tile) || t % m->tile_multiple() == 0,
                    "tile size should be a multiple of window_size (SwinIR) / 8 (Restormer)");
        IRB_REQUIRE(overlap >= 0 && overlap < t, "tile_overlap must be smaller than the tile");
        P.th = P.tw = t;
        IRB_REQUIRE((P.hp - t) / (t - overlap) + 2 <= 128 && (P.wp - t) / (t - overlap) + 2 <= 128,
                    "too many tiles per axis (max 128)");
        IRB_TRY(irb_plan_tiles(P.hp, t, overlap, P.ys, &P.ny));
        IRB_TRY(irb_plan_tiles(P.wp, t, overlap, P.xs, &P.nx));
    }
    const int nt = P.ny * P.nx;
    P.nb = pick_batch(m, nt, P.th, P.tw);
    Bump b(ws);
    P.padded = b.take<float>(size_t(P.c) * P.hp * P.wp * 4);
    P.tiles_in = b.take<float>(size_t(nt) * P.th * P.tw * P.c * 4);
    P.tiles_out = b.take<float>(size_t(nt) * P.th * P.tw * P.sf * P.sf * P.c * 4);
    P.model_ws_bytes = m->forward_workspace(P.nb, P.th, P.tw);
    P.model_ws = b.take<void>(P.model_ws_bytes);
    P.total = b.off + 256;
    return 0;
}

// Every model entry point runs on the model's device whatever device the calling thread has current (a Flask worker
// thread defaults to device 0); the caller's current device is restored on return.
struct DeviceGuard {
    int prev = -1;
    cudaError_t err = cudaSuccess;
    explicit DeviceGuard(int dev) {
        err = cudaGetDevice(&prev);
        if (err == cudaSuccess && prev != dev) err = cudaSetDevice(dev);
        else if (err == cudaSuccess) prev = -1;
    }
    ~DeviceGuard() {
        if (prev >= 0) cudaSetDevice(prev);
    }
};
#define IRB_ON_DEVICE(m)            \
    DeviceGuard _dg((m)->device);   \
    IRB_CHECK_CUDA(_dg.err)

extern "C" {

int irb_create(const irb_model_desc* desc, const void* weights_blob, size_t nbytes, int device, irb_handle** out) {
    IRB_REQUIRE(desc && out, "irb_create: NULL argument");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        set_error("no CUDA device: irb200 has no CPU fallback");
        return 2;
    }
    IRB_REQUIRE(device >= 0 && device < ndev, "irb_create: bad device index");
    DeviceGuard _dg(device);
    IRB_CHECK_CUDA(_dg.err);
    cudaDeviceProp prop;
    IRB_CHECK_CUDA(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) {
        set_error("irb200 kernels are built for sm_100a only (device is sm_" + std::to_string(prop.major) +
                  std::to_string(prop.minor) + ")");
        return 2;
    }
    IRB_REQUIRE(desc->precision == IRB_F32 || desc->precision == IRB_BF16, "irb_create: bad precision");
    Blob blob;
    IRB_TRY(blob.parse(weights_blob, nbytes));
    Model* m = nullptr;
    int rc;
    if (desc->arch == IRB_ARCH_RESTORMER || desc->arch == IRB_ARCH_XRESTORMER)
        rc = build_restormer(*desc, blob, device, &m);
    else if (desc->arch == IRB_ARCH_SWINIR || desc->arch == IRB_ARCH_HAT)
        rc = build_swinir(*desc, blob, device, &m);
    else {
        set_error("irb_create: unknown arch");
        return 1;
    }
    if (rc != 0) {
        delete m;
        return rc;
    }
    IRB_CHECK_CUDA(cudaDeviceSynchronize());
    *out = new irb_handle{m};
    return 0;
}

void irb_destroy(irb_handle* h) {
    if (!h) return;
    delete h->model;
    delete h;
}

int irb_set_option(irb_handle* h, const char* key, long long value) {
    IRB_REQUIRE(h && key, "irb_set_option: NULL argument");
    if (strcmp(key, "tile_batch") == 0) {
        IRB_REQUIRE(value >= 0, "tile_batch must be >= 0");
        h->model->tile_batch = value;
        return 0;
    }
    if (strcmp(key, "fuse_mlp") == 0) {
        h->model->fuse_mlp = value != 0;
        return 0;
    }
    if (strcmp(key, "fuse_pwdw") == 0) {
        h->model->fuse_pwdw = value != 0;
        return 0;
    }
    if (strcmp(key, "fp32_tc") == 0) {
        h->model->fp32_tc = value != 0;
        return 0;
    }
    if (strcmp(key, "strict_tile") == 0) {
        h->model->strict_tile = value != 0;
        return 0;
    }
    set_error(std::string("irb_set_option: unknown key ") + key);
    return 1;
}

int irb_scale(const irb_handle* h) { return h ? h->model->scale : 0; }

int irb_tile_batch(irb_handle* h, int n_tiles, int tile_h, int tile_w, int* out_batch) {
    IRB_REQUIRE(h && out_batch && n_tiles > 0 && tile_h > 0 && tile_w > 0, "irb_tile_batch: bad arguments");
    *out_batch = pick_batch(h->model, n_tiles, tile_h, tile_w);
    return 0;
}

int irb_workspace_bytes(irb_handle* h, int n_tiles, int tile_h, int tile_w, size_t* out_bytes) {
    IRB_REQUIRE(h && out_bytes && n_tiles > 0 && tile_h > 0 && tile_w > 0, "irb_workspace_bytes: bad arguments");
    Model* m = h->model;
    const size_t px = size_t(n_tiles) * tile_h * tile_w * m->cin * 4;
    *out_bytes = m->forward_workspace(n_tiles, tile_h, tile_w) + px + px * m->scale * m->scale + 1024;
    return 0;
}

int irb_forward_tiles(irb_handle* h, const float* d_in, float* d_out, int n_tiles, int tile_h, int tile_w,
                      void* d_workspace, size_t workspace_bytes, void* stream) {
    IRB_REQUIRE(h && d_in && d_out && d_workspace, "irb_forward_tiles: NULL argument");
    Model* m = h->model;
    IRB_ON_DEVICE(m);
    size_t need;
    IRB_TRY(irb_workspace_bytes(h, n_tiles, tile_h, tile_w, &need));
    IRB_REQUIRE(workspace_bytes >= need, "irb_forward_tiles: workspace too small");
    cudaStream_t s = (cudaStream_t)stream;
    Bump b(d_workspace);
    const size_t px = size_t(n_tiles) * tile_h * tile_w * m->cin * 4;
    float* in_l = b.take<float>(px);
    float* out_l = b.take<float>(px * m->scale * m->scale);
    const size_t mws = m->forward_workspace(n_tiles, tile_h, tile_w);
    void* mw = b.take<void>(mws);
    IRB_TRY(nchw_to_nhwc_f32(d_in, in_l, n_tiles, m->cin, tile_h, tile_w, s));
    IRB_TRY(m->forward(in_l, out_l, n_tiles, tile_h, tile_w, mw, mws, s));
    IRB_TRY(nhwc_to_nchw_f32(out_l, d_out, n_tiles, m->cin, tile_h * m->scale, tile_w * m->scale, s));
    return 0;
}

int irb_restore_workspace_bytes(irb_handle* h, int img_h, int img_w, int tile, int overlap, size_t* out_bytes) {
    IRB_REQUIRE(h && out_bytes, "irb_restore_workspace_bytes: NULL argument");
    RestorePlan P;
    IRB_TRY(make_plan(h->model, img_h, img_w, tile, overlap, nullptr, P));
    *out_bytes = P.total;
    return 0;
}

int irb_restore_image(irb_handle* h, const void* d_in, int in_fmt, int reverse_channels, int img_h, int img_w,
                      int tile, int overlap, uint8_t* d_out_u8, float* d_out_f32, void* d_workspace,
                      size_t workspace_bytes, void* stream) {
    IRB_REQUIRE(h && d_in && d_workspace, "irb_restore_image: NULL argument");
    IRB_REQUIRE(d_out_u8 || d_out_f32, "irb_restore_image: no output requested");
    IRB_REQUIRE(in_fmt == 0 || in_fmt == 1, "irb_restore_image: in_fmt must be 0 (u8 HWC) or 1 (f32 planar)");
    Model* m = h->model;
    IRB_ON_DEVICE(m);
    RestorePlan P;
    IRB_TRY(make_plan(m, img_h, img_w, tile, overlap, d_workspace, P));
    IRB_REQUIRE(workspace_bytes >= P.total, "irb_restore_image: workspace too small");
    cudaStream_t s = (cudaStream_t)stream;
    const int pad_mode = m->desc.arch == IRB_ARCH_SWINIR ? 1 : 0;
    if (in_fmt == 0)
        IRB_TRY(image_prepare(reinterpret_cast<const uint8_t*>(d_in), img_h, img_w, P.c, reverse_channels, P.padded,
                              P.hp, P.wp, pad_mode, s));
    else
        IRB_TRY(image_pad_f32(reinterpret_cast<const float*>(d_in), P.c, img_h, img_w, P.padded, P.hp, P.wp, pad_mode, s));
    IRB_TRY(tiles_gather(P.padded, P.c, P.hp, P.wp, P.ys, P.xs, P.ny, P.nx, P.th, P.tw, P.tiles_in, nullptr, 1.0f, s));
    const int nt = P.ny * P.nx;
    const size_t in_stride = size_t(P.th) * P.tw * P.c;
    const size_t out_stride = in_stride * P.sf * P.sf;
    for (int t0 = 0; t0 < nt; t0 += P.nb) {
        const int nb = nt - t0 < P.nb ? nt - t0 : P.nb;
        IRB_TRY(m->forward(P.tiles_in + t0 * in_stride, P.tiles_out + t0 * out_stride, nb, P.th, P.tw, P.model_ws,
                           P.model_ws_bytes, s));
    }
    IRB_TRY(tiles_stitch(P.tiles_out, P.c, P.th, P.tw, P.sf, P.ys, P.xs, P.ny, P.nx, P.hp, P.wp, img_h * P.sf,
                         img_w * P.sf, d_out_f32, d_out_u8, reverse_channels, s));
    return 0;
}

// ---- HATModel's tile mode (hat_model.py:16-109): crop-with-pad tiles, no blending -------------------------------
struct CropPlan {
    int hp, wp, ny, nx, ts, tp, mth, mtw, sf, c;
    float *padded, *tile_in, *tile_out;
    void* model_ws;
    size_t model_ws_bytes, total;
};
static int make_crop_plan(Model* m, int img_h, int img_w, int tile_size, int tile_pad, void* ws, CropPlan& P) {
    IRB_REQUIRE(img_h > 0 && img_w > 0, "restore: empty image");
    IRB_TRY(irb_padded_size(m->desc.arch, m->desc.arch == IRB_ARCH_XRESTORMER ? m->tile_multiple() : m->desc.window_size,
                            img_h, img_w, &P.hp, &P.wp));
    P.sf = m->scale;
    P.c = m->cin;
    if (tile_size <= 0) {  // HATModel.process: the whole padded image
        P.ts = P.hp > P.wp ? P.hp : P.wp;
        P.tp = 0;
    } else {
        // hat.yml:7-9: both must be integer multiples of the window size (the network does not pad)
        IRB_REQUIRE(tile_size % m->tile_multiple() == 0 && tile_pad >= 0 && tile_pad % m->tile_multiple() == 0,
                    "tile_size and tile_pad must be integer multiples of the window size");
        P.ts = tile_size;
        P.tp = tile_pad;
    }
    P.ny = (P.hp + P.ts - 1) / P.ts;
    P.nx = (P.wp + P.ts - 1) / P.ts;
    P.mth = P.ts + 2 * P.tp < P.hp ? P.ts + 2 * P.tp : P.hp;
    P.mtw = P.ts + 2 * P.tp < P.wp ? P.ts + 2 * P.tp : P.wp;
    Bump b(ws);
    P.padded = b.take<float>(size_t(P.c) * P.hp * P.wp * 4);
    P.tile_in = b.take<float>(size_t(P.mth) * P.mtw * P.c * 4);
    P.tile_out = b.take<float>(size_t(P.mth) * P.mtw * P.sf * P.sf * P.c * 4);
    P.model_ws_bytes = m->forward_workspace(1, P.mth, P.mtw);
    P.model_ws = b.take<void>(P.model_ws_bytes);
    P.total = b.off + 256;
    return 0;
}

int irb_restore_crop_workspace_bytes(irb_handle* h, int img_h, int img_w, int tile_size, int tile_pad, size_t* out_bytes) {
    IRB_REQUIRE(h && out_bytes, "irb_restore_crop_workspace_bytes: NULL argument");
    CropPlan P;
    IRB_TRY(make_crop_plan(h->model, img_h, img_w, tile_size, tile_pad, nullptr, P));
    *out_bytes = P.total;
    return 0;
}

int irb_restore_image_crop(irb_handle* h, const void* d_in, int in_fmt, int reverse_channels, int img_h, int img_w,
                           int tile_size, int tile_pad, uint8_t* d_out_u8, float* d_out_f32, void* d_workspace,
                           size_t workspace_bytes, void* stream) {
    IRB_REQUIRE(h && d_in && d_workspace, "irb_restore_image_crop: NULL argument");
    IRB_REQUIRE(d_out_u8 || d_out_f32, "irb_restore_image_crop: no output requested");
    IRB_REQUIRE(in_fmt == 0 || in_fmt == 1, "irb_restore_image_crop: in_fmt must be 0 (u8 HWC) or 1 (f32 planar)");
    Model* m = h->model;
    IRB_ON_DEVICE(m);
    CropPlan P;
    IRB_TRY(make_crop_plan(m, img_h, img_w, tile_size, tile_pad, d_workspace, P));
    IRB_REQUIRE(workspace_bytes >= P.total, "irb_restore_image_crop: workspace too small");
    cudaStream_t s = (cudaStream_t)stream;
    const int pad_mode = m->desc.arch == IRB_ARCH_SWINIR ? 1 : 0;
    if (in_fmt == 0)
        IRB_TRY(image_prepare(reinterpret_cast<const uint8_t*>(d_in), img_h, img_w, P.c, reverse_channels, P.padded,
                              P.hp, P.wp, pad_mode, s));
    else
        IRB_TRY(image_pad_f32(reinterpret_cast<const float*>(d_in), P.c, img_h, img_w, P.padded, P.hp, P.wp, pad_mode, s));
    const int sf = P.sf;
    for (int ty = 0; ty < P.ny; ++ty)
        for (int tx = 0; tx < P.nx; ++tx) {
            const int y0 = ty * P.ts, x0 = tx * P.ts;
            const int y1 = y0 + P.ts < P.hp ? y0 + P.ts : P.hp, x1 = x0 + P.ts < P.wp ? x0 + P.ts : P.wp;
            const int y0p = y0 - P.tp > 0 ? y0 - P.tp : 0, x0p = x0 - P.tp > 0 ? x0 - P.tp : 0;
            const int y1p = y1 + P.tp < P.hp ? y1 + P.tp : P.hp, x1p = x1 + P.tp < P.wp ? x1 + P.tp : P.wp;
            const int th = y1p - y0p, tw = x1p - x0p;
            IRB_TRY(tiles_gather(P.padded, P.c, P.hp, P.wp, &y0p, &x0p, 1, 1, th, tw, P.tile_in, nullptr, 1.0f, s));
            IRB_TRY(m->forward(P.tile_in, P.tile_out, 1, th, tw, P.model_ws, P.model_ws_bytes, s));
            IRB_TRY(tile_paste(P.tile_out, P.c, th * sf, tw * sf, (y0 - y0p) * sf, (x0 - x0p) * sf, (y1 - y0) * sf,
                               (x1 - x0) * sf, y0 * sf, x0 * sf, img_h * sf, img_w * sf, d_out_f32, d_out_u8,
                               reverse_channels, s));
        }
    return 0;
}

// ---- building blocks of irb_restore_image, exposed so that tiles can be sharded across ranks -------------------
int irb_restore_plan(irb_handle* h, int img_h, int img_w, int tile, int overlap, int* hp, int* wp, int* th, int* tw,
                     int* ys, int* ny, int* xs, int* nx) {
    IRB_REQUIRE(h && hp && wp && th && tw && ys && ny && xs && nx, "irb_restore_plan: NULL argument");
    RestorePlan P;
    IRB_TRY(make_plan(h->model, img_h, img_w, tile, overlap, nullptr, P));
    *hp = P.hp;
    *wp = P.wp;
    *th = P.th;
    *tw = P.tw;
    *ny = P.ny;
    *nx = P.nx;
    for (int i = 0; i < P.ny; ++i) ys[i] = P.ys[i];
    for (int i = 0; i < P.nx; ++i) xs[i] = P.xs[i];
    return 0;
}

int irb_tiles_prepare(irb_handle* h, const void* d_in, int in_fmt, int reverse_channels, int img_h, int img_w,
                      int tile, int overlap, float* d_padded, float* d_tiles, void* stream) {
    IRB_REQUIRE(h && d_in && d_padded && d_tiles, "irb_tiles_prepare: NULL argument");
    Model* m = h->model;
    IRB_ON_DEVICE(m);
    RestorePlan P;
    IRB_TRY(make_plan(m, img_h, img_w, tile, overlap, nullptr, P));
    cudaStream_t s = (cudaStream_t)stream;
    const int pad_mode = m->desc.arch == IRB_ARCH_SWINIR ? 1 : 0;
    if (in_fmt == 0)
        IRB_TRY(image_prepare(reinterpret_cast<const uint8_t*>(d_in), img_h, img_w, P.c, reverse_channels, d_padded,
                              P.hp, P.wp, pad_mode, s));
    else
        IRB_TRY(image_pad_f32(reinterpret_cast<const float*>(d_in), P.c, img_h, img_w, d_padded, P.hp, P.wp, pad_mode, s));
    return tiles_gather(d_padded, P.c, P.hp, P.wp, P.ys, P.xs, P.ny, P.nx, P.th, P.tw, d_tiles, nullptr, 1.0f, s);
}

int irb_forward_tiles_cl(irb_handle* h, const float* d_in, float* d_out, int n_tiles, int tile_h, int tile_w,
                         void* d_workspace, size_t workspace_bytes, void* stream) {
    IRB_REQUIRE(h && d_in && d_out && d_workspace && n_tiles > 0, "irb_forward_tiles_cl: bad argument");
    Model* m = h->model;
    IRB_ON_DEVICE(m);
    const size_t need = m->forward_workspace(n_tiles, tile_h, tile_w);
    IRB_REQUIRE(workspace_bytes >= need, "irb_forward_tiles_cl: workspace too small");
    return m->forward(d_in, d_out, n_tiles, tile_h, tile_w, d_workspace, workspace_bytes, (cudaStream_t)stream);
}

}  // extern "C"
