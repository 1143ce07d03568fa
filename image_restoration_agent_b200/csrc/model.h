// Host-side model runner plumbing: weights blob, device uploads, workspace bump allocator, Model interface.
#pragma once
#include "../../include/irb200.h"
#include "common.cuh"
#include "kernels.h"
#include <map>
#include <string>
#include <vector>

namespace irb {

// ---- IRBW v1 weights blob (written by image_restoration_agent_b200/weights.py) --------------------------
//   header : char magic[4] = "IRBW"; u32 version = 1; u32 n_tensors; u32 reserved
//   entry  : char name[96]; u32 ndim; u32 dims[4]; u64 offset (bytes from blob start, 16-aligned); u64 numel
//   data   : fp32, little endian
struct BlobTensor {
    const float* data;
    int ndim;
    long long dims[4];
    long long numel;
};
struct Blob {
    std::map<std::string, BlobTensor> t;
    int parse(const void* p, size_t nbytes);
    // fetches a tensor and checks its shape (d* = -1 matches anything)
    int get(const std::string& name, BlobTensor& out, long long d0, long long d1 = -1, long long d2 = -1,
            long long d3 = -1) const;
    bool has(const std::string& name) const { return t.count(name) != 0; }
};

struct Bump {
    char* base;
    size_t off;
    explicit Bump(void* b) : base(reinterpret_cast<char*>(b)), off(0) {}
    template <typename T = void>
    T* take(size_t bytes) {
        const size_t a = (off + 255) & ~size_t(255);
        off = a + bytes;
        return base ? reinterpret_cast<T*>(base + a) : nullptr;
    }
};

struct Model {
    irb_model_desc desc{};
    int dt = DT_BF16;  // storage / arithmetic mode
    int device = 0;
    int cin = 3;
    int scale = 1;
    long long tile_batch = 0;
    // 1: tile % window_size == 0 is required like the SwinIR script (inference.py:302); 0: the SwinIR service path
    // (model_serving.py:146-166 has no such assert: the model reflect-pads ragged tiles itself). Restormer always checks.
    int strict_tile = 1;
    int fuse_mlp = 1;   // SwinIR: fc1 -> GELU -> fc2 -> + x as one kernel (mlp_fused.cu) where the shape allows
    int fuse_pwdw = 0;
    // fp32 mode, SwinIR / HAT: linear layers and 3x3 convolutions on the tensor cores with the operands split into three
    // bf16 terms each (six bf16 products per fp32 product, fp32 accumulation); 0 = fp32 FMA kernels
    int fp32_tc = 1;  // Restormer: 1x1 -> depthwise pairs as one kernel (pwdw_fused.cu); opt-in, see DESIGN.md section 8
    std::vector<void*> dev_allocs;
    virtual ~Model();
    // model(x) on channels-last fp32 tiles: in [n,H,W,cin] -> out [n,H*s,W*s,cin]
    virtual size_t forward_workspace(int n, int H, int W) = 0;
    virtual int forward(const float* in, float* out, int n, int H, int W, void* ws, size_t ws_bytes,
                        cudaStream_t stream) = 0;
    virtual int tile_multiple() const = 0;  // 8 (Restormer) or window_size (SwinIR)
    size_t esize() const { return dt == DT_BF16 ? 2 : 4; }

    // uploads (tracked, freed by the destructor)
    int upload_f32(const float* host, size_t n, float** out);
    int upload_act(const float* host, size_t n, void** out);  // converted to the storage dtype
    int upload_bf16(const float* host, size_t n, void** out); // always bf16
    int dev_alloc(size_t bytes, void** out);
};

int build_restormer(const irb_model_desc& d, const Blob& blob, int device, Model** out);
int build_swinir(const irb_model_desc& d, const Blob& blob, int device, Model** out);

// layout helpers (tiles.cu)
int nchw_to_nhwc_f32(const float* in, float* out, int n, int c, int h, int w, cudaStream_t stream);
int nhwc_to_nchw_f32(const float* in, float* out, int n, int c, int h, int w, cudaStream_t stream);

// conv weight [N][C][3][3] -> [Npad][9*kpad] (K index = tap*kpad + c), linear [N][K] -> [Npad][Kpad]
std::vector<float> pack_conv3x3(const BlobTensor& w, int& kpad, int& npad);
std::vector<float> pack_conv3x3_split(const BlobTensor& w, int& kpad, int& npad);  // C <= 4: hi/lo input split
void permute_rows_ps_major(std::vector<float>& w, int N, int row_len, std::vector<float>* bias);
std::vector<float> pack_linear(const float* w, int N, int K, int& kpad, int& npad);
std::vector<float> pack_dw(const BlobTensor& w);  // [C][1][3][3] -> [9][C]

}  // namespace irb

struct irb_handle {
    irb::Model* model;
};
