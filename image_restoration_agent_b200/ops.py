"""Single-kernel host wrappers (torch tensors in, C-ABI calls out). Used by the parity tests and ncu captures.

Activations are channels-last: [rows, C] or [n, H, W, C] contiguous torch CUDA tensors (fp32 or bf16).
torch is used only for device memory and the current stream.
"""
from __future__ import annotations

import torch

from . import _lib
from ._lib import BF16, F32, check


def _dt(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    raise TypeError(f"irb200 supports fp32 / bf16 activations, got {t.dtype}")


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t):
    return None if t is None else t.data_ptr()


def layernorm(x, w, b=None, bias_free=False, eps=1e-5):
    rows, C = x.numel() // x.shape[-1], x.shape[-1]
    y = torch.empty_like(x)
    check(_lib.load().irb_layernorm(x.data_ptr(), _dt(x), C, y.data_ptr(), C, rows, C, w.data_ptr(), _ptr(b),
                                    int(bias_free), eps, _stream()))
    return y


def gemm1x1(x, w, bias=None, res=None, act=0, slope=0.0, batch=1, w_rows_per_batch=0, N=None):
    """x [batch*M, K], w [rows, K] (same dtype) -> [batch*M, N]."""
    K = x.shape[-1]
    M = (x.numel() // K) // batch
    N = N if N is not None else (w_rows_per_batch if w_rows_per_batch else w.shape[0])
    y = torch.empty((batch * M, N), device=x.device, dtype=x.dtype)
    check(_lib.load().irb_gemm1x1(x.data_ptr(), _dt(x), x.stride(-2), M, K, batch, w.data_ptr(), w.stride(0), N,
                                  w.shape[0], w_rows_per_batch, _ptr(bias), _ptr(res), N if res is not None else 0,
                                  act, slope, y.data_ptr(), N, _stream()))
    return y


def pack_conv3x3_weight(w: torch.Tensor, dtype, row_pad=16) -> tuple[torch.Tensor, int]:
    """nn.Conv2d weight [N, C, 3, 3] -> [N_pad, 9*kpad] with K index = (ky*3+kx)*kpad + c (zero padded)."""
    N, Cin = w.shape[0], w.shape[1]
    kpad = (Cin + 63) // 64 * 64
    npad = (N + row_pad - 1) // row_pad * row_pad
    out = torch.zeros((npad, 9, kpad), device=w.device, dtype=torch.float32)
    out[:N, :, :Cin] = w.float().permute(0, 2, 3, 1).reshape(N, 9, Cin)
    return out.reshape(npad, 9 * kpad).to(dtype).contiguous(), kpad


def conv3x3(x, w_packed, kpad, N, bias=None, res=None, act=0, slope=0.0, shuffle=0):
    """x [n, H, W, C] -> [n, H, W, N] (or the (un)shuffled shape)."""
    n, H, W, Cin = x.shape
    if shuffle == 1:
        y = torch.empty((n, 2 * H, 2 * W, N // 4), device=x.device, dtype=x.dtype)
    elif shuffle == 2:
        y = torch.empty((n, H // 2, W // 2, N * 4), device=x.device, dtype=x.dtype)
    else:
        y = torch.empty((n, H, W, N), device=x.device, dtype=x.dtype)
    check(_lib.load().irb_conv3x3(x.data_ptr(), _dt(x), x.stride(-2), n, H, W, Cin, w_packed.data_ptr(), kpad, N,
                                  w_packed.shape[0], _ptr(bias), _ptr(res), N if res is not None else 0, act, slope,
                                  shuffle, y.data_ptr(), y.shape[-1], _stream()))
    return y


def pack_dw_weight(w: torch.Tensor) -> torch.Tensor:
    """depthwise nn.Conv2d weight [C, 1, 3, 3] -> fp32 [9, C] tap-major."""
    return w.float().reshape(w.shape[0], 9).t().contiguous()


def dwconv3x3(x, w9c, sumsq_channels=0):
    n, H, W, Cc = x.shape
    y = torch.empty((n, H, W, Cc), device=x.device, dtype=x.dtype)
    ss = torch.zeros((n, sumsq_channels), device=x.device, dtype=torch.float32) if sumsq_channels else None
    check(_lib.load().irb_dwconv3x3(x.data_ptr(), _dt(x), x.stride(-2), y.data_ptr(), Cc, n, H, W, Cc, w9c.data_ptr(), _ptr(ss),
                                    sumsq_channels, _stream()))
    return (y, ss) if sumsq_channels else y


def gdfn_dw_gate(x, w9c):
    n, H, W, C2 = x.shape
    hp = C2 // 2
    y = torch.empty((n, H, W, hp), device=x.device, dtype=x.dtype)
    check(_lib.load().irb_gdfn_dw_gate(x.data_ptr(), _dt(x), C2, y.data_ptr(), hp, n, H, W, hp, w9c.data_ptr(),
                                       _stream()))
    return y


def mdta_gram(qkv, heads):
    """-> (Gpart [n, heads, nchunks, c, c], ssq_part [n, heads, nchunks, 2, c]): per-chunk partial sums."""
    n, H, W, C3 = qkv.shape
    Cc = C3 // 3
    c = Cc // heads
    lib = _lib.load()
    nch = lib.irb_mdta_nchunks(H * W, heads, n)
    G = torch.full((n, heads, nch, c, c), float("nan"), device=qkv.device, dtype=torch.float32)
    ss = torch.full((n, heads, nch, 2, c), float("nan"), device=qkv.device, dtype=torch.float32)
    check(lib.irb_mdta_gram(qkv.data_ptr(), _dt(qkv), C3, n, H * W, Cc, heads, nch, G.data_ptr(), ss.data_ptr(),
                            _stream()))
    return G, ss


def mdta_attn(Gpart, ssq_part, temperature, wproj, out_dtype):
    n, heads, nch, c, _ = Gpart.shape
    Cc = heads * c
    weff = torch.empty((n, Cc, Cc), device=Gpart.device, dtype=out_dtype)
    attn = torch.empty((n, heads, c, c), device=Gpart.device, dtype=torch.float32)
    check(_lib.load().irb_mdta_attn(Gpart.data_ptr(), ssq_part.data_ptr(), nch, temperature.data_ptr(),
                                    wproj.data_ptr(), n, Cc, heads, weff.data_ptr(),
                                    BF16 if out_dtype == torch.bfloat16 else F32, attn.data_ptr(), _stream()))
    return weff, attn


def stitch(tiles, ys, xs, sf, h, w, crop_h, crop_w, want_u8=True, reverse_channels=False):
    nt, TH, TW, c = tiles.shape
    th, tw = TH // sf, TW // sf
    out = torch.empty((c, crop_h, crop_w), device=tiles.device, dtype=torch.float32)
    u8 = torch.empty((crop_h, crop_w, c), device=tiles.device, dtype=torch.uint8) if want_u8 else None
    check(_lib.load().irb_stitch(tiles.data_ptr(), c, th, tw, sf, _lib.int_array(ys), _lib.int_array(xs), len(ys),
                                 len(xs), h, w, crop_h, crop_w, out.data_ptr(), _ptr(u8), int(reverse_channels),
                                 _stream()))
    return out, u8


def stitch_rows(tiles, ys, xs, sf, h, w, crop_h, crop_w, row0, rows, want_u8=True, want_f32=False,
                reverse_channels=False, out_u8=None):
    """Blend of output rows [row0, row0+rows) only (a rank's strip). `tiles` holds ALL tile slots [ny*nx, TH, TW, c];
    only the tile rows covering the strip are read. -> (f32 [c, rows, crop_w] | None, u8 [rows, crop_w, c] | None)."""
    nt, TH, TW, c = tiles.shape
    th, tw = TH // sf, TW // sf
    out = torch.empty((c, rows, crop_w), device=tiles.device, dtype=torch.float32) if want_f32 else None
    u8 = out_u8 if out_u8 is not None else (
        torch.empty((rows, crop_w, c), device=tiles.device, dtype=torch.uint8) if want_u8 else None)
    check(_lib.load().irb_stitch_rows(tiles.data_ptr(), c, th, tw, sf, _lib.int_array(ys), _lib.int_array(xs),
                                      len(ys), len(xs), h, w, crop_h, crop_w, row0, rows, _ptr(out), _ptr(u8),
                                      int(reverse_channels), torch.cuda.current_stream(tiles.device).cuda_stream))
    return out, u8


def window_attention(qkv, C, heads, window, shift, bias_table):
    """qkv [n, H, W, 3*heads*dstride] (q pre-scaled) -> [n, H, W, C]."""
    n, H, W, C3 = qkv.shape
    dstride = C3 // (3 * heads)
    y = torch.zeros((n, H, W, C), device=qkv.device, dtype=qkv.dtype)
    check(_lib.load().irb_window_attention(qkv.data_ptr(), _dt(qkv), C3, y.data_ptr(), C, n, H, W, C, heads, dstride,
                                           window, shift, bias_table.data_ptr(), _stream()))
    return y


def hat_attention(qkv, heads, head_dim, window, overlap_window, shift, bias_table):
    """HAT window attention (irb_hat_attention): qkv [n, H, W, 3*heads*dstride] (q pre-scaled, fp32 or bf16) ->
    [n, H, W, heads*head_dim]. overlap_window == window: (shifted-)window self-attention; larger: OCAB."""
    n, H, W, C3 = qkv.shape
    dstride = C3 // (3 * heads)
    C = heads * head_dim
    y = torch.zeros((n, H, W, C), device=qkv.device, dtype=qkv.dtype)
    check(_lib.load().irb_hat_attention(qkv.data_ptr(), _dt(qkv), C3, y.data_ptr(), C, n, H, W, heads, head_dim, dstride,
                                        window, overlap_window, shift, bias_table.data_ptr(), _stream()))
    return y


def pwdw_fused(x, stats, w1, dww, colsum=None, t=None, gate=False):
    """x bf16 [n,H,W,C]; stats fp32 [n*H*W, 2] (mu, rstd); w1 bf16 [rows, C] (LayerNorm gain folded); dww fp32
    [9, rows]; colsum / t fp32 [rows] or None (BiasFree). -> bf16 [n,H,W,rows] (plain) or [n,H,W,rows/2] (gate)."""
    n, H, W, C = x.shape
    rows = w1.shape[0]
    hp = rows // 2 if gate else 0
    y = torch.zeros((n, H, W, hp if gate else rows), device=x.device, dtype=torch.bfloat16)
    # [9, rows] tap-major -> [rows/2, 9, 2]: the 9 taps of a channel pair contiguous (the kernel's layout)
    dwp = dww.t().reshape(rows // 2, 2, 9).permute(0, 2, 1).contiguous()
    check(_lib.load().irb_pwdw_fused(int(gate), x.data_ptr(), x.stride(2), stats.data_ptr(), w1.data_ptr(), rows, hp,
                                     _ptr(colsum), _ptr(t), dwp.data_ptr(), rows, y.data_ptr(), y.shape[3],
                                     n, H, W, C, torch.cuda.current_stream(x.device).cuda_stream))
    return y


def mlp_fused(x, stats, w1, colsum1, b1, w2, b2, C, eps=1e-5):
    """In place on x bf16 [rows, ldx] (C valid columns): x += W2 . gelu(W1' . LN(x) + b1') + b2; -> (x, stats_out)."""
    rows = x.shape[0]
    hid = w1.shape[0]
    st_out = torch.full((rows, 2), float("nan"), device=x.device, dtype=torch.float32)
    check(_lib.load().irb_mlp_fused(x.data_ptr(), x.stride(0), rows, C, hid, stats.data_ptr(), w1.data_ptr(),
                                    w1.stride(0), colsum1.data_ptr(), b1.data_ptr(), w2.data_ptr(), w2.stride(0),
                                    b2.data_ptr(), st_out.data_ptr(), eps,
                                    torch.cuda.current_stream(x.device).cuda_stream))
    return x, st_out
