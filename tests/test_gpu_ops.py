"""GPU (-m gpu): every single-kernel C-ABI entry against a plain torch fp32 restatement of the same op on the same
seeded inputs (TF32 disabled). fp32 kernels must match to fp32 round-off, bf16 kernels to bf16 output rounding."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _no_tf32():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(0)


def close(got, ref, rel):
    got, ref = got.float(), ref.float()
    assert torch.isfinite(got).all()
    assert (got - ref).abs().max().item() <= rel * max(ref.abs().max().item(), 1e-6)


DT = [(torch.float32, 3e-5), (torch.bfloat16, 1.6e-2)]


@pytest.mark.parametrize("dt,tol", DT)
@pytest.mark.parametrize("C", [48, 96, 180, 192, 240, 384])
@pytest.mark.parametrize("bias_free", [False, True])
def test_layernorm(dt, tol, C, bias_free):
    from image_restoration_agent_b200 import ops
    x = (torch.randn(1537, C, device="cuda") * 2 + 0.5).to(dt)
    w = torch.randn(C, device="cuda") * 0.2 + 1
    b = torch.randn(C, device="cuda") * 0.1
    y = ops.layernorm(x, w, None if bias_free else b, bias_free=bias_free)
    xf = x.float()
    var = xf.var(-1, keepdim=True, unbiased=False)
    ref = xf / torch.sqrt(var + 1e-5) * w if bias_free else (xf - xf.mean(-1, keepdim=True)) / torch.sqrt(var + 1e-5) * w + b
    close(y, ref, tol)


@pytest.mark.parametrize("dt,tol", DT)
@pytest.mark.parametrize("M,K,N", [(300, 48, 144), (1000, 96, 288), (257, 192, 192), (512, 384, 1152), (700, 256, 96),
                                   (640, 1024, 384), (130, 184, 360), (129, 360, 180), (1, 64, 16)])
def test_gemm1x1_and_epilogues(dt, tol, M, K, N):
    from image_restoration_agent_b200 import ops
    x = torch.randn(M, K, device="cuda").to(dt)
    w = (torch.randn(N, K, device="cuda") / K ** 0.5).to(dt)
    bias = torch.randn(N, device="cuda")
    res = torch.randn(M, N, device="cuda").to(dt)
    ref = x.float() @ w.float().t()
    close(ops.gemm1x1(x, w), ref, tol)
    close(ops.gemm1x1(x, w, bias=bias, act=1), F.gelu(ref + bias), tol)
    close(ops.gemm1x1(x, w, bias=bias, act=2, slope=0.2), F.leaky_relu(ref + bias, 0.2), tol)
    close(ops.gemm1x1(x, w, res=res), ref + res.float(), tol)


@pytest.mark.parametrize("dt,tol", DT)
def test_gemm_batched_per_image_weights_on_a_channel_slice(dt, tol):
    """The MDTA apply step: A is the v slice of a qkv tensor (row stride 3C), weights differ per image."""
    from image_restoration_agent_b200 import ops
    n, P, C = 3, 200, 96
    qkv = torch.randn(n, P, 3 * C, device="cuda").to(dt)
    weff = (torch.randn(n, C, C, device="cuda") / C ** 0.5).to(dt)
    v = qkv[:, :, 2 * C:]
    y = ops.gemm1x1(v, weff.reshape(n * C, C), batch=n, w_rows_per_batch=C)
    close(y, torch.einsum("npk,nok->npo", v.float(), weff.float()).reshape(n * P, C), tol)


@pytest.mark.parametrize("dt,tol", DT)
@pytest.mark.parametrize("n,H,W,Cin,N", [(2, 32, 32, 48, 24), (1, 24, 40, 96, 192), (2, 16, 16, 384, 768),
                                         (1, 40, 24, 180, 180), (1, 64, 64, 96, 3), (1, 33, 20, 64, 64),
                                         (1, 8, 8, 60, 240), (1, 136, 8, 64, 256),
                                         # outputs <= 64 wide: the dy-reuse path (16 x 10 pixel A box per dx), three K
                                         # slabs with a partial last one, ragged in both directions, several images
                                         (1, 21, 50, 180, 60), (2, 16, 48, 240, 60), (1, 9, 17, 64, 16),
                                         # box shape by fewest boxes: 544 wide -> 32 x 4 boxes, ragged rows
                                         (1, 6, 544, 60, 180),
                                         # N <= 4: the shared-memory stencil kernel (conv_small.cu), ragged tiles / chunks
                                         (2, 37, 45, 180, 3), (1, 9, 70, 64, 1), (1, 20, 33, 64, 4), (3, 16, 32, 96, 2)])
def test_conv3x3_implicit_gemm(dt, tol, n, H, W, Cin, N):
    from image_restoration_agent_b200 import ops
    Cp = (Cin + 7) // 8 * 8
    x = torch.randn(n, H, W, Cp, device="cuda").to(dt)[..., :Cin]
    wq = (torch.randn(N, Cin, 3, 3, device="cuda") / (9 * Cin) ** 0.5).to(dt)
    wp, kpad = ops.pack_conv3x3_weight(wq, dt)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), wq.float(), padding=1)
    close(ops.conv3x3(x, wp, kpad, N), ref.permute(0, 2, 3, 1), tol)
    bias = torch.randn(N, device="cuda")
    close(ops.conv3x3(x, wp, kpad, N, bias=bias, act=2, slope=0.01),
          F.leaky_relu(ref + bias.view(1, -1, 1, 1), 0.01).permute(0, 2, 3, 1), tol)
    if N % 4 == 0:   # fused PixelShuffle(2) (restormer_arch.py:185-186, network_swinir.py:585)
        close(ops.conv3x3(x, wp, kpad, N, shuffle=1), F.pixel_shuffle(ref, 2).permute(0, 2, 3, 1), tol)
    if H % 2 == 0 and W % 2 == 0 and N >= 8:  # fused PixelUnshuffle(2) (restormer_arch.py:175-176)
        close(ops.conv3x3(x, wp, kpad, N, shuffle=2), F.pixel_unshuffle(ref, 2).permute(0, 2, 3, 1), tol)


@pytest.mark.parametrize("dt,tol", [(torch.float32, 1e-5), (torch.bfloat16, 1e-2)])
@pytest.mark.parametrize("n,H,W,Cc", [(2, 32, 32, 144), (1, 40, 24, 288), (1, 8, 8, 1152), (1, 70, 36, 576), (1, 1, 5, 48)])
def test_dwconv3x3_and_sumsq(dt, tol, n, H, W, Cc):
    from image_restoration_agent_b200 import ops
    x = torch.randn(n, H, W, Cc, device="cuda").to(dt)
    w = torch.randn(Cc, 1, 3, 3, device="cuda") / 3
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w, padding=1, groups=Cc)
    sc = Cc // 3 * 2
    y, ss = ops.dwconv3x3(x, ops.pack_dw_weight(w), sumsq_channels=sc)
    close(y, ref.permute(0, 2, 3, 1), tol)
    close(ss, (ref[:, :sc] ** 2).sum((2, 3)), 1e-4)
    close(ops.dwconv3x3(x, ops.pack_dw_weight(w)), ref.permute(0, 2, 3, 1), tol)  # bf16: the TMA-fed kernel


@pytest.mark.parametrize("n,H,W,Cc", [(3, 100, 70, 96), (2, 64, 64, 288), (1, 9, 33, 16), (2, 17, 31, 200)])
def test_dwconv3x3_tma_ragged_tiles(n, H, W, Cc):
    """bf16 TMA kernel on sizes that are not multiples of the 32x8 tile / 64-channel chunk, on a channel slice."""
    from image_restoration_agent_b200 import ops
    big = torch.randn(n, H, W, Cc + 8, device="cuda").bfloat16()
    x = big[..., :Cc]  # row stride Cc + 8
    w = torch.randn(Cc, 1, 3, 3, device="cuda") / 3
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w, padding=1, groups=Cc).permute(0, 2, 3, 1)
    close(ops.dwconv3x3(x, ops.pack_dw_weight(w)), ref, 1e-2)


@pytest.mark.parametrize("dt,tol", [(torch.float32, 1e-5), (torch.bfloat16, 1e-2)])
@pytest.mark.parametrize("n,H,W,hp", [(2, 32, 32, 128), (1, 24, 40, 256), (1, 8, 8, 1024), (1, 35, 9, 512)])
def test_gdfn_dwconv_gelu_gate(dt, tol, n, H, W, hp):
    from image_restoration_agent_b200 import ops
    x = torch.randn(n, H, W, 2 * hp, device="cuda").to(dt)
    w = torch.randn(2 * hp, 1, 3, 3, device="cuda") / 3
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w, padding=1, groups=2 * hp)
    close(ops.gdfn_dw_gate(x, ops.pack_dw_weight(w)), (F.gelu(ref[:, :hp]) * ref[:, hp:]).permute(0, 2, 3, 1), tol)


@pytest.mark.parametrize("dt,tol", [(torch.float32, 1e-4), (torch.bfloat16, 2e-3)])
@pytest.mark.parametrize("n,H,W,Cc,heads", [(2, 32, 32, 48, 1), (1, 64, 48, 96, 2), (2, 16, 16, 384, 8), (1, 50, 30, 96, 1),
                                            (1, 128, 128, 192, 4), (1, 8, 8, 96, 1)])
def test_mdta_gram_softmax_fold(dt, tol, n, H, W, Cc, heads):
    """restormer_arch.py:117-129 split into Gram (split-K over pixels) + normalise/temperature/softmax + fold into
    project_out."""
    from image_restoration_agent_b200 import ops
    qkv = torch.randn(n, H, W, 3 * Cc, device="cuda").to(dt)
    c = Cc // heads
    q = qkv[..., :Cc].float().reshape(n, H * W, heads, c)
    k = qkv[..., Cc:2 * Cc].float().reshape(n, H * W, heads, c)
    ref = torch.einsum("nphi,nphj->nhij", q, k)
    Gp, sp = ops.mdta_gram(qkv, heads)
    close(Gp.sum(2), ref, tol)
    ss = torch.stack([(q ** 2).sum(1), (k ** 2).sum(1)], 2).contiguous()  # [n, heads, 2, c]
    close(sp.sum(2), ss, tol)
    Gp2, sp2 = ops.mdta_gram(qkv, heads)
    assert torch.equal(Gp, Gp2) and torch.equal(sp, sp2)  # no atomics: bit-reproducible
    temp = torch.rand(heads, device="cuda") * 3 + 0.5
    wproj = torch.randn(Cc, Cc, device="cuda") / Cc ** 0.5
    ss = ss.unsqueeze(2).contiguous()                                        # one chunk
    weff, attn = ops.mdta_attn(ref.unsqueeze(2).contiguous(), ss, temp, wproj, torch.float32)
    qn, kn = F.normalize(q.permute(0, 2, 3, 1), dim=-1), F.normalize(k.permute(0, 2, 3, 1), dim=-1)
    a_ref = ((qn @ kn.transpose(-2, -1)) * temp.view(1, heads, 1, 1)).softmax(-1)
    close(attn, a_ref, 1e-4)
    close(weff, torch.einsum("ohi,nhij->nohj", wproj.reshape(Cc, heads, c), a_ref).reshape(n, Cc, Cc), 1e-4)
    # all-zero q row: F.normalize's eps path (max(||q||, 1e-12)) must not produce NaN
    z = torch.zeros_like(ss)
    weff0, attn0 = ops.mdta_attn(torch.zeros_like(ref).unsqueeze(2).contiguous(), z, temp, wproj, torch.float32)
    assert torch.isfinite(attn0).all() and torch.allclose(attn0, torch.full_like(attn0, 1.0 / c))


@pytest.mark.parametrize("dt,tol", [(torch.float32, 2e-5), (torch.bfloat16, 1.6e-2)])
@pytest.mark.parametrize("ws,H,W,C,heads,shift,dstride", [(8, 16, 24, 180, 6, 0, 32), (8, 16, 24, 180, 6, 4, 32),
                                                          (8, 8, 8, 180, 6, 4, 32), (7, 14, 21, 180, 6, 3, 32),
                                                          (8, 32, 16, 240, 8, 4, 32), (8, 16, 16, 60, 6, 4, 16),
                                                          (8, 16, 24, 180, 6, 4, 30)])
def test_window_attention_vs_oracle(dt, tol, ws, H, W, C, heads, shift, dstride):
    """network_swinir.py:114-145 + 216-237 + roll/partition/reverse, against the oracle's restatement."""
    from image_restoration_agent_b200 import ops
    from oracle import swinir_ref
    n, d = (3 if H == 8 else 2), C // heads  # 8x8 image x 3: an odd number of windows (half-empty last window pair)
    q = torch.randn(n, H, W, heads, d) * 0.5
    k = torch.randn(n, H, W, heads, d)
    v = torch.randn(n, H, W, heads, d)
    table = torch.randn((2 * ws - 1) ** 2, heads) * 0.5
    if dt == torch.bfloat16:
        q, k, v = (t.bfloat16().float() for t in (q, k, v))
    # oracle: roll -> window_partition -> softmax(qk^T + bias + mask) v -> window_reverse -> roll back
    def part(t):
        t = torch.roll(t, (-shift, -shift), (1, 2)) if shift else t
        return t.view(n, H // ws, ws, W // ws, ws, heads, d).permute(0, 1, 3, 5, 2, 4, 6).reshape(-1, heads, ws * ws, d)
    qw, kw, vw = part(q), part(k), part(v)
    attn = qw @ kw.transpose(-2, -1)
    bias = table[swinir_ref.relative_position_index(ws).reshape(-1)].reshape(ws * ws, ws * ws, heads).permute(2, 0, 1)
    attn = attn + bias.unsqueeze(0)
    if shift:
        mask = swinir_ref.shift_mask(H, W, ws, shift)
        nW = mask.shape[0]
        attn = (attn.view(n, nW, heads, ws * ws, ws * ws) + mask[None, :, None]).view(-1, heads, ws * ws, ws * ws)
    o = attn.softmax(-1) @ vw
    o = o.view(n, H // ws, W // ws, heads, ws, ws, d).permute(0, 1, 4, 2, 5, 3, 6).reshape(n, H, W, C)
    want = torch.roll(o, (shift, shift), (1, 2)) if shift else o
    # device layout: [n,H,W,3*heads*dstride]
    buf = torch.zeros(n, H, W, 3, heads, dstride)
    buf[..., 0, :, :d], buf[..., 1, :, :d], buf[..., 2, :, :d] = q, k, v
    qkv = buf.reshape(n, H, W, 3 * heads * dstride).to(dt).cuda()
    if (3 * heads * dstride) % 8:  # keep rows 16-byte aligned for the vectorised path
        pytest.skip("unaligned layout")
    got = ops.window_attention(qkv, C, heads, ws, shift, table.cuda())
    close(got.cpu(), want, tol)


@pytest.mark.parametrize("h,w,tile,ov,sf,c", [(720, 1280, 256, 32, 1, 3), (264, 264, 128, 32, 1, 1), (72, 88, 48, 16, 4, 3),
                                              (64, 64, 64, 32, 2, 3), (2160, 3840, 512, 64, 1, 3)])
def test_stitch_is_bit_exact_with_the_reference_accumulation(h, w, tile, ov, sf, c):
    """E += out; W += 1; E / W in the reference's loop order (inference.py:309-320, demo.py:142-153): bit-exact
    floats, bit-exact integer coverage, exact round-half-even uint8."""
    from image_restoration_agent_b200 import ops
    from oracle import tiling
    ys, xs = tiling.tile_indices(h, tile, ov), tiling.tile_indices(w, tile, ov)
    tiles = torch.rand(len(ys) * len(xs), tile * sf, tile * sf, c, device="cuda") * 1.2 - 0.1
    E = torch.zeros(c, h * sf, w * sf, device="cuda")
    Wt = torch.zeros_like(E)
    i = 0
    for y0 in ys:
        for x0 in xs:
            E[:, y0 * sf:(y0 + tile) * sf, x0 * sf:(x0 + tile) * sf] += tiles[i].permute(2, 0, 1)
            Wt[:, y0 * sf:(y0 + tile) * sf, x0 * sf:(x0 + tile) * sf] += 1
            i += 1
    cy = np.outer(tiling.coverage_count(h, tile, ov, sf), tiling.coverage_count(w, tile, ov, sf))
    assert np.array_equal(Wt[0].cpu().numpy().astype(np.int64), cy)
    ref = E / Wt
    ch, cw = h * sf - 3, w * sf - 5
    out, u8 = ops.stitch(tiles, ys, xs, sf, h, w, ch, cw)
    assert torch.equal(out, ref[:, :ch, :cw])
    ref8 = (ref[:, :ch, :cw].clamp(0, 1).permute(1, 2, 0).cpu().numpy() * 255.0).round().astype(np.uint8)
    assert np.array_equal(u8.cpu().numpy(), ref8)


@pytest.mark.parametrize("gate", [False, True])
@pytest.mark.parametrize("with_bias", [True, False])
@pytest.mark.parametrize("n,H,W,C,rows", [(2, 16, 32, 48, 144), (1, 24, 40, 96, 288), (3, 8, 72, 96, 512),
                                          (1, 40, 24, 48, 256), (2, 9, 33, 96, 288)])
def test_pwdw_fused_matches_separate_ops(gate, with_bias, n, H, W, C, rows):
    """1x1 (folded LayerNorm) -> depthwise 3x3 [-> gelu gate] in one kernel vs the torch fp32 restatement of
    restormer_arch.py:115-116 / :88-93 on the same bf16-rounded operands (incl. ragged tiles: H % 8, W % 32 != 0)."""
    from image_restoration_agent_b200 import ops
    if gate:
        rows = 256 if C == 48 else 512
    x = (torch.randn(n, H, W, C, device="cuda") * 1.5 + 0.3).bfloat16()
    xf = x.float()
    mu = xf.mean(-1)
    rstd = 1.0 / torch.sqrt(xf.var(-1, unbiased=False) + 1e-5)
    stats = torch.stack([mu, rstd], -1).reshape(-1, 2).contiguous()
    gain = torch.randn(C, device="cuda") * 0.1 + 1.0
    ln_b = torch.randn(C, device="cuda") * 0.05
    w = torch.randn(rows, C, device="cuda") / C ** 0.5
    w1 = (w * gain).bfloat16()
    colsum = w1.float().sum(1).contiguous() if with_bias else None
    t = (w @ ln_b).contiguous() if with_bias else None
    dww = (torch.randn(9, rows, device="cuda") / 3.0).contiguous()
    y = ops.pwdw_fused(x, stats, w1, dww, colsum, t, gate=gate)
    # reference: hidden rounded to bf16 (as the unfused path stores it), zero padding of the hidden tensor
    acc = xf @ w1.float().t()
    if with_bias:
        hid = rstd[..., None] * (acc - mu[..., None] * colsum) + t
    else:
        hid = rstd[..., None] * acc
    hid = hid.bfloat16().float().permute(0, 3, 1, 2)
    dw = F.conv2d(hid, dww.t().reshape(rows, 1, 3, 3), padding=1, groups=rows).permute(0, 2, 3, 1)
    ref = F.gelu(dw[..., :rows // 2]) * dw[..., rows // 2:] if gate else dw
    assert y.shape == ref.shape
    close(y, ref, 1.6e-2)


@pytest.mark.parametrize("rows,C,hid", [(1000, 180, 360), (128, 180, 360), (129, 64, 128), (4097, 96, 192), (300, 32, 64), (70000, 180, 360)])
def test_mlp_fused_matches_torch(rows, C, hid):
    """fc1 (norm2 folded) -> GELU -> fc2 -> + x in place (network_swinir.py:24-30, :277) as one kernel vs fp32 torch on the
    same bf16-rounded operands, and the LayerNorm statistics of the new rows."""
    from image_restoration_agent_b200 import ops
    ld = (C + 15) // 16 * 16
    xb = torch.zeros(rows, ld, device="cuda", dtype=torch.bfloat16)
    xb[:, :C] = (torch.randn(rows, C, device="cuda") * 1.2 + 0.2).bfloat16()
    x0 = xb[:, :C].float()
    mu = x0.mean(-1)
    rstd = 1.0 / torch.sqrt(x0.var(-1, unbiased=False) + 1e-5)
    stats = torch.stack([mu, rstd], -1).contiguous()
    gain = torch.randn(C, device="cuda") * 0.1 + 1.0
    ln_b = torch.randn(C, device="cuda") * 0.05
    W1 = torch.randn(hid, C, device="cuda") / C ** 0.5
    b1 = torch.randn(hid, device="cuda") * 0.1
    W2 = torch.randn(C, hid, device="cuda") / hid ** 0.5
    b2 = torch.randn(C, device="cuda") * 0.1
    k1 = (C + 7) // 8 * 8
    w1 = torch.zeros(hid, k1, device="cuda", dtype=torch.bfloat16)
    w1[:, :C] = (W1 * gain).bfloat16()
    cs1 = w1.float().sum(1).contiguous()
    b1f = (b1 + W1 @ ln_b).contiguous()
    k2 = (hid + 7) // 8 * 8
    w2 = torch.zeros(C, k2, device="cuda", dtype=torch.bfloat16)
    w2[:, :hid] = W2.bfloat16()
    # reference: the two-GEMM path's arithmetic (hidden rounded to bf16 between the GEMMs, tanh-form GELU)
    h = rstd[:, None] * (x0 @ w1[:, :C].float().t() - mu[:, None] * cs1) + b1f
    h = F.gelu(h, approximate="tanh").bfloat16().float()
    ref = x0 + h @ w2[:, :hid].float().t() + b2
    y, st = ops.mlp_fused(xb, stats, w1, cs1, b1f, w2, b2.contiguous(), C)
    close(y[:, :C], ref, 1.6e-2)
    assert (st[:, 0] - ref.mean(1)).abs().max().item() <= 2e-2 * max(ref.abs().max().item(), 1.0)
    r2 = 1.0 / torch.sqrt(ref.var(1, unbiased=False) + 1e-5)
    assert ((st[:, 1] - r2).abs() / r2).max().item() <= 3e-2
