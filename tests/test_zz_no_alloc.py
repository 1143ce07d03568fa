"""GPU (-m gpu), collected LAST on purpose: irb_mdta_attn_static, the entry point that
allocates nothing (its scratch is a file-scope static __device__ array, include/irb200.h), against the existing pair
irb_mdta_gram + irb_mdta_attn with caller-allocated scratch. Same kernels -> the outputs must be bit-identical; the
assertion uses the tolerance of the other MDTA op tests (1e-5 fp32, 2e-2 bf16 weights) and reports exact equality.
Two calls in a row with DIFFERENT inputs and shapes: the second must not see the first call's leftovers in the shared
array. All tensors are allocated before the calls; the entry point itself is called through the C ABI (ctypes).
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _existing(lib, check, qkv, heads, temp, wproj, dt_code, out_dtype):
    n, H, W, C3 = qkv.shape
    Cc = C3 // 3
    c = Cc // heads
    s = torch.cuda.current_stream().cuda_stream
    nch = lib.irb_mdta_nchunks(H * W, heads, n)
    G = torch.full((n, heads, nch, c, c), float("nan"), device="cuda")
    ss = torch.full((n, heads, nch, 2, c), float("nan"), device="cuda")
    weff = torch.zeros((n, Cc, Cc), device="cuda", dtype=out_dtype)
    attn = torch.zeros((n, heads, c, c), device="cuda")
    check(lib.irb_mdta_gram(qkv.data_ptr(), dt_code, C3, n, H * W, Cc, heads, nch, G.data_ptr(), ss.data_ptr(), s))
    check(lib.irb_mdta_attn(G.data_ptr(), ss.data_ptr(), nch, temp.data_ptr(), wproj.data_ptr(), n, Cc, heads,
                            weff.data_ptr(), dt_code, attn.data_ptr(), s))
    return weff, attn


def _static(lib, check, qkv, heads, temp, wproj, dt_code, weff, attn):
    n, H, W, C3 = qkv.shape
    check(lib.irb_mdta_attn_static(qkv.data_ptr(), dt_code, C3, n, H * W, C3 // 3, heads, temp.data_ptr(),
                                   wproj.data_ptr(), weff.data_ptr(), attn.data_ptr(),
                                   torch.cuda.current_stream().cuda_stream))


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_static_scratch_entry_matches_existing_entry_twice(dtype):
    from image_restoration_agent_b200 import _lib
    lib, check = _lib.load(), _lib.check
    dt_code = _lib.BF16 if dtype == torch.bfloat16 else _lib.F32
    tol = 2e-2 if dtype == torch.bfloat16 else 1e-5
    assert lib.irb_mdta_attn_static_max_bytes() == 4 << 20
    torch.manual_seed(7)
    # two different shapes back to back: (n, H, W, C, heads); the second is smaller, so stale slots of the first call
    # sit right behind the ones it uses
    for k, (n, H, W, Cc, heads) in enumerate([(2, 96, 128, 96, 2), (1, 64, 80, 48, 1)]):
        qkv = (torch.randn(n, H, W, 3 * Cc, device="cuda") * (1.0 + k)).to(dtype).contiguous()
        temp = torch.rand(heads, device="cuda") + 0.5
        wproj = torch.randn(Cc, Cc, device="cuda") / Cc ** 0.5
        want_w, want_a = _existing(lib, check, qkv, heads, temp, wproj, dt_code, dtype)
        c = Cc // heads
        got_w = torch.full((n, Cc, Cc), float("nan"), device="cuda", dtype=dtype)
        got_a = torch.full((n, heads, c, c), float("nan"), device="cuda")
        torch.cuda.synchronize()
        before = torch.cuda.memory_stats()["allocation.all.allocated"]
        _static(lib, check, qkv, heads, temp, wproj, dt_code, got_w, got_a)
        torch.cuda.synchronize()
        assert torch.cuda.memory_stats()["allocation.all.allocated"] == before   # no torch allocation inside the call
        assert torch.isfinite(got_a).all() and torch.isfinite(got_w.float()).all()
        ea = (got_a - want_a).abs().max().item()
        ew = (got_w.float() - want_w.float()).abs().max().item()
        print(f"\n[parity] irb_mdta_attn_static call {k + 1} {dtype}: attn max-abs {ea:.1e}, weff max-abs {ew:.1e}, "
              f"bit-identical {bool(torch.equal(got_a, want_a) and torch.equal(got_w, want_w))}")
        assert ea <= 1e-5 and ew <= tol


def test_static_scratch_entry_rejects_shapes_beyond_the_array():
    from image_restoration_agent_b200 import _lib
    lib = _lib.load()
    # 8 images x 8 heads x 32 chunks x (48*48 + 96) floats = 19.7 MB > 4 MiB; the size check comes before any launch
    n, H, W, Cc, heads = 8, 256, 256, 384, 8
    qkv = torch.zeros(1, device="cuda", dtype=torch.bfloat16)   # never read: the call must fail on the size check
    f = torch.zeros(1, device="cuda")
    rc = lib.irb_mdta_attn_static(qkv.data_ptr(), _lib.BF16, 3 * Cc, n, H * W, Cc, heads, f.data_ptr(), f.data_ptr(),
                                  qkv.data_ptr(), f.data_ptr(), torch.cuda.current_stream().cuda_stream)
    msg = lib.irb_last_error().decode()
    assert rc != 0 and "static array holds 4194304" in msg and "needs" in msg
    with pytest.raises(ValueError):
        _lib.check(rc)
