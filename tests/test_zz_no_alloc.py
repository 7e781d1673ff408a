"""GPU: b200_decode_noalloc, the decode entry point for hosts that cannot allocate device memory.

Some users of libb200attn.so are not PyTorch programs: a serving process in another language bound over
the C ABI, a plugin thread inside somebody else's worker, or an engine that handed every free byte of HBM
to its KV pool at start-up.  They cannot produce the workspace b200_decode asks for (no allocator, or
cudaMalloc is forbidden because it stalls the running worker, or it would simply fail).  For them the
library keeps 8 MiB of scratch inside its own image and b200_decode_noalloc takes inputs and outputs only.

What is checked here, through ctypes on the raw C ABI (no torch custom op in between):
  * its output equals b200_decode's on the same inputs, for calls in a row on DIFFERENT batches, so the
    later calls run on a scratch array still full of the previous call's partials (the array is never
    cleared; the kernels must overwrite everything they read);
  * both results are also within the suite's tolerance of the fp32 oracle (tests/util.py: 2e-2 abs, 1e-2 rel);
  * a shape that needs more than the built-in array is refused with a message that says so;
  * the call really allocates nothing: torch's allocator statistics and the driver's free-memory figure
    do not move across the calls (after the first call, which is allowed to load the module).

This file is named test_zz_no_alloc so that pytest collects it after every other test file: with `-x`
a failure here cannot hide another test.  It does not depend on what ran before it: it passes alone in
a fresh process (first call does the per-device set-up) and at the end of the whole suite.
"""
import ctypes

import pytest
import torch

from oracle import paged_attention_oracle as orc
from tests.util import check_close, dev, make_paged_case

pytestmark = pytest.mark.gpu

SCRATCH_BYTES = 8 << 20   # B200_DECODE_NOALLOC_SCRATCH_BYTES in include/b200attn.h


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _decode_both(pkg, case):
    """Run b200_decode (caller-owned workspace) and b200_decode_noalloc on one case; return both outputs."""
    lib = pkg._lib.load()
    d = dev(case)
    B, Hq, Hkv, D, ps = case["B"], case["Hq"], case["Hkv"], case["D"], case["page_size"]
    q, kc, vc = d["q"], d["k_cache"], d["v_cache"]
    scale = float(D) ** -0.5
    need = lib.b200_decode_workspace_bytes(B, Hq, Hkv, D)
    assert 0 < need <= SCRATCH_BYTES
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    out_ws = torch.full((B, Hq, D), float("nan"), dtype=q.dtype, device="cuda")
    out_na = torch.full((B, Hq, D), float("nan"), dtype=q.dtype, device="cuda")
    rc = lib.b200_decode(q.data_ptr(), q.stride(0), kc.data_ptr(), vc.data_ptr(), kc.stride(0),
                         d["kv_indptr"].data_ptr(), d["kv_indices"].data_ptr(), d["last_page_len"].data_ptr(),
                         out_ws.data_ptr(), out_ws.stride(0), ws.data_ptr(), ws.numel(), B, Hq, Hkv, D, ps, scale, 0,
                         _stream())
    pkg._lib.check(rc, "b200_decode")
    rc = lib.b200_decode_noalloc(q.data_ptr(), q.stride(0), kc.data_ptr(), vc.data_ptr(), kc.stride(0),
                                 d["kv_indptr"].data_ptr(), d["kv_indices"].data_ptr(), d["last_page_len"].data_ptr(),
                                 out_na.data_ptr(), out_na.stride(0), B, Hq, Hkv, D, ps, scale, 0, _stream())
    pkg._lib.check(rc, "b200_decode_noalloc")
    torch.cuda.synchronize()
    return out_ws, out_na


def _oracle(case):
    return orc.decode_attention(case["q"], case["k_cache"], case["v_cache"], case["kv_indptr"], case["kv_indices"],
                                case["last_page_len"], case["page_size"], return_fp32=True)


def test_noalloc_decode_matches_workspace_decode_twice_in_a_row(pkg):
    # Different batches back to back.  Both split sequences across streams (long contexts), so partials,
    # log-sum-exps and segment indices all go through the built-in array; the second batch has another head
    # layout (one KV head, 8 warps per head) and another page size, so it reuses the same bytes differently.
    first = make_paged_case([3000, 17, 900, 1, 0, 4500, 64], 32, 8, page_size=1, seed=101)
    second = make_paged_case([7000, 33, 5, 2048, 600], 4, 1, page_size=16, seed=102)
    for n, case in enumerate((first, second, first)):
        out_ws, out_na = _decode_both(pkg, case)
        assert torch.isfinite(out_na.float()).all(), f"call {n}: rows left unwritten or non-finite"
        # same kernels, same static schedule: the two entry points agree to the bit, far inside the suite's tolerance
        diff = (out_ws.float() - out_na.float()).abs().max().item()
        assert diff == 0.0, f"call {n}: b200_decode_noalloc differs from b200_decode by {diff}"
        check_close(out_na, _oracle(case), f"noalloc decode, call {n}")


def test_noalloc_decode_allocates_nothing_and_refuses_oversize(pkg):
    lib = pkg._lib.load()
    case = make_paged_case([1500, 40, 777], 16, 8, page_size=1, seed=103)
    _decode_both(pkg, case)                     # first call on this device: set-up (may load the module)
    d = dev(case)
    q, kc, vc = d["q"], d["k_cache"], d["v_cache"]
    B, Hq, Hkv, D = case["B"], case["Hq"], case["Hkv"], case["D"]
    out = torch.empty((B, Hq, D), dtype=q.dtype, device="cuda")
    torch.cuda.synchronize()
    stats0 = torch.cuda.memory_stats()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(3):
        rc = lib.b200_decode_noalloc(q.data_ptr(), q.stride(0), kc.data_ptr(), vc.data_ptr(), kc.stride(0),
                                     d["kv_indptr"].data_ptr(), d["kv_indices"].data_ptr(),
                                     d["last_page_len"].data_ptr(), out.data_ptr(), out.stride(0), B, Hq, Hkv, D, 1,
                                     float(D) ** -0.5, 0, _stream())
        assert rc == 0, lib.b200_last_error()
    torch.cuda.synchronize()
    stats1 = torch.cuda.memory_stats()
    free1, _ = torch.cuda.mem_get_info()
    assert stats1["allocation.all.allocated"] == stats0["allocation.all.allocated"]
    assert stats1["reserved_bytes.all.current"] == stats0["reserved_bytes.all.current"]
    assert free1 == free0, "device free memory moved across b200_decode_noalloc calls"
    check_close(out, _oracle(case), "noalloc decode, repeated")

    # A batch whose partials do not fit the built-in array: refused before anything is launched.
    big_B = 4096
    need = lib.b200_decode_workspace_bytes(big_B, 32, 8, 128)
    assert need > SCRATCH_BYTES
    sentinel = out.clone()
    rc = lib.b200_decode_noalloc(q.data_ptr(), q.stride(0), kc.data_ptr(), vc.data_ptr(), kc.stride(0),
                                 d["kv_indptr"].data_ptr(), d["kv_indices"].data_ptr(), d["last_page_len"].data_ptr(),
                                 out.data_ptr(), out.stride(0), big_B, 32, 8, 128, 1, 1.0, 0, _stream())
    msg = lib.b200_last_error().decode()
    assert rc != 0 and str(need) in msg and str(SCRATCH_BYTES) in msg and "b200_decode" in msg, msg
    torch.cuda.synchronize()
    assert torch.equal(out, sentinel)
