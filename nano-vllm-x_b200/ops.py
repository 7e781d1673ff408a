"""torch.library registration of the b200attn C-ABI entry points.

Each op is a thin shim: pull raw device pointers out of the tensors, pass the current CUDA
stream, call the `extern "C"` function through ctypes.  Fake (meta) implementations are
registered so torch.compile'd neighbours (the reference compiles RMSNorm / RoPE / sampler,
layers/layernorm.py:16, rotary_embedding.py:73) trace through.  No op has a CPU
implementation: a non-CUDA tensor raises.
"""
from __future__ import annotations

import torch

from . import _lib

_DTYPE_CODE = {torch.bfloat16: 0, torch.float16: 1}


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _need_cuda(*ts: torch.Tensor) -> None:
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("b200attn ops run on CUDA tensors only (there is no CPU fallback)")


def _ptr(t):
    return 0 if t is None else t.data_ptr()


# ---- (1) slot write ---------------------------------------------------------------------------------
@torch.library.custom_op("b200attn::store_kv", mutates_args=("k_cache", "v_cache"), device_types="cuda")
def store_kv(k: torch.Tensor, v: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
             slots: torch.Tensor) -> None:
    _need_cuda(k, v, k_cache, v_cache, slots)
    T, H, D = k.shape
    row = H * D
    assert k.stride(-1) == 1 and v.stride(-1) == 1
    assert k.stride(-2) == D and v.stride(-2) == D
    assert k_cache.stride(-3) == row and v_cache.stride(-3) == row
    assert k_cache.stride(-1) == 1 and k_cache.stride(-2) == D
    assert slots.numel() == T and slots.dtype == torch.int64 and slots.is_contiguous()
    assert k.dtype == k_cache.dtype == v.dtype == v_cache.dtype
    lib = _lib.load()
    rc = lib.b200_store_kv(k.data_ptr(), k.stride(0), v.data_ptr(), v.stride(0), k_cache.data_ptr(),
                           v_cache.data_ptr(), slots.data_ptr(), T, row, k.element_size(), _stream())
    _lib.check(rc, "b200_store_kv")


# ---- (2) metadata -----------------------------------------------------------------------------------
@torch.library.custom_op("b200attn::slot_table_update", mutates_args=("slot_table",), device_types="cuda")
def slot_table_update(slot_table: torch.Tensor, step_msg: torch.Tensor, B: int, n_payload: int) -> None:
    _need_cuda(slot_table, step_msg)
    assert slot_table.dtype == torch.int32 and step_msg.dtype == torch.int32
    assert slot_table.dim() == 2 and slot_table.stride(1) == 1 and step_msg.is_contiguous()
    assert step_msg.numel() >= 7 * B + n_payload
    lib = _lib.load()
    rc = lib.b200_slot_table_update(slot_table.data_ptr(), slot_table.stride(0), step_msg.data_ptr(), B,
                                    n_payload, _stream())
    _lib.check(rc, "b200_slot_table_update")


@torch.library.custom_op("b200attn::build_meta",
                         mutates_args=("qo_indptr", "kv_indptr", "kv_indices", "last_page_len", "out_slots",
                                       "scratch"),
                         device_types="cuda")
def build_meta(slot_table: torch.Tensor, step_msg: torch.Tensor, B: int, padded_bs: int, page_size: int,
               fill_tail: bool, qo_indptr: torch.Tensor | None, kv_indptr: torch.Tensor,
               kv_indices: torch.Tensor, last_page_len: torch.Tensor | None, out_slots: torch.Tensor,
               scratch: torch.Tensor) -> None:
    _need_cuda(slot_table, step_msg, kv_indptr, kv_indices, out_slots, scratch)
    assert kv_indptr.dtype == torch.int32 and kv_indices.dtype == torch.int32
    assert out_slots.dtype == torch.int64 and scratch.dtype == torch.int32
    assert kv_indptr.numel() >= padded_bs + 1 and scratch.numel() >= padded_bs + 1
    assert qo_indptr is None or (qo_indptr.dtype == torch.int32 and qo_indptr.numel() >= B + 1)
    assert last_page_len is None or (last_page_len.dtype == torch.int32 and last_page_len.numel() >= padded_bs)
    for t in (kv_indptr, kv_indices, out_slots, scratch, qo_indptr, last_page_len):
        assert t is None or t.is_contiguous()
    lib = _lib.load()
    rc = lib.b200_build_meta(slot_table.data_ptr(), slot_table.stride(0), step_msg.data_ptr(), B, padded_bs,
                             page_size, int(fill_tail), _ptr(qo_indptr), kv_indptr.data_ptr(),
                             kv_indices.data_ptr(), kv_indices.numel(), _ptr(last_page_len),
                             out_slots.data_ptr(), out_slots.numel(), scratch.data_ptr(), _stream())
    _lib.check(rc, "b200_build_meta")



@torch.library.custom_op("b200attn::step_meta",
                         mutates_args=("slot_table", "qo_indptr", "kv_indptr", "kv_indices", "last_page_len",
                                       "out_slots", "scratch"),
                         device_types="cuda")
def step_meta(slot_table: torch.Tensor, step_msg: torch.Tensor, B: int, n_payload: int, padded_bs: int,
              page_size: int, fill_tail: bool, qo_indptr: torch.Tensor | None, kv_indptr: torch.Tensor,
              kv_indices: torch.Tensor, last_page_len: torch.Tensor | None, out_slots: torch.Tensor,
              scratch: torch.Tensor) -> None:
    """slot_table_update + build_meta behind ONE dispatcher round trip (the per-step host cost of the
    metadata path is dominated by Python/dispatcher overhead, not by the three tiny kernels)."""
    _need_cuda(slot_table, step_msg, kv_indptr, kv_indices, out_slots, scratch)
    lib = _lib.load()
    st = _stream()
    if B > 0 and n_payload > 0:
        _lib.check(lib.b200_slot_table_update(slot_table.data_ptr(), slot_table.stride(0), step_msg.data_ptr(), B,
                                              n_payload, st), "b200_slot_table_update")
    _lib.check(lib.b200_build_meta(slot_table.data_ptr(), slot_table.stride(0), step_msg.data_ptr(), B, padded_bs,
                                   page_size, int(fill_tail), _ptr(qo_indptr), kv_indptr.data_ptr(),
                                   kv_indices.data_ptr(), kv_indices.numel(), _ptr(last_page_len),
                                   out_slots.data_ptr(), out_slots.numel(), scratch.data_ptr(), st), "b200_build_meta")


# ---- (3) decode -------------------------------------------------------------------------------------
def decode_workspace_bytes(B: int, Hq: int, Hkv: int, D: int) -> int:
    return int(_lib.load().b200_decode_workspace_bytes(B, Hq, Hkv, D))


@torch.library.custom_op("b200attn::decode", mutates_args=("workspace",), device_types="cuda")
def decode(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, kv_indptr: torch.Tensor,
           kv_indices: torch.Tensor, last_page_len: torch.Tensor, workspace: torch.Tensor, page_size: int,
           sm_scale: float) -> torch.Tensor:
    _need_cuda(q, k_cache, v_cache, kv_indptr, kv_indices, last_page_len, workspace)
    B, Hq, D = q.shape
    Hkv = k_cache.shape[-2]
    assert q.stride(-1) == 1 and q.stride(-2) == D
    assert k_cache.stride(-1) == 1 and k_cache.stride(-2) == D and v_cache.stride() == k_cache.stride()
    assert kv_indptr.dtype == torch.int32 and kv_indices.dtype == torch.int32 and last_page_len.dtype == torch.int32
    assert kv_indptr.numel() >= B + 1 and last_page_len.numel() >= B
    assert workspace.dtype == torch.uint8
    out = torch.empty((B, Hq, D), dtype=q.dtype, device=q.device)
    if B == 0:
        return out
    lib = _lib.load()
    rc = lib.b200_decode(q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(), k_cache.stride(-3),
                         kv_indptr.data_ptr(), kv_indices.data_ptr(), last_page_len.data_ptr(), out.data_ptr(),
                         out.stride(0), workspace.data_ptr(), workspace.numel(), B, Hq, Hkv, D, page_size,
                         float(sm_scale), _DTYPE_CODE[q.dtype], _stream())
    _lib.check(rc, "b200_decode")
    return out


@decode.register_fake
def _(q, k_cache, v_cache, kv_indptr, kv_indices, last_page_len, workspace, page_size, sm_scale):
    return torch.empty_like(q)


# ---- (4) prefill ------------------------------------------------------------------------------------
def prefill_workspace_bytes(B: int, total_q: int, Hq: int, Hkv: int, D: int) -> int:
    return int(_lib.load().b200_prefill_workspace_bytes(B, total_q, Hq, Hkv, D))


@torch.library.custom_op("b200attn::prefill", mutates_args=("workspace",), device_types="cuda")
def prefill(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, qo_indptr: torch.Tensor,
            kv_indptr: torch.Tensor, kv_indices: torch.Tensor, last_page_len: torch.Tensor,
            workspace: torch.Tensor, page_size: int, sm_scale: float, causal: bool) -> torch.Tensor:
    _need_cuda(q, k_cache, v_cache, qo_indptr, kv_indptr, kv_indices, last_page_len, workspace)
    T, Hq, D = q.shape
    Hkv = k_cache.shape[-2]
    B = qo_indptr.numel() - 1
    assert q.stride(-1) == 1 and q.stride(-2) == D
    assert k_cache.stride(-1) == 1 and k_cache.stride(-2) == D and v_cache.stride() == k_cache.stride()
    assert qo_indptr.dtype == torch.int32 and kv_indptr.dtype == torch.int32
    assert kv_indices.dtype == torch.int32 and last_page_len.dtype == torch.int32
    out = torch.empty((T, Hq, D), dtype=q.dtype, device=q.device)
    if T == 0 or B == 0:
        return out
    lib = _lib.load()
    rc = lib.b200_prefill(q.data_ptr(), q.stride(0), k_cache.data_ptr(), v_cache.data_ptr(), k_cache.stride(-3),
                          qo_indptr.data_ptr(), kv_indptr.data_ptr(), kv_indices.data_ptr(),
                          last_page_len.data_ptr(), out.data_ptr(), out.stride(0), workspace.data_ptr(),
                          workspace.numel(), B, T, Hq, Hkv, D, page_size, float(sm_scale), int(causal),
                          _DTYPE_CODE[q.dtype], _stream())
    _lib.check(rc, "b200_prefill")
    return out


@prefill.register_fake
def _(q, k_cache, v_cache, qo_indptr, kv_indptr, kv_indices, last_page_len, workspace, page_size, sm_scale,
      causal):
    return torch.empty_like(q)


# ---- (f-3) fused q/k RMSNorm + RoPE + slot write ------------------------------------------------------
@torch.library.custom_op("b200attn::qk_norm_rope_store", mutates_args=("q_out", "k_out", "k_cache", "v_cache"),
                         device_types="cuda")
def qk_norm_rope_store(qkv: torch.Tensor, positions: torch.Tensor, q_weight: torch.Tensor, k_weight: torch.Tensor,
                       eps: float, cos_sin_cache: torch.Tensor, q_out: torch.Tensor, k_out: torch.Tensor | None,
                       k_cache: torch.Tensor, v_cache: torch.Tensor, slots: torch.Tensor, Hq: int, Hkv: int,
                       D: int) -> None:
    """models/qwen3.py:123-133 + layers/attention.py:347-352 in one launch (see include/b200attn.h, f-3).
    ``q_out`` may be the q columns of ``qkv`` itself (in place)."""
    _need_cuda(qkv, positions, q_weight, k_weight, cos_sin_cache, q_out, k_cache, v_cache, slots)
    T = qkv.shape[0]
    assert qkv.dim() == 2 and qkv.shape[1] == (Hq + 2 * Hkv) * D and qkv.stride(1) == 1
    assert qkv.dtype in _DTYPE_CODE and q_out.dtype == qkv.dtype == k_cache.dtype == v_cache.dtype
    assert q_weight.dtype == qkv.dtype and k_weight.dtype == qkv.dtype and q_weight.numel() == D and k_weight.numel() == D
    assert q_weight.is_contiguous() and k_weight.is_contiguous()
    assert positions.dtype == torch.int64 and positions.numel() == T and positions.is_contiguous()
    assert slots.dtype == torch.int64 and slots.numel() == T and slots.is_contiguous()
    assert cos_sin_cache.dtype == torch.float32 and cos_sin_cache.dim() == 2 and cos_sin_cache.shape[1] == D \
        and cos_sin_cache.stride(1) == 1, "RotaryEmbedding.cos_sin_cache must be fp32 [max_pos, head_dim]"
    assert q_out.dim() == 2 and q_out.shape == (T, Hq * D) and q_out.stride(1) == 1
    assert k_out is None or (k_out.shape == (T, Hkv * D) and k_out.stride(1) == 1 and k_out.dtype == qkv.dtype)
    row = Hkv * D
    assert k_cache.stride(-3) == row and v_cache.stride(-3) == row and k_cache.stride(-1) == 1 and k_cache.stride(-2) == D
    lib = _lib.load()
    rc = lib.b200_qk_norm_rope_store(qkv.data_ptr(), qkv.stride(0), positions.data_ptr(), q_weight.data_ptr(),
                                     k_weight.data_ptr(), float(eps), cos_sin_cache.data_ptr(), cos_sin_cache.stride(0),
                                     cos_sin_cache.shape[0], q_out.data_ptr(), q_out.stride(0), _ptr(k_out),
                                     0 if k_out is None else k_out.stride(0), k_cache.data_ptr(), v_cache.data_ptr(),
                                     slots.data_ptr(), T, Hq, Hkv, D, _DTYPE_CODE[qkv.dtype], _stream())
    _lib.check(rc, "b200_qk_norm_rope_store")


# ---- (f-4) sampler tail -----------------------------------------------------------------------------
_LOGIT_CODE = {torch.bfloat16: 0, torch.float32: 2}


@torch.library.custom_op("b200attn::sample", mutates_args=(), device_types="cuda")
def sample(logits: torch.Tensor, temperatures: torch.Tensor | None, min_ps: torch.Tensor | None,
           top_ps: torch.Tensor | None, top_ks: torch.Tensor | None, noise: torch.Tensor | None,
           seed: int, offset: int) -> torch.Tensor:
    """layers/sampler.py:19-55 in one launch (see include/b200attn.h, f-4).  Argument order follows
    ``Sampler.forward``.  ``noise`` = the Exp(1) tensor torch.multinomial would draw (parity), or None
    for the in-kernel Philox stream keyed by (seed, offset)."""
    _need_cuda(logits, temperatures, min_ps, top_ps, top_ks, noise)
    assert logits.dim() == 2 and logits.stride(1) == 1 and logits.dtype in _LOGIT_CODE, "logits: [B, V] bf16 / fp32"
    B, V = logits.shape
    for t, dt in ((temperatures, torch.float32), (min_ps, torch.float32), (top_ps, torch.float32), (top_ks, torch.int64)):
        assert t is None or (t.dtype == dt and t.numel() == B and t.is_contiguous())
    assert noise is None or (noise.dtype == torch.float32 and noise.shape == logits.shape and noise.stride(1) == 1)
    out = torch.empty((B,), dtype=torch.int64, device=logits.device)
    lib = _lib.load()
    rc = lib.b200_sample(logits.data_ptr(), logits.stride(0), _LOGIT_CODE[logits.dtype], B, V, _ptr(temperatures),
                         _ptr(top_ks), _ptr(top_ps), _ptr(min_ps), _ptr(noise), 0 if noise is None else noise.stride(0),
                         seed & 0xFFFFFFFFFFFFFFFF, offset & 0xFFFFFFFFFFFFFFFF, out.data_ptr(), _stream())
    _lib.check(rc, "b200_sample")
    return out


@sample.register_fake
def _(logits, temperatures, min_ps, top_ps, top_ks, noise, seed, offset):
    return torch.empty((logits.shape[0],), dtype=torch.int64, device=logits.device)


def launch_count() -> int:
    return int(_lib.load().b200_launch_count())


def reset_launch_count() -> None:
    _lib.load().b200_reset_launch_count()


# ---- dispatcher-free calls for the layers ----------------------------------------------------------------
# A torch.library custom op costs ~25 us of dispatcher work per call; a decode step issues three of them (step_meta,
# store_kv, decode) and is itself 0.19 ms at the TP8 shapes.  Outside torch.compile tracing the layers therefore call
# the ops' Python bodies directly (same checks, same C ABI call); under tracing the registered ops are used.
_RAW = {n: getattr(globals()[n], "_init_fn", globals()[n])
        for n in ("store_kv", "slot_table_update", "build_meta", "step_meta", "decode", "prefill", "qk_norm_rope_store", "sample")}


def call(name: str, *args):
    if torch.compiler.is_compiling():
        return globals()[name](*args)
    return _RAW[name](*args)
