#!/usr/bin/env python
"""Microbenchmarks of every hot-path kernel next to the reference's GPU implementation
(FlashInfer wrappers built exactly as nanovllmx/layers/attention.py:49-59; the Triton slot-write
kernel semantics via torch index_copy as a stand-in is NOT used: FlashInfer/Triton only).
BASELINE.md M1-M3.  CUDA-event timing, inputs >> L2 or L2 flushed between iterations.

  python benchmarks/sweep.py [decode|prefill|store|qknorm|meta|all] [--no-flashinfer]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import nanovllmx_b200 as pkg

DEV = torch.device("cuda")
PEAK = 6524.6
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    PEAK_TF = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]
except Exception:
    PEAK_TF = 1590.0

_flush = None


def flush_l2():
    global _flush
    if _flush is None:
        _flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=DEV)
    _flush.zero_()


def timeit(fn, iters=20, warmup=5, flush=False):
    """Median GPU time of one call.  All iterations are enqueued before the single sync, so the GPU never
    waits for the Python dispatch of the next call (an event pair around a call issued to an IDLE GPU also
    measures the host's launch latency, ~45 us through torch.library + ctypes)."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    evs = []
    if not flush:
        fn()                                     # keeps the queue non-empty ahead of the first timed call
    for _ in range(iters):
        if flush:
            torch.cuda._sleep(300000)            # ~150 us of GPU time: lets the host enqueue the timed call first
            flush_l2()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[len(ts) // 2]


def time_graph(fn, reps=10, iters=10):
    """GPU time per call with launch/dispatch overhead removed: `reps` calls captured in one CUDA graph."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush_l2()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); g.replay(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) / reps)
    ts.sort()
    return ts[len(ts) // 2]


def decode_case(B, L, Hq, Hkv, ps, ragged=False, seed=0):
    D = 128
    rng = np.random.default_rng(seed)
    lens = rng.integers(max(1, L // 8), L + 1, B) if ragged else np.full(B, L)
    pages = -(-lens // ps)
    n_pages = int(pages.sum())
    perm = torch.from_numpy(rng.permutation(n_pages).astype(np.int32)).to(DEV)
    indptr = torch.from_numpy(np.concatenate([[0], np.cumsum(pages)]).astype(np.int32)).to(DEV)
    last = torch.from_numpy((lens - (pages - 1) * ps).astype(np.int32)).to(DEV)
    k = torch.randn((n_pages * ps, Hkv, D), device=DEV, dtype=torch.bfloat16)
    v = torch.randn_like(k)
    q = torch.randn((B, Hq, D), device=DEV, dtype=torch.bfloat16)
    bytes_ = int(lens.sum()) * 2 * Hkv * D * 2 + 2 * B * Hq * D * 2 + 4 * (n_pages + 2 * B + 1)
    return dict(q=q, k=k, v=v, indptr=indptr, idx=perm, last=last, ps=ps, B=B, Hq=Hq, Hkv=Hkv, bytes=bytes_, lens=lens)


def bench_decode(use_fi=True):
    rows = []
    cfgs = []
    for (Hq, Hkv) in [(32, 8), (16, 4), (8, 2), (4, 1)]:
        cfgs.append(("cfg5 B64 L32k", 64, 32768, Hq, Hkv, 16, False))
        cfgs.append(("cfg5 B64 L32k", 64, 32768, Hq, Hkv, 1, False))
    for (Hq, Hkv) in [(16, 8), (32, 8), (4, 1)]:
        cfgs.append(("bench-like B256 L<=2048 ragged", 256, 2048, Hq, Hkv, 1, True))
        cfgs.append(("small B8 L<=1024 ragged", 8, 1024, Hq, Hkv, 1, True))
        cfgs.append(("single B1 L4096", 1, 4096, Hq, Hkv, 1, False))
    for name, B, L, Hq, Hkv, ps, ragged in cfgs:
        c = decode_case(B, L, Hq, Hkv, ps, ragged)
        ws = torch.empty(pkg.ops.decode_workspace_bytes(B, Hq, Hkv, 128), dtype=torch.uint8, device=DEV)
        small = c["bytes"] < 512e6
        f = lambda: pkg.ops.decode(c["q"], c["k"], c["v"], c["indptr"], c["idx"], c["last"], ws, ps, 128 ** -0.5)
        ms = timeit(f, flush=small)
        row = dict(case=name, B=B, L=L, Hq=Hq, Hkv=Hkv, ps=ps, MB=c["bytes"] / 1e6, ms=ms, GBs=c["bytes"] / ms / 1e6,
                   frac_measured=c["bytes"] / ms / 1e6 / PEAK)
        if small:
            row["graph_ms"] = time_graph(f)   # working set < L2 after the first of 10 captured calls: latency, not bandwidth
        if use_fi:
            try:
                import flashinfer as fi
                fws = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device=DEV)
                w = fi.BatchDecodeWithPagedKVCacheWrapper(fws, "NHD", use_tensor_cores=False, backend="auto")
                w.plan(indptr=c["indptr"], indices=c["idx"], last_page_len=c["last"], num_qo_heads=Hq, num_kv_heads=Hkv,
                       head_dim=128, page_size=ps, q_data_type=torch.bfloat16, kv_data_type=torch.bfloat16)
                kk, vv = c["k"].view(-1, ps, Hkv, 128), c["v"].view(-1, ps, Hkv, 128)
                g = lambda: w.run(c["q"], (kk, vv))
                fms = timeit(g, flush=small)
                if small:
                    row["flashinfer_graph_ms"] = time_graph(g)
                row.update(flashinfer_ms=fms, flashinfer_GBs=c["bytes"] / fms / 1e6, speedup=fms / ms)
                t0 = time.perf_counter()
                for _ in range(20):
                    w.plan(indptr=c["indptr"], indices=c["idx"], last_page_len=c["last"], num_qo_heads=Hq,
                           num_kv_heads=Hkv, head_dim=128, page_size=ps, q_data_type=torch.bfloat16,
                           kv_data_type=torch.bfloat16)
                torch.cuda.synchronize()
                row["flashinfer_plan_host_ms"] = (time.perf_counter() - t0) / 20 * 1e3
                err = (g().float() - f().float()).abs().max().item()
                row["max_abs_diff_vs_flashinfer"] = err
            except Exception as e:
                row["flashinfer"] = f"unavailable: {type(e).__name__}: {str(e)[:100]}"
        print(json.dumps(row), flush=True)
        rows.append(row)
        del c
        torch.cuda.empty_cache()
    return rows


def prefill_case(q_lens, prefix, Hq, Hkv, ps=1, seed=0):
    D = 128
    rng = np.random.default_rng(seed)
    q_lens, prefix = np.asarray(q_lens), np.asarray(prefix)
    lens = q_lens + prefix
    pages = -(-lens // ps)
    n_pages = int(pages.sum())
    perm = torch.from_numpy(rng.permutation(n_pages).astype(np.int32)).to(DEV)
    indptr = torch.from_numpy(np.concatenate([[0], np.cumsum(pages)]).astype(np.int32)).to(DEV)
    qo = torch.from_numpy(np.concatenate([[0], np.cumsum(q_lens)]).astype(np.int32)).to(DEV)
    last = torch.from_numpy((lens - (pages - 1) * ps).astype(np.int32)).to(DEV)
    k = torch.randn((n_pages * ps, Hkv, D), device=DEV, dtype=torch.bfloat16)
    v = torch.randn_like(k)
    q = torch.randn((int(q_lens.sum()), Hq, D), device=DEV, dtype=torch.bfloat16)
    flops = 4.0 * Hq * D * float((q_lens * prefix + q_lens * (q_lens + 1) / 2).sum())
    return dict(q=q, k=k, v=v, qo=qo, indptr=indptr, idx=perm, last=last, ps=ps, Hq=Hq, Hkv=Hkv, flops=flops,
                B=len(q_lens), T=int(q_lens.sum()))


def bench_prefill(use_fi=True):
    from oracle.paged_attention_oracle import bench_workload
    prompts, _ = bench_workload()
    plens = [len(p) for p in prompts]
    cfgs = [
        ("bench prompts 64 seqs", plens[:64], [0] * 64),
        ("bench prompts 256 seqs (one step)", plens, [0] * 256),
        ("uniform 16 x 4096", [4096] * 16, [0] * 16),
        ("shared prefix 2048 + suffix 16..256 x256", list(np.random.default_rng(1).integers(16, 257, 256)), [2048] * 256),
    ]
    rows = []
    for (Hq, Hkv) in [(16, 8), (32, 8), (4, 1)]:
        for name, ql, pf in cfgs:
            c = prefill_case(ql, pf, Hq, Hkv)
            ws = torch.empty(pkg.ops.prefill_workspace_bytes(c["B"], c["T"], Hq, Hkv, 128), dtype=torch.uint8, device=DEV)
            f = lambda: pkg.ops.prefill(c["q"], c["k"], c["v"], c["qo"], c["indptr"], c["idx"], c["last"], ws, 1,
                                        128 ** -0.5, True)
            ms = timeit(f, iters=10, warmup=3)
            row = dict(case=name, Hq=Hq, Hkv=Hkv, T=c["T"], GFLOP=c["flops"] / 1e9, ms=ms, TFLOPs=c["flops"] / ms / 1e9,
                       frac_measured_bf16=c["flops"] / ms / 1e9 / PEAK_TF)
            if use_fi:
                try:
                    import flashinfer as fi
                    fws = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device=DEV)
                    w = fi.BatchPrefillWithPagedKVCacheWrapper(fws, "NHD", backend="auto")
                    w.plan(qo_indptr=c["qo"], paged_kv_indices=c["idx"], paged_kv_indptr=c["indptr"],
                           paged_kv_last_page_len=c["last"], num_qo_heads=Hq, num_kv_heads=Hkv, head_dim_qk=128,
                           head_dim_vo=128, page_size=1, causal=True, q_data_type=torch.bfloat16,
                           kv_data_type=torch.bfloat16)
                    kk, vv = c["k"].view(-1, 1, Hkv, 128), c["v"].view(-1, 1, Hkv, 128)
                    g = lambda: w.run(c["q"], (kk, vv))
                    fms = timeit(g, iters=10, warmup=3)
                    row.update(flashinfer_ms=fms, flashinfer_TFLOPs=c["flops"] / fms / 1e9, speedup=fms / ms,
                               max_abs_diff_vs_flashinfer=(g().float() - f().float()).abs().max().item())
                except Exception as e:
                    row["flashinfer"] = f"unavailable: {type(e).__name__}: {str(e)[:100]}"
            print(json.dumps(row), flush=True)
            rows.append(row)
            del c
            torch.cuda.empty_cache()
    return rows


def bench_store():
    rows = []
    for T, Hkv in [(256, 8), (256, 1), (16384, 8), (142827, 8), (142827, 1)]:
        D = 128
        qkv = torch.randn((T, (4 * Hkv + 2 * Hkv) * D), device=DEV, dtype=torch.bfloat16)
        k = qkv[:, 4 * Hkv * D:5 * Hkv * D].contiguous().view(T, Hkv, D)
        v = qkv[:, 5 * Hkv * D:].view(T, Hkv, D)
        pool = max(2 * T, 4096)
        kc = torch.empty((pool, Hkv, D), device=DEV, dtype=torch.bfloat16)
        vc = torch.empty_like(kc)
        slots = torch.randperm(pool, device=DEV)[:T]
        f = lambda: pkg.store_kvcache(k, v, kc, vc, slots)
        ms = timeit(f, flush=True)
        by = T * 2 * Hkv * D * 2 * 2 + 8 * T
        row = dict(case="store_kv", T=T, Hkv=Hkv, MB=by / 1e6, ms=ms, GBs=by / ms / 1e6, frac_measured=by / ms / 1e6 / PEAK)
        try:   # the reference's Triton kernel semantics, timed via torch index_put as the stock GPU alternative
            g = lambda: (kc.index_copy_(0, slots, k), vc.index_copy_(0, slots, v.contiguous()))
            row["torch_index_copy_ms"] = timeit(g, flush=True)
        except Exception as e:
            row["torch_index_copy"] = str(e)[:80]
        print(json.dumps(row), flush=True)
        rows.append(row)
    return rows


def bench_qknorm():
    """f-3: the fused q/k RMSNorm + RoPE + slot write kernel vs the unfused sequence of the reference
    (models/qwen3.py:123-134: q_norm, k_norm, rotary_emb as separate elementwise kernels, then the slot write),
    restated with stock torch ops (F.rms_norm + eager RoPE) + this repo's store kernel."""
    import torch.nn.functional as F
    rows = []
    D, max_pos = 128, 4096
    inv = 1.0 / (1e6 ** (torch.arange(0, D, 2, dtype=torch.float, device=DEV) / D))
    fr = torch.outer(torch.arange(max_pos, dtype=torch.float, device=DEV), inv)
    cache = torch.cat((fr.cos(), fr.sin()), -1).contiguous()
    for T, Hq, Hkv in [(256, 16, 8), (256, 32, 8), (256, 4, 1), (16384, 32, 8), (142827, 16, 8), (142827, 32, 8), (142827, 4, 1)]:
        qkv = torch.randn((T, (Hq + 2 * Hkv) * D), device=DEV, dtype=torch.bfloat16)
        qw = torch.ones(D, device=DEV, dtype=torch.bfloat16)
        kw = torch.ones(D, device=DEV, dtype=torch.bfloat16)
        pos = torch.randint(0, max_pos, (T,), device=DEV)
        pool = max(2 * T, 4096)
        kc = torch.empty((pool, Hkv, D), device=DEV, dtype=torch.bfloat16)
        vc = torch.empty_like(kc)
        slots = torch.randperm(pool, device=DEV)[:T]
        q_view = qkv[:, : Hq * D]
        f = lambda: pkg.ops.qk_norm_rope_store(qkv, pos, qw, kw, 1e-6, cache, q_view, None, kc, vc, slots, Hq, Hkv, D)
        ms = timeit(f, flush=True)
        by = 2 * T * (Hq + 2 * Hkv) * D * 2 + 16 * T
        row = dict(case="qk_norm_rope_store", T=T, Hq=Hq, Hkv=Hkv, MB=by / 1e6, ms=ms, GBs=by / ms / 1e6,
                   frac_measured=by / ms / 1e6 / PEAK)
        if T <= 1024:
            row["graph_ms"] = time_graph(f)

        def unfused():
            q, k, v = qkv.split([Hq * D, Hkv * D, Hkv * D], dim=-1)
            qh = F.rms_norm(q.reshape(T, Hq, D), (D,), qw, 1e-6)
            kh = F.rms_norm(k.reshape(T, Hkv, D), (D,), kw, 1e-6)
            cs = cache.index_select(0, pos)
            cos, sin = cs[:, None, : D // 2].to(qh.dtype), cs[:, None, D // 2:].to(qh.dtype)

            def rope(x):
                x1, x2 = x.chunk(2, -1)
                return torch.stack((x1 * cos - x2 * sin, x2 * cos + x1 * sin), -1).flatten(-2)
            qo, ko = rope(qh), rope(kh)
            pkg.store_kvcache(ko, v.view(T, Hkv, D), kc, vc, slots)
            return qo
        try:
            row["unfused_torch_eager_ms"] = timeit(unfused, flush=True)
            if T <= 1024:
                row["unfused_graph_ms"] = time_graph(unfused)
        except Exception as e:
            row["unfused"] = str(e)[:80]
        print(json.dumps(row), flush=True)
        rows.append(row)
    return rows


def bench_meta():
    """Host wall time per step of the metadata build at bench-like decode shape (B=256, L~1-2k):
    ours (one message + 3 kernels) vs the reference's algorithm (2B H2D copies + B .item() syncs),
    restated here for timing only."""
    class S:
        pass
    rng = np.random.default_rng(0)
    B = 256
    seqs = []
    for i in range(B):
        n = int(rng.integers(500, 2048))
        s = S(); s.seq_id, s.token_ids, s.kv_indices, s.cached_kv_len = f"s{i}", [0] * n, [int(x) for x in rng.integers(0, 1 << 20, n)], n - 1
        seqs.append(s)
    class Bt:
        pass
    b = Bt(); b.forward_mode, b.seqs, b.num_seqs = pkg.ForwardMode.DECODE, seqs, B
    kvp = pkg.KVCachePool(torch.bfloat16, DEV, 16, 1, 8, 128)
    be = pkg.AttentionBackend(kvp, 16, 8, 128, torch.bfloat16, DEV)
    graph = pkg.CUDAGraph()
    be.prepare_for_cuda_graph_capture(graph, 256, 4096)
    be.build_metadata_for_cuda_graph_capture(graph, 256, 4096)
    def ours_eager():
        be.build_metadata(b)
    def ours_replay():
        be.build_metadata_for_cuda_graph_replay(graph, b, 256)
    def ref_algo():   # layers/attention.py:219-275 restated (device tensors per sequence, fills, cat)
        lens = torch.tensor([len(s.kv_indices) for s in seqs], dtype=torch.int32, device=DEV)
        indptr = torch.cat([torch.zeros(1, dtype=torch.int32, device=DEV), torch.cumsum(lens, 0, dtype=torch.int32)])
        idx = [torch.tensor(s.kv_indices, dtype=torch.int32, device=DEV) for s in seqs]
        graph.get_input_buffer("paged_kv_indices_buffer").fill_(0)
        flat = torch.cat(idx)
        graph.get_input_buffer("paged_kv_indices_buffer")[:flat.numel()].copy_(flat)
        out = torch.cat([torch.tensor(s.kv_indices[s.cached_kv_len:], dtype=torch.long, device=DEV) for s in seqs])
        return indptr, out
    rows = []
    for name, fn in [("ours build_metadata (eager)", ours_eager), ("ours graph replay refill", ours_replay),
                     ("reference algorithm restated (no FlashInfer plan)", ref_algo)]:
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        n = 20
        for _ in range(n):
            fn()
        torch.cuda.synchronize()
        row = dict(case="metadata B=256 L~1.3k decode step", impl=name, wall_ms_per_step=(time.perf_counter() - t0) / n * 1e3)
        print(json.dumps(row), flush=True)
        rows.append(row)
    return rows


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("what", nargs="?", default="all")
    ap.add_argument("--no-flashinfer", action="store_true")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    res = {}
    if a.what in ("decode", "all"):
        res["decode"] = bench_decode(not a.no_flashinfer)
    if a.what in ("prefill", "all"):
        res["prefill"] = bench_prefill(not a.no_flashinfer)
    if a.what in ("store", "all"):
        res["store"] = bench_store()
    if a.what in ("qknorm", "all"):
        res["qknorm"] = bench_qknorm()
    if a.what in ("meta", "all"):
        res["meta"] = bench_meta()
    if a.out:
        json.dump(res, open(a.out, "w"), indent=1)
