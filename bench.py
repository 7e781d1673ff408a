#!/usr/bin/env python
"""bench.py — the hot-path benchmark contract.

Workload (BASELINE.json configs[4], the configuration the headline metric "paged decode-attn HBM
GB/s" is quoted on; it fits one GPU): Qwen3-8B attention shapes, ONE layer, batch 64, 32 768-token
context per sequence, page size 16, bf16, page table = seeded random permutation of the pool's
pages (fragmented), Hq/Hkv = 32/8 at TP1 and 32/N / 8/N per rank at TP N (KV-head sharding, no
collective on the path).  One "step" = one decode pass of the hot path over the
batch: slot write of the 64 new K/V rows + split-KV paged decode attention (+ merge).

  value     algorithmic GB/s (SURVEY.md 8d formula) with everything resident in HBM, whole step, K CUDA-graph replays
  e2e       same metric through the public drop-in API with q/k/v in pinned HOST memory and the output read back to the
            host every step: a real decode progression (one new token and slot per sequence per step) driven as the
            reference drives decode -- AttentionBackend.build_metadata_for_cuda_graph_replay (host Sequence lists) +
            replay of the captured Attention.forward (core/model_runner.py:299-305); --no-graph: eager build_metadata
  roofline  the dominant kernel (decode streaming kernel + its merge): its share of the step from a paired pass
            (step graph / decode-only graph replayed alternately) applied to the timed ms_per_step
  cpu_baseline  the oracle's plain-C port of the reference semantics on the host cores, on the whole workload
  prefill / gpu_reference / engine   extra keys: tensor-pipe roofline of the prefill kernel, FlashInfer on the same
            inputs, the STOCK reference engine (baseline/_ref) with and without the drop-in at tp = N

`--impl reference` times that CPU restatement only (the reference has no CPU attention of its own:
FlashInfer and Triton are CUDA-only) and prints the same JSON line shape.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np
import torch

METRIC = "paged_decode_attn_hbm_gbps"
UNIT = "GB/s"
NOMINAL_HBM_GBS = 8000.0   # north_star's "roughly 8 TB/s per B200"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--ctx", type=int, default=32768)
    ap.add_argument("--page-size", type=int, default=16)
    ap.add_argument("--cpu-sample-seqs", type=int, default=64,
                    help="sequences of the workload the CPU arm runs per pass (default: all of them)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-prefill", action="store_true", help="skip the prefill (tensor-pipe) extra key")
    ap.add_argument("--no-gpu-reference", action="store_true", help="skip the FlashInfer-on-the-same-inputs extra key")
    ap.add_argument("--no-m3", action="store_true",
                    help="skip the extra keys store_kv / meta / decode_page1 (BASELINE.md M3 and the page-size-1 headline)")
    ap.add_argument("--no-engine", action="store_true",
                    help="skip the engine extra key (stock reference engine, FlashInfer arm and b200 arm, bench.py workload)")
    ap.add_argument("--engine-model", default="8b", choices=["0.6b", "8b"])
    ap.add_argument("--engine-timeout", type=int, default=420, help="seconds per engine arm")
    ap.add_argument("--no-graph", action="store_true",
                    help="time eager launches instead of CUDA-graph replays of the step (the reference replays its decode "
                         "step from a CUDA graph: core/cuda_graph.py:27-39)")
    return ap.parse_args()


def algorithmic_bytes(B, L, Hq, Hkv, D, ps):
    """KV read once per KV head + q read + o written + index bytes."""
    kv = B * 2 * L * Hkv * D * 2
    qo = 2 * B * Hq * D * 2
    idx = 4 * (B * (-(-L // ps)) + 2 * B + 1)
    return kv + qo + idx


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """Samples SM clock + throttle reasons during the timed region (pynvml, 100 ms period)."""

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(vis.split(",")[index]) if vis and vis.split(",")[index].isdigit() else index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        nv = self.nv
        names = {"hw_slowdown": 0x8, "sw_power_cap": 0x4, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20,
                 "hw_power_brake_slowdown": 0x80}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(nv, "nvmlDeviceGetCurrentClocksEventReasons") \
                    else nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.02)

    def __enter__(self):
        if self.nv is not None:
            self._t = threading.Thread(target=self._loop, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t is not None:
            self._t.join(timeout=2)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": 0}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# -------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port on the host cores
# -------------------------------------------------------------------------------------------------------
def cpu_port_gbps(B_sample, L, Hq, Hkv, D, ps, steps=1, warmup=0):
    from oracle import paged_attention_oracle as orc
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(0)
    n_pages = B_sample * (L // ps)
    # N(0,1) bf16 for 8 sequences' worth of pages, tiled over the pool (drawing 2 x 4.3 G normals for all 64 sequences
    # takes longer than the measurement; the arithmetic has no data-dependent work and the page table is a random
    # permutation of the whole pool either way)
    chunk = min(n_pages, 8 * (L // ps)) * ps
    reps = -(-(n_pages * ps) // chunk)
    k_cache = torch.randn((chunk, Hkv, D), generator=g, dtype=torch.bfloat16).repeat(reps, 1, 1)[:n_pages * ps].contiguous()
    v_cache = torch.randn((chunk, Hkv, D), generator=g, dtype=torch.bfloat16).repeat(reps, 1, 1)[:n_pages * ps].contiguous()
    q = torch.randn((B_sample, Hq, D), generator=g).to(torch.bfloat16)
    k_new = torch.randn((B_sample, Hkv, D), generator=g).to(torch.bfloat16)
    v_new = torch.randn((B_sample, Hkv, D), generator=g).to(torch.bfloat16)
    perm = np.random.default_rng(0).permutation(n_pages).astype(np.int32)
    indptr = (np.arange(B_sample + 1) * (L // ps)).astype(np.int32)
    last = np.full(B_sample, ps, np.int32)
    new_slots = torch.from_numpy(np.array([int(perm[indptr[i + 1] - 1]) * ps + ps - 1 for i in range(B_sample)], np.int64))

    # the plain-C port (pthreads, every host core; one work item per (sequence, KV head) so the query heads of a
    # group share each K/V row): ~10x the numpy/torch restatement of the same arithmetic on the same cores
    from oracle import c_oracle
    qo = np.arange(B_sample + 1, dtype=np.int32)

    def step():
        orc.store_kvcache(k_new, v_new, k_cache, v_cache, new_slots)
        return c_oracle.paged_attention(q, k_cache, v_cache, qo, indptr, perm, last, ps, causal=False, threads=threads)

    for _ in range(warmup):
        step()
    if steps is None:                       # default bench run: about 10 s of CPU work
        t0 = time.perf_counter()
        step()
        steps = int(min(200, max(3, 10.0 / max(time.perf_counter() - t0, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    return algorithmic_bytes(B_sample, L, Hq, Hkv, D, ps) / dt / 1e9, dt, threads, steps


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # one host runs the WHOLE job's heads (the GPU arm shards them over the ranks; total work is fixed)
    Hq, Hkv, D = 32, 8, 128
    n_sample = min(args.cpu_sample_seqs, args.batch)
    val, dt, threads, _ = cpu_port_gbps(n_sample, args.ctx, Hq, Hkv, D, args.page_size,
                                        steps=max(args.steps, 1), warmup=args.warmup)
    sample = (f"{n_sample} of {args.batch} sequences x {args.ctx} tokens per step" +
              (" (the whole workload)" if n_sample >= args.batch else
               ", same shapes; ms_per_step is the time of this sample, not of the whole batch"))
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": workload_config(args, 32 // args.gpus, max(1, 8 // args.gpus), D),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, Hq, Hkv, D):
    return {"workload": "BASELINE.json configs[4]: Qwen3-8B long-context paged decode attention, one layer",
            "batch": args.batch, "context_len": args.ctx, "page_size": args.page_size, "num_qo_heads_per_rank": Hq,
            "num_kv_heads_per_rank": Hkv, "head_dim": D, "parallelism": f"tp{args.gpus} (KV-head sharding, no collective)",
            "page_table": "seeded random permutation (fragmented)",
            "l2": "inputs larger than L2 (KV pool per rank >> 126 MB)"}



# -------------------------------------------------------------------------------------------------------
# extra keys: prefill (tensor-pipe roofline), FlashInfer on the same inputs, the stock engine
# -------------------------------------------------------------------------------------------------------
def measured_tflops():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        return float(d["bf16_tflops"]), float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 2250.0, 2250.0, "fallback (nominal dense bf16 2.25 PFLOP/s)"


def prefill_flops(q_lens, prefix_lens, Hq, D):
    """4 Hq D sum_i (N_i P_i + N_i (N_i + 1) / 2); the masked half is not counted."""
    return 4.0 * Hq * D * sum(n * p + n * (n + 1) / 2 for n, p in zip(q_lens, prefix_lens))


def median_event_ms(fn, iters, stream):
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in evs:
        a.record(stream)
        fn()
        b.record(stream)
    torch.cuda.synchronize()
    return statistics.median(a.elapsed_time(b) for a, b in evs)


def make_prefill_case(q_lens, Hq, Hkv, D, devc, seed):
    """Fragmented page-size-1 pool holding exactly the prompts' tokens (the reference's only mode)."""
    T = int(sum(q_lens))
    g = torch.Generator(device=devc).manual_seed(seed)
    k_cache = torch.randn((T + 64, Hkv, D), device=devc, dtype=torch.bfloat16, generator=g)
    v_cache = torch.randn((T + 64, Hkv, D), device=devc, dtype=torch.bfloat16, generator=g)
    q = torch.randn((T, Hq, D), device=devc, dtype=torch.bfloat16, generator=g)
    perm = torch.from_numpy(np.random.default_rng(seed).permutation(T + 64)[:T].astype(np.int32)).to(devc)
    indptr = torch.from_numpy(np.concatenate([[0], np.cumsum(q_lens)]).astype(np.int32)).to(devc)
    last = torch.ones(len(q_lens), dtype=torch.int32, device=devc)
    return dict(q=q, k_cache=k_cache, v_cache=v_cache, qo_indptr=indptr, kv_indptr=indptr.clone(), kv_indices=perm,
                last_page_len=last, T=T, B=len(q_lens), q_lens=list(q_lens))


def make_shared_prefix_case(Hq, Hkv, D, devc, seed, B=256, P=2048):
    """BASELINE.json configs[3] / SURVEY.md 8(d) config 4 at the kernel: B sequences whose first P KV slots are the SAME
    P slots (one radix-tree path) + a private suffix of randint(16, 256) new tokens each; page size 1."""
    rng = np.random.default_rng(seed)
    suffix = rng.integers(16, 257, size=B)
    T = int(suffix.sum())
    n_slots = P + T + 64
    perm = rng.permutation(n_slots).astype(np.int32)
    qo = np.concatenate([[0], np.cumsum(suffix)]).astype(np.int32)
    kvp = np.concatenate([[0], np.cumsum(P + suffix)]).astype(np.int32)
    idx = np.concatenate([np.concatenate([perm[:P], perm[P + qo[i]:P + qo[i + 1]]]) for i in range(B)]).astype(np.int32)
    g = torch.Generator(device=devc).manual_seed(seed)
    rnd = lambda *shape: torch.randn(shape, device=devc, dtype=torch.bfloat16, generator=g)
    return dict(q=rnd(T, Hq, D), k_cache=rnd(n_slots, Hkv, D), v_cache=rnd(n_slots, Hkv, D),
                qo_indptr=torch.from_numpy(qo).to(devc), kv_indptr=torch.from_numpy(kvp).to(devc),
                kv_indices=torch.from_numpy(idx).to(devc), last_page_len=torch.ones(B, dtype=torch.int32, device=devc),
                T=T, B=B, q_lens=[int(x) for x in suffix], prefix_lens=[P] * B)


def prefill_cases(Hq, Hkv, D, devc):
    from oracle.paged_attention_oracle import bench_workload      # request generator of the reference's bench.py:14-29
    plens = [len(p) for p in bench_workload()[0]]
    return [("bench_prompts_256 (one prefill step of the reference bench.py workload: 256 prompts, 142 827 tokens, page 1)",
             make_prefill_case(plens, Hq, Hkv, D, devc, 11)),
            ("uniform_16x4096 (16 sequences x 4096 tokens, causal, page 1)", make_prefill_case([4096] * 16, Hq, Hkv, D, devc, 12)),
            ("shared_prefix_256x2048 (config 4: 256 sequences, the SAME 2048 cached prefix slots + randint(16,256) new tokens "
             "each, page 1)", make_shared_prefix_case(Hq, Hkv, D, devc, 13))]


def bench_prefill(pkg, Hq, Hkv, D, devc, stream, iters=10):
    burst, sustained, src = measured_tflops()
    out = {"bound": "tensor", "unit": "TFLOP/s", "peak_burst": burst, "peak_sustained": sustained, "peak_source": src, "cases": []}
    sm = D ** -0.5
    for name, c in prefill_cases(Hq, Hkv, D, devc):
        ws = torch.empty(pkg.ops.prefill_workspace_bytes(c["B"], c["T"], Hq, Hkv, D), dtype=torch.uint8, device=devc)
        run = lambda: pkg.ops.prefill(c["q"], c["k_cache"], c["v_cache"], c["qo_indptr"], c["kv_indptr"], c["kv_indices"],
                                      c["last_page_len"], ws, 1, sm, True)
        for _ in range(3):
            o = run()
        ms = median_event_ms(run, iters, stream)
        fl = prefill_flops(c["q_lens"], c.get("prefix_lens", [0] * c["B"]), Hq, D)
        tf = fl / (ms * 1e-3) / 1e12
        # spot check of the timed output: last token of the longest sequence, one head, direct fp32
        i = int(np.argmax(c["q_lens"]))
        a, b = int(c["qo_indptr"][i]), int(c["qo_indptr"][i + 1])
        h = Hq - 1
        slots = c["kv_indices"][int(c["kv_indptr"][i]):int(c["kv_indptr"][i + 1])].long()
        K, V = c["k_cache"][slots, h // (Hq // Hkv)].float(), c["v_cache"][slots, h // (Hq // Hkv)].float()
        ref = torch.softmax((K @ c["q"][b - 1, h].float()) * sm, 0) @ V
        err = float((o[b - 1, h].float() - ref).abs().max())
        out["cases"].append({"workload": name, "kernel_ms": ms, "algorithmic_tflop": fl / 1e12, "tflops": tf,
                             "frac_burst": tf / burst, "frac_sustained": tf / sustained, "self_check_max_abs_err": err,
                             "kernel": "prefill_plan_kernel + prefill_sm100_persist_kernel"})
        del ws
    return out


def bench_store_kv(pkg, Hkv, D, devc, stream, peak, iters=20):
    """BASELINE.md M3, slot write (replaces core/kv_cache.py:12-44): one prefill step of the bench workload (T = 142 827
    new tokens) and one decode step at B = 256.  Algorithmic bytes: K and V rows read once and written once + 8 B of
    slot id per token.  v is the strided view of the qkv GEMM output, as in the model (models/qwen3.py:123)."""
    out = []
    for T in (142827, 256):
        n_slots = T + 4096
        kc = torch.zeros((n_slots, Hkv, D), device=devc, dtype=torch.bfloat16)
        vc = torch.zeros_like(kc)
        kv = torch.randn((T, 2 * Hkv * D), device=devc, dtype=torch.bfloat16)
        k, v = kv[:, :Hkv * D].view(T, Hkv, D), kv[:, Hkv * D:].view(T, Hkv, D)
        slots = torch.from_numpy(np.random.default_rng(T).permutation(n_slots)[:T].astype(np.int64)).to(devc)
        run = lambda: pkg.store_kvcache(k, v, kc, vc, slots)
        for _ in range(3):
            run()
        ms = median_event_ms(run, iters, stream)
        ok = bool(torch.equal(kc[slots], k) and torch.equal(vc[slots], v))
        nbytes = T * (4 * Hkv * D * 2 + 8)
        gbs = nbytes / (ms * 1e-3) / 1e9
        out.append({"tokens": T, "kernel_ms": ms, "algorithmic_bytes": nbytes, "gbs": gbs, "frac": gbs / peak,
                    "bit_exact_self_check": ok, "kernel": "store_kv_rows_kernel"})
        del kc, vc, kv, slots
    return out


def bench_meta(pkg, Hq, Hkv, D, devc, stream, B=256, ctx=1300, steps=50):
    """BASELINE.md M3, metadata: a B = 256 decode step (mean context 1.3 k tokens, the bench workload's steady state)
    through AttentionBackend.build_metadata (eager; replaces layers/attention.py:62-149) and through
    build_metadata_for_cuda_graph_replay (the refill of the captured buffers; replaces :219-295).  host_us = wall time of
    the call on the host (packing + launches, no sync); device_us = CUDA-event time of what it enqueued."""
    pool = pkg.KVCachePool(torch.bfloat16, devc, B * (ctx + 2 * steps + 16) + 64, 1, Hkv, D)
    be = pkg.AttentionBackend(pool, Hq, Hkv, D, torch.bfloat16, devc)
    nxt = [0]

    def take(n):
        a = nxt[0]
        nxt[0] += n
        return list(range(a, a + n))
    seqs = [Seq(f"m{i}", ctx, take(ctx), ctx - 1) for i in range(B)]
    batch = Batch(pkg.ForwardMode.DECODE, seqs)
    g = pkg.CUDAGraph()
    be.prepare_for_cuda_graph_capture(g, B, ctx + 2 * steps + 16)
    be.build_metadata_for_cuda_graph_capture(g, B, ctx + 2 * steps + 16)

    def advance():
        for sq in seqs:
            sq.cached_kv_len = len(sq.kv_indices)
            sq.token_ids.append(0)
            sq.kv_indices.append(nxt[0])
            nxt[0] += 1
    res = {"batch": B, "context_len": ctx}
    for name, fn in (("eager_build_metadata", lambda: be.build_metadata(batch)),
                     ("graph_replay_refill", lambda: be.build_metadata_for_cuda_graph_replay(g, batch, B))):
        host, devt = [], []
        for i in range(steps + 5):
            advance()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(stream)
            t0 = time.perf_counter()
            fn()
            t1 = time.perf_counter()
            b.record(stream)
            b.synchronize()
            if i >= 5:
                host.append((t1 - t0) * 1e6)
                devt.append(a.elapsed_time(b) * 1e3)
        res[name] = {"host_us_median": statistics.median(host), "issue_to_done_us_median": statistics.median(devt),
                     "h2d_bytes_per_step": int(be.slot_table.last_msg_bytes)}
    res["full_resends_after_first_step"] = int(be.slot_table.full_resends - B)
    return res


def bench_decode_page1(pkg, B, L, Hq, Hkv, D, q, k_cache, v_cache, devc, stream, peak, iters=10):
    """The headline shapes at page size 1: the only page size the reference's token-granular allocator can drive
    (layers/attention.py:66, core/kv_cache.py:101-117).  Same pool and queries, slot-granular random page table."""
    n = B * L
    idx = torch.from_numpy(np.random.default_rng(1).permutation(n).astype(np.int32)).to(devc)
    indptr = (torch.arange(B + 1, dtype=torch.int64) * L).to(torch.int32).to(devc)
    last = torch.ones(B, dtype=torch.int32, device=devc)
    ws = torch.empty(pkg.ops.decode_workspace_bytes(B, Hq, Hkv, D), dtype=torch.uint8, device=devc)
    run = lambda: pkg.ops.decode(q, k_cache, v_cache, indptr, idx, last, ws, 1, D ** -0.5)
    for _ in range(3):
        o = run()
    ms = median_event_ms(run, iters, stream)
    b, h = B // 2, Hq - 1
    sl = idx[b * L:(b + 1) * L].long()
    K, V = k_cache[sl, h // (Hq // Hkv)].float(), v_cache[sl, h // (Hq // Hkv)].float()
    ref = torch.softmax((K @ q[b, h].float()) * D ** -0.5, 0) @ V
    nbytes = algorithmic_bytes(B, L, Hq, Hkv, D, 1)
    gbs = nbytes / (ms * 1e-3) / 1e9
    return {"page_size": 1, "kernel_ms": ms, "algorithmic_bytes_per_launch": nbytes, "gbs": gbs, "frac": gbs / peak,
            "frac_of_nominal_8TBs": gbs / NOMINAL_HBM_GBS, "self_check_max_abs_err": float((o[b, h].float() - ref).abs().max()),
            "kernel": "decode_kernel + decode_merge_kernel"}


def bench_flashinfer(B, L, ps, Hq, Hkv, D, q, k_cache, v_cache, kv_indptr, kv_indices, last_page_len, bytes_rank, devc, stream,
                     with_prefill=True, iters=20):
    """The reference's own arithmetic library on the SAME inputs: wrappers constructed and planned exactly as
    nanovllmx/layers/attention.py:49-59, 106-132 does ("NHD", backend="auto", decode with use_tensor_cores=False).
    FlashInfer's host plan() is outside the timed region (the reference pays it every step)."""
    import flashinfer
    res = {"library": f"flashinfer {getattr(flashinfer, '__version__', '?')}",
           "constructed_as": "nanovllmx/layers/attention.py:49-59 (NHD, backend=auto, use_tensor_cores=False)"}
    ws = torch.empty(512 * 1024 * 1024, dtype=torch.uint8, device=devc)
    w = flashinfer.BatchDecodeWithPagedKVCacheWrapper(ws, "NHD", use_tensor_cores=False, backend="auto")
    t0 = time.perf_counter()
    w.plan(indptr=kv_indptr, indices=kv_indices, last_page_len=last_page_len, num_qo_heads=Hq, num_kv_heads=Hkv, head_dim=D,
           page_size=ps, q_data_type=torch.bfloat16, kv_data_type=torch.bfloat16)
    w._sm_scale = D ** -0.5
    kc, vc = k_cache.view(-1, ps, Hkv, D), v_cache.view(-1, ps, Hkv, D)
    run = lambda: w.run(q, (kc, vc))
    for _ in range(3):
        o = run()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    w.plan(indptr=kv_indptr, indices=kv_indices, last_page_len=last_page_len, num_qo_heads=Hq, num_kv_heads=Hkv, head_dim=D,
           page_size=ps, q_data_type=torch.bfloat16, kv_data_type=torch.bfloat16)
    w._sm_scale = D ** -0.5
    plan_ms = (time.perf_counter() - t0) * 1e3
    ms = median_event_ms(run, iters, stream)
    res["decode"] = {"kernel_ms": ms, "gbps": bytes_rank / (ms * 1e-3) / 1e9, "host_plan_ms_per_step": plan_ms, "out": o}
    if with_prefill:
        res["prefill"] = []
        pw = flashinfer.BatchPrefillWithPagedKVCacheWrapper(ws, "NHD", backend="auto")
        for name, c in prefill_cases(Hq, Hkv, D, devc):
            pw.plan(qo_indptr=c["qo_indptr"], paged_kv_indices=c["kv_indices"], paged_kv_indptr=c["kv_indptr"],
                    paged_kv_last_page_len=c["last_page_len"], num_qo_heads=Hq, num_kv_heads=Hkv, head_dim_qk=D, head_dim_vo=D,
                    page_size=1, causal=True, q_data_type=torch.bfloat16, kv_data_type=torch.bfloat16)
            pw._sm_scale = D ** -0.5
            prun = lambda: pw.run(c["q"], (c["k_cache"].view(-1, 1, Hkv, D), c["v_cache"].view(-1, 1, Hkv, D)))
            for _ in range(2):
                prun()
            pms = median_event_ms(prun, 5, stream)
            fl = prefill_flops(c["q_lens"], [0] * c["B"], Hq, D)
            res["prefill"].append({"workload": name, "kernel_ms": pms, "tflops": fl / (pms * 1e-3) / 1e12})
    return res


def run_engine_arms(args, tp):
    """configs[1]/[2] of BASELINE.json: the reference's bench.py workload through the STOCK reference engine
    (baseline/_ref, real scheduler / radix cache / ModelRunner / CUDA-graph decode), once as shipped (FlashInfer +
    Triton) and once with INTEGRATION.md section 1's drop-in.  tok/s by the reference's own formula
    (bench.py:41-43: sum(max_tokens) / wall) and as really generated tokens / wall; first pass = the reference
    bench protocol (one 1-request warm-up), second pass = steady state."""
    import subprocess
    res = {"model": f"Qwen3-{args.engine_model} shapes, random init", "tp": tp, "workload": "reference bench.py:13-31 (256 requests)",
           "harness": "benchmarks/ref_engine.py (unmodified reference tree; import fix, stub tokenizer, seeded init, eager sampler)"}
    for arm, key in (("stock", "flashinfer"), ("b200", "b200")):
        out = f"/tmp/bench_engine_{arm}_tp{tp}.json"
        cmd = [sys.executable, os.path.join(ROOT, "benchmarks", "ref_engine.py"), "--model", args.engine_model, "--backend", arm,
               "--tp", str(tp), "--runs", "2", "--nccl-port", str(29620 + (0 if arm == "stock" else 10)), "--out", out]
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis:                       # the engine sets CUDA_VISIBLE_DEVICES itself (core/executor.py:27-28): hand it OUR devices
            cmd += ["--device-ids", ",".join(vis.split(",")[:tp])]
        env = {k: v for k, v in os.environ.items()
               if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT", "CUDA_VISIBLE_DEVICES", "LOCAL_WORLD_SIZE",
                            "GROUP_RANK", "ROLE_RANK", "ROLE_WORLD_SIZE", "TORCHELASTIC_RUN_ID", "GROUP_WORLD_SIZE", "ROLE_NAME")
               and not k.startswith("TORCHELASTIC")}
        try:
            r = subprocess.run(cmd, capture_output=True, text=True, timeout=args.engine_timeout, env=env, start_new_session=True)
            if r.returncode != 0:
                res[key] = {"error": (r.stderr or r.stdout)[-400:]}
                continue
            with open(out) as f:
                d = json.load(f)
            res[key] = {"first_pass": {k: d["runs"][0][k] for k in ("wall_s", "tok_s_reference_formula", "tok_s_generated")},
                        "steady_pass": {k: d["runs"][-1][k] for k in ("wall_s", "tok_s_reference_formula", "tok_s_generated")},
                        "init_s": d["init_s"]}
        except subprocess.TimeoutExpired:
            res[key] = {"error": f"timed out after {args.engine_timeout} s"}
        except Exception as e:      # the engine leg must never take the headline line down with it
            res[key] = {"error": f"{type(e).__name__}: {e}"[:400]}
    a, b = res.get("b200", {}), res.get("flashinfer", {})
    if "steady_pass" in a and "steady_pass" in b:
        res["tok_s_reference_formula"] = a["steady_pass"]["tok_s_reference_formula"]
        res["tok_s_generated"] = a["steady_pass"]["tok_s_generated"]
        res["flashinfer_tok_s"] = b["steady_pass"]["tok_s_reference_formula"]
        res["b200_over_flashinfer"] = a["steady_pass"]["tok_s_reference_formula"] / b["steady_pass"]["tok_s_reference_formula"]
    return res


# -------------------------------------------------------------------------------------------------------
# GPU arm
# -------------------------------------------------------------------------------------------------------
class Seq:
    __slots__ = ("seq_id", "token_ids", "kv_indices", "cached_kv_len")

    def __init__(self, sid, n_tok, kv, cached):
        self.seq_id, self.token_ids, self.kv_indices, self.cached_kv_len = sid, [0] * n_tok, kv, cached


class Batch:
    def __init__(self, mode, seqs):
        self.forward_mode, self.seqs, self.num_seqs = mode, seqs, len(seqs)


def run_b200_arm(args):
    import nanovllmx_b200 as pkg
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device; there is no CPU fallback")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        host_group = dist.new_group(backend="gloo")      # host-side waits (engine leg): no NCCL kernel spinning on a GPU
    tp = max(world, 1)
    assert args.gpus == tp, f"--gpus {args.gpus} but WORLD_SIZE {world}"
    Hq, Hkv, D = 32 // tp, max(1, 8 // tp), 128
    B, L, ps = args.batch, args.ctx, args.page_size
    devc = torch.device("cuda", local)
    pages_per_seq = -(-L // ps)
    n_pages = B * pages_per_seq
    # the e2e leg is a real decode progression (one new token per sequence per step): spare pages past the timed pool
    grow_pages = -(-(2 * args.steps + max(args.warmup, 3) + 4) // ps) + 1      # pipelined pass + synchronised pass of the e2e leg
    num_slots = (n_pages + B * grow_pages) * ps

    # ---- resident state: pool (one layer), page table, inputs ---------------------------------------
    torch.manual_seed(0)
    pool = pkg.KVCachePool(torch.bfloat16, devc, num_slots, 1, Hkv, D)
    pool.k_cache.normal_()
    pool.v_cache.normal_()
    perm_np = np.random.default_rng(0).permutation(n_pages).astype(np.int32)   # same table on every rank
    kv_indices = torch.from_numpy(perm_np).to(devc)
    kv_indptr = (torch.arange(B + 1, dtype=torch.int32) * pages_per_seq).to(devc)
    last_len = L - (pages_per_seq - 1) * ps
    last_page_len = torch.full((B,), last_len, dtype=torch.int32, device=devc)
    new_slots_np = np.array([int(perm_np[(i + 1) * pages_per_seq - 1]) * ps + last_len - 1 for i in range(B)], np.int64)
    new_slots = torch.from_numpy(new_slots_np).to(devc)
    qkv = torch.randn((B, (Hq + 2 * Hkv) * D), device=devc, dtype=torch.bfloat16)
    q = qkv[:, :Hq * D].contiguous().view(B, Hq, D)
    k_new = qkv[:, Hq * D:(Hq + Hkv) * D].contiguous().view(B, Hkv, D)
    v_new = qkv[:, (Hq + Hkv) * D:].view(B, Hkv, D)          # strided view, as in the reference model
    k_cache, v_cache = pool.get_kv_cache(0)
    ws = torch.empty(pkg.ops.decode_workspace_bytes(B, Hq, Hkv, D), dtype=torch.uint8, device=devc)
    sm_scale = D ** -0.5
    bytes_rank = algorithmic_bytes(B, L, Hq, Hkv, D, ps)
    stream = torch.cuda.current_stream()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device():
        pkg.store_kvcache(k_new, v_new, k_cache, v_cache, new_slots)
        return pkg.ops.decode(q, k_cache, v_cache, kv_indptr, kv_indices, last_page_len, ws, ps, sm_scale)

    # ---- (A) device-resident timing -----------------------------------------------------------------
    # Timed region = K steps, each step = slot write + decode attention.  By default the step is replayed from a
    # CUDA graph, as the reference runs its decode step (core/cuda_graph.py:27-39, core/model_runner.py:299-305):
    # at the TP4/TP8 shapes a step is 0.2-0.3 ms and eager launches through torch.library + ctypes (~45 us each)
    # leave the GPU idle between kernels.  The per-kernel times for the roofline come from an eager, event-
    # instrumented pass of the same K steps right after (CUDA events cannot be recorded inside a graph replay).
    for _ in range(max(args.warmup, 3)):
        step_device()
    barrier()
    graph = None
    launches_per_step = 0
    if not args.no_graph:
        pkg.ops.reset_launch_count()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            out = step_device()
        launches_per_step = pkg.ops.launch_count()
        for _ in range(3):
            graph.replay()
        barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
          for _ in range(args.steps)]
    pkg.ops.reset_launch_count()
    with ClockSampler(local) as clk:
        t_first = torch.cuda.Event(enable_timing=True)
        t_last = torch.cuda.Event(enable_timing=True)
        if graph is not None:
            t_first.record(stream)
            for _ in range(args.steps):
                graph.replay()
            t_last.record(stream)
            barrier()
            pkg.ops.reset_launch_count()
        e_first = torch.cuda.Event(enable_timing=True)
        e_last = torch.cuda.Event(enable_timing=True)
        e_first.record(stream)
        for a, b, c in ev:
            a.record(stream)
            pkg.store_kvcache(k_new, v_new, k_cache, v_cache, new_slots)
            b.record(stream)
            out = pkg.ops.decode(q, k_cache, v_cache, kv_indptr, kv_indices, last_page_len, ws, ps, sm_scale)
            c.record(stream)
        e_last.record(stream)
        barrier()
        if graph is None:
            t_first, t_last = e_first, e_last
    launches = pkg.ops.launch_count() if graph is None else launches_per_step * args.steps
    total_ms = t_first.elapsed_time(t_last)
    eager_ms_per_step = e_first.elapsed_time(e_last) / args.steps
    decode_ms = [b.elapsed_time(c) for _, b, c in ev]
    # The dominant kernel (decode + its merge) in the SAME timing mode as `value`: its own CUDA graph, one event pair
    # per replay, median over the K replays (eager intervals include dispatch gaps and, rarely, a stray multi-ms
    # outlier: SCALE_r01 N=1 printed a 3.1 ms mean inside a 1.28 ms step).
    paired = None
    if graph is not None:
        gdec = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gdec):
            out_dec = pkg.ops.decode(q, k_cache, v_cache, kv_indptr, kv_indices, last_page_len, ws, ps, sm_scale)
        for _ in range(3):
            gdec.replay()
        barrier()
        # PAIRED pass: step graph and decode-only graph replayed alternately, one event per boundary, so both see the same
        # clocks / power state (two separate back-to-back passes drift by ~1 % under sw_power_cap: BENCH_r01 printed a
        # kernel 2 % longer than the step that contains it).  The kernel's SHARE of the step comes from the medians of
        # this pass; roofline.kernel_ms = share x the timed ms_per_step, so it cannot exceed the step by construction.
        n_pair = max(args.steps, 20)
        pe = [torch.cuda.Event(enable_timing=True) for _ in range(2 * n_pair + 1)]
        pe[0].record(stream)
        for i in range(n_pair):
            graph.replay()
            pe[2 * i + 1].record(stream)
            gdec.replay()
            pe[2 * i + 2].record(stream)
        barrier()
        step_i = [pe[2 * i].elapsed_time(pe[2 * i + 1]) for i in range(n_pair)]
        dec_i = [pe[2 * i + 1].elapsed_time(pe[2 * i + 2]) for i in range(n_pair)]
        paired = {"step_ms_median": statistics.median(step_i), "decode_ms_median": statistics.median(dec_i), "pairs": n_pair}
        paired["decode_share_of_step"] = min(1.0, paired["decode_ms_median"] / paired["step_ms_median"])
        kernel_ms_from = ("share of the step from a paired pass (step graph / decode-only graph replayed alternately, medians) "
                          "x the timed ms_per_step")
        dec_kernel_ms = None      # filled in below, once the max-over-ranks step time is known
    else:
        dec_kernel_ms = statistics.median(decode_ms)
        kernel_ms_from = "eager pass with CUDA events around the decode launches, median"
    t = torch.tensor([total_ms], device=devc, dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms_max = float(t.item())
    ms_per_step = total_ms_max / args.steps
    value = bytes_rank * tp / (ms_per_step * 1e-3) / 1e9
    if dec_kernel_ms is None:
        dec_kernel_ms = paired["decode_share_of_step"] * (total_ms / args.steps)      # this rank's own step time
    dec_kernel_ms = min(dec_kernel_ms, ms_per_step)
    dec_ms_avg = dec_kernel_ms
    achieved = bytes_rank / (dec_ms_avg * 1e-3) / 1e9
    peak, peak_src = measured_peak()
    # quick self-check of the timed output against a direct fp32 computation of one (seq, head)
    b_chk, h_chk = B // 2, Hq - 1
    pages = kv_indices[b_chk * pages_per_seq:(b_chk + 1) * pages_per_seq].long()
    slots = (pages[:, None] * ps + torch.arange(ps, device=devc)[None]).view(-1)[:L]
    Kc = k_cache[slots, h_chk // (Hq // Hkv)].float()
    Vc = v_cache[slots, h_chk // (Hq // Hkv)].float()
    ref = torch.softmax((Kc @ q[b_chk, h_chk].float()) * sm_scale, 0) @ Vc
    err = float((out[b_chk, h_chk].float() - ref).abs().max().item())

    # ---- (B) end to end through the drop-in API with host buffers -------------------------------------
    e2e = e2e_sync = None
    if not args.no_e2e:
        be = pkg.AttentionBackend(pool, Hq, Hkv, D, torch.bfloat16, devc, page_size=ps)
        attn = pkg.Attention(Hq, D, None, Hkv, 0).to(devc)
        holder = torch.nn.ModuleList([attn])
        seqs, spare = [], []
        for i in range(B):
            pg = perm_np[i * pages_per_seq:(i + 1) * pages_per_seq].astype(np.int64)
            kv = (pg[:, None] * ps + np.arange(ps)[None]).reshape(-1)[:L].tolist()
            seqs.append(Seq(f"r{i}", L, kv, L - 1))
            first = (n_pages + i * grow_pages) * ps          # this sequence's growth pages: consecutive, page aligned
            spare.append(list(range(first + grow_pages * ps - 1, first - 1, -1)))
        if L % ps:                                           # growth continues inside the last page first
            for i, sq in enumerate(seqs):
                spare[i] += list(range(sq.kv_indices[-1] + ps - (L % ps), sq.kv_indices[-1], -1))
        batch = Batch(pkg.ForwardMode.DECODE, seqs)
        e2e_steps_done = [0]
        h_qkv = qkv.cpu().pin_memory()
        h_out = torch.empty((B, Hq * D), dtype=torch.bfloat16).pin_memory()
        d_qkv = torch.empty_like(qkv)

        # The decode step of the reference engine is a CUDA-graph replay around build_metadata_for_cuda_graph_replay
        # (core/model_runner.py:299-305, 332-417); the e2e leg drives the drop-in through that same public API:
        # prepare_for_cuda_graph_capture -> build_metadata_for_cuda_graph_capture -> capture, then per step
        # build_metadata_for_cuda_graph_replay + replay.  --no-graph uses the eager build_metadata path instead.
        e2e_graph = None
        static_out = torch.empty((B, Hq * D), dtype=torch.bfloat16, device=devc)

        def fwd_static():
            o = attn(d_qkv[:, :Hq * D], d_qkv[:, Hq * D:(Hq + Hkv) * D], d_qkv[:, (Hq + Hkv) * D:])
            static_out.copy_(o)

        if not args.no_graph:
            e2e_graph = pkg.CUDAGraph()
            ctx_cap = L + 2 * args.steps + max(args.warmup, 3) + 8
            be.prepare_for_cuda_graph_capture(e2e_graph, B, ctx_cap)
            md0 = be.build_metadata_for_cuda_graph_capture(e2e_graph, B, ctx_cap)
            with pkg.attention_kv_cache(holder, md0):
                fwd_static()
                with e2e_graph.capture(B):
                    fwd_static()
            torch.cuda.synchronize()

        def step_e2e():
            # what the scheduler does between two decode steps (core/scheduler.py:119-128, 161-171): the token of the last
            # step is now cached, one more token and one more slot per sequence
            if e2e_steps_done[0]:
                for i, sq in enumerate(seqs):
                    sq.cached_kv_len = len(sq.kv_indices)
                    sq.token_ids.append(0)
                    sq.kv_indices.append(spare[i].pop())
            e2e_steps_done[0] += 1
            d_qkv.copy_(h_qkv, non_blocking=True)
            if e2e_graph is not None:
                md = be.build_metadata_for_cuda_graph_replay(e2e_graph, batch, B)
                with pkg.attention_kv_cache(holder, md):
                    e2e_graph.replay(B)
                h_out.copy_(static_out, non_blocking=True)
            else:
                md = be.build_metadata(batch)
                with pkg.attention_kv_cache(holder, md):
                    o = attn(d_qkv[:, :Hq * D], d_qkv[:, Hq * D:(Hq + Hkv) * D], d_qkv[:, (Hq + Hkv) * D:])
                h_out.copy_(o, non_blocking=True)
            return be.slot_table.last_msg_bytes

        for _ in range(max(args.warmup, 3)):
            step_e2e()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        resends0 = be.slot_table.full_resends
        e0.record(stream)
        msg_bytes = 0
        for _ in range(args.steps):
            msg_bytes = step_e2e()
        e1.record(stream)
        barrier()
        t = torch.tensor([e0.elapsed_time(e1)], device=devc, dtype=torch.float64)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item()) / args.steps
        L_avg = L + e2e_steps_done[0] - args.steps / 2.0       # context grows by one token per step
        e2e_bytes = algorithmic_bytes(B, L_avg, Hq, Hkv, D, ps)
        e2e = {"value": e2e_bytes * tp / (e2e_ms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": e2e_ms,
               "progression": f"real decode steps: context {L} -> {L + e2e_steps_done[0] - 1} tokens, one new slot per sequence per step",
               "h2d_bytes_per_step": int(h_qkv.numel() * 2 + msg_bytes), "d2h_bytes_per_step": int(h_out.numel() * 2),
               "api": ("AttentionBackend.build_metadata_for_cuda_graph_replay(host Sequence lists) + CUDAGraph.replay of "
                       "Attention.forward (the reference's decode path, core/model_runner.py:299-305) + output read-back")
               if e2e_graph is not None else
               "AttentionBackend.build_metadata(host Sequence lists) + Attention.forward + output read-back"}
        torch.cuda.synchronize()
        # The same steps once more, but as the reference's decode loop runs them: the host WAITS for a step's output
        # before it advances the sequences (core/worker.py:104 `.tolist()` every step), so packing step n+1 cannot
        # overlap the device work of step n.  `e2e` above is the pipelined upper bound; this is the lock-step figure.
        barrier()
        t0 = time.perf_counter()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record(stream)
        for _ in range(args.steps):
            step_e2e()
            stream.synchronize()                 # h_out (pinned host memory) is complete: the step's result is on the host
        s1.record(stream)
        s1.synchronize()
        wall_ms = (time.perf_counter() - t0) * 1e3
        t = torch.tensor([max(s0.elapsed_time(s1), wall_ms)], device=devc, dtype=torch.float64)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        sync_ms = float(t.item()) / args.steps
        L_sync = L + e2e_steps_done[0] - args.steps / 2.0
        e2e_sync = {"value": algorithmic_bytes(B, L_sync, Hq, Hkv, D, ps) * tp / (sync_ms * 1e-3) / 1e9, "unit": UNIT,
                    "ms_per_step": sync_ms, "h2d_bytes_per_step": e2e["h2d_bytes_per_step"],
                    "d2h_bytes_per_step": e2e["d2h_bytes_per_step"],
                    "note": "host waits for each step's output before advancing the sequences (core/worker.py:104); "
                            "max of device-event and host wall time over the K steps"}
        sl_e = torch.tensor(seqs[b_chk].kv_indices, dtype=torch.long, device=devc)
        Ke, Ve = k_cache[sl_e, h_chk // (Hq // Hkv)].float(), v_cache[sl_e, h_chk // (Hq // Hkv)].float()
        qe = d_qkv[b_chk, h_chk * D:(h_chk + 1) * D].float()
        ref_e = torch.softmax((Ke @ qe) * sm_scale, 0) @ Ve
        e2e["self_check_max_abs_err"] = float((h_out.view(B, Hq, D)[b_chk, h_chk].to(devc).float() - ref_e).abs().max())
        e2e["full_resends_in_timed_steps"] = int(be.slot_table.full_resends - resends0)
        del be, attn, holder, seqs, batch, d_qkv
        e2e_graph = static_out = None

    # ---- (C) CPU baseline (rank 0, N=1 only) -----------------------------------------------------------
    cpu = None
    if rank == 0 and tp == 1 and not args.no_cpu_baseline:
        val, dt, threads, n_pass = cpu_port_gbps(min(args.cpu_sample_seqs, B), L, Hq, Hkv, D, ps, steps=None, warmup=1)
        cpu = {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"{min(args.cpu_sample_seqs, B)} of {B} sequences x {L} tokens" +
                         (" (the whole workload)" if args.cpu_sample_seqs >= B else " (same shapes)") +
                         f", {n_pass} passes of {dt:.2f} s"}

    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "decode_traffic.json")) as f:
            cap = json.load(f)
        # a static ncu figure: only quoted when this run launches the kernel on exactly the captured shapes
        if tp == 1 and cap.get("algorithmic_bytes_per_launch") == algorithmic_bytes(B, L, Hq, Hkv, D, ps):
            traffic = cap.get("dram_bytes_per_launch")
    except Exception:
        pass

    timed_region = "eager launches" if graph is None else "CUDA-graph replays of the step (as the reference replays decode)"

    # ---- (D) extra keys: prefill roofline, FlashInfer on the same inputs, the stock engine ---------------------
    prefill = gpu_ref = engine = None
    m3 = {}
    if rank == 0 and not args.no_m3:
        for key, fn in (("decode_page1", lambda: bench_decode_page1(pkg, B, L, Hq, Hkv, D, q, k_cache, v_cache, devc, stream, peak)),
                        ("store_kv", lambda: bench_store_kv(pkg, Hkv, D, devc, stream, peak)),
                        ("meta", lambda: bench_meta(pkg, Hq, Hkv, D, devc, stream))):
            try:
                m3[key] = fn()
            except Exception as e:
                m3[key] = {"error": f"{type(e).__name__}: {e}"[:400]}
    if rank == 0 and not args.no_prefill:
        try:
            prefill = bench_prefill(pkg, Hq, Hkv, D, devc, stream)
        except Exception as e:
            prefill = {"error": f"{type(e).__name__}: {e}"[:400]}
    if rank == 0 and not args.no_gpu_reference:
        try:
            gpu_ref = bench_flashinfer(B, L, ps, Hq, Hkv, D, q, k_cache, v_cache, kv_indptr, kv_indices, last_page_len, bytes_rank,
                                       devc, stream, with_prefill=not args.no_prefill)
            o_fi = gpu_ref["decode"].pop("out")
            gpu_ref["decode"]["max_abs_diff_vs_b200_all_rows"] = float((o_fi.float() - out.float()).abs().max())
            gpu_ref["decode"]["b200_over_flashinfer"] = gpu_ref["decode"]["kernel_ms"] / dec_ms_avg
            if prefill and "cases" in prefill:
                for mine, theirs in zip(prefill["cases"], gpu_ref.get("prefill", [])):
                    theirs["b200_over_flashinfer"] = theirs["kernel_ms"] / mine["kernel_ms"]
            del o_fi
        except Exception as e:
            gpu_ref = {"error": f"{type(e).__name__}: {e}"[:400]}
    if not args.no_engine:
        # every rank lets go of its device memory; rank 0 runs the engine (its own processes, tp = N) while the others
        # wait on the host
        del pool, k_cache, v_cache, ws
        graph = gdec = None
        import gc
        gc.collect()
        torch.cuda.empty_cache()
        if dist is not None:
            dist.barrier(group=host_group)
        if rank == 0:
            engine = run_engine_arms(args, tp)
        if dist is not None:
            dist.barrier(group=host_group)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": tp, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic", "config": workload_config(args, Hq, Hkv, D),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": "static: ncu --set full capture of this kernel at these shapes "
                         "(profiles/decode_traffic.json), not measured in this run", "peak_source": peak_src,
                         "kernel": "decode_kernel + decode_merge_kernel", "kernel_ms": dec_ms_avg,
                         "kernel_ms_le_ms_per_step": bool(dec_ms_avg <= ms_per_step),
                         "frac_note": "peak = read+write copy bandwidth (MEASURED_PEAKS.json); this kernel only reads, so a "
                                      "fraction slightly above 1 is not an error",
                         "algorithmic_bytes_per_launch": bytes_rank, "frac_of_nominal_8TBs": achieved / NOMINAL_HBM_GBS},
            "prefill": prefill, "gpu_reference": gpu_ref, "engine": engine,
            "decode_page1": m3.get("decode_page1"), "store_kv": m3.get("store_kv"), "meta": m3.get("meta"),
            "cpu_baseline": cpu, "e2e": e2e, "e2e_sync": e2e_sync, "gpu_launches": int(launches), "clocks": clk.summary(),
            "timing": {"timed_region": timed_region,
                       "eager_ms_per_step": eager_ms_per_step, "roofline_kernel_ms_from": kernel_ms_from,
                       "eager_decode_ms_median": statistics.median(decode_ms), "paired_pass": paired},
            "self_check_max_abs_err": err,
        }
        print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
