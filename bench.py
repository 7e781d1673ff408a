#!/usr/bin/env python
"""bench.py — the hot-path benchmark contract.

Workload (BASELINE.json configs[4], the configuration the headline metric "paged decode-attn HBM
GB/s" is quoted on; it fits one GPU): Qwen3-8B attention shapes, ONE layer, batch 64, 32 768-token
context per sequence, page size 16, bf16, page table = seeded random permutation of the pool's
pages (fragmented), Hq/Hkv = 32/8 at TP1 and 32/N / 8/N per rank at TP N (KV-head sharding, no
collective on the path).  One "step" = one decode pass of the hot path over the
batch: slot write of the 64 new K/V rows + split-KV paged decode attention (+ merge).

  value     algorithmic GB/s (SURVEY.md 8d formula) with everything resident in HBM, whole step
  e2e       same metric through the public drop-in API (AttentionBackend.build_metadata from host
            Sequence lists -> Attention.forward) with q/k/v in pinned HOST memory and the output
            read back to the host every step
  roofline  the dominant kernel (decode streaming kernel + its merge) timed alone with CUDA events
  cpu_baseline  the oracle's fp32 torch restatement of the reference semantics on the host cores,
            on a bounded sample (a few sequences of the same workload)

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
    ap.add_argument("--cpu-sample-seqs", type=int, default=8)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
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
    k_cache = torch.randn((n_pages * ps, Hkv, D), generator=g, dtype=torch.bfloat16)
    v_cache = torch.randn((n_pages * ps, Hkv, D), generator=g, dtype=torch.bfloat16)
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
    val, dt, threads, _ = cpu_port_gbps(args.cpu_sample_seqs, args.ctx, Hq, Hkv, D, args.page_size,
                                        steps=max(args.steps, 1), warmup=args.warmup)
    sample = f"{args.cpu_sample_seqs} of {args.batch} sequences x {args.ctx} tokens, same shapes, per step"
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
    tp = max(world, 1)
    assert args.gpus == tp, f"--gpus {args.gpus} but WORLD_SIZE {world}"
    Hq, Hkv, D = 32 // tp, max(1, 8 // tp), 128
    B, L, ps = args.batch, args.ctx, args.page_size
    devc = torch.device("cuda", local)
    pages_per_seq = -(-L // ps)
    n_pages = B * pages_per_seq
    num_slots = n_pages * ps

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
    t = torch.tensor([total_ms], device=devc, dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms_max = float(t.item())
    ms_per_step = total_ms_max / args.steps
    value = bytes_rank * tp / (ms_per_step * 1e-3) / 1e9
    dec_ms_avg = sum(decode_ms) / len(decode_ms)
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
    e2e = None
    if not args.no_e2e:
        be = pkg.AttentionBackend(pool, Hq, Hkv, D, torch.bfloat16, devc, page_size=ps)
        attn = pkg.Attention(Hq, D, None, Hkv, 0).to(devc)
        holder = torch.nn.ModuleList([attn])
        seqs = []
        for i in range(B):
            pg = perm_np[i * pages_per_seq:(i + 1) * pages_per_seq].astype(np.int64)
            kv = (pg[:, None] * ps + np.arange(ps)[None]).reshape(-1)[:L].tolist()
            seqs.append(Seq(f"r{i}", L, kv, L - 1))
        batch = Batch(pkg.ForwardMode.DECODE, seqs)
        h_qkv = qkv.cpu().pin_memory()
        h_out = torch.empty((B, Hq * D), dtype=torch.bfloat16).pin_memory()
        d_qkv = torch.empty_like(qkv)

        def step_e2e():
            d_qkv.copy_(h_qkv, non_blocking=True)
            md = be.build_metadata(batch)
            with pkg.attention_kv_cache(holder, md):
                o = attn(d_qkv[:, :Hq * D], d_qkv[:, Hq * D:(Hq + Hkv) * D], d_qkv[:, (Hq + Hkv) * D:])
            h_out.copy_(o, non_blocking=True)
            return be.slot_table.last_msg_bytes

        for _ in range(max(args.warmup, 3)):
            step_e2e()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
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
        e2e = {"value": bytes_rank * tp / (e2e_ms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": e2e_ms,
               "h2d_bytes_per_step": int(h_qkv.numel() * 2 + msg_bytes), "d2h_bytes_per_step": int(h_out.numel() * 2),
               "api": "AttentionBackend.build_metadata(host Sequence lists) + Attention.forward + output read-back"}
        ok = torch.equal(h_out.view(B, Hq, D)[b_chk, h_chk].to(devc), out[b_chk, h_chk])
        e2e["matches_device_run"] = bool(ok)

    # ---- (C) CPU baseline (rank 0, N=1 only) -----------------------------------------------------------
    cpu = None
    if rank == 0 and tp == 1 and not args.no_cpu_baseline:
        val, dt, threads, n_pass = cpu_port_gbps(args.cpu_sample_seqs, L, Hq, Hkv, D, ps, steps=None, warmup=1)
        cpu = {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
               "sample": f"{args.cpu_sample_seqs} of {B} sequences x {L} tokens (same shapes), {n_pass} passes of {dt:.2f} s"}

    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "decode_traffic.json")) as f:
            traffic = json.load(f).get("dram_bytes_per_launch") if tp == 1 else None   # captured at TP1 shapes
    except Exception:
        pass

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": tp, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic", "config": workload_config(args, Hq, Hkv, D),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "peak_source": peak_src, "kernel": "decode_kernel + decode_merge_kernel",
                         "kernel_ms": dec_ms_avg, "algorithmic_bytes_per_launch": bytes_rank,
                         "frac_of_nominal_8TBs": achieved / NOMINAL_HBM_GBS},
            "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clk.summary(),
            "timing": {"timed_region": "eager launches" if graph is None else "CUDA-graph replays of the step (as the reference replays decode)",
                       "eager_ms_per_step": eager_ms_per_step, "roofline_kernel_ms_from": "eager pass with CUDA events around the decode launches"},
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
