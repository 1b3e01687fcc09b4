#!/usr/bin/env python
"""bench.py — HNSW QPS at recall@10 >= 0.95 on the SIFT1M-shape config (BASELINE.json configs[1]).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

One "step" = one batched hnsw_search pass over a 10k-query batch (k=10) on the 1M×128 L2,
M=16, ef_construction=200 graph, at the smallest ef of the sweep {64,96,128,192,256} whose
recall@10 against exact k-NN is >= 0.95 (measured before the timed region, reported in
`config`). Under torchrun every rank holds a replica of the graph and searches its own 10k
batch (no data-path collective: weak scaling); `value` = total queries / max-over-ranks time.

JSON keys beyond the base contract: `roofline` (algorithmic bytes of the search kernel ÷ its
CUDA-event duration vs the measured HBM peak), `cpu_baseline` (the reference's own hnsw_search
on the host cores, bounded sample), `e2e` (same metric through muninn_cuda_search_batch with
pinned HOST buffers, copies inside the timed region), `gpu_launches`, `clocks`.

`--impl reference` times the reference's CPU hnsw_search (oracle/_ref when it was built, else
the oracle port) on all host threads over bounded samples of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC_NAME = "HNSW QPS at recall@10>=0.95"
EF_SWEEP = (64, 96, 128, 192, 256)

# BASELINE.json `configs` (1-based like SURVEY.md §8d). Config 2 is the bench line the driver reads;
# 3-5 are run by hand (`--config N`) and their lines are kept under profiles/. efc of configs 3/4 is
# not fixed by BASELINE.json: the values below are the ones used.
CONFIGS = {
    2: dict(name="SIFT1M-shape", n=1_000_000, dim=128, metric=0, M=16, efc=200, nq=10_000, k=10, ef_sweep=EF_SWEEP,
            data="latent16", seeds=(11, 12), scaling="weak", ef_calibrated=64),
    3: dict(name="text-embedding shape", n=1_000_000, dim=768, metric=1, M=32, efc=100, nq=10_000, k=10, ef_sweep=(128, 192, 256, 384),
            data="spectrum", seeds=(21, 22), scaling="strong", ef_calibrated=128),
    4: dict(name="Deep10M-shape", n=10_000_000, dim=96, metric=2, M=32, efc=64, nq=10_000, k=10, ef_sweep=(64, 128, 256),
            data="latent16n", seeds=(31, 32), scaling="strong", ef_calibrated=64),
    5: dict(name="exact brute-force kNN", n=6_250_000, dim=128, metric=0, M=8, efc=0, nq=10_000, k=100, ef_sweep=(),
            data="latent16", seeds=(41, 42), scaling="weak", ef_calibrated=0),
}
METRIC_TEXT = {0: "L2", 1: "cosine", 2: "inner product"}
DATA_TEXT = {"latent16": "rank-16 Gaussian latent embedded in dim-d + 0.05 isotropic noise",
             "latent16n": "rank-16 Gaussian latent embedded in dim-d + 0.05 isotropic noise, rows L2-normalised",
             "spectrum": "Gaussian with eigenvalue spectrum i^-1.5 (text-embedding-like anisotropy), rows L2-normalised"}
CACHE_DIR = os.environ.get("MUNINN_BENCH_CACHE", os.path.join(tempfile.gettempdir(), "muninn_bench_cache"))


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ───────────────────────── workload ─────────────────────────

def make_vectors(n: int, dim: int, seed: int, latent: int = 16, noise: float = 0.05, kind: str = "latent16") -> np.ndarray:
    """Synthetic rows. "latent16": a `latent`-dimensional Gaussian embedded linearly in `dim`
    dimensions plus isotropic noise (SIFT-like low intrinsic dimension, no cluster structure);
    i.i.d. uniform data in 128-d cannot reach recall 0.95 with the reference's closest-M neighbour
    selection at any ef <= 256 (SURVEY.md finding 8; calibrated on the CPU oracle).
    "latent16n": the same, rows L2-normalised (Deep1B-like, for inner product). "spectrum":
    independent Gaussian coordinates with variances i^-1.5, rows L2-normalised (text-embedding-like)."""
    A = (np.random.default_rng(999).standard_normal((latent, dim)) / np.sqrt(latent)).astype(np.float32)
    scale = (np.arange(1, dim + 1, dtype=np.float32)) ** np.float32(-0.75)
    rng = np.random.default_rng(seed)
    out = np.empty((n, dim), np.float32)
    step = 1 << 17
    for i in range(0, n, step):
        m = min(step, n - i)
        if kind == "spectrum":
            blk = rng.standard_normal((m, dim)).astype(np.float32) * scale
        else:
            z = rng.standard_normal((m, latent)).astype(np.float32)
            blk = z @ A + np.float32(noise) * rng.standard_normal((m, dim)).astype(np.float32)
        if kind != "latent16":
            blk /= np.linalg.norm(blk, axis=1, keepdims=True)
        out[i:i + m] = blk
    return out


def graph_tag(args) -> str:
    return (f"n{args.n}_d{args.dim}_m{args.metric}_M{args.M}_efc{args.efc}_b{args.build_batch}_{args.data}"
            f"_s{args.seeds[0]}")


def get_graph(args, rank: int, world: int, barrier):
    """Rank 0 builds (or loads) the graph; the other ranks load the cached copy."""
    from sqlite_muninn_b200.builder import build_graph
    from sqlite_muninn_b200.graph import FlatGraph
    os.makedirs(CACHE_DIR, exist_ok=True)
    tag = graph_tag(args)
    path = os.path.join(CACHE_DIR, tag + ".npz")
    t0 = time.time()
    if rank == 0 and not os.path.exists(path):
        base = make_vectors(args.n, args.dim, args.seeds[0], kind=args.data)
        log(f"[bench] building {tag} with {os.cpu_count()} host threads …")
        if args.device_build and args.impl != "reference":
            # same group-synchronous rule with the layer-0 candidate searches on the GPU (builder.build_graph_device):
            # the same graph wherever no two candidate distances are bit-equal
            from sqlite_muninn_b200.builder import build_graph_device
            st = {}
            g = build_graph_device(base, args.metric, args.M, args.efc, batch=args.build_batch, warm=8192, stats=st,
                                   device=int(os.environ.get("LOCAL_RANK", "0")))
            log(f"[bench] device-assisted build: {st}")
        else:
            g = build_graph(base, args.metric, args.M, args.efc, batch=args.build_batch, warm=8192)
        tmp = path + ".tmp.npz"
        np.savez(tmp, levels=g.levels, list_offsets=g.list_offsets, targets=g.targets, entry_slot=g.entry_slot,
                 max_level=g.max_level)
        os.replace(tmp, path)
        log(f"[bench] graph built in {time.time() - t0:.1f}s")
    barrier()
    base = make_vectors(args.n, args.dim, args.seeds[0], kind=args.data)
    z = np.load(path)
    g = FlatGraph(args.dim, args.metric, args.M, args.efc, np.arange(args.n, dtype=np.int64), z["levels"],
                  np.zeros(args.n, np.uint8), base, z["list_offsets"], z["targets"], int(z["entry_slot"]),
                  int(z["max_level"]))
    log(f"[bench] rank {rank}: graph ready in {time.time() - t0:.1f}s (n={g.n}, max_level={g.max_level})")
    return g


# ───────────────────────── clocks ─────────────────────────

class ClockSampler:
    """nvidia-smi sampled every 200 ms during the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc:
            time.sleep(0.25)
            self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i] == "Active"})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ───────────────────────── CPU reference leg ─────────────────────────

class CpuReference:
    """The reference's own hnsw_search on the host cores: oracle/_ref/libmuninn_ref.so (the
    unmodified sources, kind "reference") when it was built, else oracle/liboracle.so (the
    restatement with an O(1) visited set, kind "port"). Queries are partitioned over all host
    threads; each thread calls the search on the shared read-only index."""

    def __init__(self, g):
        from oracle.bindings import Oracle, Reference, have_reference
        self.threads = os.cpu_count() or 1
        t0 = time.time()
        if have_reference():
            self.kind = "reference"
            self.idx = Reference().load(g)
            self._run = lambda q, k, ef: self.idx.search_mt(q, k, ef, self.threads)
        else:
            self.kind = "port"
            self.idx = Oracle().graph(g)
            self._run = lambda q, k, ef: (lambda r: (r[0], r[1], r[2], r[4]))(
                self.idx.search(q, k, ef, threads=self.threads, want_stats=False))
        log(f"[bench] CPU {self.kind} index loaded in {time.time() - t0:.1f}s, {self.threads} threads")

    def sample_size(self, queries, k, ef, target_s: float) -> int:
        probe = queries[: min(len(queries), self.threads * 2)]
        secs = self._run(probe, k, ef)[3]
        qps = len(probe) / max(secs, 1e-9)
        return int(min(len(queries), max(self.threads * 4, qps * target_s)))

    def run(self, queries, k, ef):
        ids, dist, cnt, secs = self._run(queries, k, ef)
        return ids, secs


def recall_at_k(ids: np.ndarray, truth: np.ndarray) -> float:
    k = truth.shape[1]
    return float(np.mean([len(set(a) & set(b)) / k for a, b in zip(ids, truth)]))


def run_reference_arm(args, rank: int, world: int) -> None:
    """`--impl reference`: rank 0 alone times the CPU path; the other ranks exit 0."""
    if rank != 0:
        return
    g = get_graph(args, 0, 1, lambda: None)
    queries = make_vectors(args.nq, args.dim, args.seeds[1], kind=args.data)
    cpu = CpuReference(g)
    # the ef the GPU arm times: its cached choice when it ran first on this box, else the calibrated
    # value (the graph and the queries are deterministic, so the recall sweep always picks the same ef)
    ef = args.ef or int(os.environ.get("MUNINN_BENCH_EF", "0")) or load_chosen_ef(args) or CONFIGS[args.config]["ef_calibrated"]
    n_s = cpu.sample_size(queries, args.k, ef, args.cpu_seconds / max(1, args.steps + args.warmup) * 2)
    sample = queries[:n_s]
    for _ in range(args.warmup):
        cpu.run(sample[: max(cpu.threads * 2, n_s // 8)], args.k, ef)
    secs = 0.0
    for _ in range(args.steps):
        secs += cpu.run(sample, args.k, ef)[1]
    qps = args.steps * n_s / secs
    desc = {"value": qps, "unit": "queries/s", "cores": cpu.threads, "kind": cpu.kind,
            "sample": f"{n_s} of the {args.nq} queries per step, ef={ef}, k={args.k}, {cpu.threads} host threads"}
    print(json.dumps({
        "impl": "reference", "metric": METRIC_NAME, "value": qps, "unit": "queries/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1000.0 * secs / args.steps,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, ef, None), "cpu_baseline": desc,
        "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}), flush=True)


def chosen_ef_path(args) -> str:
    return os.path.join(CACHE_DIR, "ef_" + graph_tag(args) + ".json")


def load_chosen_ef(args):
    try:
        return int(json.load(open(chosen_ef_path(args)))["ef"])
    except Exception:
        return None


def workload_config(args, ef, recall):
    slab_mb, list_mb = args.n * args.dim * 4 / 1e6, args.n * 2 * args.M * 4 / 1e6
    per_gpu = (f"one {args.nq}-query batch per GPU" if args.scaling == "weak"
               else f"the {args.nq}-query batch split evenly across the GPUs")
    return {"workload": f"{args.cfg_name} {args.n}x{args.dim} fp32 {METRIC_TEXT[args.metric]}, M={args.M}, "
                        f"ef_construction={args.efc}, {args.nq}-query batch, k={args.k} "
                        f"(BASELINE.json configs[{args.config - 1}])",
            "ef": ef, "ef_sweep": list(args.ef_sweep), "recall_at_10": recall,
            "data_distribution": f"{DATA_TEXT[args.data]}, seeds {args.seeds[0]}/{args.seeds[1]}",
            "graph": f"reference construction rule, batch-synchronous inserts (groups of {args.build_batch}), "
                     "neighbour lists in shadow-table order",
            "l2_policy": f"inputs ({slab_mb:.0f} MB slab + {list_mb:.0f} MB lists) exceed the 126 MB L2; no flush between steps",
            "parallelism": f"graph replicated, {per_gpu}, no collective"}


# ───────────────────────── GPU arm ─────────────────────────

def algorithmic_bytes(stats: np.ndarray, dim: int, M: int, k: int) -> float:
    """B(q) = n_dist·d·4 (rows gathered) + n_expand·M0·4 (layer-0 lists)
    + n_greedy_steps·M·4 (upper-layer lists) + d·4 (query) + k·12 (results). (The 4-byte row norm
    cosine also reads per distance is left out: < 0.2 % at d = 768.)"""
    nd = stats["n_dist"].astype(np.float64).sum()
    ne = stats["n_expand"].astype(np.float64).sum()
    ng = stats["n_greedy_steps"].astype(np.float64).sum()
    nq = len(stats)
    return float(nd * dim * 4 + ne * 2 * M * 4 + ng * M * 4 + nq * (dim * 4 + k * 12))


def measured_peak_tf32():
    """Dense TF32 runs at half the bf16 rate; the bf16 burst figure is the driver-measured one."""
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]) / 2.0, "measured bf16 burst / 2"
    except Exception:
        return 1590.0 / 2.0, "fallback bf16 / 2"


def exact_summary(args, nq: int, total_ms: float, st: dict, n_rows: int | None = None) -> dict:
    """Ground-truth pass: exact k-NN through the tensor-core filter. Tensor roofline = TF32 MMA flops
    the filter ISSUES (3 per product for 3xTF32) ÷ its own CUDA-event time; algorithmic = 2·nq·n·d."""
    n_rows = n_rows or args.n
    flops = 2.0 * nq * n_rows * args.dim
    issued = flops * {0: 0, 1: 3, 2: 1}[st["mode"]]
    peak, which = measured_peak_tf32()
    out = {"ms": round(total_ms, 3), "mode": {0: "cuda-core chains", 1: "3xTF32 filter + exact re-rank", 2: "1xTF32 filter + exact re-rank"}[st["mode"]],
           "filter_ms": round(st["filter_ms"], 3), "chunks": st["chunks"],
           "survivors_per_query": round(st["candidates"] / max(nq, 1), 1), "overflow_queries": st["overflow_queries"],
           "algorithmic_TFLOPs": round(flops / max(total_ms, 1e-9) / 1e9, 1)}
    if st["mode"] == 2 and st["filter_ms"] > 0:  # small batches: one TF32 product off the raw slab, HBM-bound
        hbm, hw = measured_peak_gbs()
        ach = n_rows * args.dim * 4.0 / st["filter_ms"] / 1e6
        out["roofline"] = {"bound": "hbm", "achieved": round(ach, 1), "peak": hbm, "unit": "GB/s", "frac": round(ach / hbm, 4),
                           "peak_source": hw, "kernel": "exact_tc_filter (base rows read once / its CUDA-event time)"}
    elif st["mode"] != 0 and st["filter_ms"] > 0:
        ach = issued / st["filter_ms"] / 1e9
        out["roofline"] = {"bound": "tensor", "achieved": round(ach, 1), "peak": round(peak, 1), "unit": "TFLOP/s",
                           "frac": round(ach / peak, 4), "peak_source": which, "kernel": "exact_tc_filter",
                           "of_nominal_1100": round(ach / 1100.0, 4)}
    return out


def ncu_traffic(args, ef):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE level-0 search launch, from the `ncu --set full` capture
    of this command summarised in profiles/r1_search_kernel_v7_ncu_full.txt (6.812 GB + 0.284 GB; algorithmic bytes
    of the same launch: 8.472 GB — the 58 MB visited arena leaves the rest of the 126 MB L2 to the hot rows and lists,
    so DRAM moves 0.84x of the algorithmic bytes). Only the captured configuration has a number."""
    if (args.config, args.n, args.nq, ef) == (2, 1_000_000, 10_000, 64):
        return 7.096e9
    return None


def measured_peak_gbs():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def run_gpu_arm(args, rank: int, local_rank: int, world: int) -> None:
    import torch
    import torch.distributed as dist
    from sqlite_muninn_b200.cuda import STATS_DTYPE, HnswIndexCuda

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()

    g = get_graph(args, rank, world, barrier)
    idx = HnswIndexCuda.from_graph(g, device=local_rank)
    stream = torch.cuda.Stream()  # a real stream: the legacy default stream (0) means "handle's own" to the ABI
    torch.cuda.set_stream(stream)
    idx.set_stream(stream.cuda_stream)
    if args.teams_per_sm:
        idx.set_tuning(1, args.teams_per_sm)
    k, dim = args.k, args.dim
    if args.scaling == "weak":  # every rank searches its own batch
        queries = make_vectors(args.nq, dim, args.seeds[1] + rank, kind=args.data)
    else:  # one batch, rank r takes queries [r·nq/W, (r+1)·nq/W)
        from sqlite_muninn_b200.multi import shard_range
        lo, hi = shard_range(args.nq, rank, world)
        queries = np.ascontiguousarray(make_vectors(args.nq, dim, args.seeds[1], kind=args.data)[lo:hi])
    nq = queries.shape[0]
    nq_job = world * nq if args.scaling == "weak" else args.nq
    h_q = torch.from_numpy(queries).pin_memory()
    d_q = h_q.cuda(non_blocking=True)
    d_ids = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    d_dist = torch.empty((nq, k), dtype=torch.float32, device="cuda")
    d_cnt = torch.empty(nq, dtype=torch.int32, device="cuda")
    d_stats = torch.empty(nq * STATS_DTYPE.itemsize, dtype=torch.uint8, device="cuda")

    # ground truth (exact k-NN on the GPU, bit-exact vs the CPU oracle in tests) and ef choice
    t_idx = torch.empty((nq, k), dtype=torch.int64, device="cuda")
    t_dist = torch.empty((nq, k), dtype=torch.float32, device="cuda")
    t_cnt = torch.empty(nq, dtype=torch.int32, device="cuda")
    idx.exact_knn_device(d_q.data_ptr(), nq, k, 0, t_idx.data_ptr(), t_dist.data_ptr(), t_cnt.data_ptr())
    torch.cuda.synchronize()
    exact_ms = idx.last_kernel_ms()
    exact_st = idx.exact_stats()
    truth = t_idx.cpu().numpy()

    def search_dev(ef, stats=False):
        idx.search_device(d_q.data_ptr(), nq, k, ef, d_ids.data_ptr(), d_dist.data_ptr(), d_cnt.data_ptr(),
                          d_stats.data_ptr() if stats else None)

    sweep = []
    chosen = args.ef or None
    for ef in ([args.ef] if args.ef else args.ef_sweep):
        search_dev(ef, stats=True)
        torch.cuda.synchronize()
        ms = idx.last_kernel_ms()
        rec = recall_at_k(d_ids.cpu().numpy(), truth)
        st = d_stats.cpu().numpy().view(STATS_DTYPE)
        sweep.append({"ef": ef, "recall_at_10": round(rec, 4), "kernel_ms": round(ms, 3),
                      "qps_kernel": round(nq / ms * 1000.0), "n_dist_mean": float(st["n_dist"].mean()),
                      "n_expand_mean": float(st["n_expand"].mean()),
                      "algo_GBps": round(algorithmic_bytes(st, dim, args.M, k) / ms / 1e6, 1)})
        if chosen is None and rec >= 0.95:
            chosen = ef
    if rank == 0:
        log("[bench] ef sweep: " + json.dumps(sweep))
    if chosen is None:
        chosen = args.ef_sweep[-1]
        if rank == 0:
            log(f"[bench] NOTE: recall@10 < 0.95 at every ef of the sweep; timing ef={chosen}")
    if world > 1:  # all ranks time the same ef: rank 0's choice
        t = torch.tensor([chosen], device="cuda")
        dist.broadcast(t, 0)
        chosen = int(t.item())
    ef = chosen
    if rank == 0:
        json.dump({"ef": ef}, open(chosen_ef_path(args), "w"))
    search_dev(ef, stats=True)
    torch.cuda.synchronize()
    stats = d_stats.cpu().numpy().view(STATS_DTYPE).copy()
    recall = recall_at_k(d_ids.cpu().numpy(), truth)
    algo_bytes = algorithmic_bytes(stats, dim, args.M, k)

    # ── device-resident timing: W warm-up, K timed steps, events on the launching stream ──
    for _ in range(args.warmup):
        search_dev(ef)
    torch.cuda.synchronize()
    barrier()
    clocks = ClockSampler(local_rank)
    clocks.start()
    launches0 = idx.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # the search kernel's own launch duration, live, inside the timed region: one event pair per step on the
    # launching stream (recording is asynchronous; the pairs are read after the final synchronize). A pair spans the
    # level-0 launch and the two retry-level launches that exit at once.
    step_ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    torch.cuda.synchronize()
    e0.record(stream)
    for a, b in step_ev:
        a.record(stream)
        search_dev(ef)
        b.record(stream)
    e1.record(stream)
    torch.cuda.synchronize()
    barrier()
    total_ms = e0.elapsed_time(e1)
    launches = idx.launch_count() - launches0
    kernel_ms = [a.elapsed_time(b) for a, b in step_ev]
    # ── end-to-end: host (pinned) buffers through muninn_cuda_search_batch ──
    h_ids = torch.empty((nq, k), dtype=torch.int64).pin_memory()
    h_dist = torch.empty((nq, k), dtype=torch.float32).pin_memory()
    h_cnt = torch.empty(nq, dtype=torch.int32).pin_memory()
    for _ in range(max(1, args.warmup // 2)):
        idx.search_into(h_q.data_ptr(), nq, k, ef, h_ids.data_ptr(), h_dist.data_ptr(), h_cnt.data_ptr())
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        idx.search_into(h_q.data_ptr(), nq, k, ef, h_ids.data_ptr(), h_dist.data_ptr(), h_cnt.data_ptr())
    e2e_s = time.perf_counter() - t0
    clk = clocks.stop()
    assert np.array_equal(h_ids.numpy(), d_ids.cpu().numpy()), "host-buffer path and device path disagree"

    tmax = torch.tensor([total_ms, e2e_s * 1000.0], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms = float(tmax[0]), float(tmax[1])

    cpu_desc = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = CpuReference(g)
        n_s = cpu.sample_size(queries, k, ef, args.cpu_seconds)
        c_ids, secs = cpu.run(queries[:n_s], k, ef)
        same = float(np.mean((c_ids == h_ids.numpy()[:n_s]).all(axis=1)))
        cpu_desc = {"value": n_s / secs, "unit": "queries/s", "cores": cpu.threads, "kind": cpu.kind,
                    "sample": f"first {n_s} of the {nq} queries, ef={ef}, k={k}, {secs:.1f}s on {cpu.threads} host threads",
                    "id_exact_fraction_vs_gpu": same}
        # the same search with only the linear visited scan swapped for O(1) stamps
        # (oracle/hnsw_oracle.c, results bit-identical to the reference's) — the stronger CPU comparison
        from oracle.bindings import Oracle
        o_ids, _, _, _, o_secs = Oracle().graph(g).search(queries[:n_s], k, ef, threads=cpu.threads, want_stats=False)
        cpu_desc["modified_reference_o1_visited"] = {
            "value": n_s / max(o_secs, 1e-9), "unit": "queries/s", "cores": cpu.threads, "kind": "port",
            "ids_identical_to_reference": bool(np.array_equal(o_ids, c_ids))}
    if rank == 0:
        peak, which = measured_peak_gbs()
        k_ms = float(np.mean(kernel_ms))
        achieved = algo_bytes / k_ms / 1e6
        from sqlite_muninn_b200.cuda import gather_ceiling
        gather_peak = gather_ceiling(((dim + 3) // 4) * 16, 1 << 30, local_rank)  # measured live, after the timed region
        out = {
            "metric": METRIC_NAME, "value": nq_job * args.steps / total_ms * 1000.0, "unit": "queries/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, ef, round(recall, 4)),
            "e2e": {"value": nq_job * args.steps / e2e_ms * 1000.0, "unit": "queries/s",
                    "h2d_bytes_per_step": nq * dim * 4, "d2h_bytes_per_step": nq * k * 12 + nq * 4},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(achieved / peak, 4), "traffic": ncu_traffic(args, ef), "peak_source": which,
                         "kernel": f"hnsw_search_kernel<{METRIC_TEXT[args.metric]}>", "kernel_ms": round(k_ms, 3),
                         "of_nominal_8TBps": round(achieved / 8000.0, 4),
                         "random_row_gather_peak": {"GBps": round(gather_peak, 1), "frac": round(achieved / gather_peak, 4),
                                                    "what": f"independent random {((dim + 3) // 4) * 16}-byte row gathers from a 1 GiB slab, "
                                                            "measured in this run (muninn_cuda_gather_ceiling): the ceiling of "
                                                            "this access pattern, below the streaming-copy peak"},
                         "algorithmic_bytes_per_launch": algo_bytes,
                         "bytes_per_query": round(algo_bytes / nq, 1)},
            "cpu_baseline": cpu_desc, "clocks": clk, "ef_sweep": sweep,
            "exact_knn": exact_summary(args, nq, exact_ms, exact_st),
        }
        print(json.dumps(out), flush=True)
    idx.close()
    if world > 1:
        dist.destroy_process_group()


def run_exact_arm(args, rank: int, local_rank: int, world: int) -> None:
    """`--config 5`: exact brute-force k-NN, base rows sharded across the GPUs (args.n rows per GPU —
    8 GPUs make BASELINE.json's 50M), every GPU sees all queries; per step each rank scans its shard
    (tensor-core filter + exact re-rank), the per-shard top-k lists are exchanged with one NCCL
    all-gather per array and merged under (distance, row index) on every rank."""
    import torch
    import torch.distributed as dist
    from sqlite_muninn_b200.cuda import HnswIndexCuda
    from sqlite_muninn_b200.graph import FlatGraph
    from sqlite_muninn_b200.multi import allgather_topk_parts

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()

    n, dim, k, nq = args.n, args.dim, args.k, args.nq
    t0 = time.time()
    base = make_vectors(n, dim, args.seeds[0] + 100 * rank, kind=args.data)  # this rank's shard
    g = FlatGraph(dim, args.metric, args.M, 10, np.arange(n, dtype=np.int64), np.zeros(n, np.int32), np.zeros(n, np.uint8),
                  base, np.zeros(n + 1, np.int64), np.empty(0, np.int32), 0, 0)
    idx = HnswIndexCuda.from_graph(g, device=local_rank)
    log(f"[bench] rank {rank}: shard of {n} rows resident after {time.time() - t0:.1f}s")
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    idx.set_stream(stream.cuda_stream)
    if args.exact_mode is not None:
        idx.set_exact_mode(args.exact_mode, 0)
    queries = make_vectors(nq, dim, args.seeds[1], kind=args.data)
    h_q = torch.from_numpy(queries).pin_memory()
    d_q = h_q.cuda(non_blocking=True)
    mk = lambda *shape, dt: torch.empty(shape, dtype=dt, device="cuda")
    s_idx, s_dist, s_cnt = mk(nq, k, dt=torch.int64), mk(nq, k, dt=torch.float32), mk(nq, dt=torch.int32)
    o_idx, o_dist, o_cnt = mk(nq, k, dt=torch.int64), mk(nq, k, dt=torch.float32), mk(nq, dt=torch.int32)

    def step(dq):
        idx.exact_knn_device(dq.data_ptr(), nq, k, rank * n, s_idx.data_ptr(), s_dist.data_ptr(), s_cnt.data_ptr())
        if world == 1:
            return s_idx, s_dist, s_cnt
        pi, pd, pc = allgather_topk_parts(s_idx, s_dist, s_cnt)
        idx.topk_merge_device(world, nq, k, pi.data_ptr(), pd.data_ptr(), pc.data_ptr(), o_idx.data_ptr(), o_dist.data_ptr(),
                              o_cnt.data_ptr())
        return o_idx, o_dist, o_cnt

    for _ in range(args.warmup):
        step(d_q)
    torch.cuda.synchronize()
    st = idx.exact_stats()
    shard_ms = idx.last_kernel_ms()
    barrier()
    clocks = ClockSampler(local_rank)
    clocks.start()
    launches0 = idx.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step(d_q)
    e1.record(stream)
    torch.cuda.synchronize()
    barrier()
    total_ms = e0.elapsed_time(e1)
    launches = idx.launch_count() - launches0
    # end to end: pinned host queries in, merged lists out, every step
    h_idx = torch.empty((nq, k), dtype=torch.int64).pin_memory()
    h_dist = torch.empty((nq, k), dtype=torch.float32).pin_memory()
    d_q2 = torch.empty_like(d_q)
    barrier()
    t1 = time.perf_counter()
    for _ in range(args.steps):
        d_q2.copy_(h_q, non_blocking=True)
        ri, rd, _ = step(d_q2)
        h_idx.copy_(ri, non_blocking=True)
        h_dist.copy_(rd, non_blocking=True)
        torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t1
    clk = clocks.stop()
    tmax = torch.tensor([total_ms, e2e_s * 1000.0], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms = float(tmax[0]), float(tmax[1])

    cpu_desc = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle.bindings import Oracle
        threads = os.cpu_count() or 1
        n_s = max(threads, min(nq, int(args.cpu_seconds * threads * 2.0e9 / (3.0 * n * dim))))  # ~cpu_seconds of SSE chains
        c_idx, c_dist, c_cnt, secs = Oracle().exact_knn(base, queries[:n_s], k, args.metric, threads=threads)
        same = bool(np.array_equal(c_idx, h_idx.numpy()[:n_s]) and c_dist.tobytes() == h_dist.numpy()[:n_s].tobytes())
        cpu_desc = {"value": n_s / secs, "unit": "queries/s", "cores": threads, "kind": "port",
                    "sample": f"first {n_s} of the {nq} queries against the {n}-row shard, k={k}, {secs:.1f}s on {threads} host threads",
                    "bit_identical_to_gpu": same}
    if rank == 0:
        ex = exact_summary(args, nq, shard_ms, st, n_rows=n)
        roof = dict(ex.get("roofline") or {"bound": "tensor", "achieved": None, "peak": None, "unit": "TFLOP/s", "frac": None})
        roof["traffic"] = None
        out = {
            "metric": "exact k-NN queries/s (k=100, base sharded, top-k all-gather merge)",
            "value": nq * args.steps / total_ms * 1000.0, "unit": "queries/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32 (tf32x3 filter, f32 re-rank)", "data": "synthetic",
            "config": {"workload": f"exact brute-force kNN {world * n}x{dim} {METRIC_TEXT[args.metric]} ({n} rows per GPU x {world}), "
                                   f"k={k}, {nq}-query batch (BASELINE.json configs[4])",
                       "data_distribution": f"{DATA_TEXT[args.data]}, seeds {args.seeds[0]}+100*rank/{args.seeds[1]}",
                       "l2_policy": f"the {n * dim * 4 / 1e9:.1f} GB shard exceeds the 126 MB L2; no flush between steps",
                       "parallelism": "base rows sharded, every GPU sees all queries, one all-gather of the [nq x k] lists per step"},
            "e2e": {"value": nq * args.steps / e2e_ms * 1000.0, "unit": "queries/s", "h2d_bytes_per_step": nq * dim * 4,
                    "d2h_bytes_per_step": nq * k * 12},
            "gpu_launches": int(launches), "roofline": roof, "exact_knn": ex,
            "algorithmic_TFLOPs_job": round(2.0 * nq * world * n * dim * args.steps / total_ms / 1e9, 1),
            "cpu_baseline": cpu_desc, "clocks": clk,
        }
        print(json.dumps(out), flush=True)
    idx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", type=int, default=2, choices=sorted(CONFIGS),
                    help="BASELINE.json config (1-based); 2 = the headline bench line")
    ap.add_argument("--n", type=int, default=None, help="override the config's size (dry runs)")
    ap.add_argument("--dim", type=int, default=None)
    ap.add_argument("--M", type=int, default=None)
    ap.add_argument("--efc", type=int, default=None)
    ap.add_argument("--nq", type=int, default=None)
    ap.add_argument("--k", type=int, default=None)
    ap.add_argument("--ef", type=int, default=0, help="0 = smallest ef of the sweep with recall@10 >= 0.95")
    ap.add_argument("--build-batch", type=int, default=1024)
    ap.add_argument("--device-build", action="store_true",
                    help="build the input graph with builder.build_graph_device (outside the timed region either way)")
    ap.add_argument("--teams-per-sm", type=int, default=0)
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--per-step-kernel-time", action="store_true", help="kept for compatibility: per-step event pairs are always recorded")
    ap.add_argument("--exact-mode", type=int, default=None, help="config 5: 0 cuda-core, 1 3xTF32 (default), 2 1xTF32")
    args = ap.parse_args()
    cfg = CONFIGS[args.config]
    for key in ("n", "dim", "M", "efc", "nq", "k"):
        if getattr(args, key) is None:
            setattr(args, key, cfg[key])
    args.metric, args.data, args.seeds, args.scaling = cfg["metric"], cfg["data"], cfg["seeds"], cfg["scaling"]
    args.ef_sweep, args.cfg_name = cfg["ef_sweep"], cfg["name"]
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    import __graft_entry__ as ge
    if rank == 0:
        ge.build()
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
    elif args.config == 5:
        run_exact_arm(args, rank, local_rank, world)
    else:
        run_gpu_arm(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
