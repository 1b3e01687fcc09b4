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
CACHE_DIR = os.environ.get("MUNINN_BENCH_CACHE", os.path.join(tempfile.gettempdir(), "muninn_bench_cache"))


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ───────────────────────── workload ─────────────────────────

def make_vectors(n: int, dim: int, seed: int, latent: int = 16, noise: float = 0.05) -> np.ndarray:
    """SIFT-like synthetic rows: a `latent`-dimensional Gaussian embedded linearly in `dim`
    dimensions plus isotropic noise (low intrinsic dimension, no cluster structure). i.i.d.
    uniform data in 128-d cannot reach recall 0.95 with the reference's closest-M neighbour
    selection at any ef <= 256 (SURVEY.md finding 8; calibrated on the CPU oracle)."""
    A = (np.random.default_rng(999).standard_normal((latent, dim)) / np.sqrt(latent)).astype(np.float32)
    rng = np.random.default_rng(seed)
    out = np.empty((n, dim), np.float32)
    step = 1 << 17
    for i in range(0, n, step):
        m = min(step, n - i)
        z = rng.standard_normal((m, latent)).astype(np.float32)
        out[i:i + m] = z @ A + np.float32(noise) * rng.standard_normal((m, dim)).astype(np.float32)
    return out


def get_graph(args, rank: int, world: int, barrier):
    """Rank 0 builds (or loads) the graph; the other ranks load the cached copy."""
    from sqlite_muninn_b200.builder import build_graph
    from sqlite_muninn_b200.graph import FlatGraph
    os.makedirs(CACHE_DIR, exist_ok=True)
    tag = f"n{args.n}_d{args.dim}_M{args.M}_efc{args.efc}_b{args.build_batch}_s11"
    path = os.path.join(CACHE_DIR, tag + ".npz")
    t0 = time.time()
    if rank == 0 and not os.path.exists(path):
        base = make_vectors(args.n, args.dim, 11)
        log(f"[bench] building {tag} with {os.cpu_count()} host threads …")
        g = build_graph(base, 0, args.M, args.efc, batch=args.build_batch, warm=8192)
        tmp = path + ".tmp.npz"
        np.savez(tmp, levels=g.levels, list_offsets=g.list_offsets, targets=g.targets, entry_slot=g.entry_slot,
                 max_level=g.max_level)
        os.replace(tmp, path)
        log(f"[bench] graph built in {time.time() - t0:.1f}s")
    barrier()
    base = make_vectors(args.n, args.dim, 11)
    z = np.load(path)
    g = FlatGraph(args.dim, 0, args.M, args.efc, np.arange(args.n, dtype=np.int64), z["levels"],
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
    queries = make_vectors(args.nq, args.dim, 12)
    cpu = CpuReference(g)
    ef = args.ef or int(os.environ.get("MUNINN_BENCH_EF", "0")) or load_chosen_ef(args) or 128
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
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, ef, None), "cpu_baseline": desc,
        "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}), flush=True)


def chosen_ef_path(args) -> str:
    return os.path.join(CACHE_DIR, f"ef_n{args.n}_d{args.dim}_M{args.M}_efc{args.efc}_b{args.build_batch}.json")


def load_chosen_ef(args):
    try:
        return int(json.load(open(chosen_ef_path(args)))["ef"])
    except Exception:
        return None


def workload_config(args, ef, recall):
    return {"workload": f"SIFT1M-shape {args.n}x{args.dim} fp32 L2, M={args.M}, ef_construction={args.efc}, "
                        f"{args.nq}-query batch, k={args.k} (BASELINE.json configs[1])",
            "ef": ef, "ef_sweep": list(EF_SWEEP), "recall_at_10": recall,
            "data_distribution": "rank-16 Gaussian latent embedded in 128-d + 0.05 isotropic noise, seeds 11/12",
            "graph": f"reference construction rule, batch-synchronous inserts (groups of {args.build_batch}), "
                     "neighbour lists in shadow-table order",
            "l2_policy": "inputs (512 MB slab + 128 MB lists) exceed the 126 MB L2; no flush between steps",
            "parallelism": f"graph replicated, one {args.nq}-query batch per GPU, no collective"}


# ───────────────────────── GPU arm ─────────────────────────

def algorithmic_bytes(stats: np.ndarray, dim: int, M: int, k: int) -> float:
    """B(q) = n_dist·d·4 (rows gathered) + n_expand·M0·4 (layer-0 lists)
    + n_greedy_steps·M·4 (upper-layer lists) + d·4 (query) + k·12 (results)."""
    nd = stats["n_dist"].astype(np.float64).sum()
    ne = stats["n_expand"].astype(np.float64).sum()
    ng = stats["n_greedy_steps"].astype(np.float64).sum()
    nq = len(stats)
    return float(nd * dim * 4 + ne * 2 * M * 4 + ng * M * 4 + nq * (dim * 4 + k * 12))


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
    nq, k, dim = args.nq, args.k, args.dim
    queries = make_vectors(nq, dim, 12 + rank)  # every rank searches its own batch
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
    truth = t_idx.cpu().numpy()

    def search_dev(ef, stats=False):
        idx.search_device(d_q.data_ptr(), nq, k, ef, d_ids.data_ptr(), d_dist.data_ptr(), d_cnt.data_ptr(),
                          d_stats.data_ptr() if stats else None)

    sweep = []
    chosen = args.ef or None
    for ef in ([args.ef] if args.ef else EF_SWEEP):
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
        chosen = EF_SWEEP[-1]
        log(f"[bench] WARNING: recall@10 < 0.95 at every ef of the sweep; timing ef={chosen}")
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
    kernel_ms = []
    torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(args.steps):
        search_dev(ef)
        if args.per_step_kernel_time:
            kernel_ms.append(idx.last_kernel_ms())
    e1.record(stream)
    torch.cuda.synchronize()
    barrier()
    total_ms = e0.elapsed_time(e1)
    launches = idx.launch_count() - launches0
    if not kernel_ms:  # one extra untimed step measures the kernel's own duration live
        for _ in range(3):
            search_dev(ef)
            kernel_ms.append(idx.last_kernel_ms())
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
    if rank == 0:
        peak, which = measured_peak_gbs()
        k_ms = float(np.mean(kernel_ms))
        achieved = algo_bytes / k_ms / 1e6
        out = {
            "metric": METRIC_NAME, "value": world * nq * args.steps / total_ms * 1000.0, "unit": "queries/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, ef, round(recall, 4)),
            "e2e": {"value": world * nq * args.steps / e2e_ms * 1000.0, "unit": "queries/s",
                    "h2d_bytes_per_step": nq * dim * 4, "d2h_bytes_per_step": nq * k * 12 + nq * 4},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(achieved / peak, 4), "traffic": None, "peak_source": which,
                         "kernel": "hnsw_search_kernel<L2>", "kernel_ms": round(k_ms, 3),
                         "algorithmic_bytes_per_launch": algo_bytes,
                         "bytes_per_query": round(algo_bytes / nq, 1)},
            "cpu_baseline": cpu_desc, "clocks": clk, "ef_sweep": sweep, "exact_knn_ms": round(exact_ms, 2),
        }
        print(json.dumps(out), flush=True)
    idx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=1_000_000)
    ap.add_argument("--dim", type=int, default=128)
    ap.add_argument("--M", type=int, default=16)
    ap.add_argument("--efc", type=int, default=200)
    ap.add_argument("--nq", type=int, default=10_000)
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--ef", type=int, default=0, help="0 = smallest ef of the sweep with recall@10 >= 0.95")
    ap.add_argument("--build-batch", type=int, default=1024)
    ap.add_argument("--teams-per-sm", type=int, default=0)
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--per-step-kernel-time", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    import __graft_entry__ as ge
    if rank == 0:
        ge.build()
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
    else:
        run_gpu_arm(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
