#!/usr/bin/env python
"""bench.py — rollouts/sec of the Monte-Carlo leaf evaluator on B200 (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W            our CUDA path (one rank per GPU)
  python bench.py --impl reference ...                      the reference's own CPU code

A "step" is one pass of the hot path over one batch: 2^24 RandomTree rollouts (depth_bound 10,
3x3 actions, 1 chance transition, root prng{0}; BASELINE.json configs[1]) per GPU, consumed in
index order by ONE model prng{0} exactly as MonteCarloModel::inference(batch) does
(monte-carlo-model.hh:58-67).  With N GPUs the global batch is N * 2^24 rollouts of that one
stream; rank r takes the r-th slice (its engine is jumped to the slice start on the host), no
data-path collective, and a final NCCL all-reduce of the win/draw/loss/step sums.

One JSON line on stdout (rank 0).  `value` = rollouts/s with everything resident in HBM;
`e2e` = the same through the C ABI with host buffers (D2H of per-rollout payoffs inside).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

WORKLOADS = {
    # name: (kind, rows, cols, transitions, depth, log2 n per gpu, algorithmic int32 ops/step, ops/rollout setup)
    # traffic: dram__bytes_read.sum + dram__bytes_write.sum of k_stream2 from the ncu --set full capture of this build
    # (profiles/r1i_ncu_full_summary.csv, 2^24 rollouts): 1.53 MB read, 0 written — under ncu the 16.8 MB of payoff
    # bytes were still in L2 when the kernel ended; they reach HBM later (or go straight to the host when zero-copy).
    "tree_d10_3x3x1": dict(kind="tree", rows=3, cols=3, T=1, depth=10, log2n=24, ops_step=318.0, ops_exec=76.0,
                           kernel="k_stream2<TREE,u8,tab_smem>", traffic=1528320.0),
    "tree_d20_9x9x3": dict(kind="tree", rows=9, cols=9, T=3, depth=20, log2n=24, ops_step=4920.0, ops_exec=80.0,
                           kernel="k_stream2<TREE,u8,tab_smem>"),
    # distinct roots (state_seed_i = i): every rollout owns a private state engine, so the
    # kernel pays the 623-step seeding chain per rollout.  ops_setup = 2492, ops_exec per ROLLOUT here:
    # chain 397*3 + lock-step 227*(3+3+5) + tail 56*(3+5) + play 10*45.
    "tree_d10_3x3x1_dr": dict(kind="tree", rows=3, cols=3, T=1, depth=10, log2n=24, ops_step=318.0, ops_setup=2492.0,
                              ops_exec=458.6, dr=True, kernel="k_lane_pre + k_lane<TREE,S>"),
    # config B with distinct roots and distinct transition seeds -> k_seed_pre + k_warp (private
    # 624-word engines refilled ~19.5 times per rollout).  2^20 rollouts per GPU (the CPU needs ~0.3 ms per rollout).
    # ops_setup = 2492 (state seeding) + 1592 (transition prng); ops_exec per rollout = 430 warp instructions per step
    # of which 55 % are int32 ALU/FMA-pipe instructions (ncu r1e, r1g) x 32 lanes = 7150 int ops per step.
    "tree_d20_9x9x3_dr": dict(kind="tree", rows=9, cols=9, T=3, depth=20, log2n=20, ops_step=4920.0, ops_setup=4084.0,
                              ops_exec=7150.0, dr=True, kernel="k_seed_pre + k_warp<TREE,S>"),
    "mold_2_10": dict(kind="mold", rows=2, cols=2, T=1, depth=10, log2n=24, ops_step=104.0, ops_exec=70.0,
                      kernel="k_stream2<MOLD,u8>"),
}
DEFAULT_WORKLOAD = "tree_d10_3x3x1"
WORDS_PER_STEP = 4  # two random_int per step, two 32-bit draws each (random.hh:26,28)


def describe(w, n):
    if w["kind"] == "mold":
        return f"MoldState n_actions={w['rows']} max_depth={w['depth']}, batch {n} per GPU, mode S (one model prng{{0}})"
    root = ("DISTINCT roots state_seed_i = i, transition_seed_i = i*0x9E3779B97F4A7C15 (SURVEY M1-DR)" if w.get("dr")
            else "root prng{0}")
    return (f"RandomTree depth_bound={w['depth']} {w['rows']}x{w['cols']} actions {w['T']} chance transition(s), "
            f"{root}, batch {n} (2^{w['log2n']}) per GPU, mode S (one model prng{{0}} consumed in index order)")


class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.rows = []
        self.proc = None
        self.gpu = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        inside = [r for (t, r) in self.rows if t0 - 0.05 <= t <= t1 + 0.15] or [r for (_, r) in self.rows]
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in inside:
            try:
                sm.append(float(r[1]))
                mx = float(r[2])
                for name, val in zip(names, r[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
            except (ValueError, IndexError):
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return json.loads(p.read_text()), "measured"
        except ValueError:
            pass
    return {"hbm_gbs": 6650.0}, "fallback"


# ----------------------------------------------------------------------- reference arm


def cpu_reference_rate(w, seconds_target=12.0, threads=None):
    """Time the reference's own MonteCarloModel (oracle/_ref, compiled from the reference's
    headers) on this box's host cores; falls back to the C restatement when absent."""
    from oracle import loader

    threads = threads or os.cpu_count() or 1
    ref = loader.load_reference()
    kind_id = 0 if w["kind"] == "tree" else 1
    if ref is not None:
        def run(n):
            wdl = np.zeros(3, np.uint64)
            steps = np.zeros(1, np.uint64)
            sec = ref.ref_time_rollouts(kind_id, n, threads, 0, w["depth"], w["rows"], w["cols"], w["T"], 0,
                                        wdl.ctypes.data, steps.ctypes.data)
            return sec
        kind = "reference"
    else:
        orc = loader.load_oracle()
        threads = 1

        def run(n):
            t0 = time.perf_counter()
            if kind_id == 0:
                loader.random_tree_rollouts(orc, n, rows=w["rows"], cols=w["cols"], transitions=w["T"], depth=w["depth"],
                                            want_traj=False)
            else:
                loader.mold_rollouts(orc, n, n_actions=w["rows"], depth=w["depth"], want_traj=False)
            return time.perf_counter() - t0
        kind = "port"
    probe_n = 2000 * threads
    sec = max(run(probe_n), 1e-6)
    n = int(min(max(probe_n, probe_n * seconds_target / sec), 5e9))
    sec = run(n)
    return dict(rate=n / sec, n=n, seconds=sec, cores=threads, kind=kind)


def run_reference_arm(args, w):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    per_step = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    rates = []
    info = None
    for i in range(args.warmup + args.steps):
        info = cpu_reference_rate(w, seconds_target=per_step)
        if i >= args.warmup:
            rates.append(info)
    tot_n = sum(r["n"] for r in rates)
    tot_s = sum(r["seconds"] for r in rates)
    value = tot_n / tot_s
    n_gpu = 1 << w["log2n"]
    line = {
        "impl": "reference", "metric": "rollouts/sec", "value": value, "unit": "rollouts/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot_s / max(1, len(rates)),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": describe(w, n_gpu)},
        "rollout_steps_per_sec": value * w["depth"],
        "cpu_baseline": {"value": value, "unit": "rollouts/s", "cores": info["cores"], "kind": info["kind"],
                         "sample": f"{rates[0]['n']} rollouts of the same workload per step, {info['cores']} threads "
                                   f"each with its own Model copy and a contiguous slice"},
        "e2e": {"value": value, "unit": "rollouts/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------ search workload (M4)

SEARCH_BIN = ROOT / "tests" / "cpp" / "_bin" / "bench_search"


def run_search_binary(impl, gpu_id, seed, leaves, iters, warmup):
    """tests/cpp/bench_search.cc: the reference's unmodified OffPolicy<Exp3<...>> over the reference model
    (impl 'ref') or the drop-in GPU model (impl 'gpu'); prebuilt next to the reference headers."""
    if not SEARCH_BIN.exists():
        raise RuntimeError(f"{SEARCH_BIN} missing: build it with `python tests/cpp/build.py` next to the reference")
    res = subprocess.run([str(SEARCH_BIN), impl, str(gpu_id), str(seed), str(leaves), str(iters), str(warmup)],
                         capture_output=True, text=True, timeout=1200)
    if res.returncode != 0:
        raise RuntimeError(f"bench_search {impl} failed: {res.stderr[-2000:]}")
    return json.loads(res.stdout.strip().splitlines()[-1])


def run_search(args):
    """BASELINE.json configs[4] / SURVEY 8(d) M4: OffPolicy<Exp3<MonteCarloModel<RandomTree>>> on M1's root, 4096
    leaves per batched inference, ROOT-PARALLEL over the GPUs: rank r searches its own tree with device prng{r},
    root statistics (exp3.hh:9-16) are summed over ranks (the path's only collective).  A step = one learner
    iteration (4096 tree descents on the host + one batched leaf evaluation on the GPU + the tree update)."""
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    reference = args.impl == "reference"
    if reference and rank != 0:
        return
    if world > 1 and not reference:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    leaves = 4096
    if not reference:
        torch.cuda.set_device(local)
        # parity first: equal seeds and iteration counts must give the same tree (ref on the host cores, gpu here)
        if rank == 0:
            a = run_search_binary("ref", 0, 0, leaves, 3, 0)
            b = run_search_binary("gpu", local, 0, leaves, 3, 0)
            keys = ("nodes", "visits", "value_total_row", "row_visits", "col_visits")
            assert all(a[k] == b[k] for k in keys), ("GPU search diverged from the reference", a, b)
        if world > 1:
            dist.barrier()
    sampler = ClockSampler(local)
    if not reference:
        sampler.start()
    t0 = time.time()
    out = run_search_binary("ref" if reference else "gpu", local, rank, leaves, args.steps, args.warmup)
    t1 = time.time()
    sec = out["seconds"]
    stats = [out["nodes"], out["visits"], int(round(2 * out["value_total_row"])), *out["row_visits"], *out["col_visits"]]
    if not reference:
        clocks = sampler.stop(t0, t1)
        t = torch.tensor([sec], dtype=torch.float64, device="cuda")
        tot = torch.tensor(stats, dtype=torch.int64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        sec = float(t.item())
        stats = tot.cpu().tolist()
    if rank == 0:
        n = world if not reference else 1
        value = leaves * args.steps * n / sec
        line = {
            "metric": "search leaves/sec", "value": value, "unit": "leaves/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * sec / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": "OffPolicy<Exp3<MonteCarloModel<RandomTree>>> (off-policy.hh:83-109) on root "
                                   "RandomTree{prng{0},10,3,3,1}, 4096 leaves per batched inference, gamma 0.01, "
                                   "model prng{0}, device prng{rank}; one search tree per GPU (root-parallel), root "
                                   "statistics summed over ranks",
                       "parallelism": f"root-parallel x{n}", "l2": "inputs differ every step (the tree grows)"},
            "root_stats_sum": {"nodes": stats[0], "visits": stats[1], "value_total_row_x2": stats[2],
                               "row_visits": stats[3:6], "col_visits": stats[6:9]},
        }
        if reference:
            line["impl"] = "reference"
            line["cpu_baseline"] = {"value": value, "unit": "leaves/s", "cores": 1, "kind": "reference",
                                    "sample": f"{args.steps} learner iterations of 4096 leaves, the reference's own "
                                              "single-threaded OffPolicy + MonteCarloModel"}
            line["e2e"] = {"value": value, "unit": "leaves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
        else:
            # the whole search IS the end-to-end path: host tree walk, H2D of the SoA leaf batch, rollouts, D2H of payoffs
            line["e2e"] = {"value": value, "unit": "leaves/s", "h2d_bytes_per_step": leaves * 26 + 624 * 4 + 8,
                           "d2h_bytes_per_step": leaves + 64 + 624 * 4}
            line["gpu_launches"] = None
            line["clocks"] = clocks
            line["note"] = ("host-bound: 4096 tree descents and state copies per step run on ONE host thread in the "
                            "reference's OffPolicy; the GPU evaluates the 4096 leaves in one call. parity: ref == gpu "
                            "root statistics checked for 3 learner iterations before timing")
        print(json.dumps(line), flush=True)
    if world > 1 and not reference:
        dist.barrier()
        dist.destroy_process_group()


# ------------------------------------------------------------------------------ our arm


def run_ours(args, w):
    import torch
    import torch.distributed as dist

    from pinyon_b200 import Context, Engine, sharding

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    ctx = Context(local)
    if args.streams_per_sm:
        ctx.set_option("streams_per_sm", args.streams_per_sm)
    if args.no_zero_copy:
        ctx.set_option("zero_copy", 0)
    n = 1 << (args.log2n or w["log2n"])
    depth = w["depth"]

    # this rank's slice of the one global model stream: rollouts [rank*n, (rank+1)*n)
    shard = sharding.plan(n * world, world, rank, depth)
    assert shard.count == n
    base_engine = Engine(0).discard(shard.engine_offset)

    payoff_d = torch.empty(n, dtype=torch.uint8, device=dev)
    sums_d = torch.zeros(8, dtype=torch.int64, device=dev)
    payoff_h = ctx.pinned(n, np.uint8)
    sums_h = np.zeros(8, np.uint64)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)

    dr_dev, dr_host = {}, {}
    if w.get("dr"):
        first = shard.first
        idx = np.arange(first, first + n, dtype=np.uint64)
        ts = idx * np.uint64(0x9E3779B97F4A7C15)
        dr_host = dict(state_seed=ctx.pinned(n, np.uint64), transition_seed=ctx.pinned(n, np.uint64))
        dr_host["state_seed"][:] = idx
        dr_host["transition_seed"][:] = ts
        dr_dev = {k: torch.from_numpy(v.view(np.int64)).to(dev) for k, v in dr_host.items()}

    def call(payoff, sums):
        eng = base_engine.copy()
        if w["kind"] == "tree":
            extra = dr_dev if isinstance(payoff, torch.Tensor) else dr_host
            ctx.rollout_random_tree(n, rows=w["rows"], cols=w["cols"], transitions=w["T"], depth=depth, engine=eng,
                                    payoff=payoff, sums=sums, **extra)
        else:
            ctx.rollout_mold(n, n_actions=w["rows"], depth=depth, engine=eng, payoff=payoff, sums=sums)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing (value)
    for _ in range(args.warmup):
        flush.zero_()
        call(payoff_d, sums_d)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    # nvidia-smi takes up to a second to attach to the driver, and launches stall while it does: wait for its first
    # sample so that only steady 100 ms polling overlaps the timed region (a 1.66 ms outlier step was seen otherwise)
    t_wait = time.time()
    while not sampler.rows and time.time() - t_wait < 3.0:
        time.sleep(0.05)
    time.sleep(0.1)
    launches0 = ctx.launches
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps)]
    barrier()
    t_wall0 = time.time()
    kernel_ms = []
    for i in range(args.steps):
        flush.zero_()
        starts[i].record(stream)
        call(payoff_d, sums_d)
        ends[i].record(stream)
        kernel_ms.append(ctx.last_kernel_ms)  # the rollout kernel alone: CUDA events on the launching stream
    barrier()
    t_wall1 = time.time()
    launches = ctx.launches - launches0
    dev_ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
    clocks = sampler.stop(t_wall0, t_wall1)
    sums_host = sums_d.cpu().numpy().view(np.uint64).copy()

    # ---- end-to-end timing through the C ABI with host buffers
    for _ in range(min(args.warmup, 3)):
        call(payoff_h, sums_h)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        call(payoff_h, sums_h)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    assert np.array_equal(sums_h, sums_host), "host and device paths disagree"
    assert int(sums_h[:3].sum()) == n

    t = torch.tensor([dev_ms, e2e_s * 1e3], dtype=torch.float64, device=dev)
    tot = torch.from_numpy(sums_host.astype(np.int64)).to(dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)  # the path's only collective: W/D/L/step sums over NVLink
    dev_ms, e2e_ms = t.tolist()
    totals = tot.cpu().numpy()

    if rank == 0:
        peaks, peak_kind = measured_peaks()
        total_rollouts = float(n) * world * args.steps
        value = total_rollouts / (dev_ms * 1e-3)
        e2e = total_rollouts / (e2e_ms * 1e-3)
        # integer-issue peaks measured in this run (pcu_measure_int_peak): LOP3 only (ALU pipe), and
        # LOP3:IMAD 1:1 (ALU + FMA pipes together: the issue-rate ceiling of sm_100 for int32 work)
        int_peak_alu = ctx.int_peak(0, 5)
        int_peak = max(ctx.int_peak(5, 5), ctx.int_peak(2, 5), int_peak_alu)
        ms_step = dev_ms / args.steps
        k_ms = float(np.mean(kernel_ms))
        ops_launch = (w["ops_step"] * depth + w.get("ops_setup", 0.0)) * n   # SURVEY.md §8(d) algorithmic figure
        exe_launch = w["ops_exec"] * depth * n          # what the kernel really has to do
        achieved = ops_launch / (k_ms * 1e-3)
        executed = exe_launch / (k_ms * 1e-3)
        hbm_bytes = float(n)  # one payoff byte per rollout; inputs are batch-uniform scalars
        h2d = 624 * 4 + 8
        if w.get("dr"):
            hbm_bytes = float(n) * (1 + 8 + depth + 8)  # payoff + state seed + action codes + x[397] write/read
            h2d += 8 * n * (2 if w["T"] > 1 else 1)  # state seeds (+ transition seeds when transitions > 1)
        note = ("achieved = algorithmic int32 ops (SURVEY.md 8(d): %.0f per rollout-step, counted for a PRIVATE state "
                "stream) / the rollout kernel's own CUDA-event time; peak = int-issue microbenchmark measured in this "
                "run (LOP3 only %.1f, LOP3:IMAD 1:1 %.1f Gop/s). " % (w["ops_step"], int_peak_alu / 1e9, int_peak / 1e9))
        if w.get("dr"):
            note += ("DISTINCT-ROOT run: every rollout seeds and advances its own state engine (+%.0f ops per rollout "
                     "for the seeding chains, included in achieved); `executed` counts the %.0f ops per rollout the "
                     "kernels really issue. kernel_ms brackets the seeding pre-kernel + the rollout kernel (%s); the "
                     "model-stream code generator (k_gen_codes2 + jump-ahead) runs before it and is in ms_per_step only"
                     % (w["ops_setup"], w["ops_exec"] * depth, w["kernel"]))
        else:
            note += ("SHARED-STREAM run: every rollout of the batch shares the root's state seed, so the kernel serves "
                     "the state stream from a shared-memory table and only has to generate the model stream: "
                     "`executed` counts those %.0f ops per step (4 words x (4 twist + 8 temper) + 2 u64 mod n x 9 + "
                     "table walk) and is the figure that measures the kernel; frac > executed.frac is the shortcut's "
                     "gain, not kernel quality" % w["ops_exec"])
        line = {
            "metric": "rollouts/sec", "value": value, "unit": "rollouts/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": describe(w, n), "l2": "flushed between steps by a 256 MiB device write",
                       "parallelism": f"dp{world}: rollout-index shards of one model stream, NCCL all-reduce of sums only",
                       "streams_per_sm": args.streams_per_sm or ("auto: 8 from 2.5e8 rollout-steps per call, else 4")},
            "rollout_steps_per_sec": value * depth,
            "e2e": {"value": e2e, "unit": "rollouts/s", "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": n + 64 + 624 * 4, "rollout_steps_per_sec": e2e * depth},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "int_issue", "kernel": w["kernel"], "kernel_ms": k_ms,
                         "kernel_share_of_step": k_ms / ms_step,
                         "achieved": achieved / 1e9, "peak": int_peak / 1e9, "unit": "Gop/s",
                         "frac": achieved / int_peak,
                         "traffic": w.get("traffic") if n == (1 << 24) else None,
                         "executed": {"ops_per_step": w["ops_exec"], "achieved": executed / 1e9,
                                      "frac": executed / int_peak, "frac_of_alu_pipe": executed / int_peak_alu},
                         "note": note,
                         "hbm": {"achieved": hbm_bytes / (k_ms * 1e-3) / 1e9, "peak": peaks["hbm_gbs"],
                                 "unit": "GB/s", "frac": hbm_bytes / (k_ms * 1e-3) / 1e9 / peaks["hbm_gbs"],
                                 "peak_source": peak_kind}},
            "sums": {"row_wins": int(totals[0]), "draws": int(totals[1]), "row_losses": int(totals[2]),
                     "steps": int(totals[3])},
        }
        if world == 1 and not args.no_cpu:
            cb = cpu_reference_rate(w, seconds_target=12.0)
            line["cpu_baseline"] = {"value": cb["rate"], "unit": "rollouts/s", "cores": cb["cores"], "kind": cb["kind"],
                                    "sample": f"{cb['n']} rollouts of the same workload in {cb['seconds']:.1f} s"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    ctx.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS) + ["search_offpolicy"])
    ap.add_argument("--log2n", type=int, default=0, help="override rollouts per GPU (log2)")
    ap.add_argument("--streams-per-sm", type=int, default=0, help="model-stream generator warps per SM (default 4)")
    ap.add_argument("--no-zero-copy", action="store_true", help="e2e: stage payoffs in HBM and copy them back")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.workload == "search_offpolicy":
        return run_search(args)
    w = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference_arm(args, w)
    else:
        run_ours(args, w)


if __name__ == "__main__":
    main()
