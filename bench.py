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
    # roofline.traffic comes from profiles/traffic.json (an ncu --set full capture of THESE sources) or is null
    "tree_d10_3x3x1": dict(kind="tree", rows=3, cols=3, T=1, depth=10, log2n=24, ops_step=318.0, ops_exec=76.0,
                           kernel="k_stream2<TREE,u8,tab_smem>"),
    "tree_d20_9x9x3": dict(kind="tree", rows=9, cols=9, T=3, depth=20, log2n=24, ops_step=4920.0, ops_exec=80.0,
                           kernel="k_stream2<TREE,u8,tab_smem>"),
    # distinct roots (state_seed_i = i): every rollout owns a private state engine, so the
    # kernel pays the 623-step seeding chain per rollout.  ops_setup = 2492, ops_exec per ROLLOUT here:
    # chain 397*3 + lock-step 227*(3+3+5) + tail 56*(3+5) + play 10*45.
    "tree_d10_3x3x1_dr": dict(kind="tree", rows=3, cols=3, T=1, depth=10, log2n=24, ops_step=318.0, ops_setup=2492.0,
                              ops_exec=458.6, dr=True, kernel="k_lane_pre + k_lane<TREE,S>"),
    # config B with distinct roots and distinct transition seeds -> k_warp2 (private 624-word engines
    # seeded and advanced in shared memory, ~20 refills per rollout).  2^20 rollouts per GPU (the CPU needs ~0.3 ms per
    # rollout).  ops_setup = 2492 (state seeding) + 1592 (transition prng); ops_exec per rollout = 427 warp instructions
    # per step of which 51 % are int32 ALU/FMA-pipe instructions (ncu r2u) x 32 lanes = 6980 int ops per step.
    "tree_d20_9x9x3_dr": dict(kind="tree", rows=9, cols=9, T=3, depth=20, log2n=20, ops_step=4920.0, ops_setup=4084.0,
                              ops_exec=6980.0, dr=True, kernel="k_warp2<1> (seeders + players)"),
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


def run_search_binary(impl, gpu_id, seed, leaves, iters, warmup, *extra):
    """tests/cpp/bench_search.cc: the reference's unmodified OffPolicy<Exp3<...>> over the reference model
    (impl 'ref') or the drop-in GPU model (impl 'gpu'); prebuilt next to the reference headers."""
    if not SEARCH_BIN.exists():
        raise RuntimeError(f"{SEARCH_BIN} missing: build it with `python tests/cpp/build.py` next to the reference")
    res = subprocess.run([str(SEARCH_BIN), impl, str(gpu_id), str(seed), str(leaves), str(iters), str(warmup),
                          *map(str, extra)], capture_output=True, text=True, timeout=1200)
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

CSRC = ROOT / "pinyon_b200" / "csrc"


def source_hash():
    """Hash of the kernel sources: an ncu capture under profiles/ speaks for this build only if it recorded the same one."""
    import hashlib

    h = hashlib.sha256()
    for path in sorted(CSRC.glob("*")):
        if path.suffix in {".cu", ".cuh", ".h", ".cpp"}:
            h.update(path.read_bytes())
    return h.hexdigest()[:16]


def measured_traffic(workload):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from an `ncu --set full` capture
    summarised in profiles/traffic.json by tools/ncu_summary.py — only if it was taken on these very sources."""
    path = ROOT / "profiles" / "traffic.json"
    if not path.exists():
        return None, None
    try:
        rec = json.loads(path.read_text()).get(workload)
    except ValueError:
        return None, None
    if not rec or rec.get("source_hash") != source_hash():
        return None, None
    return float(rec["dram_bytes"]), rec.get("capture")


class Rig:
    """One rank: its GPU, its context, the NCCL group, and the timing protocol of the bench contract."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist

        from pinyon_b200 import Context

        self.torch, self.dist, self.args = torch, dist, args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        t0 = time.perf_counter()
        self.ctx = Context(self.local)
        self.create_ms = 1e3 * (time.perf_counter() - t0)
        if args.streams_per_sm:
            self.ctx.set_option("streams_per_sm", args.streams_per_sm)
        if args.rb:
            self.ctx.set_option("rb", args.rb)
        if args.no_zero_copy:
            self.ctx.set_option("zero_copy", 0)
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)  # > 126 MB L2
        self.stream = torch.cuda.current_stream()
        self.ctx.set_stream(self.stream.cuda_stream)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, values, op="max"):
        t = self.torch.tensor(values, dtype=self.torch.float64 if op == "max" else self.torch.int64, device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.SUM)
        return t.cpu().tolist()

    def close(self):
        if self.world > 1:
            self.dist.barrier()
            self.dist.destroy_process_group()
        self.ctx.close()


def measure(rig, w, n, steps, warmup, packed=False, sample_clocks=False):
    """W warm-up steps, then exactly `steps` timed steps of workload `w` with n rollouts on this rank: device-resident
    (CUDA events, L2 flushed between steps) and end to end through the C ABI with host buffers; max over ranks."""
    torch, ctx = rig.torch, rig.ctx
    from pinyon_b200 import Engine, sharding

    depth = w["depth"]
    shard = sharding.plan(n * rig.world, rig.world, rig.rank, depth)  # this rank's slice of the ONE global model stream
    base_engine = Engine(0).discard(shard.engine_offset)
    payoff_d = torch.empty(n, dtype=torch.uint8, device=rig.dev)
    sums_d = torch.zeros(8, dtype=torch.int64, device=rig.dev)
    payoff_h = ctx.pinned((n + 3) // 4 if packed else n, np.uint8)
    sums_h = np.zeros(8, np.uint64)
    dr_dev, dr_host = {}, {}
    if w.get("dr"):
        idx = np.arange(shard.first, shard.first + n, dtype=np.uint64)
        dr_host = dict(state_seed=ctx.pinned(n, np.uint64), transition_seed=ctx.pinned(n, np.uint64))
        dr_host["state_seed"][:] = idx
        dr_host["transition_seed"][:] = idx * np.uint64(0x9E3779B97F4A7C15)
        dr_dev = {k: torch.from_numpy(v.view(np.int64)).to(rig.dev) for k, v in dr_host.items()}

    def call(payoff, sums, host):
        eng = base_engine.copy()
        out = dict(payoff_packed=payoff) if (host and packed) else dict(payoff=payoff)
        if w["kind"] == "tree":
            ctx.rollout_random_tree(n, rows=w["rows"], cols=w["cols"], transitions=w["T"], depth=depth, engine=eng, sums=sums,
                                    **out, **(dr_host if host else dr_dev))
        else:
            ctx.rollout_mold(n, n_actions=w["rows"], depth=depth, engine=eng, sums=sums, **out)

    t_cold = time.perf_counter()
    call(payoff_d, sums_d, False)
    torch.cuda.synchronize()
    cold_ms = 1e3 * (time.perf_counter() - t_cold)  # first call of this shape: jump-family build, table build, allocations
    cold = dict(first_call_ms=cold_ms, family_build_ms=ctx.stat("last_family_build_ms"),
                table_build_ms=ctx.stat("last_table_build_ms"))
    for _ in range(max(warmup - 1, 0)):
        rig.flush.zero_()
        call(payoff_d, sums_d, False)
    rig.barrier()
    sampler = None
    if sample_clocks:
        sampler = ClockSampler(rig.local)
        sampler.start()
        # nvidia-smi takes up to a second to attach to the driver, and launches stall while it does: wait for its first
        # sample so that only steady 100 ms polling overlaps the timed region
        t_wait = time.time()
        while not sampler.rows and time.time() - t_wait < 3.0:
            time.sleep(0.05)
        time.sleep(0.1)
    launches0 = ctx.launches
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]
    rig.barrier()
    t_wall0 = time.time()
    kernel_ms = []
    for i in range(steps):
        rig.flush.zero_()
        starts[i].record(rig.stream)
        call(payoff_d, sums_d, False)
        ends[i].record(rig.stream)
        kernel_ms.append(ctx.last_kernel_ms)  # the rollout kernel alone: CUDA events on the launching stream
    rig.barrier()
    t_wall1 = time.time()
    launches = ctx.launches - launches0
    dev_ms = sum(s.elapsed_time(e) for s, e in zip(starts, ends))
    clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
    sums_dev = sums_d.cpu().numpy().view(np.uint64).copy()
    path = ctx.last_path
    geometry = dict(streams=int(ctx.stat("last_streams")), chunk=int(ctx.stat("last_chunk")))

    # ---- end to end through the C ABI with host buffers
    for _ in range(min(warmup, 3)):
        call(payoff_h, sums_h, True)
    rig.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        call(payoff_h, sums_h, True)
    torch.cuda.synchronize()
    e2e_ms = 1e3 * (time.perf_counter() - t0)
    assert np.array_equal(sums_h, sums_dev), "host and device paths disagree"
    assert int(sums_h[:3].sum()) == n

    dev_ms, e2e_ms = rig.reduce([dev_ms, e2e_ms], "max")
    totals = rig.reduce(sums_dev.astype(np.int64).tolist(), "sum")  # the path's only collective: counters over NVLink
    h2d = 624 * 4 + 8
    if w.get("dr"):
        h2d += 8 * n * (2 if w["T"] > 1 else 1)  # state seeds (+ transition seeds when transitions > 1)
    return dict(n=n, steps=steps, dev_ms=dev_ms, e2e_ms=e2e_ms, kernel_ms=float(np.mean(kernel_ms)), launches=int(launches),
                clocks=clocks, totals=totals, path=path, geometry=geometry, cold=cold, h2d=h2d,
                d2h=(len(payoff_h)) + 64 + 624 * 4)


def roofline(w, m, int_peak, int_peak_alu, peaks, peak_kind, workload_name):
    depth, n, k_ms = w["depth"], m["n"], m["kernel_ms"]
    ops_launch = (w["ops_step"] * depth + w.get("ops_setup", 0.0)) * n   # SURVEY.md §8(d) algorithmic figure
    exe_launch = w["ops_exec"] * depth * n                                 # what the kernel really has to do
    achieved = ops_launch / (k_ms * 1e-3)
    executed = exe_launch / (k_ms * 1e-3)
    hbm_bytes = float(n)  # one payoff byte per rollout; inputs are batch-uniform scalars
    if w.get("dr"):
        hbm_bytes = float(n) * (1 + 8 + depth + 8)  # payoff + state seed + action codes + x[397] write/read
    note = ("achieved = algorithmic int32 ops (SURVEY.md 8(d): %.0f per rollout-step, counted for a PRIVATE state "
            "stream) / the rollout kernel's own CUDA-event time; peak = int-issue microbenchmark measured in this "
            "run (LOP3 only %.1f, LOP3:IMAD 1:1 %.1f Gop/s). " % (w["ops_step"], int_peak_alu / 1e9, int_peak / 1e9))
    if w.get("dr"):
        note += ("DISTINCT-ROOT run: every rollout seeds and advances its own state engine (+%.0f ops per rollout "
                 "for the seeding chains, included in achieved); `executed` counts the %.0f ops per rollout the "
                 "kernels really issue. kernel_ms brackets the seeding and the rollouts (%s); the "
                 "model-stream code generator (k_gen_codes2 + jump-ahead) runs before it and is in ms_per_step only"
                 % (w["ops_setup"], w["ops_exec"] * depth, w["kernel"]))
    else:
        note += ("SHARED-STREAM run: every rollout of the batch shares the root's state seed, so the kernel serves "
                 "the state stream from a shared-memory table and only has to generate the model stream: frac is "
                 "computed on the algorithmic figure as the contract asks and exceeds what the kernel executes; "
                 "`executed` counts the %.0f ops per step it really issues (4 words x (4 twist + 8 temper) + 2 u64 "
                 "mod n x 9 + table walk) and executed.frac / frac_of_alu_pipe are the figures that measure the "
                 "kernel" % w["ops_exec"])
    traffic, capture = measured_traffic(workload_name)
    return {"bound": "int_issue", "kernel": w["kernel"], "kernel_ms": k_ms,
            "kernel_share_of_step": k_ms / (m["dev_ms"] / m["steps"]),
            "achieved": achieved / 1e9, "peak": int_peak / 1e9, "unit": "Gop/s", "frac": achieved / int_peak,
            "traffic": traffic, "traffic_source": capture,
            "executed": {"ops_per_step": w["ops_exec"], "achieved": executed / 1e9, "frac": executed / int_peak,
                         "frac_of_alu_pipe": executed / int_peak_alu},
            "note": note,
            "hbm": {"achieved": hbm_bytes / (k_ms * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                    "frac": hbm_bytes / (k_ms * 1e-3) / 1e9 / peaks["hbm_gbs"], "peak_source": peak_kind}}


# Secondary workloads measured in the same run (BASELINE configs[2..4] and SURVEY 8(d)'s distinct-root runs): name,
# workload key, log2 of the TOTAL batch (strong scaling: divided over the ranks) or None for the workload's per-GPU size.
EXTRA = [
    ("B_tree_d20_9x9x3_2p26", "tree_d20_9x9x3", 26),
    ("M_mold_2_10_2p28", "mold_2_10", 28),
    ("M1_DR_tree_d10_3x3x1", "tree_d10_3x3x1_dr", None),
    ("M2_DR_tree_d20_9x9x3", "tree_d20_9x9x3_dr", None),
]


def single_process_line(args, w, world, depth):
    """The same global batch through ONE context over all N GPUs (pcu_create_multi: shards, per-shard jump slices and the
    NCCL all-reduce live inside the library) — what `Types::Model model{prng{0}}` + model.inference(batch) gets."""
    import torch

    from pinyon_b200 import Context, Engine

    n = (1 << (args.log2n or w["log2n"])) * world
    ctx = Context(list(range(world)))
    payoff = ctx.pinned(n, np.uint8)
    packed = ctx.pinned((n + 3) // 4, np.uint8)
    sums = np.zeros(8, np.uint64)

    def call(**out):
        eng = Engine(0)
        if w["kind"] == "tree":
            ctx.rollout_random_tree(n, rows=w["rows"], cols=w["cols"], transitions=w["T"], depth=depth, engine=eng, sums=sums, **out)
        else:
            ctx.rollout_mold(n, n_actions=w["rows"], depth=depth, engine=eng, sums=sums, **out)

    res = {}
    for name, out in (("bytes", dict(payoff=payoff)), ("packed2", dict(payoff_packed=packed)), ("sums_only", dict())):
        for _ in range(3):
            call(**out)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            call(**out)
        sec = time.perf_counter() - t0
        res[name] = n * args.steps / sec
    path = ctx.last_path
    nccl = ctx.stat("nccl_calls")
    ctx.close()
    torch.cuda.synchronize()
    return {"e2e_rollouts_per_s": res["bytes"], "e2e_packed2_rollouts_per_s": res["packed2"],
            "e2e_sums_only_rollouts_per_s": res["sums_only"], "rollouts_per_step": n, "path": path, "nccl_all_reduces": nccl,
            "note": "one process, one pcu_create_multi context over all GPUs, host buffers (pinned), wall clock around "
                    "the calls; the rank-per-GPU figures above are the bench contract's form"}


def run_ours(args, w, workload_name):
    rig = Rig(args)
    world, rank = rig.world, rig.rank
    n = 1 << (args.log2n or w["log2n"])
    depth = w["depth"]
    m = measure(rig, w, n, args.steps, args.warmup, sample_clocks=True)
    packed_m = measure(rig, w, n, max(3, args.steps // 2), 3, packed=True) if not w.get("dr") else None

    extras = {}
    if not args.no_extra:
        for name, key, log2total in EXTRA:
            if key == workload_name:
                continue
            we = WORKLOADS[key]
            ne = (1 << log2total) // world if log2total else 1 << we["log2n"]
            em = measure(rig, we, ne, max(3, min(args.steps, 5)), 3)
            extras[name] = (we, em, "strong" if log2total else "weak", key)

    search = None
    if not args.no_extra and SEARCH_BIN.exists():
        try:
            rig.barrier()
            mine = run_search_binary("gpu", rig.local, rank, 4096, 10, 2)
            sec = rig.reduce([mine["seconds"]], "max")[0]
            search = {"metric": "search leaves/sec", "value": 4096 * 10 * world / sec, "unit": "leaves/s",
                      "config": "OffPolicy<Exp3<GpuModel<RandomTree 3x3x1 d10>>>, 4096 leaves per batched inference, "
                                "one tree per GPU (root-parallel)", "ms_per_step": 1e3 * sec / 10}
            if rank == 0:
                ref = run_search_binary("ref", 0, 0, 4096, 5, 1)
                search["reference_single_thread_leaves_per_s"] = ref["leaves_per_s"]
            # the same search root-parallel on this rank's share of the host threads, all over this rank's GPU
            cores = max(1, (os.cpu_count() or 1) // world)
            rig.barrier()
            mt = run_search_binary("gpu_op_mt", rig.local, 1000 * rank, 4096, 6, 1, cores)
            sec = rig.reduce([mt["seconds"]], "max")[0]
            search["root_parallel"] = {"value": 4096 * 6 * cores * world / sec, "unit": "leaves/s",
                                       "host_threads_per_gpu": cores, "trees_per_gpu": cores,
                                       "config": "one unmodified OffPolicy search per host thread, 4096 leaves per "
                                                 "batched inference, the threads of a rank share its GPU context"}
            if rank == 0:
                refm = run_search_binary("ref_op_mt", 0, 0, 4096, 3, 1, os.cpu_count() or 1)
                search["root_parallel"]["reference_all_threads_leaves_per_s"] = refm["leaves_per_s"]
                search["root_parallel"]["reference_threads"] = refm["threads"]
        except Exception as exc:  # the search line is extra evidence; the headline must still print
            search = {"error": str(exc)[:300]}
    batched = None
    if not args.no_extra and SEARCH_BIN.exists():
        # BASELINE configs[4] as written: TreeBandit<Exp3> with GPU batched leaf evaluation, 4096 leaves per batch,
        # root-parallel.  UNMODIFIED TreeBandit searches run in fibers on this rank's share of the host threads; their
        # single-leaf inferences rendezvous in the batching adapter (include/pinyon_cuda/batcher.hh).
        try:
            cores = max(1, (os.cpu_count() or 1) // world)
            fibers = max(1, (3 * 4096 + cores - 1) // cores)  # three batches in flight per GPU
            iters = 150
            rig.barrier()
            mine = run_search_binary("gpu_tb_batched", rig.local, 1000003 * rank, 4096, iters, 0, cores, fibers)
            sec = rig.reduce([mine["seconds"]], "max")[0]
            tot = rig.reduce([mine["trees"] * iters, mine["gpu_calls"], mine["gpu_leaves"]], "sum")
            batched = {"metric": "search leaves/sec", "value": tot[0] / sec, "unit": "leaves/s",
                       "config": "TreeBandit<Exp3<BatchedMonteCarloModel<RandomTree 3x3x1 d10>>> unmodified, root-parallel: "
                                 f"{mine['trees']} trees per GPU in fibers on {cores} host threads, {iters} iterations each, "
                                 "4096 leaves per GPU call (model-rng mode W)",
                       "host_threads_per_gpu": cores, "trees_per_gpu": mine["trees"], "gpu_calls": tot[1],
                       "leaves_per_gpu_call": tot[2] / max(1, tot[1]), "seconds": sec}
            if rank == 0:
                ref1 = run_search_binary("ref_tb", 0, 0, 0, 20000, 2000)
                refm = run_search_binary("ref_tb_mt", 0, 0, 0, iters, 0, os.cpu_count() or 1, 64)
                batched["reference_tree_bandit_one_thread_leaves_per_s"] = ref1["leaves_per_s"]
                batched["reference_tree_bandit_all_threads_leaves_per_s"] = refm["leaves_per_s"]
                batched["reference_threads"] = refm["threads"]
                batched["vs_reference_one_thread"] = batched["value"] / ref1["leaves_per_s"]
                batched["vs_reference_all_threads"] = batched["value"] / refm["leaves_per_s"]
        except Exception as exc:
            batched = {"error": str(exc)[:300]}

    single = None
    if world > 1 and not args.no_extra:
        # the other ranks must leave their GPUs IDLE meanwhile: they wait on the rendezvous store (a host-side wait),
        # not in an NCCL barrier, whose kernel would spin on the very GPUs rank 0's context is about to use
        from torch.distributed.distributed_c10d import _get_default_store

        rig.barrier()
        store = _get_default_store()
        if rank == 0:
            try:
                single = single_process_line(args, w, world, depth)
            except Exception as exc:
                single = {"error": str(exc)[:300]}
            store.set("pinyon_single_process_done", "1")
        else:
            store.wait(["pinyon_single_process_done"])
        rig.barrier()

    if rank == 0:
        ctx = rig.ctx
        peaks, peak_kind = measured_peaks()
        # integer-issue peaks measured in this run (pcu_measure_int_peak): LOP3 only (ALU pipe), and
        # LOP3:IMAD 1:1 (ALU + FMA pipes together: the issue-rate ceiling of sm_100 for int32 work)
        int_peak_alu = ctx.int_peak(0, 5)
        int_peak = max(ctx.int_peak(5, 5), ctx.int_peak(2, 5), int_peak_alu)
        total_rollouts = float(n) * world * args.steps
        value = total_rollouts / (m["dev_ms"] * 1e-3)
        e2e = total_rollouts / (m["e2e_ms"] * 1e-3)
        line = {
            "metric": "rollouts/sec", "value": value, "unit": "rollouts/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": m["dev_ms"] / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": describe(w, n), "l2": "flushed between steps by a 256 MiB device write",
                       "parallelism": f"dp{world}: rollout-index shards of one model stream, NCCL all-reduce of sums only",
                       "streams_per_sm": args.streams_per_sm or ("auto: 8 from 2.5e8 rollout-steps per call, else 4"),
                       "kernel_path": m["path"], "model_streams": m["geometry"]["streams"],
                       "rollouts_per_stream": m["geometry"]["chunk"]},
            "rollout_steps_per_sec": value * depth,
            "e2e": {"value": e2e, "unit": "rollouts/s", "h2d_bytes_per_step": m["h2d"], "d2h_bytes_per_step": m["d2h"],
                    "rollout_steps_per_sec": e2e * depth},
            "gpu_launches": m["launches"],
            "clocks": m["clocks"],
            "roofline": roofline(w, m, int_peak, int_peak_alu, peaks, peak_kind, workload_name),
            "sums": {"row_wins": int(m["totals"][0]), "draws": int(m["totals"][1]), "row_losses": int(m["totals"][2]),
                     "steps": int(m["totals"][3])},
            "cold_call": dict(m["cold"], context_create_ms=rig.create_ms,
                              note="first call of this batch shape in a fresh context (jump-polynomial family of a new "
                                   "stride, state-stream table, buffers); pcu_warmup pays it ahead of time; not in "
                                   "ms_per_step"),
        }
        if packed_m:
            pe = float(n) * world * packed_m["steps"] / (packed_m["e2e_ms"] * 1e-3)
            line["e2e_packed2"] = {"value": pe, "unit": "rollouts/s", "d2h_bytes_per_step": packed_m["d2h"],
                                   "note": "same call with pcu_out.payoff_packed2 (2 bits per rollout) instead of bytes"}
        wl = {}
        for name, (we, em, scaling, key) in extras.items():
            tot = float(em["n"]) * world * em["steps"]
            rf = roofline(we, em, int_peak, int_peak_alu, peaks, peak_kind, key)
            wl[name] = {"workload": describe(we, em["n"]), "scaling": scaling, "rollouts_per_step_all_gpus": em["n"] * world,
                        "value": tot / (em["dev_ms"] * 1e-3), "unit": "rollouts/s",
                        "rollout_steps_per_sec": tot / (em["dev_ms"] * 1e-3) * we["depth"],
                        "e2e": tot / (em["e2e_ms"] * 1e-3), "ms_per_step": em["dev_ms"] / em["steps"],
                        "kernel_ms": em["kernel_ms"], "kernel_path": em["path"], "model_streams": em["geometry"]["streams"],
                        "gpu_launches": em["launches"],
                        "roofline": {"frac": rf["frac"], "executed": rf["executed"], "kernel": rf["kernel"]},
                        "cold_first_call_ms": em["cold"]["first_call_ms"]}
        if search is not None:
            wl["search_offpolicy"] = search
        if batched is not None:
            wl["search_treebandit_batched"] = batched
        if wl:
            line["workloads"] = wl
        if single is not None:
            line["single_process_multi_gpu"] = single
        if world == 1 and not args.no_cpu:
            cb = cpu_reference_rate(w, seconds_target=12.0)
            c1 = cpu_reference_rate(w, seconds_target=5.0, threads=1)
            line["cpu_baseline"] = {"value": cb["rate"], "unit": "rollouts/s", "cores": cb["cores"], "kind": cb["kind"],
                                    "sample": f"{cb['n']} rollouts of the same workload in {cb['seconds']:.1f} s",
                                    "one_thread": {"value": c1["rate"], "cores": 1,
                                                   "sample": f"{c1['n']} rollouts in {c1['seconds']:.1f} s"}}
        print(json.dumps(line), flush=True)
    rig.close()


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
    ap.add_argument("--no-extra", action="store_true", help="headline workload only (no `workloads` object): ncu runs")
    ap.add_argument("--rb", type=int, default=0, help="rollouts per consumer batch of the fused kernel (32, 64, 128)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.workload == "search_offpolicy":
        return run_search(args)
    w = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference_arm(args, w)
    else:
        run_ours(args, w, args.workload)


if __name__ == "__main__":
    main()
