#!/usr/bin/env python
"""bench.py -- headline benchmark of python_motion_planning_b200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c5|c4|c3|c2] [--impl reference]

One "step" is one pass of the hot path over one batch of synthetic input.  Headline workload
(default, every N): **c5** = BASELINE.json configs[4], the full cost-to-go field of a 1536^3
6-connected unit-weight voxel grid (u32), source (1,1,1).  N = 1: the whole grid on one GPU.
N = 2/4/8 (torchrun): the SAME grid cut into N z-slabs, one per GPU, halo exchange on the device over
NVLink peer memory (pmp_bfs_slab_*) -- strong scaling.  Every timed configuration is checked by the
local-optimality certificate (pmp_verify_field: 0 violations) before the line is printed.
Sub-records in the same line (`extras`): c4 = configs[3], one A* query on a 512^3 26-connected grid
(rank 0; its cost is asserted against the oracle's known answer, tests/golden/bench_sizes.json) and
c3 = configs[2], 4096 Dijkstra queries on a 1024^2 8-connected grid, sharded over the ranks.
`--workload c4|c3|c2` makes one of those the headline instead (profiling runs).

Metric: edge relaxations / s = |motions| * N_settled / t, in GTEPS.
`--impl reference` times the CPU restatement of the reference (oracle/, all host threads) on a
bounded sample of the same workload; rank 0 only.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ALG_BYTES = {"f64": 16.125, "u32": 8.125}   # read g + write g + 1 occupancy bit per settled cell
REL = 1e-9
DEFAULT_N = {"c5": 1536, "c4": 512, "c3": 1024, "c2": 0}


def workload_name(wl, n, nq=4096, world=1):
    if wl == "c5":
        par = "single GPU" if world == 1 else f"{world} z-slabs, device-side halo exchange over NVLink peer memory"
        return f"c5: full cost-to-go field, {n}^3 6-connected unit weights (u32), {par}"
    if wl == "c4":
        return f"c4: single-query A* (euclidean), {n}^3 26-connected voxel grid, random boxes"
    if wl == "c3":
        return f"c3: batched Dijkstra, {nq} queries, {n}^2 8-connected, 30% random obstacles"
    return "c2: examples/3d_example.py defaults, Grid3D(26,20,16) 'door', A*"


def measured_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def golden():
    try:
        return json.load(open(os.path.join(ROOT, "tests", "golden", "bench_sizes.json")))
    except Exception:
        return {}


class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region: NVML (nvidia_ml_py) polled by
    a thread every few ms plus one explicit sample per step while the GPU is busy; nvidia-smi as
    the fallback when NVML cannot be loaded."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"),
               (0x4, "sw_power_cap"))

    def __init__(self, index, uuid=None):
        self.sm, self.mx, self.bits, self.rows = [], None, 0, []
        self.proc, self.nvml, self.h, self.stop_flag = None, None, None, False
        self.lock = threading.Lock()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            try:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid if isinstance(uuid, bytes) else str(uuid).encode()) \
                    if uuid else pynvml.nvmlDeviceGetHandleByIndex(index)
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.mx = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.th = threading.Thread(target=self._poll, daemon=True)
            self.th.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def sample(self):
        """one NVML sample now (called from the timed loop while kernels are in flight)"""
        if self.nvml is None:
            return
        try:
            c = float(self.nvml.nvmlDeviceGetClockInfo(self.h, self.nvml.NVML_CLOCK_SM))
            try:
                r = int(self.nvml.nvmlDeviceGetCurrentClocksEventReasons(self.h))
            except Exception:
                r = int(self.nvml.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
            with self.lock:
                self.sm.append(c)
                self.bits |= r
        except Exception:
            pass

    def _poll(self):
        while not self.stop_flag:
            self.sample()
            time.sleep(0.005)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.nvml is not None:
            self.stop_flag = True
            self.th.join(timeout=1)
            reasons = sorted(n for b, n in self.REASONS if self.bits & b)
            return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.mx,
                    "reasons": reasons, "samples": len(self.sm), "source": "nvml"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvidia-smi"}


# ------------------------------------------------------------------------------------------------
# workloads: each builder returns a dict
#   run(step) -> my kernel launches of the step;  events: list of (start, stop) around the search phase
#   account() -> (edges, settled cells, algorithmic bytes, check value)   [after a run]
#   verify()  -> dict of asserted facts (raises AssertionError on a wrong answer)
#   e2e(step) -> (check, h2d bytes, d2h bytes): the same through the public API with HOST buffers
# ------------------------------------------------------------------------------------------------
def timed(torch, events, fn):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fn()
    e1.record()
    events.append((e0, e1))


def build_c5(args, ctx):
    torch, pmp, workloads = ctx["torch"], ctx["pmp"], ctx["workloads"]
    from python_motion_planning_b200 import engine, _lib, slab
    from python_motion_planning_b200.environment import motion_table
    rank, world, dist = ctx["rank"], ctx["world"], ctx["dist"]
    n = args.n or DEFAULT_N["c5"]
    blk, _ = workloads.replicated_boxes3d(base=128, reps=max(1, n // 128))
    src = (1, 1, 1)
    events = []
    cfg = {"grid": [n, n, n], "connectivity": 6, "occupancy_input": "bit rows (Grid3D.update_packed), 1 bit per cell"}
    if world > 1:
        solver = slab.BfsSlabSolver(lambda z0, z1: workloads.replicated_packed(blk, (n, n, n), z0, z1), (n, n, n),
                                    rank, world, dist, packed=True)
        sl = solver.slab
        host_rows = torch.from_numpy(workloads.replicated_packed(blk, (n, n, n), solver.z0, solver.z1).view(np.int32)).pin_memory()

        def run(step):
            l0 = sl.launches
            timed(torch, events, lambda: solver.solve(src))
            return sl.launches - l0

        def account():
            bad, reached, pend = solver.verify(src)          # global numbers (all-reduced)
            assert bad == 0 and pend == 0, f"certificate: {bad} violations, {pend} pending"
            solver.check()
            npl = sl.dg.padded[0] * sl.dg.padded[1]
            nzl = solver.z1 - solver.z0
            local = int(((sl.field[:nzl * npl] != -1)).sum().item())
            # algorithmic halo bytes of this rank: one 4-byte read per reached cell of each neighbour's boundary plane
            # = (by symmetry of the exchange) the reached cells of the own boundary planes that face another slab
            hb = 0
            if rank > 0:
                hb += int((sl.field[:npl] != -1).sum().item()) * 4
            if rank < world - 1:
                hb += int((sl.field[(nzl - 1) * npl:nzl * npl] != -1).sum().item()) * 4
            cfg["algorithmic_halo_bytes_rank0"] = hb
            return local * 6, local, local * ALG_BYTES["u32"], float(reached)

        def verify():
            return {"certificate_violations": 0, "pending": 0}

        def e2e(step):
            sl.env.update_packed(host_rows)                  # H2D of this rank's bit rows from pinned memory
            sl.env.device_grid()
            solver.solve(src)
            far = int(sl.field[0].item())                    # D2H of one field word
            return float(far), host_rows.numel() * 4 + 12, 4

        cfg.update(slabs=world, planes_per_slab=solver.z1 - solver.z0)
        return dict(run=run, events=events, account=account, verify=verify, e2e=e2e, dtype="u32", queries=1,
                    scaling="strong", parallelism=f"z-slabs x{world}", name=workload_name("c5", n, world=world), cfg=cfg,
                    stats=lambda: sl.stats_host(), K=6, summed_account=True)
    rows = workloads.replicated_packed(blk, (n, n, n), 0, n)
    host_rows = torch.from_numpy(rows.view(np.int32)).pin_memory()
    env = pmp.Grid3D(n, n, n, walls=False)
    env.motions = [m for m in env.motions if m.g == 1]
    env.update_packed(host_rows)
    dg = env.device_grid()
    dg.set_motions(*motion_table(env, 3))
    ws = engine.get_workspace(dg, 1, _lib.W_U32)

    def run(step):
        engine.upload_queries(ws, [src], None, 3)
        timed(torch, events, lambda: engine.sssp(ws, _lib.DIJKSTRA, 0, full_field=True, have_goals=False))
        return 4     # init, k_bfs, fill of untouched tiles, finish

    def account():
        settled = int(engine.count_settled(ws, bounded=False).sum().item())
        return settled * 6, settled, settled * ALG_BYTES["u32"], float(settled)

    def verify():
        out = torch.zeros(3, dtype=torch.int64, device=dg.device)
        s = torch.tensor(list(src), dtype=torch.int32, device=dg.device)
        _lib.check(dg.L.pmp_verify_field(dg.handle, _lib.W_U32, ws.g.data_ptr(), s.data_ptr(), 0, n, out.data_ptr(),
                                         dg.stream()), "pmp_verify_field")
        bad, reached, pend = (int(v) for v in out.cpu().tolist())
        engine.check_field_status(ws)
        assert bad == 0 and pend == 0, f"certificate: {bad} violations, {pend} pending"
        return {"certificate_violations": bad, "pending": pend, "reached": reached}

    def e2e(step):
        env.update_packed(host_rows)                          # H2D of the bit rows from pinned memory
        wsx, st = pmp.cost_field(env, src, weight="u32", to_host=False)
        far = int(wsx.g[0][-1].item())                        # D2H of one field word
        return float(far), host_rows.numel() * 4 + 12, 4

    return dict(run=run, events=events, account=account, verify=verify, e2e=e2e, dtype="u32", queries=1,
                scaling="strong", parallelism="z-slabs x1 (whole grid on one GPU)", name=workload_name("c5", n), cfg=cfg,
                stats=lambda: ws.stats.cpu().numpy(), K=6)


def build_c4(args, ctx, n=None, small=False):
    torch, pmp, workloads = ctx["torch"], ctx["pmp"], ctx["workloads"]
    from python_motion_planning_b200 import engine, _lib
    from python_motion_planning_b200.environment import motion_table
    events = []
    if small:
        occ, dims = workloads.door3d()
        start, goal = (1, 1, 1), (24, 18, 14)
        name, gold = workload_name("c2", 0), None
    else:
        n = n or args.n or DEFAULT_N["c4"]
        occ = workloads.boxes3d(n, seed=0)
        dims, start, goal = (n, n, n), (1, 1, 1), (n - 2, n - 2, n - 2)
        workloads.seal(occ)
        name, gold = workload_name("c4", n), golden().get(f"c4_{n}")
    env = pmp.Grid3D(*dims, walls=False)
    host_occ = torch.from_numpy(occ.view(np.uint8)).pin_memory()
    env.update(host_occ)
    dg = env.device_grid()
    dg.set_motions(*motion_table(env, 3))
    ws = engine.get_workspace(dg, 1, _lib.W_F64)
    path_cap = int(min(dg.pcells, 6 * sum(dims) + 4096))

    def run(step):
        engine.upload_queries(ws, [start], [goal], 3)
        timed(torch, events, lambda: engine.sssp(ws, _lib.ASTAR, 0))
        engine.extract_paths(ws, path_cap)
        return 4   # init, k_sssp, finish, extract

    def account():
        settled = int(engine.count_settled(ws, bounded=True).sum().item())
        return settled * 26, settled, settled * ALG_BYTES["f64"], float(ws.cost[0].item())

    def verify():
        engine.check_field_status(ws)
        cost = float(ws.cost[0].item())
        if gold:
            want = float.fromhex(gold["cost_hex"])
            assert abs(cost - want) <= REL * want, f"c4 cost {cost!r} != oracle's {want!r}"
            return {"cost": cost, "oracle_cost": want, "rel_err": abs(cost - want) / want}
        return {"cost": cost, "oracle_cost": None}

    def e2e(step):
        env.update(host_occ)                      # H2D of the occupancy from pinned memory, one byte per cell
        cost, path, expand = pmp.AStar3D(start, goal, env).plan()   # H2D query, D2H cost + path
        return cost, host_occ.numel() + 24, 32 + 4 * len(path)

    host_rows = torch.from_numpy(workloads.pack_rows(occ).view(np.int32)).pin_memory()

    def e2e_packed():
        """the same for a caller that holds the map as bits (Grid3D.update_packed: 1/8 of the upload)"""
        ts = []
        for _ in range(5):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            env.update_packed(host_rows)
            cost, path, expand = pmp.AStar3D(start, goal, env).plan()
            ts.append((time.perf_counter() - t0) * 1e3)
        env.update(host_occ)
        return {"ms_per_step": float(np.median(ts)), "h2d_bytes_per_step": host_rows.numel() * 4 + 24, "cost": cost}

    return dict(run=run, events=events, account=account, verify=verify, e2e=e2e, e2e_packed=e2e_packed, name=name, dtype="f64", queries=1,
                scaling="weak", parallelism="independent replicas", stats=lambda: ws.stats.cpu().numpy(), K=26,
                cfg={"grid": list(dims), "connectivity": 26, "heuristic": "euclidean", "free_cells": int((~occ).sum())})


def build_c3(args, ctx, nq=None):
    torch, pmp, workloads = ctx["torch"], ctx["pmp"], ctx["workloads"]
    from python_motion_planning_b200 import engine, _lib, batch
    from python_motion_planning_b200.environment import motion_table
    rank, world = ctx["rank"], ctx["world"]
    events = []
    n = (args.n if args.workload == "c3" else 0) or DEFAULT_N["c3"]
    nq_all = nq or args.queries or 4096
    occ, starts, goals = workloads.random2d(n, nq=nq_all, seed=0)
    q0, q1 = batch.shard_range(nq_all, rank, world)          # the queries of this rank: no communication
    starts, goals = starts[q0:q1], goals[q0:q1]
    env = pmp.Grid(n, n)
    env.update(occ)
    dg = env.device_grid()
    dg.set_motions(*motion_table(env, 2))
    ws = engine.get_workspace(dg, q1 - q0, _lib.W_F64)
    path_cap = int(min(dg.pcells, 8 * n + 4096))
    gold = golden().get("c3_1024_first64") if n == 1024 and nq_all == 4096 else None

    def run(step):
        engine.upload_queries(ws, starts, goals, 2)
        timed(torch, events, lambda: engine.sssp(ws, _lib.DIJKSTRA, 0))
        engine.extract_paths(ws, path_cap)
        return 4

    def account():
        settled = int(engine.count_settled(ws, bounded=True, use_h=False).sum().item())
        c = ws.cost.cpu().numpy()
        return settled * 8, settled, settled * ALG_BYTES["f64"], float(np.nanmean(c[np.isfinite(c)]))

    def verify():
        engine.check_field_status(ws)
        if gold and rank == 0 and q1 - q0 >= 64:
            want = np.array([float.fromhex(h) for h in gold["cost_hex"]])
            got = ws.cost.cpu().numpy()[:64]
            reach = np.isfinite(want)
            assert np.array_equal(np.isfinite(got), reach), "c3: reachability differs from the oracle"
            assert np.allclose(got[reach], want[reach], rtol=REL, atol=0), "c3: costs differ from the oracle"
            return {"queries_checked_vs_oracle": 64}
        return {"queries_checked_vs_oracle": 0}

    def e2e(step):
        res = pmp.plan_batch(env, starts, goals, algo="dijkstra", want_paths=True)
        return float(np.isfinite(res.costs).sum()), res.h2d_bytes, res.d2h_bytes

    return dict(run=run, events=events, account=account, verify=verify, e2e=e2e, dtype="f64", queries=q1 - q0,
                scaling="strong" if world > 1 else "weak", parallelism=f"queries sharded x{world}, no communication",
                stats=lambda: ws.stats.cpu().numpy(), K=8, name=workload_name("c3", n, nq_all),
                cfg={"grid": [n, n], "connectivity": 8, "queries": nq_all, "queries_this_rank": q1 - q0})


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle (C restatement of the reference) on the host cores
# ------------------------------------------------------------------------------------------------
def cpu_sample(wl, args, workloads, budget_s=12.0):
    """bounded sample of the workload on all host threads; returns dict for the JSON line"""
    from oracle import oracle
    threads = oracle.max_threads()
    t0 = time.perf_counter()
    if wl in ("c4", "c2"):
        if wl == "c4":
            n = min(args.n or 512, args.cpu_n or 192)
            occ = workloads.boxes3d(n, seed=0)
            sample = f"{threads} A* queries (one per host thread) on a {n}^3 copy of the workload (same generator; the reference's data structures do not hold 512^3), start (1,1,1), goals on the far faces"
            rng = np.random.default_rng(5)
            free_far = np.argwhere(~occ[n // 2:, n // 2:, n // 2:]) + n // 2
            goals = free_far[rng.integers(0, len(free_far), size=threads)]
            starts = np.tile(np.array([[1, 1, 1]]), (threads, 1))
        else:
            occ, _ = workloads.door3d()
            sample = f"{threads * 8} repeats of the c2 query, one per host thread"
            starts = np.tile(np.array([[1, 1, 1]]), (threads * 8, 1))
            goals = np.tile(np.array([[24, 18, 14]]), (threads * 8, 1))
        K, algo = 26, oracle.ASTAR
    elif wl == "c3":
        n = (args.n if args.workload == "c3" else 0) or 1024
        occ, starts, goals = workloads.random2d(n, nq=max(threads, 16), seed=0)
        sample = f"{len(starts)} of the Dijkstra queries on the {n}^2 grid, one per host thread"
        K, algo = 8, oracle.DIJKSTRA
    else:
        blk, _ = workloads.replicated_boxes3d(base=128, reps=1)
        occ = workloads.replicated_slab(blk, (128, 128, 128), 0, 128)
        starts = np.tile(np.array([[1, 1, 1]]), (threads, 1))
        goals = np.zeros((threads, 3), dtype=np.int32)      # an obstacle: runs to exhaustion
        sample = (f"{threads} full fields (Dijkstra3D to exhaustion, 6 motions) on the 128^3 block the workload "
                  f"replicates, one per host thread")
        K, algo = 6, oracle.DIJKSTRA
    motions = oracle.MOTIONS_3D if occ.ndim == 3 else oracle.MOTIONS_2D
    mg = oracle.motion_costs(motions)
    if K == 6:
        keep = mg == 1.0
        motions, mg = motions[keep], mg[keep]
    reps, edges, t_total = 0, 0, 0.0
    while True:
        t1 = time.perf_counter()
        cost, nexp, plen, flags = oracle.plan_batch(occ, starts, goals, algo=algo, motions=motions,
                                                    motion_g=mg, n_threads=threads)
        t_total += time.perf_counter() - t1
        edges += int(nexp.sum()) * K
        reps += 1
        if t_total > budget_s or reps >= 64:
            break
    v = edges / t_total / 1e9
    return {"value": v, "unit": "GTEPS", "cores": threads, "kind": "port", "per_thread_gteps": v / threads,
            "sample": sample + f"; {reps} repetition(s), {t_total:.1f} s",
            "queries_per_s": reps * len(starts) / t_total, "seconds": t_total, "reps": reps,
            "edges_per_rep": edges // reps, "queries_per_rep": len(starts), "setup_s": time.perf_counter() - t0 - t_total}


def reference_python_sample():
    """The UNMODIFIED reference planners (staged under the git-ignored oracle/_ref by oracle/stage_ref.py),
    single-threaded as the reference is, `plan()` only as examples/benchmark.py:79-83 times it:
    C1 README AStar and C2 3d_example AStar3D, median of 10 / 5.  None when the staging is absent."""
    try:
        from oracle import ref_python
        return ref_python.time_c1_c2()
    except Exception as e:       # not staged on this box
        return {"kind": "reference-python", "unavailable": str(e)[:200]}


# ------------------------------------------------------------------------------------------------
# measurement of one workload: `value` leg (inputs resident in HBM) and `e2e` leg (host buffers)
# ------------------------------------------------------------------------------------------------
def measure(W, ctx, steps, warmup, dist, sample_clocks=True):
    torch, flush = ctx["torch"], ctx["flush"]

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(warmup):
        W["run"](i)
    torch.cuda.synchronize()
    edges, settled, abytes, check = W["account"]()
    facts = W["verify"]()
    sampler = ClockSampler(ctx["local_rank"], ctx["uuid"]) if sample_clocks else None
    barrier()
    evs, launches = [], 0
    n0 = len(W["events"])
    for i in range(steps):
        flush.fill_(i & 0xFF)                      # L2 flush between timed iterations
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        launches += W["run"](i)
        e1.record()
        if sampler:
            sampler.sample()                       # the step's kernels are still in flight
        evs.append((e0, e1))
    barrier()
    clocks = sampler.stop() if sampler else None
    t_dev = sum(a.elapsed_time(b) for a, b in evs) / 1e3
    kms = [a.elapsed_time(b) for a, b in W["events"][n0:]]
    t_kernel = (sum(kms) / len(kms) / 1e3) if kms else t_dev / steps
    st = W["stats"]() if "stats" in W else None
    # ---- end-to-end leg through the public API with host buffers ----
    for i in range(2):
        W["e2e"](i)
    barrier()
    t0 = time.perf_counter()
    h2d = d2h = 0
    for i in range(steps):
        _, h2d, d2h = W["e2e"](i)
    barrier()
    t_e2e = time.perf_counter() - t0
    tt = torch.tensor([t_dev, t_e2e, t_kernel], dtype=torch.float64, device="cuda")
    ee = torch.tensor([float(edges), float(settled), float(abytes)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)      # max over ranks
        dist.all_reduce(ee, op=dist.ReduceOp.SUM)      # whole-job work
    return dict(t_dev=float(tt[0]), t_e2e=float(tt[1]), t_kernel=float(tt[2]), t_kernel_local=t_kernel,
                edges=float(ee[0]), settled=float(ee[1]), abytes=float(ee[2]), abytes_local=abytes, check=check,
                facts=facts, clocks=clocks, launches=launches, h2d=int(h2d), d2h=int(d2h), stats=st, steps=steps)


def kernel_stats(st, dtype, clocks):
    st = [int(v) for v in st]
    if dtype == "u32":      # pmp_bfs (warp-per-tile level solver): see include/pmp_b200.h
        mhz = ((clocks or {}).get("sm_mhz") or 1965.0)
        na = max(st[1], 1)
        return {"solver": "k_bfs (one warp per tile)", "tile_activations": st[1], "idle_activations": st[13],
                "cells_lowered": st[2], "edges_relaxed": st[2] * 6, "levels_swept": st[12], "passes": st[14],
                "requeues": st[5], "remote_pushes": st[7], "kernel_ms": st[8] / 1e6,
                "rank0_busy_ms": (st[16] - st[9]) / 1e6 if len(st) > 17 and st[16] else None,
                "nvlink_bytes_read_rank0": st[17] * 128 if len(st) > 17 else 0,
                "us_per_activation_per_warp": {"wait_for_work": st[4] / na / mhz, "setup": st[6] / na / mhz,
                                               "sweep": st[11] / na / mhz, "extract_store": st[3] / na / mhz,
                                               "publish": st[0] / na / mhz}}
    mhz = ((clocks or {}).get("sm_mhz") or 1965.0)
    nt = max(st[1], 1)
    return {"rounds": st[0], "tile_activations": st[1], "cells_expanded": st[2], "edges_relaxed": st[3],
            "worklist_max": st[4], "sub_iterations": st[12],
            "us_per_activation": {"load": st[5] / nt / mhz, "settle": st[6] / nt / mhz, "publish": st[7] / nt / mhz},
            "kernel_ms": st[8] / 1e6, "cta0_ms": {"scan": st[9] / 1e6, "process": st[10] / 1e6, "barrier_wait": st[11] / 1e6}}


def sub_record(W, m, world):
    return {"workload": W["name"], "dtype": W["dtype"], "parallelism": W["parallelism"],
            "value": m["edges"] * m["steps"] / m["t_dev"] / 1e9, "unit": "GTEPS",
            "ms_per_step": m["t_dev"] / m["steps"] * 1e3, "search_ms": m["t_kernel"] * 1e3,
            "queries_per_s": W["queries"] * m["steps"] / m["t_dev"] if W["scaling"] == "weak" or world == 1
            else None,
            "e2e_ms_per_step": m["t_e2e"] / m["steps"] * 1e3,
            "e2e_value": m["edges"] * m["steps"] / m["t_e2e"] / 1e9,
            "settled_cells_per_step": m["settled"], "check_value": m["check"], "verified": m["facts"],
            "roofline_frac": m["abytes_local"] / m["t_kernel_local"] / 1e9 / measured_peak()[0],
            "h2d_bytes_per_step": m["h2d"], "d2h_bytes_per_step": m["d2h"]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c5", choices=["c5", "c4", "c3", "c2"])
    ap.add_argument("--n", type=int, default=0, help="grid edge (default: the BASELINE.json size)")
    ap.add_argument("--queries", type=int, default=0)
    ap.add_argument("--cpu-n", type=int, default=0, help="grid edge of the CPU sample (c4)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extras", action="store_true", help="skip the c4 / c3 sub-records")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    from python_motion_planning_b200 import workloads
    n_head = args.n or DEFAULT_N[args.workload]

    if args.impl == "reference":
        if rank != 0:
            return
        ts = []
        out = None
        for i in range(args.warmup + args.steps):
            out = cpu_sample(args.workload, args, workloads, budget_s=0.0)     # one repetition per step
            if i >= args.warmup:
                ts.append(out)
        secs = sum(o["seconds"] for o in ts)
        edges = sum(o["edges_per_rep"] for o in ts)
        val = edges / secs / 1e9
        line = {"impl": "reference", "metric": "edge relaxations/s", "value": val, "unit": "GTEPS",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": secs / max(len(ts), 1) * 1e3, "higher_is_better": True,
                "scaling": "strong" if args.workload == "c5" else "weak", "vs_baseline": None,
                "dtype": "f64" if args.workload != "c5" else "u32",
                "data": "synthetic", "config": {"workload": workload_name(args.workload, n_head, world=args.gpus),
                                                  "sample": out["sample"]},
                "cpu_baseline": {"value": val, "unit": "GTEPS", "cores": out["cores"], "kind": "port",
                                 "per_thread_gteps": val / out["cores"], "sample": out["sample"]},
                "queries_per_s": sum(o["queries_per_rep"] for o in ts) / secs,
                "e2e": {"value": val, "unit": "GTEPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return

    import torch
    import python_motion_planning_b200 as pmp
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device; there is no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    try:
        uuid = "GPU-" + str(torch.cuda.get_device_properties(local_rank).uuid)
    except Exception:
        uuid = None
    ctx = dict(torch=torch, pmp=pmp, workloads=workloads, rank=rank, world=world, dist=dist, local_rank=local_rank,
               uuid=uuid, flush=torch.empty(256 << 20, dtype=torch.uint8, device="cuda"))   # > 126 MB L2
    solo = dict(ctx, rank=0, world=1, dist=None)

    builders = {"c5": build_c5, "c4": build_c4, "c3": build_c3, "c2": lambda a, c: build_c4(a, c, small=True)}
    head_ctx = ctx if args.workload in ("c5", "c3") else solo     # c4 / c2 do not shard: replicas
    W = builders[args.workload](args, head_ctx)
    m = measure(W, head_ctx, args.steps, args.warmup, dist)
    # c4 / c2 replicas: every rank ran the same query on its own GPU
    rep = world if args.workload in ("c4", "c2") else 1

    extras = {}
    if not args.no_extras and args.workload == "c5":
        del W["run"], W["e2e"]                     # release nothing yet; the field stays for kernel_stats
        torch.cuda.empty_cache()
        if rank == 0:
            W4 = build_c4(args, solo, n=512)
            extras["c4"] = sub_record(W4, measure(W4, solo, max(3, args.steps // 2), 3, None, sample_clocks=False), 1)
            extras["c4"]["e2e_packed_input"] = W4["e2e_packed"]()
            del W4
            torch.cuda.empty_cache()
        if dist is not None:
            dist.barrier()
        W3 = build_c3(args, ctx)
        m3 = measure(W3, ctx, 3, 3, dist, sample_clocks=False)
        extras["c3"] = sub_record(W3, m3, world)
        extras["c3"]["queries_per_s"] = 4096 * m3["steps"] / m3["t_dev"]
        extras["c3"]["e2e_queries_per_s"] = 4096 * m3["steps"] / m3["t_e2e"]
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peak()
    steps = args.steps
    value = m["edges"] * rep * steps / m["t_dev"] / 1e9 if rep > 1 else m["edges"] * steps / m["t_dev"] / 1e9
    achieved = m["abytes_local"] / m["t_kernel_local"] / 1e9      # this rank's dominant kernel
    traffic = None
    prof = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(prof):
        try:
            traffic = json.load(open(prof)).get(f"{args.workload}_{n_head}" if args.workload != "c2" else "c2")
            if isinstance(traffic, dict):
                traffic = traffic.get("bytes")
        except Exception:
            pass
    dom = "k_bfs" if W["dtype"] == "u32" else "k_sssp"
    line = {
        "metric": "edge relaxations/s", "value": value, "unit": "GTEPS", "n_gpus": world,
        "steps": steps, "warmup": args.warmup, "ms_per_step": m["t_dev"] / steps * 1e3,
        "higher_is_better": True, "scaling": W["scaling"], "vs_baseline": None, "dtype": W["dtype"],
        "data": "synthetic",
        "config": dict(W["cfg"], workload=W["name"], l2="flushed between timed steps (256 MiB write); field > L2",
                       parallelism=W["parallelism"] if rep == 1 else f"independent replicas x{world}",
                       settled_cells_per_step=m["settled"] * rep, check_value=m["check"], verified=m["facts"]),
        "queries_per_s": W["queries"] * rep * steps / m["t_dev"],
        "clocks": m["clocks"],
        "e2e": {"value": m["edges"] * rep * steps / m["t_e2e"] / 1e9, "unit": "GTEPS",
                "h2d_bytes_per_step": m["h2d"], "d2h_bytes_per_step": m["d2h"],
                "ms_per_step": m["t_e2e"] / steps * 1e3,
                "queries_per_s": W["queries"] * rep * steps / m["t_e2e"]},
        "gpu_launches": m["launches"],
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                     "kernel": f"{dom}: the search phase of a step, CUDA events on its stream "
                               f"(workspace reset + init + {dom} + finish; {dom} is > 95 % of it); rank 0's share at N > 1",
                     "kernel_ms": m["t_kernel_local"] * 1e3,
                     "algorithmic_bytes_per_step": m["abytes_local"],
                     "traffic_source": "profiles/traffic.json: dram__bytes_read+write of one ncu --set full capture of this "
                                       "workload, per launch (static; the date is in the file)" if traffic else None},
    }
    if m["stats"] is not None:
        line["kernel_stats"] = kernel_stats(m["stats"], W["dtype"], m["clocks"])
    if extras:
        line["extras"] = extras
    if not args.no_cpu and world == 1:
        line["cpu_baseline"] = cpu_sample(args.workload, args, workloads)
        line["cpu_baseline_reference_python"] = reference_python_sample()
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
