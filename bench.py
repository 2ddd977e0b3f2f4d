#!/usr/bin/env python
"""bench.py — body-substeps/s of the rigid-body step on many_pyramids (BASELINE.json configs[1]).

  python bench.py --gpus N --steps K --warmup W          B200 arm: libb2cuda.so through the C-ABI
  python bench.py --impl reference ...                    reference arm: the CPU restatement (oracle/) of the C# step

One JSON line on stdout (rank 0). N>1 (torchrun): every rank steps its own many_pyramids world (400 independent
pyramid worlds per GPU, no data-path collective; NCCL only reduces the timing and batch statistics) -> weak scaling.
The line's "batched" object is BASELINE.json configs[4] at every N: 8192 independent worlds behind b2cuda_batch_*, the world
range split over the N ranks (strong scaling), world-steps/s.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

DT, SUBSTEPS = 1.0 / 60.0, 4
WORKLOAD = "many_pyramids"
# algorithmic bytes (SURVEY.md §8d / DESIGN.md): per constraint per pass and per body per substep
B_WARM, B_SOLVE, B_RELAX, B_IV, B_IP, B_STORE = 140, 260, 260, 64, 48, 72


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=60)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=WORKLOAD)
    ap.add_argument("--cpu-steps", type=int, default=120, help="bounded CPU-baseline sample (steps after 20 warm-up)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--worlds", type=int, default=8192, help="batched sub-record (and the batched_pyramids workload): total independent worlds, sharded over ranks")
    ap.add_argument("--no-batched", action="store_true", help="skip the batched-worlds sub-record")
    return ap.parse_args()


def build_scene(world, name, worlds=None, rank=0, world_size=1):
    from box2d_b200 import scenes, batch
    if name == "batched_pyramids":
        b, e = batch.shard_range(worlds, rank, world_size)
        scenes.batched_pyramids(world, e - b, first=b)
    else:
        scenes.SCENES[name](world)


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self.stop_flag = index, [], set(), None, False

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown", nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                     nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown", nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
            while not self.stop_flag:
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                for bit, n in names.items():
                    if r & bit:
                        self.reasons.add(n)
                time.sleep(0.05)
        except Exception as e:  # NVML missing: report that instead of inventing numbers
            self.reasons.add("nvml_unavailable:" + type(e).__name__)

    def result(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def workload_config(workload, bodies):
    """`config` of the JSON line: the SAME dictionary in both arms (the driver compares them)."""
    scene = {"many_pyramids": "20x20 grid of independent 10-row box pyramids on static grounds, one world",
             "large_pyramid": "100-row box pyramid on a static ground", "tumbler": "20 000 mixed circles/capsules/boxes in a rotating kinematic box",
             "joint_grid": "100x100 revolute-joint grid plus contacts", "batched_pyramids": "10-row pyramids 15 m apart in one world"}.get(workload, workload)
    return {"workload": workload, "scene": scene, "dt": DT, "substeps": SUBSTEPS, "bodies": int(bodies),
            "l2": "b200 arm: 384 MB flush buffer written between timed steps, per-step CUDA-event times summed; reference arm: CPU, not applicable",
            "host_input": "none per step: WorldAPI.Step(dt, substeps) takes two scalars, the scene is uploaded once before the first step (h2d_bytes_per_step = 0 by construction)"}


def dotnet_probe():
    """the reference itself (C#, .NET 9) is timed only where a .NET SDK exists on the box; this image has none."""
    import shutil
    for exe in ("dotnet", "mono", "csc"):
        if shutil.which(exe):
            return exe
    return None


def cpu_oracle_run(workload, warm, steps):
    """Single-threaded CPU restatement of the reference step (oracle/, kind "port"): the reference's WorkerCount=1 path."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_backend import OracleBackend
    from box2d_b200 import api, scenes
    w = api.World(OracleBackend(), scenes.world_def())
    build_scene(w, workload)
    for _ in range(warm):
        w.Step(DT, SUBSTEPS)
    bodies = w.GetCounters().awakeBodyCount
    t0 = time.perf_counter()
    for _ in range(steps):
        w.Step(DT, SUBSTEPS)
    sec = time.perf_counter() - t0
    w.Destroy()
    return bodies * SUBSTEPS * steps / sec, sec, bodies


def reference_arm(args, rank):
    if rank != 0:
        return
    steps = max(1, min(args.steps, args.cpu_steps))
    warm = max(3, min(args.warmup, 20))
    value, sec, bodies = cpu_oracle_run(args.workload, warm, steps)
    runtime = dotnet_probe()
    sample = f"{steps} steps of {args.workload} ({bodies} bodies) after {warm} warm-up steps, 1 thread"
    line = {"impl": "reference", "metric": "body-substeps/sec on many_pyramids", "value": value, "unit": "body-substeps/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": warm, "ms_per_step": 1000.0 * sec / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(args.workload, bodies),
            "reference_runtime": (f"{runtime} found, but building /root/reference needs the source tree and NuGet, neither is on this box" if runtime
                                  else "unavailable on this box: no dotnet / mono / csc (probed at run time); the CPU restatement (oracle/) of the reference's single-threaded step is timed instead"),
            "cpu_baseline": {"value": value, "unit": "body-substeps/s", "cores": 1, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "body-substeps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


KERNELS_PER_STEP = 3    # k_step_front (cooperative), k_solver_islands, k_finalize_phase (cooperative)
MIN_TIMED_S, MAX_REPEATS = 0.5, 25


def median(xs):
    s = sorted(xs)
    return s[len(s) // 2] if len(s) % 2 else 0.5 * (s[len(s) // 2 - 1] + s[len(s) // 2])


def measured_traffic(workload):
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture (profiles/traffic.json), or None."""
    try:
        rec = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[workload]
        return rec["dram_bytes_per_launch"], rec["source"]
    except Exception:
        return None, None


def batched_record(args, rank, world_size, local_rank, dist, barrier):
    """BASELINE.json configs[4]: `--worlds` independent 10-row pyramids behind ONE b2cuda_batch handle per rank, the world range split
    over the ranks (strong scaling), stepped by one launch per step; NCCL reduces the batch statistics from device memory."""
    import torch
    from box2d_b200 import api, scenes, batch
    b0, b1 = batch.shard_range(args.worlds, rank, world_size)
    n = b1 - b0
    bt = api.Batch(api.CudaBackend(device=local_rank), n, scenes.world_def())
    rec = api.Snapshot()                                                   # one recorded 10-row pyramid, replayed into every world by ONE C-ABI call each
    scenes.add_pyramid(rec, 10)
    for w in bt.worlds:
        rec.upload(w)
    K, W = args.steps, max(args.warmup, 3)
    for _ in range(W):
        bt.Step(DT, SUBSTEPS)
    bt.Synchronize()
    barrier()
    reps, total = [], 0.0
    while total < MIN_TIMED_S * 1000.0 and len(reps) < MAX_REPEATS:      # the K-step region, repeated; the working set (GBs) is far larger than L2
        bt.TimerStart()
        for _ in range(K):
            bt.Step(DT, SUBSTEPS)
        ms = bt.TimerStop()
        reps.append(ms)
        total += ms
    if dist is not None:                                                  # every rank repeats the same number of times
        t = torch.tensor([len(reps)], dtype=torch.int64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        reps = reps[:int(t.item())]
    barrier()
    dev = torch.zeros(8, dtype=torch.float64, device="cuda")
    t0 = time.perf_counter()
    for _ in range(K):                                                    # end to end: step + the statistics of all worlds read back by the host
        bt.Step(DT, SUBSTEPS)
        st = bt.Stats(dev.data_ptr())
    e2e_ms = (time.perf_counter() - t0) * 1000.0
    torch.cuda.synchronize()
    (ms_max, e2e_max), _ = batch.reduce_timing([median(reps), e2e_ms], 0.0, dist, "cuda")
    if dist is not None:                                                  # the only collective: batch statistics, reduced from device memory
        sums, mx, mn = dev[0:5].clone(), dev[5:6].clone(), dev[6:7].clone()
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(mn, op=dist.ReduceOp.MIN)
        flags = dev[7:8].clone()
        dist.all_reduce(flags, op=dist.ReduceOp.MAX)
        dev = torch.cat([sums, mx, mn, flags])
    v = dev.cpu().tolist()
    bt.Destroy()
    if v[7] != 0.0:
        raise SystemExit(f"bench.py: batched worlds report device error flags {v[7]}")
    return {"workload": f"{args.worlds} independent 10-row pyramids (b2cuda_batch_*), world range split over {world_size} GPU(s)", "worlds": args.worlds,
            "worlds_per_gpu": n, "scaling": "strong", "steps": K, "warmup": W, "repeats": len(reps), "ms_per_step": ms_max / K,
            "world_steps_per_s": args.worlds * K / (ms_max * 1e-3), "body_substeps_per_s": v[0] * SUBSTEPS * K / (ms_max * 1e-3),
            "e2e_world_steps_per_s": args.worlds * K / (e2e_max * 1e-3), "e2e_d2h_bytes_per_step": 64, "gpu_launches_per_step": 3, "gpu_launches_note": "front, warp-per-world island solver, finalize: three kernels replayed as ONE cuda graph launch per step",
            "l2": "per-GPU working set (every world's own trees, contacts and solver arrays) exceeds the 126 MB L2 many times over",
            "dram_traffic_bytes_per_launch": (measured_traffic("batched_8192")[0] if (args.worlds == 8192 and world_size == 1) else None),
            "stats": {"awake_bodies": int(v[0]), "sum_v2": v[1], "touching_contacts": int(v[2]), "max_speed": v[5], "min_y": v[6]}}


def b200_arm(args, rank, world_size, local_rank):
    import torch
    from box2d_b200 import api, scenes
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; libb2cuda has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world_size > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    w = api.World(api.CudaBackend(device=local_rank), scenes.world_def())
    w.EnableGraphLaunch(True)   # b2cuda_world_enable_graph_launch: the step's three kernels replayed as one graph launch (back to back, free of host launch jitter)
    build_scene(w, args.workload, args.worlds, rank, world_size)
    flush = torch.empty(96 * 1024 * 1024, dtype=torch.float32, device="cuda")   # 384 MB > 126 MB L2
    K, W = args.steps, max(args.warmup, 3)
    for _ in range(W):
        w.Step(DT, SUBSTEPS)
    w.Synchronize()
    c = w.GetCounters()
    bodies, touching = c.awakeBodyCount, c.touchingContactCount
    w.GetProfile()

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def same_repeats(n):
        if dist is None:
            return n
        t = torch.tensor([n], dtype=torch.int64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return int(t.item())

    # ---- timed region 1: device-resident, L2 flushed between steps; per-step CUDA-event time on the world's stream. The K-step region is
    #      repeated until at least MIN_TIMED_S of step time has been measured; the reported figure is the median region. ----
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    phase_names = ("pairs", "collide", "prepareConstraints", "solveConstraints", "transforms", "treeRebuild")
    regions, total_ms = [], 0.0
    while total_ms < MIN_TIMED_S * 1000.0 and len(regions) < MAX_REPEATS:
        step_ms, phases, done = 0.0, [0.0] * 6, 0
        while done < K:
            n = min(32, K - done)
            for _ in range(n):
                flush.zero_()
                torch.cuda.synchronize()
                w.TimerStart()                 # a CUDA event pair on the world's stream around the step's launches (TimerStop waits for the second)
                w.Step(DT, SUBSTEPS)
                step_ms += w.TimerStop()
            p = w.GetProfile()                 # phase and kernel times of the same steps: %globaltimer stamps the kernels write (no events between the kernels)
            for i, f in enumerate(phase_names):
                phases[i] += getattr(p, f) * n
            done += n
        regions.append((step_ms, phases))
        total_ms += step_ms
    regions = sorted(regions[:same_repeats(len(regions))], key=lambda r: r[0])
    step_ms, phases = regions[len(regions) // 2]
    solver_ms = phases[3]
    repeats = len(regions)
    barrier()
    # ---- timed region 2: back-to-back steps, state L2-resident as in real use (reported beside the flushed number) ----
    warm_runs = []
    for _ in range(min(repeats, 5)):
        w.TimerStart()
        for _ in range(K):
            w.Step(DT, SUBSTEPS)
        warm_runs.append(w.TimerStop())
    warm_ms = median(warm_runs)
    barrier()
    # ---- timed region 3: end to end through the C-ABI with host buffers: step + bulk body read-back + contact events ----
    w.EnableGraphLaunch(False)                 # a caller that reads results after every step is latency-bound: direct launches start sooner than a graph launch
    for _ in range(max(3, min(W, 8))):         # untimed warm-up of the read-back path itself (first call packs all bodies, loads the kernel)
        w.Step(DT, SUBSTEPS)
        w.GetBodyMoveEvents(copy=False)
        w.GetContactEvents()
    w.Synchronize()
    e2e_runs, d2h = [], 0
    for _ in range(same_repeats(min(repeats, 9))):
        barrier()
        t0 = time.perf_counter()
        for _ in range(K):
            w.Step(DT, SUBSTEPS)
            b = w.GetBodyMoveEvents(copy=False)    # WorldAPI.GetBodyEvents: 24 B per awake body, filled by FinalizeBodies, one D2H into pinned memory behind the step
            ev = w.GetContactEvents()
            d2h = b.nbytes + sum(e.nbytes for e in ev)
        torch.cuda.synchronize()
        e2e_runs.append(time.perf_counter() - t0)
    e2e_s = median(e2e_runs)
    sampler.stop_flag = True
    sampler.join(timeout=2)
    c2 = w.GetCounters()
    if c2.errorFlags & ~16:
        raise SystemExit(f"bench.py: device error flags 0x{c2.errorFlags:x}")
    from box2d_b200 import batch
    (step_ms_max, warm_ms_max, e2e_ms_max), total_units = batch.reduce_timing([step_ms, warm_ms, e2e_s * 1000.0], bodies * SUBSTEPS * K, dist, "cuda")
    bodies_read = w.GetBodyEvents(copy=False)
    stats = batch.reduce_stats(batch.batch_stats(bodies_read[bodies_read["flags"] != 0]), dist, "cuda")   # NCCL is used for batch statistics only
    w.Destroy()
    del flush
    batched = None
    if not args.no_batched and args.workload == WORKLOAD:
        batched = batched_record(args, rank, world_size, local_rank, dist, barrier)
    if rank != 0:
        return
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6455.9))
    algo_bytes = SUBSTEPS * (bodies * (B_IV + B_IP) + touching * (B_WARM + B_SOLVE + B_RELAX)) + touching * B_STORE
    avg_solver_ms = solver_ms / K
    achieved = algo_bytes / (avg_solver_ms * 1e-3) / 1e9 if avg_solver_ms > 0 else 0.0
    traffic, traffic_src = measured_traffic(args.workload)
    line = {"metric": "body-substeps/sec on many_pyramids", "value": total_units / (step_ms_max * 1e-3), "unit": "body-substeps/s", "n_gpus": world_size,
            "steps": K, "warmup": W, "ms_per_step": step_ms_max / K, "higher_is_better": True, "scaling": "strong" if args.workload == "batched_pyramids" else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": workload_config(args.workload, bodies),
            "detail": {"bodies_per_gpu": bodies, "touching_contacts_per_gpu": touching, "parallelism": f"replicas x{world_size} (no data-path collective)",
                       "launch": "one cuda graph launch per step for value (b2cuda_world_enable_graph_launch), direct launches for e2e",
                       "timer": "CUDA event pair on the world's stream around each step; b2cuda_world_timer_start holds the stream 60 us before its event so the step's launches are queued when the region opens (device time, not host enqueue latency)",
                       "repeats": repeats, "timed_step_ms_total": total_ms, "protocol": f"the {K}-step region repeated {repeats}x (>= {MIN_TIMED_S} s of step time or {MAX_REPEATS} regions); value, ms_per_step and phase_ms are the median region"},
            "value_l2_resident": total_units / (warm_ms_max * 1e-3), "ms_per_step_l2_resident": warm_ms_max / K,
            "phase_ms": dict(zip(("pairs", "collide", "color_sort_prepare", "solver_kernel", "finalize_refit", "tree_rebuild_in_collide"), [x / K for x in phases])),
            "roofline": {"bound": "hbm", "kernel": "k_solver_islands (+ grid-path solver inside k_step_front)", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "frac_is": "ALGORITHMIC bytes (SURVEY 8d per-constraint / per-body figures) over the measured copy peak; the lanes and body rows stay on chip, so physical DRAM traffic (`traffic`) is far below it and the kernel is bound by its barrier / dependent-issue chain, not by HBM",
                         "traffic": traffic, "traffic_source": traffic_src,
                         "algorithmic_bytes_per_launch": algo_bytes, "avg_launch_ms": avg_solver_ms,
                         "avg_launch_ms_is": "end of k_step_front to start of k_finalize_phase, %globaltimer stamps the two kernels write on the world's stream during the timed steps (brackets the solver launch like an event pair, without the ~10 us per step that four cudaEventRecord calls cost); value / ms_per_step are CUDA-event pairs around each step",
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback"},
            "e2e": {"value": total_units / (e2e_ms_max * 1e-3), "unit": "body-substeps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": int(d2h),
                    "note": "Step(dt,substeps) has no per-step host input; scene uploaded once at the first step; each step reads back the body move events (transform + id of every awake body, the reference's GetBodyEvents payload) and the contact begin/end/hit events to host"},
            "gpu_launches": (KERNELS_PER_STEP + (1 if args.workload == "batched_pyramids" else 0)) * K * repeats, "clocks": sampler.result(),
            "batch_stats": {"bodies": int(stats[0]), "sum_v2": float(stats[1]), "max_speed": float(stats[2]), "min_y": float(stats[3])}}
    if batched is not None:
        line["batched"] = batched
    if world_size == 1 and not args.no_cpu:
        cpu_warm = min(W, 60)
        v, sec, nb = cpu_oracle_run(args.workload, cpu_warm, args.cpu_steps)
        line["cpu_baseline"] = {"value": v, "unit": "body-substeps/s", "cores": 1, "kind": "port",
                                "sample": f"{args.cpu_steps} steps of {args.workload} ({nb} bodies) after {cpu_warm} warm-up steps, 1 thread, {sec:.1f} s"}
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        reference_arm(args, rank)
        return
    b200_arm(args, rank, world_size, local_rank)
    if world_size > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
