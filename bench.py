#!/usr/bin/env python
"""bench.py — body-substeps/s of the rigid-body step on many_pyramids (BASELINE.json configs[1]).

  python bench.py --gpus N --steps K --warmup W          B200 arm: libb2cuda.so through the C-ABI
  python bench.py --impl reference ...                    reference arm: the CPU restatement (oracle/) of the C# step

One JSON line on stdout (rank 0). N>1 (torchrun): every rank steps its own many_pyramids world (400 independent
pyramid worlds per GPU, no data-path collective; NCCL only reduces the timing and batch statistics) -> weak scaling.
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
    ap.add_argument("--worlds", type=int, default=8192, help="batched_pyramids: total independent worlds, sharded over ranks")
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
    sample = f"{steps} steps of {args.workload} ({bodies} bodies) after {warm} warm-up steps, 1 thread"
    line = {"impl": "reference", "metric": "body-substeps/sec on many_pyramids", "value": value, "unit": "body-substeps/s", "n_gpus": args.gpus,
            "steps": steps, "warmup": warm, "ms_per_step": 1000.0 * sec / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": {"workload": args.workload, "dt": DT, "substeps": SUBSTEPS, "bodies": bodies,
                                                               "note": "C# runtime absent on this box: CPU restatement (oracle/) of the reference's single-threaded step"},
            "cpu_baseline": {"value": value, "unit": "body-substeps/s", "cores": 1, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "body-substeps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


KERNELS_PER_STEP = 3    # k_step_front (cooperative), k_solver_islands, k_finalize_phase (cooperative)


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
    if world_size > 1:          # several ranks share one host: one graph launch per step keeps the kernels back to back under launch jitter
        w.EnableGraphLaunch(True)
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

    # ---- timed region 1: device-resident, L2 flushed between steps; per-step CUDA-event time on the world's stream ----
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    step_ms, solver_ms, phases = 0.0, 0.0, [0.0] * 6
    done = 0
    while done < K:
        n = min(32, K - done)
        for _ in range(n):
            flush.zero_()
            torch.cuda.synchronize()
            w.Step(DT, SUBSTEPS)
            w.Synchronize()
        p = w.GetProfile()
        step_ms += p.step * n
        solver_ms += p.solveConstraints * n
        for i, f in enumerate(("pairs", "collide", "prepareConstraints", "solveConstraints", "transforms", "treeRebuild")):
            phases[i] += getattr(p, f) * n
        done += n
    barrier()
    # ---- timed region 2: back-to-back steps, state L2-resident as in real use (reported beside the flushed number) ----
    w.TimerStart()
    for _ in range(K):
        w.Step(DT, SUBSTEPS)
    warm_ms = w.TimerStop()
    barrier()
    # ---- timed region 3: end to end through the C-ABI with host buffers: step + bulk body read-back + contact events ----
    if world_size > 1:
        w.EnableGraphLaunch(False)             # a caller that reads results after every step is latency-bound: direct launches start ~7 us sooner
    for _ in range(max(3, min(W, 8))):         # untimed warm-up of the read-back path itself (first call packs all bodies, loads the kernel)
        w.Step(DT, SUBSTEPS)
        w.GetBodyMoveEvents(copy=False)
        w.GetContactEvents()
    w.Synchronize()
    barrier()
    t0 = time.perf_counter()
    d2h = 0
    for _ in range(K):
        w.Step(DT, SUBSTEPS)
        b = w.GetBodyMoveEvents(copy=False)    # WorldAPI.GetBodyEvents: 24 B per awake body, filled by FinalizeBodies, one D2H into pinned memory behind the step
        ev = w.GetContactEvents()
        d2h = b.nbytes + sum(e.nbytes for e in ev)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    sampler.stop_flag = True
    sampler.join(timeout=2)
    prof = w.GetProfile()
    c2 = w.GetCounters()
    if c2.errorFlags & ~16:
        raise SystemExit(f"bench.py: device error flags 0x{c2.errorFlags:x}")
    from box2d_b200 import batch
    (step_ms_max, warm_ms_max, e2e_ms_max), total_units = batch.reduce_timing([step_ms, warm_ms, e2e_s * 1000.0], bodies * SUBSTEPS * K, dist, "cuda")
    bodies_read = w.GetBodyEvents(copy=False)
    stats = batch.reduce_stats(batch.batch_stats(bodies_read[bodies_read["flags"] != 0]), dist, "cuda")   # NCCL is used for batch statistics only
    if rank != 0:
        w.Destroy()
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
    line = {"metric": "body-substeps/sec on many_pyramids", "value": total_units / (step_ms_max * 1e-3), "unit": "body-substeps/s", "n_gpus": world_size,
            "steps": K, "warmup": W, "ms_per_step": step_ms_max / K, "higher_is_better": True, "scaling": "strong" if args.workload == "batched_pyramids" else "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": args.workload, "dt": DT, "substeps": SUBSTEPS, "bodies_per_gpu": bodies, "touching_contacts_per_gpu": touching,
                       "worlds_per_gpu": "20x20 independent 10-row pyramids in one device world", "parallelism": f"replicas x{world_size} (no data-path collective)", "launch": "cuda graph per step for value, direct launches for e2e" if world_size > 1 else "direct launches",
                       "l2": "384 MB flush buffer written between timed steps; per-step CUDA-event time summed"},
            "value_l2_resident": total_units / (warm_ms_max * 1e-3), "ms_per_step_l2_resident": warm_ms_max / K,
            "phase_ms": dict(zip(("pairs", "collide", "color_sort_prepare", "solver_kernel", "finalize_refit", "tree_rebuild_in_collide"), [x / K for x in phases])),
            "roofline": {"bound": "hbm", "kernel": "k_solver_islands (+ grid-path solver inside k_step_front)", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": 9350912 if args.workload == "many_pyramids" else None,   # dram read+write per launch, profiles/r01d_summary.txt
                        
                         "algorithmic_bytes_per_launch": algo_bytes, "avg_launch_ms": avg_solver_ms,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback"},
            "e2e": {"value": total_units / (e2e_ms_max * 1e-3), "unit": "body-substeps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": int(d2h),
                    "note": "Step(dt,substeps) has no per-step host input; scene uploaded once at the first step; each step reads back the body move events (transform + id of every awake body, the reference's GetBodyEvents payload) and the contact begin/end/hit events to host"},
            "gpu_launches": (KERNELS_PER_STEP + (1 if args.workload == "batched_pyramids" else 0)) * K, "clocks": sampler.result(),
            "batch_stats": {"bodies": int(stats[0]), "sum_v2": float(stats[1]), "max_speed": float(stats[2]), "min_y": float(stats[3])}}
    if world_size == 1 and not args.no_cpu:
        cpu_warm = min(W, 60)
        v, sec, nb = cpu_oracle_run(args.workload, cpu_warm, args.cpu_steps)
        line["cpu_baseline"] = {"value": v, "unit": "body-substeps/s", "cores": 1, "kind": "port",
                                "sample": f"{args.cpu_steps} steps of {args.workload} ({nb} bodies) after {cpu_warm} warm-up steps, 1 thread, {sec:.1f} s"}
    print(json.dumps(line), flush=True)
    w.Destroy()


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
