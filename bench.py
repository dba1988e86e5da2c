"""bench.py — throughput of the off-grid attention hot path on B200 (driver contract: one JSON line on stdout).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch B] [--workload NAME]

A "step" is one forward+backward pass of the workload over one synthetic batch (random-init weights, random images,
no dataset).  `value` = images/s with the batch already resident in HBM; `e2e` = the same metric through the
public API with the batch in pinned HOST memory (host->device copy of the inputs and a device->host read of the
loss inside the timed region).  `roofline` describes the dominant CLUSTEN kernel, timed live with CUDA events;
`cpu_baseline` is the oracle (the reference's pure-PyTorch gather path, oracle/) on the host cores over a bounded
sample.  `--impl reference` times that CPU path alone (rank 0 only).

Workloads (see `bench_workloads.py`): the default is the one BASELINE.json's metric is quoted on.
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

_OUT = sys.stdout   # main() replaces it with a private duplicate of the original stdout

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "ml-autofocusformermod_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, index, period=0.05):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop_evt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksEventReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksEventReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksEventReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksEventReasonSwPowerCap: "sw_power_cap",
                 nv.nvmlClocksEventReasonHwPowerBrakeSlowdown: "hw_power_brake",
                 nv.nvmlClocksEventReasonApplicationsClocksSetting: "applications_clocks_setting"}
        while not self._stop_evt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                break
            time.sleep(self.period)

    def finish(self):
        self._stop_evt.set()
        self.join(timeout=2)
        return {"sm_mhz": (statistics.median(self.samples) if self.samples else None),
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons), "samples": len(self.samples)}


def physical_gpu_index(local_rank):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except Exception:
            return local_rank
    return local_rank


def timed_loop(fn, steps, world, device):
    """barrier + synchronize, K steps between CUDA events on the current stream, synchronize; max over ranks (ms)."""
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        fn(i)
    e1.record()
    torch.cuda.synchronize(device)
    ms = e0.elapsed_time(e1)
    if world > 1:
        dist.barrier()
        t = torch.tensor([ms], device=device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms


def run_reference(args, rank, world):
    """The reference's own CPU implementation of the path (oracle port: pure-PyTorch gather expressions of
    clusten/test_*_kernel.py), all host threads, bounded sample per step.  Rank 0 only."""
    if rank != 0:
        return
    import bench_workloads as W
    wl = W.get(args.workload)
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sample_b = args.ref_batch or getattr(wl, "ref_batch", 4)
    ref = wl.cpu_reference(sample_b)
    for _ in range(max(1, min(args.warmup, 2))):
        ref()
    steps = max(1, min(args.steps, args.ref_steps))
    t0 = time.perf_counter()
    for _ in range(steps):
        ref()
    dt = time.perf_counter() - t0
    v = sample_b * steps / dt
    line = {"impl": "reference", "metric": wl.metric, "value": v, "unit": wl.unit, "n_gpus": args.gpus,
            "steps": steps, "warmup": min(args.warmup, 2), "ms_per_step": dt / steps * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(wl.config(sample_b, 1), note="CPU oracle port of the reference path; bounded sample"),
            "cpu_baseline": {"value": v, "unit": wl.unit, "cores": cores, "kind": "port",
                             "sample": "%d images per step x %d steps, fp32, torch %s, %d threads" %
                                       (sample_b, steps, torch.__version__, cores)},
            "e2e": {"value": v, "unit": wl.unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), file=_OUT, flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=0, help="images per GPU (0 = the workload's default)")
    ap.add_argument("--workload", default="")
    ap.add_argument("--ref-batch", type=int, default=0, help="images per CPU-reference step (0 = the workload's: 4 unless it says otherwise)")
    ap.add_argument("--ref-steps", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    # the contract is ONE JSON line on stdout.  Libraries write to file descriptor 1 behind Python's back (NCCL prints
    # its version banner there at NCCL_DEBUG=VERSION whatever NCCL_DEBUG_FILE says), so fd 1 is pointed at stderr for
    # the whole run and the JSON line goes to a private duplicate of the original stdout.
    global _OUT
    sys.stdout.flush()
    _OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU path for --impl ours)")
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group(backend="nccl", device_id=device)

    import bench_workloads as W
    from affb200 import _native as N
    wl = W.get(args.workload)
    B = args.batch or wl.default_batch
    torch.manual_seed(0 + rank)  # per-rank seed as the reference does (main.py:428)
    run = wl.build(B, device, world)

    # ---- warm-up (also builds the per-stage index plans) ----------------------------------------------------
    host = [run.make_host_batch(i) for i in range(2)]
    dev_batch = run.to_device(host[0])
    torch.cuda.synchronize(device)
    for i in range(args.warmup):
        run.step(dev_batch)
    torch.cuda.synchronize(device)

    # ---- device-resident timing -----------------------------------------------------------------------------
    sampler = ClockSampler(physical_gpu_index(local_rank))
    sampler.start()
    l0 = N.LIB.affb200_launch_count()
    cuprof = os.environ.get("AFFB200_CUPROF") == "1"   # `ncu --profile-from-start off`: only the timed steps are captured
    if cuprof:
        torch.cuda.profiler.start()
    ms = timed_loop(lambda i: run.step(dev_batch), args.steps, world, device)
    if cuprof:
        torch.cuda.profiler.stop()
    launches = (N.LIB.affb200_launch_count() - l0) // max(args.steps, 1)
    if getattr(run, "gstep", None) is not None:
        launches = run.gstep.kernels_per_step  # recorded once at capture, launched by every graph replay
    clocks = sampler.finish()
    imgs = B * world * args.steps
    value = imgs / (ms * 1e-3)

    # ---- end-to-end timing: pinned host batch -> device, step, loss -> host ---------------------------------
    loss_host = torch.empty((), dtype=torch.float32).pin_memory()

    # Every step's batch is copied from pinned host memory inside the timed region; batch i+1 travels on a copy stream
    # while step i computes (affb200/prefetch.py), the first batch of the region is copied before its step with nothing
    # to hide behind.  AFFB200_PREFETCH=0: copy on the compute stream right before the step, as the reference's loop does.
    prefetch = os.environ.get("AFFB200_PREFETCH", "1") != "0"
    from affb200.prefetch import DevicePrefetcher
    pf = DevicePrefetcher(device)

    def make_e2e(n_steps):
        def e2e_step(i):
            if not prefetch:
                d = run.to_device(host[i % len(host)])
                loss = run.step(d)
            else:
                if i == 0:
                    pf.submit(host[0])
                if i + 1 < n_steps:
                    pf.submit(host[(i + 1) % len(host)])
                d = pf.next()
                loss = run.step(d)
                pf.release()
            loss_host.copy_(loss.detach().float(), non_blocking=True)
        return e2e_step

    warm = make_e2e(2)
    for i in range(2):
        warm(i)
    ms_e2e = timed_loop(make_e2e(args.steps), args.steps, world, device)
    e2e_value = imgs / (ms_e2e * 1e-3)
    h2d = sum(t.numel() * t.element_size() for t in host[0])

    graph_fallbacks = None
    if getattr(run, "gstep", None) is not None:
        # every replay is verified between the two graphs (4-byte read); a replay whose speculated index plans did not
        # hold is redone eagerly inside the timed region.  Reported so that a slow run can be told from a clean one.
        graph_fallbacks = run.gstep.fallbacks

    # ---- roofline of the dominant kernel, timed live --------------------------------------------------------
    peak, peak_src = peaks()
    roof = run.roofline(peak, peak_src) if rank == 0 else None
    if world > 1:
        dist.barrier()

    cpu_base = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        torch.set_num_threads(cores)
        ref_b = args.ref_batch or getattr(wl, "ref_batch", 4)
        ref = wl.cpu_reference(ref_b)
        ref()
        t0 = time.perf_counter()
        n = 0
        while n < args.ref_steps and (time.perf_counter() - t0) < 30.0:
            ref()
            n += 1
        dt = time.perf_counter() - t0
        cpu_base = {"value": ref_b * n / dt, "unit": wl.unit, "cores": cores, "kind": "port",
                    "sample": "%d images per step x %d steps of the same workload, fp32, %d torch threads" %
                              (ref_b, n, cores)}

    if rank == 0:
        line = {"metric": wl.metric, "value": value, "unit": wl.unit, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": wl.dtype_name, "data": "synthetic",
                "config": wl.config(B, world),
                "e2e": {"value": e2e_value, "unit": wl.unit, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                        "ms_per_step": ms_e2e / args.steps,
                        "input_pipeline": "double-buffered copy stream" if prefetch else "copy on the compute stream"},
                "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "cpu_baseline": cpu_base}
        if graph_fallbacks is not None:
            line["config"]["graph_fallbacks"] = graph_fallbacks
        if isinstance(roof, dict) and roof.get("fused_stages") is not None:
            line["config"]["fused_stages"] = roof.pop("fused_stages")   # which attention kernels each stage ran
        print(json.dumps(line), file=_OUT, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
