#!/usr/bin/env python
"""bench.py -- CANNet step throughput on B200 (contract: see the task statement / DESIGN.md section 6).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload train|infer] [--impl reference]

One JSON line on rank 0.  ``value`` = whole-job images/sec with inputs resident in HBM;
``e2e`` = the same metric through the public API (``model(img)`` / ``train_step``) with pinned
host inputs copied in and the result read back every step; ``roofline`` = the dominant kernel
(tcgen05 implicit-GEMM conv) against the measured bf16 peak; ``cpu_baseline`` = the oracle's
torch-CPU port of the reference on this box's host cores (bounded sample).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "can-distributed-pytorch_b200"))

WORKLOADS = {
    # BASELINE.json configs[3] (the config the metric "train imgs/sec @3x768x1024" is quoted on)
    "train": dict(name="train_768x1024_b4_bf16", batch=4, H=768, W=1024, train=True),
    # BASELINE.json configs[1]
    "infer": dict(name="infer_768x1024_b16_bf16", batch=16, H=768, W=1024, train=False),
}
FWD_GFLOP_768 = 733.36  # SURVEY.md section 8d, algorithmic, per image


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("CAN_B200_WORKLOAD", "auto"))
    ap.add_argument("--batch", type=int, default=0, help="override per-GPU batch")
    ap.add_argument("--precision", default="bf16")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def pick_workload(args):
    name = args.workload
    if name == "auto":
        from can_b200 import engine
        name = "train" if hasattr(engine, "backward") else "infer"
    wl = dict(WORKLOADS[name])
    if args.batch:
        wl["batch"] = args.batch
        wl["name"] = wl["name"].replace("_b4_", f"_b{args.batch}_").replace("_b16_", f"_b{args.batch}_")
    return wl


# ------------------------------------------------------------------------------------------
class ClockSampler:
    """SM clock / throttle reasons sampled DURING the timed region: NVML in a side thread (every 10 ms), falling
    back to an `nvidia-smi -lms 100` child process when the NVML binding is missing."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
    BITS = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}

    def __init__(self, index):
        self.index, self.rows, self.proc, self.nvml, self.stop_flag = index, [], None, None, False
        self.samples, self.mask, self.max_mhz = [], 0, None

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(visible.split(",")[self.index]) if visible and visible.replace(",", "").isdigit() else self.index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
            self.t = threading.Thread(target=self._poll, daemon=True)
            self.t.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _poll(self):
        nv = self.nvml
        while not self.stop_flag:
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM)))
                self.mask |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
            except Exception:
                pass
            time.sleep(0.01)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.nvml is not None:
            self.stop_flag = True
            self.t.join(timeout=1)
            sm = sorted(self.samples)
            return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz,
                    "reasons": [n for n, bit in self.BITS.items() if self.mask & bit], "samples": len(sm), "source": "nvml"}
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = sorted(float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 3 + i and r[3 + i].lower() == "active" for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm), "source": "nvidia-smi"}


def measured_peaks():
    """(burst bf16 TFLOP/s, sustained bf16 TFLOP/s, HBM GB/s, where the numbers come from)"""
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        burst = d.get("bf16_tflops")
        return burst, d.get("bf16_tflops_sustained", burst), d.get("hbm_gbs"), "measured (MEASURED_PEAKS.json)"
    return 1650.0, 1400.0, 6650.0, "fallback (B200_PROFILING.md)"


# ------------------------------------------------------------------------------------------
def cpu_reference_images_per_sec(wl, n_timed, warm, batch, budget_s):
    """The reference's CPU PyTorch path (oracle torch-CPU port; /root/reference is not on this box) on `batch` images of
    the workload's shape per step, all host threads.  Stops early when `budget_s` of wall time would be exceeded (never
    below two timed steps) and reports the steps it actually ran."""
    import torch
    from oracle import cannet_torch_cpu as T
    from oracle import harness
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    sd = harness.make_state_dict("he")
    x = harness.make_input(batch, wl["H"], wl["W"])
    gt = harness.make_gt(batch, wl["H"], wl["W"])
    bufs = {}
    times = []
    t_start = time.perf_counter()
    warm_run = 0
    while len(times) < n_timed:
        t0 = time.perf_counter()
        if wl["train"]:
            T.train_step(x, gt, sd, bufs, lr=1e-10)
        else:
            with torch.no_grad():
                T.forward(x, sd)
        dt = time.perf_counter() - t0
        if warm_run < warm:
            warm_run += 1
            warm = min(warm, max(1, int(0.2 * budget_s / dt)))   # slow steps: warm-up takes at most a fifth of the budget
        else:
            times.append(dt)
        if len(times) >= 2 and (time.perf_counter() - t_start) + dt > budget_s:
            break
    n_run = len(times)
    times.sort()
    med = times[len(times) // 2]
    what = "fwd+bwd+SGD" if wl["train"] else "forward"
    sample = f"{warm_run}+{n_run} steps of {batch} x {wl['H']}x{wl['W']} image(s) ({what}), fp32, median"
    return dict(ips=batch / med, ms=1e3 * med, cores=cores, threads=torch.get_num_threads(), sample=sample, steps=n_run,
                warmup=warm_run)


def run_reference(args, wl):
    """Reference arm: the same config (per-GPU batch, H, W) through the reference's CPU path, rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = cpu_reference_images_per_sec(wl, n_timed=max(2, args.steps), warm=max(1, args.warmup), batch=wl["batch"],
                                     budget_s=float(os.environ.get("CAN_B200_REF_BUDGET_S", "150")))
    line = {
        "impl": "reference", "metric": "images_per_sec", "value": r["ips"], "unit": "img/s", "n_gpus": args.gpus,
        "steps": r["steps"], "warmup": r["warmup"], "steps_requested": args.steps, "warmup_requested": args.warmup,
        "ms_per_step": r["ms"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["name"], "per_gpu_batch": wl["batch"], "global_batch": wl["batch"], "H": wl["H"],
                   "W": wl["W"], "parallelism": "cpu (one process, rank 0 only; other ranks exit)",
                   "note": "reference CPU path = oracle torch-CPU port; steps capped by a wall-time budget"},
        "cpu_baseline": {"value": r["ips"], "unit": "img/s", "cores": r["threads"], "kind": "port", "sample": r["sample"]},
        "e2e": {"value": r["ips"], "unit": "img/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------
def main():
    args = parse()
    wl = pick_workload(args)
    if args.impl == "reference":
        run_reference(args, wl)
        return

    import torch
    import torch.distributed as dist
    from can_b200 import ops
    from can_b200.model import CANNet
    from oracle import harness  # weights/input generators only (test infrastructure, not on the timed path)

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    B, H, W = wl["batch"], wl["H"], wl["W"]
    model = CANNet(load_weights=True)
    model.load_state_dict(harness.make_state_dict("he"))
    model = model.to(dev)
    model.precision = args.precision

    # resident inputs: distinct seed per rank
    x_host = harness.make_input(B, H, W, seed=1234 + rank).pin_memory()
    gt_host = harness.make_gt(B, H, W, seed=4321 + rank).pin_memory()
    x_dev, gt_dev = x_host.to(dev), gt_host.to(dev)

    # parity of the benchmarked model on the benchmarked shape (outside every timed region): the first image's
    # density map against the port run on this box's CPU (rank 0 only)
    parity = None
    if rank == 0 and os.environ.get("CAN_B200_BENCH_PARITY", "1") != "0":
        from oracle import cannet_torch_cpu as T
        torch.set_num_threads(os.cpu_count() or 1)
        with torch.no_grad():
            ref0 = T.forward(x_host[:1], harness.make_state_dict("he"))
            got0 = model(x_dev[:1]).cpu()
        parity = {"rel_l2": harness.rel_l2(got0, ref0), "rel_count": harness.rel_count(got0, ref0),
                  "against": "oracle torch-CPU port (fp32), first image of the benchmarked batch, same weights",
                  "tolerance": {"rel_count": 1e-2 if args.precision == "bf16" else 1e-3,
                                "rel_l2": 3e-2 if args.precision == "bf16" else 1e-4}}
        parity["ok"] = bool(parity["rel_l2"] < parity["tolerance"]["rel_l2"] and
                            parity["rel_count"] < parity["tolerance"]["rel_count"])

    if wl["train"]:
        from can_b200.utils.train_eval_utils import FusedTrainer
        # lr: the reference's default 1e-7 diverges to NaN within three steps on these random-init
        # (He) weights -- in the reference's own CPU path too (tools/gpu_check_train_losses.py) -- so the
        # synthetic bench uses 1e-10 (x world size, train.py:25); the arithmetic per step is identical.
        trainer = FusedTrainer(model, lr=1e-10 * world, momentum=0.95)

        def step_resident():
            return trainer.step(x_dev, gt_dev)

        from can_b200.utils.train_eval_utils import DevicePrefetcher

        loss_host = [torch.empty(1).pin_memory(), torch.empty(1).pin_memory()]
        loss_done = [torch.cuda.Event(), torch.cuda.Event()]

        def e2e_batches(n):
            # every step: H2D of that step's pinned-host inputs (on the prefetch stream, overlapped with the previous
            # step's compute) + D2H read of the step's loss.  The loss of step k is copied to pinned memory
            # asynchronously and read on the host while step k + 1 is being enqueued (as train_one_epoch reads its finite
            # flag): every step's loss is read inside the timed region, none of the reads idles the device.
            k = 0
            for xd, gd in DevicePrefetcher(((x_host, gt_host) for _ in range(n)), dev):
                loss = trainer.step(xd, gd)
                loss_host[k & 1].copy_(loss, non_blocking=True)
                loss_done[k & 1].record()
                if k > 0:
                    loss_done[(k - 1) & 1].synchronize()
                    yield float(loss_host[(k - 1) & 1])
                k += 1
            if k > 0:
                loss_done[(k - 1) & 1].synchronize()
                yield float(loss_host[(k - 1) & 1])

        def step_e2e(n=1):
            for _ in e2e_batches(n):
                pass
    else:
        def step_resident():
            with torch.no_grad():
                out = model(x_dev)
                return ops.mse_count(out, gt_dev, want_grad=False)[1]

        from can_b200.utils.train_eval_utils import DevicePrefetcher

        def step_e2e(n=1):
            with torch.no_grad():
                for xd, gd in DevicePrefetcher(((x_host, gt_host) for _ in range(n)), dev):
                    ops.mse_count(model(xd), gd, want_grad=False)[1].cpu()   # D2H read of the per-image counts

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    rank_ms = []   # device time of the last timed region on every rank (N > 1)

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            every = [torch.zeros(1, device=dev) for _ in range(world)]
            dist.all_gather(every, torch.tensor([ms], device=dev))
            rank_ms[:] = [float(v) for v in every]
            ms = max(rank_ms)
        return ms

    for _ in range(max(3, args.warmup)):
        step_resident()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ops.LAUNCHES = 0
    ms = timed(step_resident, args.steps)
    launches = ops.LAUNCHES
    clocks = sampler.stop()
    ms_ranks = [v / args.steps for v in rank_ms]
    ips = world * B * args.steps / (ms * 1e-3)

    step_e2e(2)
    ms_e2e = timed(lambda: step_e2e(args.steps), 1)
    ips_e2e = world * B * args.steps / (ms_e2e * 1e-3)
    h2d = x_host.numel() * 4 + gt_host.numel() * 4
    d2h = 4 if wl["train"] else 4 * B

    # roofline of the dominant kernel: a second pass of the same steps with an event pair around
    # every conv_tc / wgrad launch (kept out of the timed region so it cannot perturb `value`)
    ops.PROFILE = []
    if wl["train"]:
        trainer.use_graph = False   # the event pairs live in the Python wrappers: this pass runs the same kernels eagerly
    else:
        os.environ["CAN_B200_GRAPH_INFER"] = "0"   # same for the inference graph of CANNet.forward
    # (every kernel alone on the device: the weight gradients stay on the main stream for this pass, else a data gradient's
    # event pair would also time the weight-gradient kernel running beside it on the side stream)
    side_env = os.environ.get("CAN_B200_SIDE")
    os.environ["CAN_B200_SIDE"] = "0"
    torch.cuda.synchronize()
    prof_sampler = ClockSampler(local_rank)
    prof_sampler.start()
    for _ in range(min(args.steps, 5)):
        step_resident()
    torch.cuda.synchronize()
    prof_clocks = prof_sampler.stop()
    if side_env is None:
        os.environ.pop("CAN_B200_SIDE", None)
    else:
        os.environ["CAN_B200_SIDE"] = side_env
    prof, ops.PROFILE = ops.PROFILE, None
    npass = max(1, min(args.steps, 5))
    k_ms = sum(rec[0].elapsed_time(rec[1]) for rec in prof)
    k_flop = sum(rec[2] for rec in prof)
    peak_burst, peak_sust, peak_hbm, peak_src = measured_peaks()
    # which peak applies: a run whose SM clock stayed near its maximum (short, not yet power-capped) is measured against
    # the burst figure, a run whose clock dropped under sustained load against the sustained one
    # (the clock that counts is the one sampled during the pass the kernel times come from)
    pc = prof_clocks if prof_clocks.get("sm_mhz") else clocks
    near_max = bool(pc.get("sm_mhz") and pc.get("sm_max_mhz") and pc["sm_mhz"] >= 0.9 * pc["sm_max_mhz"])
    peak_tf = peak_burst if near_max else peak_sust
    peak_kind = "burst" if near_max else "sustained"
    # dominant kernel = conv_tc_kernel<256,1,STORE> on the 512->512 3x3 layers at 96x128 (conv4_2, conv4_3,
    # backend.2, backend.4 and their data gradients): algorithmic FLOPs per launch / mean launch duration
    dom = [rec for rec in prof if rec[3].startswith("conv m0 512->512 t9")]
    dom_ms = sum(rec[0].elapsed_time(rec[1]) for rec in dom)
    dom_flop = sum(rec[2] for rec in dom)
    achieved = (dom_flop / 1e12) / (dom_ms * 1e-3) if dom_ms > 0 else 0.0
    traffic = None
    for tname in ("r02_dominant_kernel_traffic.json", "r01_dominant_kernel_traffic.json"):
        tpath = os.path.join(ROOT, "profiles", tname)
        if traffic is None and os.path.exists(tpath):  # dram__bytes_read.sum + dram__bytes_write.sum of one launch, ncu --set full
            traffic = json.load(open(tpath)).get(f"batch{B}")
    all_tf = (k_flop / 1e12) / (k_ms * 1e-3) if k_ms > 0 else 0.0
    step_tf = (k_flop / npass / 1e12) / (ms / args.steps * 1e-3)   # all tensor FLOPs of a step / the timed step
    roofline = {"bound": "tensor", "kernel": "conv_tc_kernel<256,1,STORE,PAIR=2> 512->512 3x3 @96x128 (tcgen05 cta_group::2 implicit GEMM, M=256 per pair, A from one halo tile per channel chunk)",
                "achieved": achieved, "peak": peak_tf, "unit": "TFLOP/s", "frac": achieved / peak_tf if peak_tf else None,
                "peak_kind": peak_kind, "frac_burst": achieved / peak_burst, "frac_sustained": achieved / peak_sust,
                "traffic": traffic, "peak_source": peak_src,
                "whole_step": {"achieved": step_tf, "frac": step_tf / peak_tf, "frac_burst": step_tf / peak_burst,
                               "what": "algorithmic FLOPs of every tensor-core launch of a step / the timed step"},
                "flop_per_launch": dom_flop / max(1, len(dom)), "us_per_launch": 1e3 * dom_ms / max(1, len(dom)),
                "launches_per_step": len(dom) // npass, "share_of_step": (dom_ms / npass) / (ms / args.steps),
                "all_tensor_kernels": {"achieved": all_tf, "frac_burst": all_tf / peak_burst,
                                       "launches_per_step": len(prof) // npass,
                                       "share_of_step": (k_ms / npass) / (ms / args.steps)},
                "timing": "CUDA events around every launch in a separate eager pass of the same steps, all kernels on one stream",
                "clocks_in_timing_pass": {k: pc.get(k) for k in ("sm_mhz", "sm_max_mhz", "reasons")}}

    final = step_resident()
    torch.cuda.synchronize()
    final_ok = bool(torch.isfinite(final).all())

    # data parallel: every rank must hold the same parameters after the same steps (checksum over the flat arena)
    ranks_in_sync = None
    if world > 1 and wl["train"]:
        cs = torch.stack([trainer.arena.double().sum(), trainer.arena.double().abs().sum()])
        lo, hi = cs.clone(), cs.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        ranks_in_sync = bool(torch.equal(lo, hi))

    # the fp32-parity path (the one that meets the 1e-4 / 1e-3 tolerance) on the same workload, a short separate run
    fp32_path = None
    if args.precision == "bf16" and os.environ.get("CAN_B200_BENCH_FP32", "1") != "0":
        m32 = CANNet(load_weights=True)
        m32.load_state_dict(harness.make_state_dict("he"))
        m32 = m32.to(dev)
        m32.precision = "fp32"
        if wl["train"]:
            from can_b200.utils.train_eval_utils import FusedTrainer as _FT
            t32 = _FT(m32, lr=1e-10 * world, momentum=0.95)
            f32_step = lambda: t32.step(x_dev, gt_dev)
        else:
            def f32_step():
                with torch.no_grad():
                    return ops.mse_count(m32(x_dev), gt_dev, want_grad=False)[1]
        for _ in range(3):
            f32_step()
        n32 = 5
        ms32 = timed(f32_step, n32)
        fp32_path = {"value": world * B * n32 / (ms32 * 1e-3), "unit": "img/s", "steps": n32, "warmup": 3,
                     "ms_per_step": ms32 / n32, "dtype": "f32(split-bf16: hi+lo planes, 3 products)"}
        del m32
    if rank == 0:
        line = {
            "metric": "images_per_sec", "value": ips, "unit": "img/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": args.precision if args.precision == "bf16" else "f32(split-bf16)",
            "data": "synthetic",
            "config": {"workload": wl["name"], "per_gpu_batch": B, "global_batch": B * world, "H": H, "W": W,
                       "l2_policy": "inputs+activations per step exceed the 126 MB L2 (no explicit flush)",
                       "parallelism": f"dp{world}"},
            "clocks": clocks,
            "e2e": {"value": ips_e2e, "unit": "img/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": launches,
            "roofline": roofline,
            "model_tflops": ips * FWD_GFLOP_768 * (3 if wl["train"] else 1) / 1e3 / world,
            "result_finite": final_ok,
            "parity": parity,
            "fp32_path": fp32_path,
            "ranks_in_sync": ranks_in_sync,
            "ms_per_step_ranks": ms_ranks or None,
            ("final_loss" if wl["train"] else "final_counts_sum"): float(final.sum()),
        }
        if world == 1 and not args.no_cpu_baseline:
            # bounded sample: ONE image of the workload's shape per step (the reference arm times the full batch)
            r = cpu_reference_images_per_sec(wl, n_timed=7, warm=2, batch=1, budget_s=40.0)
            line["cpu_baseline"] = {"value": r["ips"], "unit": "img/s", "cores": r["threads"], "kind": "port",
                                    "sample": r["sample"]}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
