#!/usr/bin/env python
"""Benchmark of the worker-side inference hot path (BASELINE.json metric: images/sec at
"batch 400" for AlexNet and ResNet18).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle port)

A step = one 400-image query through the hot path (uint8 HWC 224x224 images -> top-1 index and
probability).  Primary workload (value / e2e / roofline): BASELINE.json configs[1], AlexNet batch
400 over range 0-3999; the same measurements for ResNet18 are reported under "resnet18".
With N > 1 (torchrun, one rank per GPU) every rank runs its own 400-image sub-range per step (the
coordinator's range split gives each of N ranks 400 of a 400*N-image query: weak scaling) and the
per-image results are gathered to rank 0 over NCCL inside the timed step.

  value : images/s, inputs already resident in HBM, CUDA-event timed on the launching stream,
          barrier + synchronize on both sides, max over ranks.
  e2e   : images/s through the C-ABI call with HOST buffers (idn_infer_u8hwc): pinned uint8 in,
          H2D + kernels + D2H of the results inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "-distributed-machine-learning-system_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

BATCH = 400
FLOP_PER_IMG = {"alexnet": 1428376960, "resnet18": 3628146688}
PRE_BYTES_PER_IMG = 451584                                            # 224x224 source, bf16 out
STEP_FLOPS = {   # algorithmic FLOPs per image of each tensor-core plan step (2*M*N*K, logical)
    "alexnet": {"features.0+relu": 140553600, "features.3+relu": 447897600,
                "features.6+relu": 224280576, "features.8+relu": 299040768,
                "features.10+relu": 199360512, "classifier.1+relu": 75497472,
                "classifier.4+relu": 33554432, "classifier.6": 8192000},
    "resnet18": dict(
        [("conv1+bn1+relu", 236027904)] +
        [(f"layer1.{b}.{c}", 231211008) for b in (0, 1) for c in ("conv1", "conv2+add")] +
        [(f"layer{l}.0.conv1", 115605504) for l in (2, 3, 4)] +
        [(f"layer{l}.0.downsample", 12845056) for l in (2, 3, 4)] +
        [(f"layer{l}.{b}.{c}", 231211008) for l in (2, 3, 4)
         for (b, c) in ((0, "conv2+add"), (1, "conv1"), (1, "conv2+add"))]),
}
WORKLOAD = {"alexnet": "AlexNet batch 400 over range 0-3999 (uint8 224x224 HWC -> top-1 idx/prob)",
            "resnet18": "ResNet18 batch 400 over range 0-79999 sharded by rank (uint8 224x224 HWC -> top-1)"}
# dram__bytes_read+write per launch of the dominant kernel from the committed `ncu --set full`
# capture (profiles/), keyed by (model, step); None until a capture exists for that kernel.
NCU_TRAFFIC_BYTES: dict = {
    # profiles/r01f_alexnet_pair_ncu_full.csv: conv_pair_kernel<192,5>, 43.77 MB read + 59.92 MB written
    ("alexnet", "features.3+relu"): 103693568,
    # profiles/r01c_stem_pool2 (stem_kernel with the max-pool folded in): 326.9 MB read + 113.8 MB written
    ("resnet18", "conv1+bn1+relu"): 440713728,
}


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(hbm=float(p["hbm_gbs"]), tf_burst=float(p["bf16_tflops"]),
                    tf_sustained=float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), source="measured")
    # /opt/skills/guides/B200_PROFILING.md fallback
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


class ClockSampler:
    """Samples SM clock and throttle reasons of one GPU while the timed region runs (NVML)."""
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap",
               0x8: "hw_slowdown", 0x10: "sync_boost", 0x20: "sw_thermal_slowdown",
               0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nv = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self._h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self._nv = None

    def _run(self):
        nv = self._nv
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self._h, nv.NVML_CLOCK_SM)))
                mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self._h))
                for bit, name in self.REASONS.items():
                    if mask & bit and name != "gpu_idle":
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.002)

    def __enter__(self):
        if self._nv is not None:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread is not None:
            self._thread.join()

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def make_batches(n_batches: int, seed: int):
    from idunno_b200.synth import noise_images
    return [noise_images(BATCH, seed=seed + b) for b in range(n_batches)]


def bench_model(model, rank, world, device, steps, warmup, peaks, dist, do_cpu_baseline):
    import torch
    from idunno_b200 import Engine
    from idunno_b200.synth import make_state_dict

    sd = make_state_dict(model, 0, "default")
    eng = Engine(model, sd, device=device, max_batch=BATCH)
    stream = torch.cuda.Stream(device=device)      # a real (non-NULL) stream: NULL selects the handle's own
    torch.cuda.set_stream(stream)
    eng.set_stream(stream.cuda_stream)

    n_pool = 4                                     # 4 x 60.2 MB of distinct uint8 input > 126 MB L2
    host_batches = make_batches(n_pool, seed=100 * rank + (0 if model == "alexnet" else 50))
    img_bytes = 224 * 224 * 3
    offsets = (np.arange(BATCH, dtype=np.int64) * img_bytes)
    hw = np.tile(np.asarray([[224, 224]], dtype=np.int32), (BATCH, 1))
    dev_batches = [torch.from_numpy(b.reshape(-1)).to(f"cuda:{device}") for b in host_batches]
    pinned = [torch.from_numpy(b.reshape(-1)).pin_memory() for b in host_batches]
    d_idx = torch.zeros(BATCH, dtype=torch.int32, device=f"cuda:{device}")
    d_prob = torch.zeros(BATCH, dtype=torch.float32, device=f"cuda:{device}")
    gathered, pending_gathers = None, []
    if world > 1:     # two sets: the gather of step i is still in flight while step i+1 computes
        gathered = [[torch.zeros(BATCH, 2, dtype=torch.float32, device=f"cuda:{device}") for _ in range(world)]
                    for _ in range(2)]

    def drain_gathers(keep=0):
        while len(pending_gathers) > keep:
            pending_gathers.pop(0).wait()          # the timing stream waits for that collective

    def step_device(i):
        b = dev_batches[i % n_pool]
        eng.infer_device(b.data_ptr(), offsets, hw, d_idx.data_ptr(), d_prob.data_ptr())
        if world > 1:   # result return: (idx, prob) per image to every rank's view (8 bytes per image, NCCL).
            # Asynchronous: the collective starts when this step's kernels are done, the next step's kernels
            # do not wait for it; every gather is waited for before the timed region ends.
            mine = torch.stack([d_idx.to(torch.float32), d_prob], dim=1)
            pending_gathers.append(dist.all_gather(gathered[i % 2], mine, async_op=True))
            drain_gathers(keep=1)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, k):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ev0.record(stream)
        for i in range(k):
            fn(i)
        drain_gathers()
        ev1.record(stream)
        barrier()
        ms = ev0.elapsed_time(ev1)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=f"cuda:{device}")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    for i in range(max(warmup, 3)):
        step_device(i)
    drain_gathers()
    eng.sync()
    l0 = eng.kernel_launches
    with ClockSampler(device) as clk:
        ms = timed(step_device, steps)
    eng.sync()                                      # also surfaces a stalled-pipeline report
    if world > 1:                                   # the last gather really carries this rank's last results
        last = gathered[(steps - 1) % 2][rank]
        assert torch.equal(last, torch.stack([d_idx.to(torch.float32), d_prob], dim=1)), "gathered results differ"
    launches = eng.kernel_launches - l0
    value = BATCH * world * steps / (ms / 1e3)

    # ---- end to end through the host-buffer C-ABI calls: every step copies its own 400 images from
    #      pinned host memory and reads its (idx, prob) back; up to three steps are in flight
    #      (idn_submit_u8hwc / idn_collect), so the copy of step i+1 overlaps the kernels of step i
    host_submit = [0.0]
    def run_e2e(k):
        pending = []
        for i in range(k):
            if len(pending) == 3:                  # IDN_MAX_IN_FLIGHT
                eng.collect(pending.pop(0))
            h0 = time.perf_counter()
            pending.append(eng.submit_packed(offsets, hw, pinned[i % n_pool].data_ptr()))
            host_submit[0] += time.perf_counter() - h0
        while pending:
            eng.collect(pending.pop(0))
    run_e2e(3)
    host_submit[0] = 0.0
    e2e_steps = max(steps, 100)     # ~0.1-0.2 s: filling/draining the 3-deep pipeline is < 2 %, host jitter averages out
    barrier()
    t0 = time.perf_counter()
    run_e2e(e2e_steps)
    barrier()
    ms_e2e = (time.perf_counter() - t0) * 1e3
    if world > 1:
        t = torch.tensor([ms_e2e], dtype=torch.float64, device=f"cuda:{device}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_e2e = float(t.item())
    e2e_value = BATCH * world * e2e_steps / (ms_e2e / 1e3)
    # the one-call form (idn_infer_u8hwc, copy then kernels then read-back, nothing overlapped)
    def step_host(i):
        eng.infer_packed(None, offsets, hw, packed_ptr=pinned[i % n_pool].data_ptr())
    step_host(0)
    ms_sync = timed(step_host, 5)
    e2e_sync_value = BATCH * world * 5 / (ms_sync / 1e3)

    # ---- per-kernel durations (CUDA events around every plan step, on the same stream)
    eng.set_timing(True)
    acc = {}
    reps = 5
    for i in range(reps + 1):
        step_device(i)
        drain_gathers()
        eng.sync()
        if i == 0:
            continue                               # first timed pass creates the events
        for name, t in eng.step_times_ms():
            acc[name] = acc.get(name, 0.0) + t / reps
    eng.set_timing(2)                              # one more pass with the pipeline-wait counters on
    step_device(0)
    drain_gathers()
    eng.sync()
    waits = {name: [int(x) for x in w] for name, w in eng.step_wait_cycles() if int(w.sum()) > 0}
    eng.set_timing(False)
    flops = STEP_FLOPS[model]
    tc_steps = {k: v for k, v in acc.items() if k in flops}
    dom = max(tc_steps, key=tc_steps.get)
    achieved = flops[dom] * BATCH / (tc_steps[dom] * 1e-3) / 1e12
    peak = peaks["tf_sustained"]
    roofline = {"bound": "tensor", "kernel": f"tcgen05 implicit-GEMM conv [{dom}]", "achieved": round(achieved, 2),
                "peak": peak, "unit": "TFLOP/s", "frac": round(achieved / peak, 4),
                "traffic": NCU_TRAFFIC_BYTES.get((model, dom)),
                "peak_source": f"MEASURED_PEAKS.json bf16 sustained ({peaks['source']}); the kernel is timed inside a "
                               "full step, so the sustained figure is the denominator (a short kernel can exceed it)",
                "frac_of_burst": round(achieved / peaks["tf_burst"], 4) if peaks.get("tf_burst") else None,
                "kernel_ms": round(tc_steps[dom], 4), "step_share": round(tc_steps[dom] / sum(acc.values()), 4)}
    whole = value / world * FLOP_PER_IMG[model] / 1e12
    pre_ms = acc.get("preprocess")
    out = {
        "value": round(value, 1), "ms_per_step": round(ms / steps, 4),
        "e2e": {"value": round(e2e_value, 1), "unit": "images/s",
                "h2d_bytes_per_step": BATCH * img_bytes + BATCH * 24,
                "d2h_bytes_per_step": BATCH * 8, "steps": e2e_steps,
                "how": "idn_submit_u8hwc/idn_collect, up to three 400-image steps in flight, host wall clock",
                "single_call_value": round(e2e_sync_value, 1),
                "host_ms_per_submit": round(host_submit[0] / e2e_steps * 1e3, 3)},
        "roofline": roofline,
        "whole_net": {"tflops": round(whole, 2), "frac_of_measured_sustained": round(whole / peak, 4),
                      "frac_of_measured_burst": round(whole / peaks["tf_burst"], 4),
                      "frac_of_spec_2250": round(whole / 2250.0, 4)},
        "preprocess": None if not pre_ms else {
            "ms": round(pre_ms, 4), "gbps": round(PRE_BYTES_PER_IMG * BATCH / (pre_ms * 1e-3) / 1e9, 1),
            "frac_of_measured_hbm": round(PRE_BYTES_PER_IMG * BATCH / (pre_ms * 1e-3) / 1e9 / peaks["hbm"], 4)},
        "kernel_ms": {k: round(v, 4) for k, v in acc.items()},
        "wait_cycles": waits,
        "gpu_launches": int(launches), "clocks": clk.summary(),
    }
    if do_cpu_baseline:
        out["cpu_baseline"] = cpu_baseline(model, sd, host_batches[0], sample=96 if model == "alexnet" else 64)
    eng.close()
    return out


def cpu_baseline(model, sd, images, sample):
    """The oracle (CPU port of the reference's per-image loop, alexnet_resnet.py:57-87) timed on this
    box's host cores on a bounded sample of the same workload."""
    import torch
    from oracle import ref_path
    imgs = [images[i] for i in range(sample)]
    ref_path.infer_arrays(model, sd, imgs[:4], library_preprocess=True)   # warm-up
    t0 = time.time()
    ref_path.infer_arrays(model, sd, imgs, library_preprocess=True)
    dt = time.time() - t0
    return {"value": round(sample / dt, 2), "unit": "images/s", "cores": torch.get_num_threads(),
            "host_cpus": os.cpu_count(), "kind": "port",
            "sample": f"first {sample} images of one 400-image query, batch-1 loop as the reference "
                      f"(preprocess + forward + softmax/top-1), model preloaded"}


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path (oracle port; the
    reference file itself cannot travel to the GPU box).  Rank 0 only."""
    if rank != 0:
        return
    import torch
    from idunno_b200.synth import make_state_dict
    from oracle import ref_path
    # torchrun exports OMP_NUM_THREADS=1; the reference arm uses every host core it can
    torch.set_num_threads(len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))
    per_step = 32
    res = {}
    for model in ("alexnet", "resnet18"):
        sd = make_state_dict(model, 0, "default")
        batch = make_batches(1, seed=0 if model == "alexnet" else 50)[0]
        imgs = [batch[i] for i in range(per_step)]
        for _ in range(max(1, min(args.warmup, 3))):
            ref_path.infer_arrays(model, sd, imgs, library_preprocess=True)
        t0 = time.time()
        for _ in range(args.steps):
            ref_path.infer_arrays(model, sd, imgs, library_preprocess=True)
        dt = time.time() - t0
        res[model] = (per_step * args.steps / dt, dt / args.steps * 1e3)
    v, ms = res["alexnet"]
    cores = torch.get_num_threads()
    sample = (f"each step = {per_step} images of a 400-image query, batch-1 loop as the reference "
              f"(alexnet_resnet.py:48-90), model preloaded, {cores} torch threads")
    line = {"impl": "reference", "metric": "images/sec", "value": round(v, 2), "unit": "images/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": WORKLOAD["alexnet"], "tensor_batch": 1, "weights": "random-init seed 0"},
            "cpu_baseline": {"value": round(v, 2), "unit": "images/s", "cores": cores,
                             "host_cpus": os.cpu_count(), "kind": "port", "sample": sample},
            "e2e": {"value": round(v, 2), "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "resnet18": {"value": round(res["resnet18"][0], 2), "ms_per_step": round(res["resnet18"][1], 3)}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--models", default="alexnet,resnet18")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    if not os.path.exists(ge.LIB):
        ge.build()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the hot path has no CPU fallback)")
    torch.cuda.set_device(local)
    numa_cores = []
    if world > 1 and os.environ.get("IDN_NO_NUMA_BIND") != "1":
        from idunno_b200 import bind_host_thread_to_gpu
        numa_cores = bind_host_thread_to_gpu(local)      # before any page-locked buffer is allocated
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    peaks = load_peaks()
    results = {}
    for model in args.models.split(","):
        results[model] = bench_model(model, rank, world, local, args.steps, args.warmup, peaks, dist,
                                     do_cpu_baseline=(rank == 0 and world == 1))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    prim = "alexnet" if "alexnet" in results else next(iter(results))
    r = results[prim]
    line = {"metric": "images/sec", "value": r["value"], "unit": "images/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": r["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic",
            "config": {"workload": WORKLOAD[prim], "tensor_batch": BATCH,
                       "images_per_step": BATCH * world, "weights": "random-init seed 0 (torchvision init)",
                       "l2": "inputs cycle through 4 distinct 60.2 MB uint8 batches per rank (241 MB > 126 MB L2); "
                             "weights + activations per step also exceed L2",
                       "parallelism": (f"range-sharded dp{world}, asynchronous NCCL all_gather of (idx, prob) per step, all waited for "
                                       "inside the timed region") if world > 1 else "1 GPU",
                       "host_binding": (f"each rank pinned to the {len(numa_cores)} cores NVML reports next to its GPU"
                                        if numa_cores else "none")},
            "roofline": r["roofline"], "e2e": r["e2e"], "clocks": r["clocks"],
            "gpu_launches": r["gpu_launches"], "whole_net": r["whole_net"], "preprocess": r["preprocess"],
            "kernel_ms": r["kernel_ms"], "wait_cycles": r["wait_cycles"]}
    if "cpu_baseline" in r:
        line["cpu_baseline"] = r["cpu_baseline"]
    for other in results:
        if other != prim:
            line[other] = dict(results[other], workload=WORKLOAD[other])
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
