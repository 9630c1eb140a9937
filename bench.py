#!/usr/bin/env python
"""Benchmark of the worker-side inference hot path (BASELINE.json metric: images/sec at
"batch 400" for AlexNet and ResNet18, with the fraction of the tensor roofline).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's own CPU implementation

A step = one 400-image query through the hot path (uint8 HWC 224x224 images -> top-1 index and
probability).  Headline workload (value / e2e / roofline): BASELINE.json configs[1], AlexNet batch 400
over range 0-3999; ResNet18 (configs[2]) rides in the same line under "resnet18" (value, e2e, roofline
fraction, and the 0-79999 range split over the ranks by the reference's rule).  With N > 1 (torchrun, one
rank per GPU) every rank runs its own 400-image sub-range per step (weak scaling) and per-image results
are gathered to rank 0 over NCCL inside the timed step.

  value : images/s, inputs resident in HBM, CUDA-event timed on the launching stream, barrier +
          synchronize on both sides, max over ranks.
  e2e   : images/s through the C-ABI calls with HOST buffers: pinned uint8 in, H2D + kernels + D2H of
          the results inside the timed region (idn_submit_u8hwc / idn_collect, three in flight);
          single_call_value is ONE blocking idn_infer_u8hwc per 400 images, the reference's call shape.
The line is kept under 1400 characters; per-kernel times, pipeline counters and the other legs go to
gpurun_out/bench_detail.json.
"""
from __future__ import annotations

import argparse
import csv
import json
import os
import re
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "-distributed-machine-learning-system_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

BATCH = int(os.environ.get("IDN_BENCH_BATCH", "400"))   # experiments only: the contract workload is 400
IMG_BYTES = 224 * 224 * 3
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
WORKLOAD = {"alexnet": "AlexNet batch 400 over range 0-3999", "resnet18": "ResNet18 batch 400 over range 0-79999"}
# `ncu --set full` summaries under profiles/ (tools/ncu_summary.py) from which roofline.traffic is read:
# the row whose kernel name matches the kernel the dominant step launches (idn_step_kernel).
TRAFFIC_CAPTURES = {"alexnet": ["profiles/r02_alexnet_ncu_full.csv", "profiles/r01f_alexnet_pair_ncu_full.csv"],
                    "resnet18": ["profiles/r02_resnet18_ncu_full.csv", "profiles/r02b_resnet18_preprocess_stem_head_ncu_full.csv"]}
DETAIL_PATH = os.path.join(ROOT, "gpurun_out", "bench_detail.json")


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(hbm=float(p["hbm_gbs"]), tf_burst=float(p["bf16_tflops"]),
                    tf_sustained=float(p.get("bf16_tflops_sustained", p["bf16_tflops"])), source="measured")
    # /opt/skills/guides/B200_PROFILING.md fallback
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback")


def traffic_from_capture(model: str, kernel: str):
    """(bytes per launch, file) of `kernel` from the newest committed capture that holds it, else (None, None)."""
    for rel in TRAFFIC_CAPTURES.get(model, []):
        path = os.path.join(ROOT, rel)
        if not os.path.exists(path):
            continue
        with open(path) as f:
            rows = list(csv.reader(f))
        hdr = rows[0]
        try:
            kn, rd, wr = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        except ValueError:
            continue
        unit = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}
        for r in rows[2:]:
            if len(r) > wr and re.search(r"\b" + re.escape(kernel), r[kn]):
                return int(float(r[rd]) * unit.get(rows[1][rd], 1e6) + float(r[wr]) * unit.get(rows[1][wr], 1e6)), os.path.basename(rel)
    return None, None


class ClockSampler:
    """Samples SM clock, power and throttle reasons of one GPU while a timed region runs (NVML)."""
    REASONS = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap",
               0x8: "hw_slowdown", 0x10: "sync_boost", 0x20: "sw_thermal_slowdown",
               0x40: "hw_thermal_slowdown", 0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index: int):
        self.samples, self.power, self.reasons, self.max_mhz = [], [], set(), None
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
                self.power.append(nv.nvmlDeviceGetPowerUsage(self._h) / 1000.0)
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

    def summary(self, power=False):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        out = {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}
        if power and self.power:
            out["power_w_max"] = round(max(self.power), 1)
        return out


def make_batches(n_batches: int, seed: int):
    from idunno_b200.synth import noise_images
    return [noise_images(BATCH, seed=seed + b) for b in range(n_batches)]


def bench_model(model, rank, world, device, args, peaks, dist):
    import torch
    from idunno_b200 import Engine, sharding
    from idunno_b200.synth import make_state_dict

    steps, warmup = args.steps, max(args.warmup, 3)
    dev = f"cuda:{device}"
    sd = make_state_dict(model, 0, "default")
    eng = Engine(model, sd, device=device, max_batch=BATCH)
    stream = torch.cuda.Stream(device=device)      # a real (non-NULL) stream: NULL selects the handle's own
    torch.cuda.set_stream(stream)
    eng.set_stream(stream.cuda_stream)

    n_pool = 4                                     # 4 x 60.2 MB of distinct uint8 input > 126 MB L2
    host_batches = make_batches(n_pool, seed=100 * rank + (0 if model == "alexnet" else 50))
    offsets = (np.arange(BATCH, dtype=np.int64) * IMG_BYTES)
    hw = np.tile(np.asarray([[224, 224]], dtype=np.int32), (BATCH, 1))
    dev_batches = [torch.from_numpy(b.reshape(-1)).to(dev) for b in host_batches]
    pinned = [torch.from_numpy(b.reshape(-1)).pin_memory() for b in host_batches]
    # results of a step: [idx int32 x 400 | prob fp32 x 400] in ONE buffer, so that the gather needs no eager op
    res = torch.zeros(2 * BATCH, dtype=torch.int32, device=dev)
    d_idx, d_prob = res[:BATCH], res[BATCH:].view(torch.float32)
    gathered, pending_gathers = None, []
    if world > 1:     # two sets: the gather of step i is still in flight while step i+1 computes
        gathered = [[torch.zeros(2 * BATCH, dtype=torch.int32, device=dev) for _ in range(world)] for _ in range(2)]

    def drain_gathers(keep=0):
        while len(pending_gathers) > keep:
            pending_gathers.pop(0).wait()          # the timing stream waits for that collective

    def step_device(i):
        eng.infer_device(dev_batches[i % n_pool].data_ptr(), offsets, hw, d_idx.data_ptr(), d_prob.data_ptr())
        if world > 1:   # result return of the query: 8 bytes per image to rank 0's view (NCCL), asynchronous --
            # the next step's kernels do not wait for it; every gather is waited for inside the timed region
            pending_gathers.append(dist.all_gather(gathered[i % 2], res, async_op=True))
            drain_gathers(keep=1)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world > 1:
            t = torch.tensor([x], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item())
        return x

    def timed(fn, k):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ev0.record(stream)
        for i in range(k):
            fn(i)
        drain_gathers()
        ev1.record(stream)
        barrier()
        return max_over_ranks(ev0.elapsed_time(ev1))

    for i in range(warmup):
        step_device(i)
    drain_gathers()
    eng.sync()
    l0 = eng.kernel_launches
    with ClockSampler(device) as clk:
        ms = timed(step_device, steps)
    eng.sync()                                      # also surfaces a stalled-pipeline report
    if world > 1:                                   # the last gather really carries this rank's last results
        assert torch.equal(gathered[(steps - 1) % 2][rank], res), "gathered results differ"
    launches = eng.kernel_launches - l0
    value = BATCH * world * steps / (ms / 1e3)
    detail = {"ms_per_step": ms / steps}

    # ---- sustained leg: the same step back to back for >= args.sustain seconds, clocks and power sampled
    sustained = None
    if args.sustain > 0:
        k = max(steps, int(args.sustain * 1e3 / (ms / steps)) + 1)
        with ClockSampler(device) as sclk:
            ms_s = timed(step_device, k)
        v = BATCH * world * k / (ms_s / 1e3)
        tf = v / world * FLOP_PER_IMG[model] / 1e12
        sustained = {"value": round(v, 1), "seconds": round(ms_s / 1e3, 2), "tflops_per_gpu": round(tf, 1),
                     "frac_of_sustained_peak": round(tf / peaks["tf_sustained"], 4), "clocks": sclk.summary(power=True)}

    # ---- end to end through the host-buffer C-ABI calls: every step copies its own 400 images from
    #      pinned host memory and reads its (idx, prob) back; up to three steps are in flight
    def run_e2e(k):
        pending = []
        for i in range(k):
            if len(pending) == 3:                  # IDN_MAX_IN_FLIGHT
                eng.collect(pending.pop(0))
            pending.append(eng.submit_packed(offsets, hw, pinned[i % n_pool].data_ptr()))
        while pending:
            eng.collect(pending.pop(0))
    run_e2e(3)
    e2e_steps = max(steps, 200)     # ~0.2-0.4 s: filling/draining the 3-deep pipeline is < 1 %, host jitter averages out
    barrier()
    h2d0 = eng.h2d_bytes
    t0 = time.perf_counter()
    run_e2e(e2e_steps)
    barrier()
    ms_e2e = max_over_ranks((time.perf_counter() - t0) * 1e3)
    e2e_value = BATCH * world * e2e_steps / (ms_e2e / 1e3)
    # image bytes the library actually queued per step: whole images are handed over, the rows the centre crop never
    # reads are not copied (include/idunno_b200.h, idn_h2d_bytes)
    h2d_step = (eng.h2d_bytes - h2d0) // e2e_steps
    # the one-call form: ONE blocking idn_infer_u8hwc per 400-image shard, the reference's call shape
    def step_host(i):
        eng.infer_packed(None, offsets, hw, packed_ptr=pinned[i % n_pool].data_ptr())
    step_host(0)
    step_host(1)
    barrier()
    t0 = time.perf_counter()
    for i in range(10):
        step_host(i)
    barrier()
    single = BATCH * world * 10 / max_over_ranks(time.perf_counter() - t0)
    # copy-only ceiling of this box's host link: the same pinned buffers, nothing else running
    copy_stream = torch.cuda.Stream(device=device)
    sink = torch.empty(BATCH * IMG_BYTES, dtype=torch.uint8, device=dev)
    def copies(k):
        with torch.cuda.stream(copy_stream):
            for i in range(k):
                sink.copy_(pinned[i % n_pool], non_blocking=True)
        copy_stream.synchronize()
    copies(3)
    barrier()
    t0 = time.perf_counter()
    copies(40)
    dt_copy = time.perf_counter() - t0
    barrier()
    # aggregate rate while every rank copies: all bytes over the SLOWEST rank's time (summing per-rank rates would
    # credit a rank that finishes early with bandwidth the others then inherit)
    h2d_gbs = world * 40 * BATCH * IMG_BYTES / max_over_ranks(dt_copy) / 1e9
    ceiling = h2d_gbs * 1e9 / (h2d_step / BATCH)       # images/s if the link did nothing but carry the bytes a step copies
    e2e = {"value": round(e2e_value, 1), "unit": "images/s", "h2d_bytes_per_step": int(h2d_step) + BATCH * 24,
           "d2h_bytes_per_step": BATCH * 8, "single_call_value": round(single, 1),
           "h2d_ceiling_gbs": round(h2d_gbs, 1), "frac_of_h2d_ceiling": round(e2e_value / ceiling, 3)}

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
    eng.set_timing(2)                              # one more pass with the pipeline counters on
    step_device(0)
    drain_gathers()
    eng.sync()
    waits = {name: [int(x) for x in w] for name, w in eng.step_wait_cycles() if int(w.sum()) > 0}
    eng.set_timing(False)
    kernels = dict(eng.step_kernels())
    flops = STEP_FLOPS[model]
    tc_steps = {k: v for k, v in acc.items() if k in flops}
    dom = max(tc_steps, key=tc_steps.get)
    achieved = flops[dom] * BATCH / (tc_steps[dom] * 1e-3) / 1e12
    traffic, traffic_file = traffic_from_capture(model, kernels[dom])
    # A sub-millisecond kernel timed with events inside a ~10 ms region runs at burst clocks: the burst figure
    # is its denominator; against the sustained figure a good kernel reads above 1.
    roofline = {"bound": "tensor", "kernel": f"{kernels[dom].replace(', ', ',')} {dom}", "achieved": round(achieved, 1),
                "peak": peaks["tf_burst"], "unit": "TFLOP/s", "frac": round(achieved / peaks["tf_burst"], 4),
                "frac_of_sustained": round(achieved / peaks["tf_sustained"], 4), "peak_src": peaks["source"] + " burst",
                "traffic": traffic, "traffic_src": traffic_file,
                "kernel_ms": round(tc_steps[dom], 4), "step_share": round(tc_steps[dom] / sum(acc.values()), 4)}
    whole = value / world * FLOP_PER_IMG[model] / 1e12
    pre_ms = acc.get("preprocess")
    detail.update({
        "kernel_ms": {k: round(v, 4) for k, v in acc.items()}, "kernels": kernels, "wait_cycles": waits,
        "whole_net": {"tflops_per_gpu": round(whole, 1), "frac_of_burst": round(whole / peaks["tf_burst"], 4),
                      "frac_of_sustained_peak_in_a_burst_run": round(whole / peaks["tf_sustained"], 4),
                      "frac_of_spec_2250": round(whole / 2250.0, 4)},
        "preprocess": None if not pre_ms else {
            "ms": round(pre_ms, 4), "gbps": round(PRE_BYTES_PER_IMG * BATCH / (pre_ms * 1e-3) / 1e9, 1),
            "frac_of_measured_hbm": round(PRE_BYTES_PER_IMG * BATCH / (pre_ms * 1e-3) / 1e9 / peaks["hbm"], 4)},
        "sustained": sustained})

    # ---- BASELINE.json configs[2] as worded: ResNet18 over range 0-79999, split over the ranks by the
    #      reference's rule (mp4_machinelearning.py:516-536), ONE gather of (idx, prob) at the end: strong scaling
    range_leg = None
    if model == "resnet18" and args.range_images > 0:
        total = args.range_images
        shard = [s for s in sharding.split_range(0, total - 1, list(range(world))) if s[0] == rank][0]
        count = shard[2] - shard[1] + 1
        cmax = -(-total // world) + 1
        mine = torch.zeros(2 * cmax, dtype=torch.int32, device=dev)
        i_all, p_all = mine[:cmax], mine[cmax:].view(torch.float32)
        everyone = [torch.zeros(2 * cmax, dtype=torch.int32, device=dev) for _ in range(world)] if world > 1 else None

        def run_range():
            for b, i0 in enumerate(range(0, count, BATCH)):
                n = min(BATCH, count - i0)
                eng.infer_device(dev_batches[b % n_pool].data_ptr(), offsets[:n], hw[:n],
                                 i_all.data_ptr() + 4 * i0, p_all.data_ptr() + 4 * i0)
            if world > 1:
                dist.all_gather(everyone, mine)
        ms_r = timed(lambda i: run_range(), 1)
        ms_r = min(ms_r, timed(lambda i: run_range(), 1))
        range_leg = {"images": total, "value": round(total / (ms_r / 1e3), 1), "ms": round(ms_r, 2),
                     "scaling": "strong", "split": "reference rule, one gather at the end"}

    eng.close()
    return {"value": round(value, 1), "ms_per_step": round(ms / steps, 4), "e2e": e2e, "roofline": roofline,
            "gpu_launches": int(launches), "clocks": clk.summary(), "range_leg": range_leg, "detail": detail,
            "sd": sd, "images": host_batches[0]}


# ------------------------------------------------------------------------------------- CPU legs
def cpu_legs(model, sd, images, n_faithful=400):
    """BASELINE.md section 3, on this box's host cores:
    (i)  faithful: the reference's own deeplearning() over range 0-399 (oracle/_ref: the unmodified module),
         batch-1 loop, model constructed inside the call as the reference does (alexnet_resnet.py:16-22);
    (ii) best effort: the same torchvision model preloaded, ONE [400,3,224,224] fp32 forward, the
         reference's preprocess timed separately."""
    import torch
    import torchvision
    from oracle import ref_path, run_ref
    cores = torch.get_num_threads()
    out = {"unit": "images/s", "cores": cores, "host_cpus": os.cpu_count()}
    imgs = {i: images[i] for i in range(n_faithful)}
    if run_ref.available():
        with tempfile.TemporaryDirectory() as d:
            rr = run_ref.ReferenceRun(d, {model: sd})
            rr.write_images(model, imgs)
            rr.deeplearning(model, 0, 3)                                    # warm caches / lazy imports
            res, secs = rr.deeplearning(model, 0, n_faithful - 1)
        out.update(value=round(len(res) / secs, 2), kind="reference",
                   sample=f"unmodified reference deeplearning() range 0-{n_faithful - 1}, {secs:.1f} s")
    else:
        t0 = time.time()
        ref_path.infer_arrays(model, sd, [images[i] for i in range(64)], library_preprocess=True)
        secs = time.time() - t0
        out.update(value=round(64 / secs, 2), kind="port",
                   sample=f"oracle port (no oracle/_ref), 64 images, model preloaded, {secs:.1f} s")
    net = getattr(torchvision.models, model)(weights=None)
    net.load_state_dict(sd)
    net.eval()
    t0 = time.time()
    x = torch.from_numpy(np.stack([ref_path.preprocess_library(images[i]) for i in range(BATCH)]))
    t_pre = time.time() - t0
    with torch.no_grad():
        net(x[:16])
        t0 = time.time()
        net(x)
        t_fwd = time.time() - t0
    out["best_effort"] = {"value": round(BATCH / t_fwd, 1), "what": "preloaded model, one [400,3,224,224] fp32 forward",
                          "preprocess_images_per_s": round(BATCH / t_pre, 1)}
    return out


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path on this box's host cores, all
    the threads it can use.  A step = one deeplearning() call over a 32-image shard (model construction
    inside the call, as the reference does).  oracle/_ref (the unmodified module, byte-compiled) when it
    travelled here, else the oracle port.  Rank 0 only."""
    if rank != 0:
        return
    import torch
    from idunno_b200.synth import make_state_dict
    from oracle import ref_path, run_ref
    # torchrun exports OMP_NUM_THREADS=1; the reference arm uses every host core it can
    torch.set_num_threads(len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))
    per_step = 80                                   # the reference's own shard: a 400-image query over its 5 workers
    use_ref = run_ref.available()
    res = {}
    with tempfile.TemporaryDirectory() as d:
        sds = {m: make_state_dict(m, 0, "default") for m in ("alexnet", "resnet18")}
        rr = run_ref.ReferenceRun(d, sds) if use_ref else None
        for model in ("alexnet", "resnet18"):
            batch = make_batches(1, seed=0 if model == "alexnet" else 50)[0]
            imgs = [batch[i] for i in range(per_step)]
            if use_ref:
                rr.write_images(model, dict(enumerate(imgs)))
                step = lambda: rr.deeplearning(model, 0, per_step - 1)
            else:
                step = lambda: ref_path.infer_arrays(model, sds[model], imgs, library_preprocess=True)
            for _ in range(max(1, min(args.warmup, 3))):
                step()
            t0 = time.time()
            for _ in range(args.steps):
                step()
            dt = time.time() - t0
            res[model] = (per_step * args.steps / dt, dt / args.steps * 1e3)
    v, ms = res["alexnet"]
    cores = torch.get_num_threads()
    kind = "reference" if use_ref else "port"
    sample = (f"each step = one deeplearning() over a {per_step}-image shard of the 400-image query, batch-1 loop, "
              f"model built inside the call (alexnet_resnet.py:16-22), {cores} torch threads")
    line = {"impl": "reference", "metric": "images/sec", "value": round(v, 2), "unit": "images/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": WORKLOAD["alexnet"], "tensor_batch": 1, "weights": "random-init seed 0"},
            "cpu_baseline": {"value": round(v, 2), "unit": "images/s", "cores": cores,
                             "host_cpus": os.cpu_count(), "kind": kind, "sample": sample},
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
    ap.add_argument("--sustain", type=float, default=2.0, help="seconds of the back-to-back leg per model (0: skip)")
    ap.add_argument("--range-images", type=int, default=80000, help="ResNet18 range leg (BASELINE configs[2]); 0: skip")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline legs")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    # libraries (NCCL's version banner, ...) must not write into the one-line result: park stdout on stderr
    real_stdout = os.dup(1)
    os.dup2(2, 1)
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
        results[model] = bench_model(model, rank, world, local, args, peaks, dist)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    prim = "alexnet" if "alexnet" in results else next(iter(results))
    r = results[prim]
    cpu = {}
    if world == 1 and not args.no_cpu:
        torch.set_num_threads(len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))
        for m, rm in results.items():
            cpu[m] = cpu_legs(m, rm["sd"], rm["images"])
    line = {"metric": "images/sec", "value": r["value"], "unit": "images/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": r["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic",
            "config": {"workload": WORKLOAD[prim], "tensor_batch": BATCH, "images_per_step": BATCH * world,
                       "weights": "random-init seed 0", "l2": "inputs 4x60MB per rank > L2",
                       "parallelism": f"range-sharded dp{world}, NCCL gather of (idx, prob) per step" if world > 1 else "1 GPU"},
            "roofline": {k: v for k, v in r["roofline"].items() if k not in ("kernel_ms", "step_share")},
            "e2e": r["e2e"], "clocks": r["clocks"], "gpu_launches": r["gpu_launches"]}
    if prim in cpu:
        c = cpu[prim]
        line["cpu_baseline"] = {"value": c["value"], "unit": c["unit"], "cores": c["cores"], "kind": c["kind"],
                                "sample": c["sample"], "best_effort_value": c["best_effort"]["value"]}
    if "resnet18" in results and prim != "resnet18":
        rr = results["resnet18"]
        line["resnet18"] = {"value": rr["value"], "ms_per_step": rr["ms_per_step"], "e2e": rr["e2e"]["value"],
                            "single_call": rr["e2e"]["single_call_value"], "roofline_frac": rr["roofline"]["frac"]}
        if rr["range_leg"]:                          # BASELINE configs[2] as worded: fixed range, strong scaling
            line["resnet18"]["range80k"] = rr["range_leg"]["value"]
            line["resnet18"]["range80k_ms"] = rr["range_leg"]["ms"]
    if r["detail"].get("sustained"):
        s = r["detail"]["sustained"]
        line["sustained"] = {"value": s["value"], "s": s["seconds"], "frac_of_sustained_peak": s["frac_of_sustained_peak"],
                             "sm_mhz": s["clocks"].get("sm_mhz")}
    detail = {"line": line, "host_binding": numa_cores, "cpu_baseline": cpu,
              "models": {m: dict(rm["detail"], roofline=rm["roofline"], e2e=rm["e2e"], range_leg=rm["range_leg"],
                                 workload=WORKLOAD[m]) for m, rm in results.items()}}
    try:
        os.makedirs(os.path.dirname(DETAIL_PATH), exist_ok=True)
        with open(DETAIL_PATH if world == 1 else DETAIL_PATH.replace(".json", f"_{world}gpu.json"), "w") as f:
            json.dump(detail, f, indent=1)
    except OSError:
        pass
    text = json.dumps(line)
    for drop in (("roofline", "peak_src"), ("config", "images_per_step"), ("sustained", "s"), ("config", "l2"),
                 ("roofline", "traffic_src"), ("sustained",)):   # the driver's record keeps 1500 characters
        if len(text) <= 1400:
            break
        if len(drop) == 1:
            line.pop(drop[0], None)
        elif drop[0] in line:
            line[drop[0]].pop(drop[1], None)
        text = json.dumps(line)
    sys.stdout.flush()
    os.dup2(real_stdout, 1)
    print(text, flush=True)


if __name__ == "__main__":
    main()
