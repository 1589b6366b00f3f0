#!/usr/bin/env python
"""bench.py -- the headline measurement: params/s of the multi-tensor fused-optimizer step.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

N = 1 (default): BASELINE.json configs[1] -- amp unscale + overflow check (multi_tensor_scale, in
place, 1/65536) followed by the FusedAdam fp32 step (multi_tensor_adam, AdamW mode) over the
synthetic ResNet-50 parameter list (161 tensors, 25,557,032 fp32 params), on one B200.
N > 1: BASELINE.json configs[3] -- the ZeRO-style sharded step of DistributedFusedAdam on the
Llama-3-8B shape list (bf16 grads, fp32 master): reduce-scatter -> shard-local fused Adam ->
all-gather over NCCL, one process per GPU (launched by torch.distributed.run).

`value` is whole-job params/s with everything resident in HBM, timed with CUDA events, max over
ranks.  `e2e` is the same step through the public Python API with the step's gradients copied
from pinned host memory and the overflow flag read back inside the timed region.  `roofline` is
for the dominant kernel (the Adam read-modify-write, 28 B/param).  `cpu_baseline` / `--impl
reference` time the reference's python-only build path (apex/amp/scaler.py:6-23,85-100 restated
in oracle/mt_oracle.py + stock torch.optim.Adam) on this box's host cores.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (os.path.join(ROOT, "rocm-apex_b200"), ROOT):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

import param_shapes  # noqa: E402

LOSS_SCALE = 65536.0
ADAM = dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, wd=0.0)
CHUNK = 2048 * 32
METRIC = "fused_adam_step_params_per_s"
UNIT = "params/s"


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for n, val in zip(names, r[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's python-only build path on the host cores
# ------------------------------------------------------------------------------------------------
def gpu_numa_cpus(dev_index):
    """CPUs of the NUMA node the GPU hangs off (sysfs), or None when the box does not say."""
    try:
        pr = torch.cuda.get_device_properties(dev_index)
        bdf = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % bdf).read())
        if node < 0:
            raise ValueError("no NUMA node recorded for %s" % bdf)
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        return (cpus & os.sched_getaffinity(0)) or None
    except Exception:
        pass
    try:    # containers often hide sysfs NUMA data; the driver still knows which CPUs sit next to the GPU
        import pynvml
        pynvml.nvmlInit()
        pr = torch.cuda.get_device_properties(dev_index)
        bdf = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bdf.encode())
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = {64 * w + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1}
        near = cpus & os.sched_getaffinity(0)
        return near if (near and near != os.sched_getaffinity(0)) else None
    except Exception:
        return None


def pinned_near_gpu(numel, dtype, dev_index):
    """Pinned host tensor whose pages sit on the GPU's own NUMA node: allocated and first touched while this thread
    is confined to that node's CPUs (H2D from the far socket runs at ~30 GB/s instead of ~55 on these boxes); the
    affinity is put back afterwards so that the CPU-baseline leg still sees every core."""
    old = os.sched_getaffinity(0)
    near = gpu_numa_cpus(dev_index)
    try:
        if near:
            os.sched_setaffinity(0, near)
        t = torch.empty(numel, dtype=dtype, pin_memory=True)   # cudaHostAlloc faults the pages in here, on this thread
    finally:
        os.sched_setaffinity(0, old)
    return t


def single_config(nparams, ntensors):
    return {"workload": "configs[1]: multi_tensor_scale unscale+overflow check, then FusedAdam fp32 step (multi_tensor_adam, AdamW) "
                        "over the ResNet-50 param list, 1xB200",
            "tensors": ntensors, "params": nparams, "chunk_size": CHUNK,
            "l2": "inputs larger than L2 (working set 409 MB g/p/m/v vs 126 MB L2); no flush between steps",
            "algorithmic_bytes_per_step": 36 * nparams}


def cpu_reference_run(steps, warmup, max_seconds=30.0, shapes=None):
    from oracle import mt_oracle as O
    if shapes is None:
        shapes = param_shapes.resnet50()
    gen = torch.Generator().manual_seed(0)
    params = [torch.nn.Parameter(torch.randn(s, generator=gen) * 0.02) for s in shapes]
    grads0 = [torch.randn(s, generator=gen) * LOSS_SCALE for s in shapes]
    opt = torch.optim.Adam(params, lr=ADAM["lr"], betas=(ADAM["beta1"], ADAM["beta2"]), eps=ADAM["eps"],
                           weight_decay=ADAM["wd"])
    for p in params:
        p.grad = torch.empty_like(p)
    nparams = sum(p.numel() for p in params)

    def step():
        for p, g in zip(params, grads0):
            p.grad.copy_(g)          # "backward" leaves scaled grads; not part of the reference path cost
        t0 = time.perf_counter()
        grads = [p.grad for p in params]
        overflow = O.unscale_python(grads, grads, LOSS_SCALE, dynamic=True)   # apex/amp/scaler.py:85-100
        if not overflow:
            opt.step()
        return time.perf_counter() - t0

    for _ in range(warmup):
        step()
    times = []
    t_begin = time.perf_counter()
    for _ in range(steps):
        times.append(step())
        if time.perf_counter() - t_begin > max_seconds:
            break
    ms = 1e3 * sum(times) / len(times)
    return {"value": nparams / (ms * 1e-3), "ms_per_step": ms, "steps": len(times), "cores": torch.get_num_threads(),
            "host_cpus": os.cpu_count(), "nparams": nparams}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = max(args.gpus, int(os.environ.get("WORLD_SIZE", "1")))
    if world > 1:
        # the sharded workload (Llama-3-8B list) does not fit a few-minute CPU run: time the python-only
        # unscale + torch.optim.Adam on a bounded sample of it (the four attention projections of layer 0)
        full = param_shapes.llama3_8b()
        shapes = full[1:5]
        r = cpu_reference_run(args.steps, args.warmup, max_seconds=120.0, shapes=shapes)
        sample = "%d steps over a %d-parameter sample (q/k/v/o projections of layer 0) of the Llama-3-8B list on host CPU; " \
                 "value = sample params / step time" % (r["steps"], r["nparams"])
        config = {"workload": "configs[3]: DistributedFusedAdam sharded step on the llama3_8b shape list (reference arm: python-only "
                              "unscale + torch.optim.Adam on a bounded sample, host CPU)",
                  "tensors": len(full), "params": param_shapes.numel(full), "parallelism": "zero-dp%d" % world}
        scaling = "strong"
    else:
        r = cpu_reference_run(args.steps, args.warmup, max_seconds=120.0)
        sample = "%d full steps of the ResNet-50 list (161 tensors, %d params) on host CPU" % (r["steps"], r["nparams"])
        config = single_config(r["nparams"], 161)
        scaling = "weak"
    line = {
        "impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": world,
        "steps": r["steps"], "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True,
        "scaling": scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config,
        "cpu_baseline": {"value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "sample": sample,
                         "host_cpus": r["host_cpus"]},
        "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------------
# our arm, N = 1: configs[1]
# ------------------------------------------------------------------------------------------------
def make_resnet_state(dev, seed=0):
    shapes = param_shapes.resnet50()
    g = torch.Generator(device=dev).manual_seed(seed)
    params = [torch.randn(s, generator=g, device=dev) * 0.02 for s in shapes]
    grads = [torch.randn(s, generator=g, device=dev) * LOSS_SCALE for s in shapes]
    m = [torch.zeros_like(p) for p in params]
    v = [torch.zeros_like(p) for p in params]
    return shapes, params, grads, m, v


def run_single(args):
    import _b200mt as mt
    import amp_C
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
    torch.cuda.set_device(dev)
    lib = mt.lib
    shapes, params, grads, m, v = make_resnet_state(dev)
    nparams = sum(p.numel() for p in params)
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    inv_scale = 1.0 / LOSS_SCALE

    # C-ABI descriptors built once: this is what a C host keeps between steps
    L_scale = mt.pack([grads, grads], 2)
    L_adam = mt.pack([grads, params, m, v], 4)
    stream = torch.cuda.current_stream(dev)
    sp = stream.cuda_stream
    fp = flag.data_ptr()
    step_no = [0]
    # grads are re-scaled in place each step (x 1/65536 then Adam); to keep them in range over many steps the
    # "backward" is emulated by scaling them back up with our own kernel OUTSIDE the per-kernel roofline events
    # but it is NOT part of the step: instead we alternate the unscale factor (1/65536, then 65536) so that the
    # gradient magnitudes stay put while each step still does one full unscale pass + one full Adam pass.
    def step(events=None):
        step_no[0] += 1
        s = inv_scale if (step_no[0] & 1) else LOSS_SCALE
        flag.zero_()
        mt.check(lib.b200mt_scale(L_scale.ref(), CHUNK, fp, s, sp))
        if events is not None:
            events[0].record(stream)
        mt.check(lib.b200mt_adam(L_adam.ref(), CHUNK, fp, ADAM["lr"], ADAM["beta1"], ADAM["beta2"], ADAM["eps"],
                                 step_no[0], 1, 1, ADAM["wd"], 1.0, sp))
        if events is not None:
            events[1].record(stream)

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    sampler = ClockSampler(dev.index)
    sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    launches0 = lib.b200mt_launch_count()
    torch.cuda.synchronize()
    e0.record(stream)
    for i in range(args.steps):
        step()
    e1.record(stream)
    torch.cuda.synchronize()
    launches = lib.b200mt_launch_count() - launches0
    total_ms = e0.elapsed_time(e1)
    # second pass over the same K steps with an event on either side of the Adam launch (the events
    # cost ~4 us of stream time per step, so they are kept out of the pass that defines `value`)
    for i in range(args.steps):
        step(ev[i])
    torch.cuda.synchronize()
    clocks = sampler.stop()
    adam_ms = [a.elapsed_time(b) for a, b in ev]
    adam_avg = sum(adam_ms) / len(adam_ms)
    ms_per_step = total_ms / args.steps
    value = nparams / (ms_per_step * 1e-3)
    assert int(flag) == 0

    # ---- e2e: public API (amp_C drop-in entry points), grads arrive from pinned host memory each step,
    # the overflow flag is read back (the one D2H sync amp does per iteration, apex/amp/scaler.py:209)
    host_grads = pinned_near_gpu(nparams, torch.float32, dev.index)
    host_grads.normal_(0, 1).mul_(LOSS_SCALE)
    flat = torch.empty(nparams, dtype=torch.float32, device=dev)
    views, off = [], 0
    for s_ in shapes:
        n = 1
        for d in s_:
            n *= d
        views.append(flat[off:off + n].view(s_))
        off += n
    e2e_steps = max(3, min(args.steps, 30))
    # The copy is cut into E2E_SEGMENTS pieces on tensor boundaries, issued on a copy stream; the unscale + Adam
    # launches of a piece follow its copy, so all but the last piece's compute hides under the PCIe transfer
    # (tools/e2e_probe.py: copy alone 1.85 ms = 55 GB/s, unsegmented 2.01 ms, 4 segments 1.94 ms).
    E2E_SEGMENTS = 4
    numels = [p.numel() for p in params]
    offs = [0]
    for n in numels:
        offs.append(offs[-1] + n)
    bounds = [0]
    for i in range(len(numels)):
        if offs[i + 1] >= nparams * len(bounds) / E2E_SEGMENTS and len(bounds) < E2E_SEGMENTS:
            bounds.append(i + 1)
    bounds.append(len(numels))
    copy_stream = torch.cuda.Stream(device=dev)
    seg_ready = [torch.cuda.Event() for _ in range(len(bounds) - 1)]

    def e2e_step(k):
        flag.zero_()
        copy_stream.wait_stream(stream)                  # the previous step is done with the gradient buffer
        with torch.cuda.stream(copy_stream):
            for s_ in range(len(bounds) - 1):
                a, b = offs[bounds[s_]], offs[bounds[s_ + 1]]
                flat[a:b].copy_(host_grads[a:b], non_blocking=True)
                seg_ready[s_].record(copy_stream)
        for s_ in range(len(bounds) - 1):
            stream.wait_event(seg_ready[s_])
            sl = slice(bounds[s_], bounds[s_ + 1])
            amp_C.multi_tensor_scale(CHUNK, flag, [views[sl], views[sl]], inv_scale)
            amp_C.multi_tensor_adam(CHUNK, flag, [views[sl], params[sl], m[sl], v[sl]], ADAM["lr"], ADAM["beta1"],
                                    ADAM["beta2"], ADAM["eps"], k, 1, 1, ADAM["wd"])
        return flag.item()

    for k in range(3):
        e2e_step(k + 1)
    torch.cuda.synchronize()
    # The host's PCIe complex is shared with other tenants of the box: the same pinned buffer copies at 55 GB/s one
    # moment and 35 GB/s the next (tools/e2e_probe.py).  Time E2E_BLOCKS blocks of e2e_steps steps and report the
    # median block, all blocks listed beside it.
    E2E_BLOCKS = 5
    e2e_blocks = []
    for blk in range(E2E_BLOCKS):
        e0.record(stream)
        for k in range(e2e_steps):
            ov = e2e_step(k + 4)
        e1.record(stream)
        torch.cuda.synchronize()
        e2e_blocks.append(e0.elapsed_time(e1) / e2e_steps)
        assert ov == 0
    e2e_ms = statistics.median(e2e_blocks)

    # ---- the same device-resident step through the drop-in Python entry points (amp_C.*): what the
    # reference's optimizers call.  Host time per call is measured separately (no sync inside).
    def api_step(k):
        flag.zero_()
        amp_C.multi_tensor_scale(CHUNK, flag, [grads, grads], inv_scale if (k & 1) else LOSS_SCALE)
        amp_C.multi_tensor_adam(CHUNK, flag, [grads, params, m, v], ADAM["lr"], ADAM["beta1"], ADAM["beta2"],
                                ADAM["eps"], k, 1, 1, ADAM["wd"])

    for k in range(1, 5):
        api_step(k)
    torch.cuda.synchronize()
    api_steps = max(10, min(args.steps, 100))
    e0.record(stream)
    t_host0 = time.perf_counter()
    for k in range(api_steps):
        api_step(k + 5)
    t_host1 = time.perf_counter()
    e1.record(stream)
    torch.cuda.synchronize()
    api_ms = e0.elapsed_time(e1) / api_steps
    api_host_us = (t_host1 - t_host0) / api_steps * 1e6

    # ---- the mixed-precision single-pass Adam the north star names as the target kernel: fp16 gradients read
    # alongside fp32 master weights and moments, fp16 model copy written in the same pass (28 B/param)
    # fp16 gradients carry a loss scale that keeps them finite in fp16 (1024, not the fp32 leg's 65536)
    mixed_scale = 1024.0
    g16 = [(torch.randn_like(g) * mixed_scale).half() for g in grads]
    assert all(bool(torch.isfinite(g).all()) for g in g16[:8])
    p16 = [torch.empty_like(g) for g in g16]
    L_mixed = mt.pack([g16, params, m, v, p16], 5)
    inv_scale = 1.0 / mixed_scale
    for k in range(3):
        mt.check(lib.b200mt_adam(L_mixed.ref(), CHUNK, fp, ADAM["lr"], ADAM["beta1"], ADAM["beta2"], ADAM["eps"], k + 1, 1, 1,
                                 ADAM["wd"], inv_scale, sp))
    mixed_steps = max(10, min(args.steps, 100))
    torch.cuda.synchronize()
    e0.record(stream)
    for k in range(mixed_steps):
        mt.check(lib.b200mt_adam(L_mixed.ref(), CHUNK, fp, ADAM["lr"], ADAM["beta1"], ADAM["beta2"], ADAM["eps"], k + 4, 1, 1,
                                 ADAM["wd"], inv_scale, sp))
    e1.record(stream)
    torch.cuda.synchronize()
    mixed_ms = e0.elapsed_time(e1) / mixed_steps

    # ---- the same two launches (+ flag clear) replayed from a CUDA graph: what is left of the step when the
    # launch gaps go.  Auxiliary only: a captured multi_tensor_adam carries its host-side step number, so the bias
    # correction does not advance between replays (the byte traffic and arithmetic are those of a real step).
    graph_ms = None
    try:
        gs = torch.cuda.Stream(device=dev)
        gs.wait_stream(stream)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=gs):
            flag.zero_()
            mt.check(lib.b200mt_scale(L_scale.ref(), CHUNK, fp, 1.0, gs.cuda_stream))
            mt.check(lib.b200mt_adam(L_adam.ref(), CHUNK, fp, ADAM["lr"], ADAM["beta1"], ADAM["beta2"], ADAM["eps"],
                                     1000, 1, 1, ADAM["wd"], 1.0, gs.cuda_stream))
        stream.wait_stream(gs)
        for _ in range(3):
            graph.replay()
        torch.cuda.synchronize()
        e0.record(stream)
        for _ in range(mixed_steps):
            graph.replay()
        e1.record(stream)
        torch.cuda.synchronize()
        graph_ms = e0.elapsed_time(e1) / mixed_steps
    except Exception as e:      # reported, not fatal: the headline numbers above do not depend on it
        sys.stderr.write("bench: CUDA-graph leg failed: %r\n" % (e,))

    peak, peak_src = peaks()
    adam_bytes = 28.0 * nparams
    achieved = adam_bytes / (adam_avg * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "adam_traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get("dram_bytes_per_launch")
        except Exception:
            traffic = None
    if os.environ.get("B200MT_BENCH_SKIP_CPU"):   # profiling runs only: keep the command short under ncu
        cpu = {"value": None, "cores": 0, "steps": 0, "host_cpus": os.cpu_count(), "ms_per_step": None}
    else:
        cpu = cpu_reference_run(steps=40, warmup=2, max_seconds=20.0)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": single_config(nparams, len(shapes)),
        "step_gbs": 36.0 * nparams / (ms_per_step * 1e-3) / 1e9,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "kernel": "mt_stream_kernel<AdamOp<float,float,float,false>>",
                     "bytes_per_launch": adam_bytes, "avg_launch_ms": adam_avg, "peak_source": peak_src,
                     "frac_of_8TBs_nominal": achieved / 8000.0},
        "cpu_baseline": {"value": cpu["value"], "unit": UNIT, "cores": cpu["cores"], "kind": "port",
                         "sample": "%d steps of the same ResNet-50 list on host CPU (python-only unscale + torch.optim.Adam)" % cpu["steps"],
                         "host_cpus": cpu["host_cpus"], "ms_per_step": cpu["ms_per_step"]},
        "e2e": {"value": nparams / (e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": 4 * nparams, "d2h_bytes_per_step": 4,
                "ms_per_step": e2e_ms, "steps": e2e_steps,
                "segments": len(bounds) - 1, "pcie_floor_ms_at_55GBs": 4 * nparams / 55e9 * 1e3,
                "blocks_ms_per_step": e2e_blocks, "statistic": "median of %d blocks of %d steps" % (E2E_BLOCKS, e2e_steps),
                "api": "pinned-host gradients copied in %d segments on a copy stream; per segment amp_C.multi_tensor_scale + "
                       "amp_C.multi_tensor_adam (drop-in entry points); flag.item() at the end" % (len(bounds) - 1)},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "mixed_precision_adam": {"what": "single-pass multi_tensor_adam [g fp16 (unscale fused), p/m/v fp32, p_copy fp16] on the same list, "
                                         "back-to-back launches", "ms_per_launch": mixed_ms, "value": nparams / (mixed_ms * 1e-3), "unit": UNIT,
                                 "achieved_gbs": 28.0 * nparams / (mixed_ms * 1e-3) / 1e9,
                                 "frac": 28.0 * nparams / (mixed_ms * 1e-3) / 1e9 / peak},
        "cuda_graph_replay": {"ms_per_step": graph_ms, "value": (nparams / (graph_ms * 1e-3)) if graph_ms else None, "unit": UNIT,
                              "what": "flag clear + b200mt_scale + b200mt_adam captured once with torch.cuda.graph and replayed "
                                      "(auxiliary: the captured Adam keeps one step number)"},
        "python_api": {"value": nparams / (api_ms * 1e-3), "unit": UNIT, "ms_per_step": api_ms, "host_us_per_step": api_host_us,
                       "native_packer": mt._fast_pack is not None,
                       "what": "same resident step through amp_C.multi_tensor_scale + amp_C.multi_tensor_adam (2 x list packing per step)"},
        "table_uploads": int(lib.b200mt_table_uploads()), "table_hits": int(lib.b200mt_table_hits()),
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device. This path has no CPU fallback (use --impl reference for the CPU arm).")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 or world > 1:
        from bench_sharded import run_sharded
        run_sharded(args)
    else:
        run_single(args)


if __name__ == "__main__":
    main()
