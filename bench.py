#!/usr/bin/env python
"""bench.py -- throughput of the sketch-rnn hot path on B200 (driver contract).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--precision f32|bf16]
  python bench.py --impl reference ...      # the CPU restatement of the reference graph, timed

Workloads (BASELINE.json configs):
  train_h512_l3_b800  configs[3]: 3-layer LSTM rnn_size 512, 24 mixtures, global batch 800,
                      seq 300, data-parallel over N GPUs (default; metric train stroke-steps/s)
  train_kanji         configs[0]: the reference default, 2-layer LSTM 256, batch 100, seq 300
  train_gru / train_rnn   configs[2]
  sample_8192         configs[4]: 8192 concurrent sketches x 300 steps, sharded over N GPUs
  sample_10           configs[1]: 10 sketches x 300 steps (sample.py's size; bit-exact fp32 persistent kernel, 1 GPU)

One "step" = one full training step (forward, MDN loss, BPTT, global-norm clip, Adam; plus the
gradient all-reduce when N > 1) over one synthetic batch, or one whole sampling call.  Rank 0
prints ONE JSON line.  Inputs are synthetic (the dataset and checkpoint blobs are absent from the
reference mount) with the shapes and statistics of SURVEY.md 8(d).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

# The CPU legs use every host core.  torch.distributed.run exports OMP_NUM_THREADS=1 to its workers when more than
# one is launched, and the BLAS behind NumPy reads that variable when NumPy is first imported: put the real core
# count back BEFORE the import (the thread count actually used is printed with every CPU number).
if "reference" in sys.argv or os.environ.get("WORLD_SIZE", "1") == "1":
    for _v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[_v] = str(os.cpu_count() or 1)

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    "train_h512_l3_b800": dict(kind="train", model="lstm", H=512, L=3, M=24, B=800, T=300),
    "train_kanji": dict(kind="train", model="lstm", H=256, L=2, M=24, B=100, T=300),
    "train_gru": dict(kind="train", model="gru", H=256, L=2, M=24, B=100, T=300),
    "train_rnn": dict(kind="train", model="rnn", H=256, L=2, M=24, B=100, T=300),
    "sample_8192": dict(kind="sample", model="lstm", H=256, L=2, M=24, B=8192, T=300),
    "sample_10": dict(kind="sample", model="lstm", H=256, L=2, M=24, B=10, T=300),
    "train_h512_l3_b100": dict(kind="train", model="lstm", H=512, L=3, M=24, B=100, T=300),  # one rank's share at 8 GPUs
}


def fwd_flops_per_stroke_step(w):
    """sum_l 2*(I_l+H)*G + 2*H*NOUT."""
    ng = {"lstm": 4, "gru": 3, "rnn": 1}[w["model"]]
    H, L, NOUT = w["H"], w["L"], 3 + 6 * w["M"]
    f = 2 * H * NOUT
    for l in range(L):
        I = 5 if l == 0 else H
        f += 2 * (I + H) * ng * H
    return f


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tc_burst=d["bf16_tflops"], tc_sustained=d["bf16_tflops_sustained"],
                    src="measured")
    return dict(hbm=6650.0, tc_burst=1590.0, tc_sustained=1400.0, src="fallback")


def ncu_traffic(workload, precision, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel`, from the `ncu --set full` captures
    summarised under profiles/ (profiles/ncu_traffic.json names the capture each figure comes from)."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(p):
        return None
    e = json.load(open(p)).get("%s/%s/%s" % (workload, precision, kernel))
    return float(e["dram_bytes_per_launch"]) if e else None


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons sampled DURING the timed region: NVML (a few microseconds per sample, so even a
    30 ms region gets several) with nvidia-smi as the fallback."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}

    def __init__(self, index):
        threading.Thread.__init__(self, daemon=True)
        self.index = index
        self.samples = []   # (sm_mhz, max_mhz, set(reasons))
        self.stop_flag = False
        self.nv = None
        try:
            import pynvml
            import torch
            pynvml.nvmlInit()
            pr = torch.cuda.get_device_properties(index)
            try:
                bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
                self.handle = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
            except Exception:
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nv = pynvml
        except Exception:
            self.nv = None

    def _nvml_sample(self):
        nv = self.nv
        sm = float(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM))
        try:
            mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
        except Exception:
            mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
        self.samples.append((sm, self.max_mhz, {n for b, n in self.REASONS.items() if mask & b}))

    def _smi_sample(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                             capture_output=True, text=True, timeout=5).stdout.strip()
        if out:
            f = [v.strip() for v in out.split(",")]
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            self.samples.append((float(f[0]), float(f[1]), {n for n, v in zip(names, f[2:6]) if v.lower().startswith("active")}))

    def run(self):
        while not self.stop_flag:
            try:
                if self.nv is not None:
                    self._nvml_sample()
                else:
                    self._smi_sample()
            except Exception:
                pass
            time.sleep(0.005 if self.nv is not None else 0.2)

    def summary(self):
        self.stop_flag = True
        sm = [v[0] for v in self.samples]
        mx = max([v[1] for v in self.samples], default=0)
        reasons = set()
        for v in self.samples:
            reasons |= v[2]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm), "source": "nvml" if self.nv is not None else "nvidia-smi"}


# --------------------------------------------------------------------------------------------
# CPU arm: the oracle (NumPy restatement of the reference graph) on the host cores
# --------------------------------------------------------------------------------------------
CPU_SAMPLE_ROWS = 100   # batch rows of the workload the CPU legs time (config 4 has 800; 100 rows x 300 steps ~ 8 s)


def blas_threads():
    """Threads the BLAS behind NumPy will really use (not what we hoped for)."""
    try:
        import threadpoolctl
        n = [i.get("num_threads", 1) for i in threadpoolctl.threadpool_info() if i.get("user_api") == "blas"]
        return int(max(n)) if n else 1
    except Exception:
        return None


def cpu_baseline(w, inputs=None):
    """Times the oracle on a bounded sample of the workload and returns the `cpu_baseline` object.
    inputs = (flat_params, x_rows, y_rows, mask_rows) from the GPU arm: the CPU step then runs on exactly those rows
    with exactly those weights, and its losses come back under "losses" for bench.py's loss check."""
    from oracle import sketch_oracle as so
    cfg = so.Config(model=w["model"], rnn_size=w["H"], num_layers=w["L"], num_mixture=w["M"],
                    batch_size=w["B"], seq_length=w["T"], keep_prob=0.8)
    nthr = blas_threads()
    if w["kind"] == "train":
        if inputs is not None:
            flat, x, y, mask = inputs
            params = so.unflatten(cfg, np.asarray(flat, dtype=np.float32))
            Bs = x.shape[0]
        else:
            Bs = min(w["B"], CPU_SAMPLE_ROWS)
            params = so.init_params(cfg, seed=0)
            x, y = so.synthetic_batch(Bs, w["T"], seed=1)
            mask = so.synthetic_dropout_mask(cfg, Bs, w["T"])
        opt = so.AdamState(cfg)
        st = so.zero_state(cfg, Bs)
        t0 = time.time()
        losses, _, _ = so.train_step(cfg, params, opt, x, y, st, 0.005, mask)
        dt = time.time() - t0
        out = {"value": Bs * w["T"] / dt, "unit": "stroke-steps/s", "cores": nthr or 1, "kind": "port",
               "sample": "1 oracle train step (fwd + BPTT + clip + Adam) on %d of %d batch rows x %d steps, NumPy fp32, "
                         "%s BLAS threads on %d host cores" % (Bs, w["B"], w["T"], nthr, os.cpu_count() or 1),
               "host_cores": os.cpu_count() or 1, "blas_threads": nthr, "sample_seconds": dt,
               "losses": [float(v) for v in losses]}
        if Bs < w["B"]:
            out["extrapolated_from_rows"] = Bs   # the rate of a Bs-row step stands for the B-row step
        return out
    from oracle import sampler_ref
    params = so.init_params(cfg, seed=0)
    ns, steps = 2, w["T"]
    u = np.random.RandomState(1).rand(ns, steps, 4).astype(np.float32)
    t0 = time.time()
    for i in range(ns):
        sampler_ref.sample(cfg, params, u[i], steps, 0.1, 0.1, False)
    dt = time.time() - t0
    return {"value": ns / dt, "unit": "sketches/s", "cores": 1, "kind": "port",
            "sample": "%d sketches x %d steps through oracle/sampler_ref.c (scalar C, 1 thread)" % (ns, steps),
            "host_cores": os.cpu_count() or 1, "sample_seconds": dt, "extrapolated_from_rows": ns}


def run_reference(args, w):
    """--impl reference: the reference's own path cannot run (Python 2 + TensorFlow 0.x); the
    timed stand-in is the oracle port of its graph on the host cores, same config and metric.
    Each "step" is one bounded sample (CPU_SAMPLE_ROWS batch rows of the workload, all T timesteps)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    vals = []
    cb = None
    nwarm = 1 if args.warmup > 0 else 0
    for i in range(nwarm + args.steps):
        cb = cpu_baseline(w)
        if i >= nwarm:
            vals.append(cb["value"])
    value = float(np.median(vals))
    cb["value"] = value
    cb.pop("losses", None)
    metric = "train_stroke_steps_per_sec" if w["kind"] == "train" else "sampled_sketches_per_sec"
    line = {"impl": "reference", "metric": metric, "value": value, "unit": cb["unit"], "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup,
            # one step of the FULL workload at the sampled rate (each timed step is the bounded sample in `sample`)
            "ms_per_step": (w["B"] * (w["T"] if w["kind"] == "train" else 1)) / value * 1e3,
            "ms_per_sampled_step": float(np.median([cb["sample_seconds"]])) * 1e3,
            "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, w),
            "cpu_baseline": cb,
            "e2e": {"value": value, "unit": cb["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def workload_config(args, w):
    c = {"workload": args.workload, "cell": w["model"], "rnn_size": w["H"], "num_layers": w["L"],
         "num_mixture": w["M"], "seq_length": w["T"], "precision": args.precision,
         "cache": "per-step activations (%.0f MB) exceed the 126 MB L2; no explicit flush" % act_mb(w)}
    if w["kind"] == "train":
        c.update({"global_batch": w["B"], "keep_prob": 0.8, "grad_clip": 5.0,
                  "parallelism": "dp%d (batch rows sharded, fp32 gradient all-reduce)" % args.gpus})
    else:
        c.update({"n_sketches": w["B"], "temperature": 0.1, "stop_if_eoc": False,
                  "parallelism": "sketches sharded over %d GPU(s), no communication" % args.gpus})
    return c


def act_mb(w):
    ng = {"lstm": 4, "gru": 3, "rnn": 1}[w["model"]]
    return w["B"] * w["T"] * w["L"] * (ng + 3) * w["H"] * 4 / 1e6


# --------------------------------------------------------------------------------------------
# loss check: the GPU arm's losses on its first rows against the oracle on the same rows and weights
# --------------------------------------------------------------------------------------------
def gpu_slice_losses(h, w, x_rows, y_rows, mask_rows):
    """cost / cost_shape / cost_pen of the handle on `x_rows` from a zero state with an injected dropout mask."""
    import torch
    R, T = x_rows.shape[0], x_rows.shape[1]
    st = torch.zeros((w["L"], R, h.S), dtype=torch.float32, device="cuda:%d" % h.device)
    losses, _ = h.loss_grad(x_rows, y_rows, st, dropout_mask=mask_rows, loss_rows=R * T)
    return [float(v) for v in losses.cpu().numpy()]


def compare_losses(gpu, oracle, precision):
    tol = 1e-4 if precision == "f32" else 2e-2
    rel = max(abs(g - o) / abs(o) for g, o in zip(gpu, oracle))
    return {"gpu": gpu, "oracle": oracle, "rel_err": rel, "tolerance": tol, "ok": bool(rel < tol)}


def loss_check(w, precision, rows=CPU_SAMPLE_ROWS):
    """Stand-alone form of the check run_ours makes (used by tests/): build the workload's handle, take the losses
    of its first `rows` rows on the GPU, run the oracle on the same rows and weights, compare."""
    from sketch_rnn_b200 import _lib, utils
    R = min(rows, w["B"])
    h = _lib.Handle(model=w["model"], rnn_size=w["H"], num_layers=w["L"], num_mixture=w["M"], batch_size=R,
                    seq_length=w["T"], keep_prob=0.8, precision=precision)
    flat = utils.default_init(h, seed=0)
    h.set_params(flat)
    x, y = utils.synthetic_batch(w["B"], w["T"], seed=1)
    mask = utils.synthetic_dropout_mask(R, w["T"], w["H"], 0.8, seed=2)
    gpu = gpu_slice_losses(h, w, x[:R], y[:R], mask)
    h.close()
    cb = cpu_baseline(w, inputs=(flat, x[:R], y[:R], mask))
    out = compare_losses(gpu, cb["losses"], precision)
    out["rows"] = R
    return out


# --------------------------------------------------------------------------------------------
# sampling block of the default line (BASELINE configs 2 and 5 next to the training metric)
# --------------------------------------------------------------------------------------------
def sampling_block(world, rank, local):
    """sketches/s of Model.sample's replacement: config 5 (8192 sketches x 300 steps sharded over the ranks) on the
    bit-exact fp32 kernel and on the bf16 tensor-core sampler, and config 2's size (10 sketches, rank 0 only).
    CUDA-event times, max over ranks."""
    import torch
    import torch.distributed as dist
    from sketch_rnn_b200 import _lib, utils
    dev = torch.device("cuda:%d" % local)
    out = {}

    def timed(h, n, iters, warm, **kw):
        for i in range(warm):
            h.sample(n, 300, 0.1, 0.1, False, seed=i, **kw)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(iters):
            h.sample(n, 300, 0.1, 0.1, False, seed=100 + i, **kw)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n = 8192 // world
    for prec, key, iters in (("bf16", "config5_bf16", 5), ("f32", "config5_exact", 2)):
        h = _lib.Handle(model="lstm", rnn_size=256, num_layers=2, num_mixture=24, precision=prec, device=local)
        h.set_params(utils.default_init(h, seed=0))
        l0 = h.launch_count()
        ms = timed(h, n, iters, 2, sketch0=rank * n)
        out[key] = {"sketches_per_sec": 8192 / (ms * 1e-3), "ms": ms, "n_sketches": 8192, "steps": 300,
                    "sketches_per_gpu": n, "bit_exact": prec == "f32",
                    "launches_per_call": (h.launch_count() - l0) // (iters + 2),
                    "kernel": "sample_kernel (persistent, fp32, bit-exact)" if prec == "f32" else
                              "sample_tp_kernel (persistent tcgen05 kernel: 64 sketches per CTA for all steps, h operands "
                              "resident in shared memory, weights streamed by TMA, draw in the epilogue)"}
        if prec == "f32" and rank == 0:
            # config 2's size: sample.py draws 10 sketches from one model
            ms10 = None
            for i in range(2):
                h.sample(10, 300, 0.1, 0.1, False, seed=i)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(5):
                h.sample(10, 300, 0.1, 0.1, False, seed=50 + i)
            e1.record()
            torch.cuda.synchronize()
            ms10 = e0.elapsed_time(e1) / 5
            out["config2_size"] = {"sketches_per_sec": 10 / (ms10 * 1e-3), "ms": ms10, "n_sketches": 10, "steps": 300,
                                   "bit_exact": True, "launches_per_call": 1,
                                   "kernel": "sample_cluster_kernel (fp32 weights resident in a 16-CTA cluster, bit-exact)",
                                   "note": "synthetic weights (save/kanji/model.ckpt-0 is absent); "
                                   "fixed 300 steps, no early stop"}
        h.close()
    flop = 1.0 * fwd_flops_per_stroke_step(WORKLOADS["sample_8192"]) * 8192 * 300
    out["config5_bf16"]["tflops"] = flop / (out["config5_bf16"]["ms"] * 1e-3) / 1e12 / world
    return out


# --------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------
def run_ours(args, w):
    import torch
    import torch.distributed as dist
    from sketch_rnn_b200 import _lib, dp, utils   # the GPU arm never touches oracle/

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda:%d" % local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda:%d" % local)
    assert args.gpus == world, "--gpus %d but WORLD_SIZE=%d (launch with torchrun)" % (args.gpus, world)
    assert w["B"] % world == 0
    Bl = w["B"] // world
    T = w["T"]
    h = _lib.Handle(model=w["model"], rnn_size=w["H"], num_layers=w["L"], num_mixture=w["M"],
                    batch_size=Bl if w["kind"] == "train" else 0, seq_length=T if w["kind"] == "train" else 0,
                    keep_prob=0.8, precision=args.precision, device=local)
    h.set_params(utils.default_init(h, seed=0))
    S = h.S
    peaks = load_peaks()
    clocks = ClockSampler(local)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    lr = 0.005
    timing = None
    if w["kind"] == "train":
        xg, yg = utils.synthetic_batch(w["B"], T, seed=1)
        x_host = np.ascontiguousarray(xg[rank * Bl:(rank + 1) * Bl])
        y_host = np.ascontiguousarray(yg[rank * Bl:(rank + 1) * Bl])
        x = torch.from_numpy(x_host).to(dev)
        y = torch.from_numpy(y_host).to(dev)
        state = torch.zeros((w["L"], Bl, S), dtype=torch.float32, device=dev)
        loss_rows = w["B"] * T
        check_inputs = check_gpu = None
        if rank == 0 and world == 1 and not args.no_cpu:
            # loss check, GPU half: the first rows of the batch with the initial weights and an injected mask; the
            # cpu_baseline leg below runs the oracle on exactly these and the two sets of losses are compared
            R = min(Bl, CPU_SAMPLE_ROWS)
            mask_rows = utils.synthetic_dropout_mask(R, T, w["H"], 0.8, seed=2)
            check_inputs = (h.get_params(), x_host[:R], y_host[:R], mask_rows)
            check_gpu = gpu_slice_losses(h, w, x_host[:R], y_host[:R], mask_rows)

        def step(i):
            nonlocal state
            losses, state = dp.dp_train_step(h, x, y, state, lr, 1000 + i, rank * Bl, loss_rows, timing=timing)
            return losses

        units_per_step = w["B"] * T
        unit, metric = "stroke-steps/s", "train_stroke_steps_per_sec"
    else:
        def step(i):
            return h.sample(Bl, T, 0.1, 0.1, False, seed=7 + i, sketch0=rank * Bl)[1]

        units_per_step = w["B"]
        unit, metric = "sketches/s", "sampled_sketches_per_sec"

    for i in range(args.warmup):
        step(i)
    barrier()
    l0 = h.launch_count()
    clocks.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    last = None
    for i in range(args.steps):
        last = step(args.warmup + i)
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    launches = h.launch_count() - l0
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    clk = clocks.summary()
    ms_per_step = ms / args.steps
    value = units_per_step / (ms_per_step * 1e-3)

    # per-category kernel time, measured live with CUDA events around every launch (separate pass,
    # so the event overhead is not inside the headline number)
    h.profile_enable(True)
    nprof = min(args.steps, 3)
    if w["kind"] == "train" and world > 1:
        timing = {}
    for i in range(nprof):
        step(args.warmup + args.steps + i)
    prof = h.profile_read()
    h.profile_enable(False)
    kern_ms = {k: v[0] / nprof for k, v in prof.items() if v[1] > 0}
    kern_cnt = {k: v[1] // nprof for k, v in prof.items() if v[1] > 0}
    if timing:
        kern_ms["allreduce"] = timing["allreduce_ms"] / nprof   # NCCL, not one of this library's kernels
        kern_cnt["allreduce"] = 2
    timing = None

    # e2e through the host-buffer entry point: H2D of the step's inputs + D2H of its result inside
    e2e = None
    if w["kind"] == "train":
        # The same host-buffer step at every N (dp.HostStep: pinned host batch -> device, loss_grad, all-reduce when
        # N > 1, clip + Adam, losses and final state back to pinned host memory), max over ranks.  At N = 1 the
        # C-ABI's own host entry point skrnn_train_step_host is timed beside it (`c_abi_value`).
        xp, yp = torch.from_numpy(x_host).pin_memory(), torch.from_numpy(y_host).pin_memory()
        hs = dp.HostStep(h, Bl, T, loss_rows, row0=rank * Bl)
        ne = max(3, min(args.steps, 10))
        for i in range(2):
            hs(xp, yp, lr, 4000 + i)
        barrier()
        t0 = time.perf_counter()
        for i in range(ne):
            hs(xp, yp, lr, 5000 + i)
        barrier()
        dt = torch.tensor([(time.perf_counter() - t0) / ne], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": units_per_step / float(dt.item()), "unit": unit,
               "h2d_bytes_per_step": int(hs.h2d_bytes) * world, "d2h_bytes_per_step": int(hs.d2h_bytes) * world,
               "path": "sketch_rnn_b200.dp.HostStep (same call at every N)"}
        if world == 1:
            st_host = torch.zeros((w["L"], Bl, S), dtype=torch.float32).pin_memory().numpy()
            xh, yh = xp.numpy(), yp.numpy()
            for i in range(2):
                h.train_step_host(xh, yh, st_host, lr, seed=i)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for i in range(ne):
                h.train_step_host(xh, yh, st_host, lr, seed=10 + i)
            torch.cuda.synchronize()
            e2e["c_abi_value"] = units_per_step / ((time.perf_counter() - t0) / ne)
            e2e["c_abi_path"] = "skrnn_train_step_host"
    else:
        # results land in pinned host buffers (what a serving caller would keep around), two sets: the copy of call i
        # runs on its own stream while call i + 1 samples, and the host waits for call i's strokes before it issues
        # call i + 2 -- every call's strokes and lengths are copied and waited for inside the timed region
        sh = [torch.empty((Bl, T, 5), dtype=torch.float32).pin_memory() for _ in range(2)]
        lh = [torch.empty((Bl,), dtype=torch.int32).pin_memory() for _ in range(2)]
        copy_stream = torch.cuda.Stream(device=dev)
        main_stream = torch.cuda.current_stream(dev)
        done = [None, None]
        ne = 6

        def call(i):
            k = i & 1
            if done[k] is not None:
                done[k].synchronize()   # the host has call i - 2's result before its buffers are reused
            strokes, lengths, _ = h.sample(Bl, T, 0.1, 0.1, False, seed=100 + i, sketch0=rank * Bl)
            ready = torch.cuda.Event()
            ready.record(main_stream)
            with torch.cuda.stream(copy_stream):
                copy_stream.wait_event(ready)
                sh[k].copy_(strokes, non_blocking=True)
                lh[k].copy_(lengths, non_blocking=True)
                strokes.record_stream(copy_stream)
                lengths.record_stream(copy_stream)
                done[k] = torch.cuda.Event()
                done[k].record(copy_stream)

        call(0)
        torch.cuda.synchronize()
        done = [None, None]
        barrier()
        t0 = time.perf_counter()
        for i in range(ne):
            call(i)
        for k in range(2):
            done[k].synchronize()
        barrier()
        dt = torch.tensor([(time.perf_counter() - t0) / ne], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        e2e = {"value": units_per_step / float(dt.item()), "unit": unit, "h2d_bytes_per_step": 0,
               "d2h_bytes_per_step": int(sh[0].numel() * 4 + Bl * 4) * world,
               "path": "Handle.sample -> pinned host buffers, double-buffered: the copy of call i overlaps call i + 1"}

    if rank == 0:
        fl = fwd_flops_per_stroke_step(w)
        if w["kind"] == "train":
            flops_step = 3.0 * fl * Bl * T   # per GPU
        else:
            flops_step = 1.0 * fl * Bl * T
        # dominant kernel by measured time; its roofline = algorithmic work of ONE launch / its mean duration
        own = {k: v for k, v in kern_ms.items() if k != "allreduce"}
        dom = max(own, key=own.get) if own else None
        # bf16 sampling of many LSTM sketches runs per-step tcgen05 GEMMs (csrc/sampler_tc.cu)
        samp_tc = (w["kind"] != "train" and args.precision == "bf16" and w["model"] == "lstm" and Bl >= 64
                   and os.environ.get("SKRNN_SAMPLER_EXACT", "0") != "1")
        tc = args.precision == "bf16" and (w["kind"] == "train" or samp_tc)
        ng = {"lstm": 4, "gru": 3, "rnn": 1}[w["model"]]
        H, L, NOUT = w["H"], w["L"], 3 + 6 * w["M"]
        rows = Bl * T
        per_launch_flops = {
            # one launch = all T steps of one layer: h[B,H] x Wh[H,G] per step (BPTT: d(pre)[B,G] x Wh^T)
            "rnn_fwd": 2.0 * rows * H * ng * H, "rnn_bwd": 2.0 * rows * H * ng * H,
            # GEMM launches differ in shape: report the mean over the step's launches
            "gemm_tc": None, "gemm_f32": None,
            "sampler": 1.0 * fl * rows,
        }
        gemm_total = (flops_step - (3.0 if w["kind"] == "train" else 1.0) * 2.0 * rows * H * ng * H * L) \
            if w["kind"] == "train" else (flops_step if samp_tc else 0.0)
        n_dom = max(1, kern_cnt.get(dom, 1)) if dom else 1
        if dom in ("gemm_tc", "gemm_f32"):
            fpl = gemm_total / n_dom
        elif dom in per_launch_flops and per_launch_flops[dom]:
            fpl = per_launch_flops[dom]
        else:
            fpl = flops_step / n_dom
        dur_ms = kern_ms[dom] / n_dom if dom else ms_per_step
        peak = peaks["tc_sustained"] if tc and dom != "sampler" else 2 * 128 * 148 * 1.965e-3
        ach = fpl / (dur_ms * 1e-3) / 1e12
        whole = flops_step / (ms_per_step * 1e-3) / 1e12
        # rnn_size 256 x 2 layers: the whole call is ONE persistent tcgen05 kernel (csrc/sampler_tp.cu), no per-step GEMMs
        samp_tp = samp_tc and dom == "sampler" and "gemm_tc" not in kern_ms
        if samp_tp:
            dur_ms = kern_ms["sampler"]   # the persistent kernel plus microseconds of weight-image packing
            ach = flops_step / (dur_ms * 1e-3) / 1e12
            roof = {"bound": "tensor", "achieved": ach, "peak": peaks["tc_sustained"], "unit": "TFLOP/s",
                    "frac": ach / peaks["tc_sustained"], "traffic": None,
                    "kernel": "sample_tp_kernel (persistent: every step of 64 sketches per CTA)", "launch_ms": dur_ms,
                    "launches_per_step": 1, "algorithmic_flops_per_launch": flops_step,
                    "per_timestep_us": dur_ms * 1e3 / T,
                    "peak_source": "MEASURED_PEAKS.json bf16 sustained (%s)" % peaks["src"],
                    "note": "M = 64 products (half the tensor pipe's rows) on 128 of 148 SMs, a step is a chain of dependent "
                            "chunks, and each CTA streams the stack's 1.6 MB of weights from L2 every step: see DESIGN.md "
                            "section 4",
                    "whole_step_tflops": whole, "per_kernel_ms": kern_ms, "per_kernel_launches": kern_cnt}
        elif samp_tc and dom == "sampler":
            # cell-update + draw kernels between the GEMMs: streaming, HBM/L2 bound.  Bytes per sketch-step:
            # per layer gates 4H fp32 read, c fp32 read+write, h bf16 written twice; logits NOUT fp32 read,
            # point written (fp32 strokes + fp32/bf16 feedback)
            bytes_step = rows * (L * (ng * H * 4 + 2 * H * 4 + 2 * H * 2) + NOUT * 4 + 5 * 4 * 2 + 10)
            ach = bytes_step / n_dom / (dur_ms * 1e-3) / 1e9
            roof = {"bound": "hbm", "achieved": ach, "peak": peaks["hbm"], "unit": "GB/s", "frac": ach / peaks["hbm"],
                    "traffic": None, "kernel": "sampler (cell update + draw)", "launch_ms": dur_ms,
                    "launches_per_step": n_dom, "algorithmic_bytes_per_launch": bytes_step / n_dom,
                    "peak_source": "MEASURED_PEAKS.json HBM copy (%s)" % peaks["src"],
                    "whole_step_tflops": whole, "per_kernel_ms": kern_ms, "per_kernel_launches": kern_cnt}
        else:
            roof = {"bound": "tensor", "achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak,
                    "traffic": ncu_traffic(args.workload, args.precision, dom) if world == 1 else None,
                    "kernel": dom, "launch_ms": dur_ms, "launches_per_step": n_dom,
                    "algorithmic_flops_per_launch": fpl,
                    # recurrent kernels: one launch = T dependent steps; their per-step latency is the figure to watch
                    "per_timestep_us": (dur_ms * 1e3 / T) if dom in ("rnn_fwd", "rnn_bwd") else None,
                    "peak_source": ("MEASURED_PEAKS.json bf16 sustained (%s)" % peaks["src"]) if (tc and dom != "sampler") else
                    "fp32 CUDA-core nominal 148 SM x 128 lanes x 2 x 1.965 GHz (this kernel runs no tensor-core math)",
                    "note": "persistent recurrent kernels are bound by per-timestep latency and the L2 exchange of h / d(h), "
                            "not by the tensor pipe: see DESIGN.md section 4 and profiles/",
                    "whole_step_tflops": whole, "per_kernel_ms": kern_ms, "per_kernel_launches": kern_cnt}
        cb = None
        lcheck = None
        if world == 1 and not args.no_cpu:
            cb = cpu_baseline(w, inputs=check_inputs if w["kind"] == "train" else None)
            if w["kind"] == "train":
                lcheck = compare_losses(check_gpu, cb.pop("losses"), args.precision)
                lcheck["rows"] = int(check_inputs[1].shape[0])
        line = {"metric": metric, "value": value, "unit": unit, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
                "scaling": "weak" if args.weak else "strong", "vs_baseline": None,
                "dtype": args.precision if (w["kind"] == "train" or samp_tc) else "f32", "data": "synthetic",
                "config": workload_config(args, w), "clocks": clk, "e2e": e2e,
                "gpu_launches": int(launches), "roofline": roof, "cpu_baseline": cb,
                "last_loss": [float(v) for v in last.float().cpu().numpy()[:3]] if w["kind"] == "train" else None,
                "loss_check": lcheck}
    h.close()
    if w["kind"] == "train" and args.workload == "train_h512_l3_b800" and not args.no_sampling:
        samp = sampling_block(world, rank, local)
        if rank == 0:
            line["sampling"] = samp
    if rank == 0:
        print(json.dumps(line))
        if line.get("loss_check") and not line["loss_check"]["ok"]:
            sys.stderr.write("bench.py: GPU losses disagree with the oracle on the checked rows: %r\n" % (line["loss_check"],))
            sys.exit(3)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="train_h512_l3_b800", choices=sorted(WORKLOADS))
    ap.add_argument("--precision", default="bf16", choices=["f32", "bf16"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (and the loss check that rides on it)")
    ap.add_argument("--no-sampling", action="store_true", help="skip the sampling block of the default line")
    ap.add_argument("--weak", action="store_true",
                    help="weak scaling: the workload's batch is the PER-GPU batch (default: it is the global batch)")
    args = ap.parse_args()
    w = dict(WORKLOADS[args.workload])
    if args.weak:
        w["B"] = w["B"] * int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, w)
    else:
        run_ours(args, w)


if __name__ == "__main__":
    main()
