#!/usr/bin/env python
"""bench.py -- headline benchmark of the scale-hyperprior hot path on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload train|train_ssim|eval]

Own arm (default): one "step" = one pass of the hot path over one batch of synthetic input:
  train       psnr_256 training step = Compressor2018 forward (noise quantisation) + RD loss + backward
              (+ gradient all-reduce when N>1), 64 crops of 256x256 per GPU, bf16 operands / fp32 accumulate
              (BASELINE.json configs[1], the configuration the metric is quoted on)
  train_ssim  the same with the ssim_64 recipe (MS_SSIMLoss, log scale)  (configs[3])
  eval        forward in rounding mode + bpp / PSNR / MS-SSIM(dB) on 768x512 images, sharded per image (configs[4])
`value` is whole-job throughput with inputs resident in HBM; `e2e` is the same metric through the public
module API with HOST (pinned) inputs copied in every step and the loss read back every step.

Reference arm (--impl reference): the reference's CPU implementation of the same step, timed on this box's
host cores with all threads.  The reference is pure Python/torch and is not present on the GPU box, so the
arm runs the oracle port (oracle/hyperprior_oracle.py, pinned bit-exactly to the reference modules by
tests/golden/make_golden.py) -- kind "port".

One JSON line on stdout (rank 0).
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
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = {"train": "train crops/s (256x256, fwd+bwd)", "train_ssim": "train crops/s (256x256, fwd+bwd)",
          "eval": "768x512 eval images/s"}
UNIT = {"train": "crops/s", "train_ssim": "crops/s", "eval": "images/s"}
RECIPE = {"train": "psnr_256", "train_ssim": "ssim_64", "eval": "psnr_256"}
L2_BYTES = 126 * 1024 * 1024

# algorithmic GEMM flops per 256x256 crop: forward 24.567 GFLOP, train step 73.23 GFLOP
TRAIN_GFLOP_PER_CROP = 73.23
EVAL_GFLOP_PER_IMAGE = 147.40


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="train", choices=list(METRIC))
    ap.add_argument("--no-graphs", action="store_true", help="eval workload: eager launches instead of a CUDA graph")
    ap.add_argument("--batch", type=int, default=None, help="crops (images) per GPU per step")
    ap.add_argument("--size", type=int, nargs=2, default=None, metavar=("H", "W"))
    ap.add_argument("--mode", default="bf16", choices=["bf16", "fp32", "bf16_simt"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--kernel-events", default="after", choices=["timed", "after"],
                    help="per-kernel-class event pairs: inside the timed region, or over the same number of steps right "
                         "after it (an event record between two kernels serialises their programmatic dependent launch)")
    ap.add_argument("--cpu-sample", type=int, default=0,
                    help="crops per CPU step (0: the reference arm uses the full per-GPU batch, the cpu_baseline leg 16)")
    ap.add_argument("--sustained", type=int, default=200, help="steps of the sustained sub-record (train workload)")
    ap.add_argument("--quick", action="store_true", help="headline measurement only (no extra / sustained / trainer records)")
    return ap.parse_args()


def workload_shape(args):
    if args.workload == "eval":
        return (args.batch or 3), tuple(args.size or (512, 768))
    return (args.batch or 64), tuple(args.size or (256, 256))


# ------------------------------------------------------------------------------------------
# clocks: sampled DURING the timed region (B200_PROFILING.md "clocks line")
# ------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0, period=0.1):
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._th = None
        self._nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nvml = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(index)
        except Exception:
            self._nvml = None

    def _sample_nvml(self):
        n = self._nvml
        self.samples.append(float(n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM)))
        if self.max_mhz is None:
            self.max_mhz = float(n.nvmlDeviceGetMaxClockInfo(self._h, n.NVML_CLOCK_SM))
        r = n.nvmlDeviceGetCurrentClocksEventReasons(self._h) if hasattr(
            n, "nvmlDeviceGetCurrentClocksEventReasons") else n.nvmlDeviceGetCurrentClocksThrottleReasons(self._h)
        for name, bit in (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20),
                          ("sw_power_cap", 0x4)):
            if r & bit:
                self.reasons.add(name)

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                              str(self.index)], capture_output=True, text=True, timeout=10).stdout.strip()
        f = [v.strip() for v in out.split(",")]
        self.samples.append(float(f[0]))
        self.max_mhz = float(f[1])
        for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[2:6]):
            if v.lower().startswith("active"):
                self.reasons.add(name)

    def _run(self):
        while not self._stop.is_set():
            try:
                self._sample_nvml() if self._nvml else self._sample_smi()
            except Exception:
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        self._th = threading.Thread(target=self._run, daemon=True)
        self._th.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._th.join(timeout=15)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    # B200_PROFILING.md fallback
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1590.0, "source": "fallback"}


def ncu_traffic(*kernel_keys):
    """DRAM bytes per launch (read + write) of a kernel class from the committed ncu capture of this same command
    (profiles/traffic.json, written by scripts/ncu_summarise.py from `ncu --metrics dram__bytes_*`): the launch-
    weighted average over the kernels of the class, or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    if not os.path.exists(p):
        return None
    try:
        with open(p) as f:
            t = json.load(f)
        tot = n = 0.0
        for k in kernel_keys:
            d = t.get(k)
            if d:
                tot += d["avg_bytes_per_launch"] * d["launches"]
                n += d["launches"]
        return tot / n if n else None
    except Exception:
        return None


# ------------------------------------------------------------------------------------------
# CPU legs: the reference's own arithmetic on the host cores (oracle port; checker/baseline only)
# ------------------------------------------------------------------------------------------
def cpu_step_fn(workload, n, hw):
    """Returns (step callable, units per step).  train*: forward + RD loss + backward of `n` crops in fp32
    (what engine/trainer.py:179-188 times); eval: forward in rounding mode + psnr + MS-SSIM of `n` images."""
    import torch
    from oracle import hyperprior_oracle as O
    torch.manual_seed(0)
    if workload == "train_ssim":
        cfg = O.default_cfg(lmbda=64.0, loss_names=("MS_SSIMLoss",), log_scale=True)
    else:
        cfg = O.default_cfg(lmbda=256.0)
    params = O.leaf_params(O.init_params(cfg, seed=0))
    H, W = hw
    x = torch.rand(n, 3, H, W)

    def train_step():
        for p in params.values():
            p.grad = None
        _, losses = O.compressor2018_forward(params, x, cfg, True)
        losses["total_loss"].backward()
        return float(losses["total_loss"].detach())

    def eval_step():
        with torch.no_grad():
            xt, losses = O.compressor2018_forward(params, x, cfg, False)
            ps = O.psnr(xt * 255, x * 255)
            ms = O.ms_ssim_metric(xt * 255, x * 255)
        return float(losses["bpp"]) + float(ps.sum()) + float(ms.sum())

    return (eval_step if workload == "eval" else train_step), n


def time_torch_gpu(workload, n, hw, steps, warmup):
    """The same oracle port in eager PyTorch (cuDNN / ATen, fp32 tensors, torch's default TF32 flags) on cuda:0: what a user of the reference runs
    on this GPU today (BASELINE.md "second baseline, same box").  A reported baseline like `cpu_baseline`; the
    product path never touches it.  Returns None when the port cannot run on the device."""
    import torch
    from oracle import hyperprior_oracle as O
    try:
        cfg = O.default_cfg(lmbda=256.0)     # torch's default precision flags, as the reference runs: cuDNN TF32 allowed
        torch.manual_seed(0)
        params = {k: v.detach().cuda().requires_grad_(True) for k, v in O.init_params(cfg, seed=0).items()}
        x = torch.rand(n, 3, *hw, device="cuda")

        def step():
            for p in params.values():
                p.grad = None
            if workload == "eval":
                with torch.no_grad():
                    return O.compressor2018_forward(params, x, cfg, False)[1]["bpp"]
            losses = O.compressor2018_forward(params, x, cfg, True)[1]
            losses["total_loss"].backward()
            return losses["total_loss"].detach()
        for _ in range(warmup):
            step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        return {"value": n / (ms / 1e3), "ms_per_step": ms, "units": n}
    except Exception as e:           # e.g. a CPU-only helper tensor inside the port
        return {"value": None, "error": f"{type(e).__name__}: {str(e)[:160]}"}


def time_cpu(workload, n, hw, steps, warmup):
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    fn, units = cpu_step_fn(workload, n, hw)
    for _ in range(warmup):
        fn()
    t0 = time.perf_counter()
    for _ in range(steps):
        fn()
    dt = time.perf_counter() - t0
    return {"value": units * steps / dt, "ms_per_step": 1e3 * dt / steps, "cores": torch.get_num_threads(),
            "units": units}


def run_reference(args):
    """--impl reference: the reference's CPU path (oracle port) on this box's host cores, on the own arm's config:
    every step is the full per-GPU batch of the workload (64 crops / 3 images), --steps and --warmup honoured; only
    if the probe step says the run would not end within the arm's time budget are the steps cut (and the line says so)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # rank 0 alone runs it
    n, hw = workload_shape(args)
    if args.cpu_sample:
        n = args.cpu_sample
    steps, warmup = args.steps, args.warmup
    budget_s = float(os.environ.get("HPC_REF_BUDGET_S", "300"))
    probe = time_cpu(args.workload, n, hw, 1, 1)                 # (counts as the first warm-up step)
    per = probe["ms_per_step"] / 1e3
    cut = None
    if per * (steps + warmup) > budget_s:
        warmup_run = min(warmup, 1)
        steps_run = max(1, min(steps, int((budget_s - per * warmup_run) / per)))
        cut = f"cut from {steps} steps / {warmup} warm-ups to {steps_run} / {warmup_run + 1}: {per:.1f} s per step, {budget_s:.0f} s budget"
    else:
        steps_run, warmup_run = steps, max(warmup - 1, 0)
    r = time_cpu(args.workload, n, hw, steps_run, warmup_run)
    sample = (f"{steps_run} step(s) of {n} x {hw[0]}x{hw[1]} synthetic "
              f"{'crops fwd+loss+bwd' if args.workload != 'eval' else 'images fwd+metrics'} in fp32, torch-CPU, "
              f"{r['cores']} threads, after {warmup_run + 1} warm-up step(s)" + (f" ({cut})" if cut else ""))
    line = {
        "impl": "reference", "metric": METRIC[args.workload], "value": r["value"], "unit": UNIT[args.workload],
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": DATA_NOTE,
        "config": config_dict(args),
        "cpu_baseline": {"value": r["value"], "unit": UNIT[args.workload], "cores": r["cores"], "kind": "port",
                         "sample": sample},
        "e2e": {"value": r["value"], "unit": UNIT[args.workload], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "steps_timed": steps_run,
    }
    emit(line)


DATA_NOTE = "synthetic (torch.rand images, random-init weights of the reference architecture)"


def config_dict(args, workload=None):
    """Names the workload; a function of the command line only, so both arms print the same `config`."""
    wl = workload or args.workload
    a = argparse.Namespace(**vars(args))
    a.workload = wl
    if wl != args.workload:
        a.batch = a.size = None
    per_gpu, hw = workload_shape(a)
    names = {"train": "psnr_256 training step (fwd+bwd, noise quantisation), 256x256 synthetic crops",
             "train_ssim": "ssim_64 MS-SSIM-distortion training step (fwd+bwd), 256x256 synthetic crops",
             "eval": "eval sweep: 768x512 Kodak-shaped synthetic images, rounding quantisation, bpp/PSNR/MS-SSIM"}
    return {"workload": names[wl], "recipe": RECIPE[wl], "per_gpu_batch": per_gpu,
            "global_batch": per_gpu * args.gpus, "height": hw[0], "width": hw[1], "channels": 192,
            "parallelism": f"dp{args.gpus}" if wl != "eval" else f"per-image x{args.gpus}",
            "l2": "GPU arm: inputs rotate over resident batches totalling > 126 MB (L2); one step streams > 1 GB of "
                  "activations.  CPU arm: not applicable",
            "timing": "GPU arm: CUDA events on the launching stream, max over ranks; CPU arm: perf_counter",
            "step": ("one pass of the hot path over one per-GPU batch: zero_grad, forward, RD loss, backward, gradient "
                     "all-reduce for N>1; every weight is marked updated before each step, so the bf16 operand copies of "
                     "all weights are re-packed in every step as in real training; optimizer step excluded (SURVEY 8d), "
                     "reported separately as `with_optimizer`") if wl != "eval" else
                    "one pass over one per-GPU batch of images: forward (rounding), bpp, PSNR, MS-SSIM; per-image results "
                    "all-gathered inside the step for N>1"}


# ------------------------------------------------------------------------------------------
# own arm
# ------------------------------------------------------------------------------------------
class Ctx:
    pass


def setup_dist():
    import torch
    import torch.distributed as dist
    c = Ctx()
    c.rank = int(os.environ.get("RANK", "0"))
    c.world = int(os.environ.get("WORLD_SIZE", "1"))
    c.local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the hot path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(c.local)
    c.dev = torch.device("cuda", c.local)
    if c.world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # (NCCL prints its version banner on file descriptor 1; main() has pointed fd 1 at stderr for the whole run
        # and emit() writes the one JSON line to the saved stdout)
        dist.init_process_group("nccl", device_id=c.dev)
    return c


def measure(args, c, wl, steps, warmup, with_clocks=False, with_optimizer=False, sustained_steps=0,
            trainer_sequence=False):
    """Time one workload on this rank's GPU.  Returns a dict (see main() for how it becomes the JSON line)."""
    import torch
    import torch.distributed as dist
    import image_compression_b200 as pkg
    from image_compression_b200 import _lib, hotpath as hp, sharding
    from image_compression_b200.utils.metric import image_metrics

    rank, world, dev = c.rank, c.world, c.dev
    a = argparse.Namespace(**vars(args))
    a.workload = wl
    if wl != args.workload:
        a.batch = a.size = None
    B, (H, W) = workload_shape(a)
    training = wl != "eval"
    torch.manual_seed(0)                       # identical random-init weights on every rank
    cfg = pkg.get_recipe_cfg(RECIPE[wl])
    model = pkg.build_model(cfg).to(dev)
    model.train(training)
    sharding.broadcast_parameters(model)
    torch.manual_seed(1000 + rank)             # per-rank inputs and in-kernel Philox noise streams
    n_rot = max(2, -(-int(1.6 * L2_BYTES) // (B * 3 * H * W * 4)))   # resident input batches: > L2 in total
    host = [torch.rand(B, 3, H, W).pin_memory() for _ in range(n_rot)]
    resident = [h.to(dev) for h in host]
    params = [p for p in model.parameters()]

    def eval_eager(x):
        with torch.no_grad():
            xt, losses = model(x)
            ps, ms = image_metrics(xt, x)
        return losses["bpp"], ps, ms

    graphed = None
    if not training and not args.no_graphs:
        # evaluation is launch bound (3 images per step, ~60 short kernels): replay the step as one CUDA graph
        from image_compression_b200.utils.graphs import GraphedCallable
        graphed = GraphedCallable(eval_eager, resident[0])
    table = torch.empty((B, 2), device=dev)

    def step(x):
        if training:
            # what engine/trainer.py:176-194 does around the model, minus the optimizer arithmetic: zero_grad
            # (set_to_none), forward, backward.  The model's GradientReducer writes the gradients into its flat buffer
            # and (N>1) all-reduces them from inside the backward.  bump_versions stands in for the optimizer's update
            # of the weights: the packed bf16 operand copies of ALL weights are rebuilt in every step.
            for p in params:
                p.grad = None
            hp.bump_versions(params)
            _, losses = model(x)
            losses["total_loss"].backward()
            return losses["total_loss"]
        out = graphed(x) if graphed is not None else eval_eager(x)
        if world > 1:           # per-image sharding: the per-image results of all ranks, in image order, on every rank
            table[:, 0].copy_(out[1]); table[:, 1].copy_(out[2])
            sharding.gather_per_image(table, B * world, rank, world)
        return out

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, nsteps):
        """EXACTLY `nsteps` calls bracketed by barrier+synchronize; device time, max over ranks."""
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(nsteps):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        sync_all()
        return float(ms)

    # ---- resident-input throughput (`value`) ------------------------------------------------
    for i in range(warmup):
        step(resident[i % n_rot])
    events_in_timed = args.kernel_events == "timed"
    clocks = ClockSampler(c.local) if with_clocks else None
    if clocks:
        clocks.__enter__()
    n0 = _lib.lib().hpc_launch_count()
    if events_in_timed:
        _lib.prof_enable(True)
    ms_total = timed(lambda i: step(resident[i % n_rot]), steps)
    _lib.prof_enable(False)
    launches = _lib.lib().hpc_launch_count() - n0
    ms_events = ms_total
    if not events_in_timed:
        _lib.prof_enable(True)
        ms_events = timed(lambda i: step(resident[i % n_rot]), steps)
        _lib.prof_enable(False)
    if clocks:
        clocks.__exit__()
    if graphed is not None:
        # kernels inside a replayed graph are neither counted nor event-timed by the library: take the launch
        # count and the kernel-class times from the same number of eager steps, outside the timed region
        sync_all()
        n0 = _lib.lib().hpc_launch_count()
        _lib.prof_enable(True)
        for i in range(steps):
            eval_eager(resident[i % n_rot])
        torch.cuda.synchronize()
        _lib.prof_enable(False)
        launches = _lib.lib().hpc_launch_count() - n0
    prof = {name: _lib.prof_read(k) for k, name in enumerate(_lib.PROF_NAMES)}
    units = B * world * steps
    r = dict(wl=wl, rank=rank, world=world, B=B, H=H, W=W, steps=steps, warmup=warmup, ms_total=ms_total,
             ms_events=ms_events, value=units / (ms_total / 1e3), launches=launches, prof=prof,
             clocks=clocks.summary() if clocks else None, n_rot=n_rot, graphed=graphed is not None)

    # ---- end to end through the module API with host inputs (`e2e`) --------------------------
    # Every step's input is copied host->device INSIDE the timed region (on a side stream, one step ahead of the
    # kernels: what a prefetching loader does) and every step's result is copied device->host and read by the host
    # inside the timed region.  The host reads the loss of step i-1 after it has enqueued step i (deferred logging),
    # so the device never idles while Python enqueues the next step; the last loss is read after the final
    # synchronize of the timed region.  Three staging buffers: the copy for step i+1 overwrites the input of step
    # i-2, whose loss the host has already read.
    copy_stream = torch.cuda.Stream(device=dev)
    stages = [torch.empty(B, 3, H, W, device=dev) for _ in range(3)]
    copied = [torch.cuda.Event() for _ in range(3)]
    nres = 1 if training else 1 + 2 * B
    res_host = [torch.empty(nres, dtype=torch.float32).pin_memory() for _ in range(2)]
    res_ready = [torch.cuda.Event(), torch.cuda.Event()]
    seen = []

    def h2d_async(i):
        with torch.cuda.stream(copy_stream):
            stages[i % 3].copy_(host[i % n_rot], non_blocking=True)
            copied[i % 3].record(copy_stream)

    def read_result(i):
        res_ready[i & 1].synchronize()
        seen.append(float(res_host[i & 1][0]))        # the host consumes the value

    def make_e2e(nsteps):
        def e2e_step(i):
            if i == 0:
                h2d_async(0)
            torch.cuda.current_stream().wait_event(copied[i % 3])
            if i + 1 < nsteps:
                h2d_async(i + 1)
            out = step(stages[i % 3])
            dst = res_host[i & 1]
            if training:
                dst.copy_(out.detach().reshape(1), non_blocking=True)        # D2H of the loss
            else:
                bpp, ps, ms = out
                dst[:1].copy_(bpp.reshape(1), non_blocking=True)             # D2H of bpp + per-image metrics
                dst[1:1 + B].copy_(ps, non_blocking=True)
                dst[1 + B:1 + 2 * B].copy_(ms, non_blocking=True)
            res_ready[i & 1].record()
            if i > 0:
                read_result(i - 1)
            if i + 1 == nsteps:
                read_result(i)
        return e2e_step
    nw = max(2, warmup // 2)
    warm = make_e2e(nw)
    for i in range(nw):
        warm(i)
    e2e_ms = timed(make_e2e(steps), steps)
    r.update(e2e_value=units / (e2e_ms / 1e3), e2e_ms=e2e_ms, h2d=B * 3 * H * W * 4,
             d2h=4 if training else 4 * (1 + 2 * B),
             grad_bytes=model.gradient_reducer().fg.nbytes() if training else 0)

    # ---- the same end-to-end loop fed by data_utils.Batch: uint8 pixels over PCIe, normalised on the GPU (8f rank 4) ----
    # The Batch objects are collated ahead of time (the DataLoader workers' job in the reference as well: the collate
    # function runs there); the timed region holds, per step, Batch.to(device) -- the uint8 H2D copy and the
    # crop/normalise kernel, one step ahead on the copy stream -- the step itself and the loss read-back.
    if training:
        from image_compression_b200 import data_utils as DU
        mean, std = tuple(cfg.DATA.MEAN), tuple(cfg.DATA.STD)
        rng_u8 = np.random.RandomState(1234 + rank)
        u8_host = [[("%d" % k, DU.U8Image(rng_u8.randint(0, 256, size=(H, W, 3)).astype(np.uint8), mean, std))
                    for k in range(B)] for _ in range(min(n_rot, 3))]
        ready = [DU.Batch(items, cfg).pin_memmory() for items in u8_host]
        ring = [None, None, None]

        def u8_async(i):
            b = ready[i % len(ready)]
            with torch.cuda.stream(copy_stream):
                b.to(dev)
                copied[i % 3].record(copy_stream)
            ring[i % 3] = b.imgs              # kept alive until step i's loss has been read (three steps later)

        def make_u8(nsteps):
            def u8_step(i):
                if i == 0:
                    u8_async(0)
                torch.cuda.current_stream().wait_event(copied[i % 3])
                if i + 1 < nsteps:
                    u8_async(i + 1)
                out = step(ring[i % 3])
                res_host[i & 1].copy_(out.detach().reshape(1), non_blocking=True)
                res_ready[i & 1].record()
                if i > 0:
                    read_result(i - 1)
                if i + 1 == nsteps:
                    read_result(i)
            return u8_step
        warm = make_u8(nw)
        for i in range(nw):
            warm(i)
        k0 = _lib.lib().hpc_launch_count()
        u8_ms = timed(make_u8(steps), steps)
        r["e2e_u8_batch"] = {
            "value": units / (u8_ms / 1e3), "unit": UNIT[wl], "ms_per_step": u8_ms / steps,
            "h2d_bytes_per_step": B * 3 * H * W, "d2h_bytes_per_step": 4,
            "launches_per_step": (_lib.lib().hpc_launch_count() - k0) / steps,
            "what": "the e2e loop fed through data_utils.Batch (the reference's collate object, data_utils/batch.py:19-29): "
                    "uint8 HWC pixels from pinned host memory, ToTensor/ChannelFirst/Normalize by hpc_u8_crop_normalize "
                    "inside Batch.to(device), bit-identical to the reference's fp32 batch"}
        del ring, ready, u8_host

    # ---- a sustained run: >= 200 steps back to back (clocks settle under the power cap) -----------------------
    if sustained_steps:
        cs = ClockSampler(c.local)
        cs.__enter__()
        ms_s = timed(lambda i: step(resident[i % n_rot]), sustained_steps)
        cs.__exit__()
        r["sustained"] = {"steps": sustained_steps, "ms_per_step": ms_s / sustained_steps,
                          "value": B * world * sustained_steps / (ms_s / 1e3), "unit": UNIT[wl], "clocks": cs.summary()}

    # ---- the full training iteration: + fused clip + AdamW ----------------------------------
    if training and with_optimizer:
        from torch.nn.utils import clip_grad_value_
        from image_compression_b200 import solver
        opt = solver.make_optimizer(cfg, model)

        def full_step(i):
            opt.zero_grad(set_to_none=True)
            _, losses = model(resident[i % n_rot])
            losses["total_loss"].backward()
            opt.step()                  # fused clip_grad_value_ + AdamW; bumps the weight versions -> re-pack next step
        for i in range(3):
            full_step(i)
        ms_full = timed(full_step, steps)
        k0 = _lib.lib().hpc_launch_count()
        _lib.prof_enable(True)
        ms_fused = timed(lambda i: opt.step(), steps)
        _lib.prof_enable(False)
        k1 = _lib.lib().hpc_launch_count()
        oms, on, owork = _lib.prof_read(_lib.PROF_NAMES.index("optim"))
        # eager torch on a shadow copy: clip_grad_value_ + torch.optim.AdamW over the same 50 single-tensor groups
        shadow = [torch.nn.Parameter(p.detach().clone()) for p in model.parameters()]
        for q, p in zip(shadow, model.parameters()):
            q.grad = p.grad.detach().clone()
        ropt = torch.optim.AdamW([{"params": [q], "lr": g["lr"], "weight_decay": g["weight_decay"]}
                                  for q, g in zip(shadow, opt.param_groups)], cfg.SOLVER.BASE_LR, eps=cfg.SOLVER.EPS)

        def ref_step(_):
            clip_grad_value_(shadow, cfg.SOLVER.GRAD_CLIP)
            ropt.step()
        for _ in range(3):
            ref_step(0)
        ms_torch = timed(ref_step, steps)
        r["with_optimizer"] = {
            "value": units / (ms_full / 1e3), "unit": UNIT[wl], "ms_per_step": ms_full / steps,
            "what": "zero_grad + forward + backward (+ all-reduce) + FusedAdamW.step() (fused clip_grad_value_ + AdamW, "
                    "one launch) per step: the whole of engine/trainer.py:176-194",
            "optimizer_only_ms": ms_fused / steps, "optimizer_kernel_ms": oms / max(on, 1),
            "optimizer_kernel_gbs": owork / max(oms, 1e-9) / 1e6, "optimizer_launches_per_step": (k1 - k0) / steps,
            "bytes_per_param": 28, "params": sum(p.numel() for p in model.parameters()),
            "torch_eager_optimizer_ms": ms_torch / steps}

    # ---- the reference trainer's own call sequence around the model (tests/trainer_replay.py) ------------------
    if training and trainer_sequence:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from trainer_replay import Monitor, train_step
        from image_compression_b200 import solver, utils as U
        dp = torch.nn.DataParallel(model, device_ids=[c.local])
        opt = solver.make_optimizer(cfg, dp)
        mon = Monitor(model.loss_names, {"psnr": U.psnr, "ms_ssim": U.MS_SSIM()})

        def seq_step(i):
            torch.cuda.current_stream().wait_event(copied[i % 3])
            h2d_async(i + 1)            # the loader's prefetch of the next batch overlaps this step (as in the e2e leg)
            train_step(dp, opt, stages[i % 3], mon, grad_clip=cfg.SOLVER.GRAD_CLIP, it=1)
        h2d_async(0)
        for i in range(3):
            seq_step(i)
        ms_seq = timed(lambda i: seq_step(i + 3), steps)
        torch.cuda.synchronize()
        r["trainer_sequence"] = {
            "value": units / (ms_seq / 1e3), "unit": UNIT[wl], "ms_per_step": ms_seq / steps,
            "what": "engine/trainer.py:176-208 replayed call for call around nn.DataParallel(model): forward, .mean(), "
                    "isnan check, backward, clip_grad_value_, FusedAdamW.step, zero_grad, Monitor.update_metric "
                    "(PSNR + 5-level MS-SSIM of the batch) and update_loss (float() of 4 loss values); inputs copied "
                    "from pinned host memory every step"}
    del model, graphed
    torch.cuda.empty_cache()
    return r


def record_of(r, peaks, args, headline=False):
    """JSON sub-record of one measured workload (roofline of its dominant kernel class, kernel classes, e2e)."""
    wl, steps = r["wl"], r["steps"]
    B, H, W, world = r["B"], r["H"], r["W"], r["world"]
    ms, n, flops = r["prof"]["igemm_fwd_umma"]
    roofline = None
    if n:
        ach = flops / (ms * 1e-3) / 1e12
        peak = peaks["bf16_tflops_sustained"]   # kernel timed inside a long step -> sustained figure
        roofline = {"bound": "tensor", "kernel": "igemm_fwd_umma_pair2_kernel + igemm_fwd_umma_pair_kernel + igemm_fwd_umma_kernel "
                    "(one class: the CTA-pair (one or two M tiles per CTA) and single-CTA builds of the implicit-GEMM "
                    "forward engine)", "achieved": ach, "peak": peak,
                    "unit": "TFLOP/s", "frac": ach / peak,
                    "traffic": ncu_traffic("igemm_fwd_umma_pair2_kernel", "igemm_fwd_umma_pair_kernel", "igemm_fwd_umma_kernel") if wl == "train" else None,
                    "peak_source": f"{peaks['source']} (bf16 sustained)", "launches": n,
                    "avg_launch_ms": ms / n, "share_of_step": ms / r["ms_events"],
                    "algorithmic_flops_per_launch": flops / n}
    kernels = {}
    for name, (kms, kn, work) in r["prof"].items():
        if not kn:
            continue
        gemm = name.startswith("igemm")
        rate = work / (kms * 1e-3) / (1e12 if gemm else 1e9)
        pk = peaks["bf16_tflops_sustained"] if gemm else peaks["hbm_gbs"]
        kernels[name] = {"ms_per_step": kms / steps, "launches_per_step": kn / steps,
                         "achieved": rate, "unit": "TFLOP/s" if gemm else "GB/s", "frac": rate / pk,
                         "share_of_step": kms / r["ms_events"]}
    # the entropy path as one figure (north_star: >= 70 % of HBM): K4 + K5 + K6 together
    ent = [r["prof"][k] for k in ("conditional", "factorized", "distortion") if r["prof"][k][1]]
    entropy_path = None
    if ent:
        ems, ework = sum(e[0] for e in ent), sum(e[2] for e in ent)
        entropy_path = {"ms_per_step": ems / steps, "achieved": ework / (ems * 1e-3) / 1e9, "unit": "GB/s",
                        "frac": ework / (ems * 1e-3) / 1e9 / peaks["hbm_gbs"], "target_frac": 0.70,
                        "kernels": "factorized + conditional + distortion (K4 + K5 + K6)"}
    gflop_unit = TRAIN_GFLOP_PER_CROP if wl != "eval" else EVAL_GFLOP_PER_IMAGE
    if (H, W) != ((256, 256) if wl != "eval" else (512, 768)):
        gflop_unit *= (H * W) / ((256 * 256) if wl != "eval" else (512 * 768))
    step_tflops = r["value"] / world * gflop_unit / 1e3
    rec = {
        "metric": METRIC[wl], "value": r["value"], "unit": UNIT[wl], "n_gpus": world, "steps": steps,
        "warmup": r["warmup"], "ms_per_step": r["ms_total"] / steps,
        "e2e": {"value": r["e2e_value"], "unit": UNIT[wl], "h2d_bytes_per_step": r["h2d"],
                "d2h_bytes_per_step": r["d2h"], "ms_per_step": r["e2e_ms"] / steps},
        "gpu_launches": int(r["launches"]),
        "roofline": roofline,
        "kernels": kernels,
        "entropy_path": entropy_path,
        "step_gemm_tflops_per_gpu": step_tflops,
        "step_frac_of_tensor_peak": step_tflops / peaks["bf16_tflops_sustained"],
        "kernel_events": ("event pairs around every kernel class inside the timed region"
                          if args.kernel_events == "timed" else
                          "event pairs around every kernel class over %d further steps right after the timed "
                          "region (%.3f ms/step with the events; an event record between two kernels "
                          "serialises their programmatic dependent launch, so `value` is timed without them)"
                          % (steps, r["ms_events"] / steps)),
    }
    if r["graphed"]:
        rec["cuda_graph"] = ("each step replays one captured graph of the package's kernels; kernel-class times and "
                             "gpu_launches come from the same number of eager steps run after the timed region")
    for k in ("sustained", "with_optimizer", "trainer_sequence", "e2e_u8_batch"):
        if r.get(k):
            rec[k] = r[k]
    if not headline:
        rec["config"] = config_dict(args, wl)
    return rec


_REAL_STDOUT = None


def capture_stdout():
    """Point file descriptor 1 at stderr for the whole run: native libraries (NCCL's version banner, torchrun
    children) write to fd 1 behind Python's back, and stdout must carry exactly ONE JSON line."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    capture_stdout()
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    import torch
    c = setup_dist()
    import image_compression_b200 as pkg
    from image_compression_b200 import _lib
    _lib.lib()  # fail loudly here if the extension is missing
    pkg.set_compute_mode(args.mode)
    wl = args.workload
    full = wl == "train" and not args.quick
    r = measure(args, c, wl, args.steps, args.warmup, with_clocks=True, with_optimizer=wl != "eval",
                sustained_steps=args.sustained if full else 0, trainer_sequence=full)
    extras = {}
    if full:
        # the rest of BASELINE's metric in front of the driver: C4 (ssim_64 training) and C5 (768x512 eval sweep) from
        # short runs after the headline timing, at the same N
        for w2 in ("train_ssim", "eval"):
            extras[w2] = measure(args, c, w2, max(5, min(args.steps, 10)), 3)
    if c.rank != 0:
        _finish_dist()
        return
    peaks = measured_peaks()
    B, H, W, world = r["B"], r["H"], r["W"], r["world"]
    line = {"higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": {"bf16": "bf16", "fp32": "f32", "bf16_simt": "bf16"}[args.mode], "data": DATA_NOTE,
            "config": config_dict(args), "clocks": r["clocks"]}
    line.update(record_of(r, peaks, args, headline=True))
    if extras:
        line["extra"] = {k: record_of(v, peaks, args) for k, v in extras.items()}
    if world == 1 and not args.no_cpu_baseline and wl in ("train", "eval"):
        g = time_torch_gpu(wl, B, (H, W), 5, 2)
        if g is not None:
            g.update(unit=UNIT[wl], kind="port", sample=f"5 steps of {B} x {H}x{W} after 2 warm-ups, eager torch on the same B200 "
                     "(fp32 tensors, cuDNN, default flags: cudnn.allow_tf32=%s, matmul.allow_tf32=%s): the oracle port of "
                     "the reference modules, " % (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
                     + ("fwd+loss+bwd" if wl == "train" else "forward in rounding mode"))
            line["torch_gpu_baseline"] = g
    if world == 1 and not args.no_cpu_baseline:
        n = (args.cpu_sample or 16) if wl != "eval" else 1
        probe = time_cpu(wl, n, (H, W), 1, 1)
        steps = max(1, min(5, int(20.0 / max(probe["ms_per_step"] / 1e3, 1e-3))))
        cb = time_cpu(wl, n, (H, W), steps, 0)
        line["cpu_baseline"] = {
            "value": cb["value"], "unit": UNIT[wl], "cores": cb["cores"], "kind": "port",
            "sample": f"{steps} step(s) of {n} x {H}x{W} synthetic "
                      f"{'crops, fwd+loss+bwd' if wl != 'eval' else 'image, fwd+metrics'}, fp32 torch-CPU oracle port of "
                      f"the reference modules, {cb['cores']} threads, after 2 warm-up steps"}
    emit(line)
    _finish_dist()


def _finish_dist():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
