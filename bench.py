#!/usr/bin/env python
"""bench.py — GRCN frames/sec on B200 (BASELINE.json metric), one process per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--mode inference|train] [--impl ours|reference]

Default workload = BASELINE.json configs[1]: gaze_grcn inference, 256 clips x T=35 per GPU, 7x7 grid,
hidden 256, bf16 operands / fp32 accumulate, synthetic C3D conv5b features [B,T,1024,7,7] and
random-init weights (reference init).  A "step" is one forward pass of the whole hot path
(pack -> projection -> input convs -> T recurrent steps -> composed head -> softmax+loss) over one batch.
Both numbers go THROUGH THE MODEL CLASS (models/gaze_grcn.py GazePredictionGRCN): `value` = model.run() on resident
placeholders; `e2e` = model.prefetch / acquire / run from pinned HOST tensors, H2D of the features + gt and D2H of the
gaze maps + loss inside the timed region.  The default (no flags) line also carries the sub-records `train` (config 3's
per-GPU share: fwd + BPTT + NCCL all-reduce + clip + Adam, 128 clips/GPU), `config4` (gaze_lstm 14x14 T=128) and
`metrics` (saliency suite at 360x480).
`--impl reference`: the oracle's torch-CPU fp32 restatement of the reference graph (TensorFlow 1 cannot be
installed here) on the host cores, a bounded sample of the same workload per step.
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

# torchrun exports OMP_NUM_THREADS=1 to every rank before Python starts, and the OpenMP / oneDNN pools are sized from it
# at import time.  The reference arm (rank 0 alone does CPU work) is meant to use every host core: lift the cap before
# torch is imported.
if "reference" in sys.argv and os.environ.get("RANK", "0") == "0":
    try:
        _ncores = len(os.sched_getaffinity(0))
    except AttributeError:
        _ncores = os.cpu_count() or 1
    for _k in ("OMP_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[_k] = str(_ncores)

import numpy as np
import torch


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="inference", choices=["inference", "train"])
    ap.add_argument("--model", default="gaze_grcn")
    ap.add_argument("--batch", type=int, default=256, help="clips per GPU")
    ap.add_argument("--T", type=int, default=35)
    ap.add_argument("--hidden", type=int, default=256)
    ap.add_argument("--grid", type=int, default=7)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="launch every kernel from the host instead of one CUDA graph")
    ap.add_argument("--cpu-clips", type=int, default=0, help="clips per CPU pass (0: sized to fill the host cores)")
    ap.add_argument("--features", default="bf16", choices=["bf16", "fp32"],
                    help="dtype of the fed host features of the headline e2e (results are bit-identical; the fp32-fed e2e is "
                         "reported beside it)")
    ap.add_argument("--train-batch", type=int, default=128, help="clips per GPU of the `train` sub-record (config 3)")
    ap.add_argument("--no-sub", action="store_true", help="headline record only (no train / config4 / metrics sub-records)")
    ap.add_argument("--entry", default="c3d", choices=["c3d", "emb"],
                    help="c3d: reference entry [B,T,1024,G,G]; emb: already-projected features [B,T,G,G,512] (SURVEY 8d 'L1')")
    ap.add_argument("--head", default=None, choices=["deconv", "fc"], help="gaze-map head (default: the model's own)")
    return ap.parse_args()


class ClockSampler(object):
    """SM clock and throttle reasons sampled DURING the timed region (B200_PROFILING.md's clocks line).  The sampler is
    started before the warm-up so it is already producing rows when the timed region opens (an `nvidia-smi -lms` child
    needs > 100 ms to print its first row, longer than ten inference steps); rows are time-stamped and only those inside
    [mark(), stop()] are reported.  Source: NVML (the library nvidia-smi reads) polled every 5 ms from a thread, else an
    `nvidia-smi --query-gpu ... -lms 20` child."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    BITS = [0x8, 0x40, 0x20, 0x4]      # nvmlClocksThrottleReason{HwSlowdown, HwThermalSlowdown, SwThermalSlowdown, SwPowerCap}
    # the remaining NVML reason bits, reported too so that a clock below max is never left unexplained
    NVML_NAMES = NAMES + ["applications_clocks_setting", "sync_boost", "hw_power_brake_slowdown", "display_clock_setting"]
    NVML_BITS = BITS + [0x2, 0x10, 0x80, 0x100]

    def __init__(self, index=0):
        self.index, self.rows, self.proc, self.source = index, [], None, None
        self.t0, self.run = None, False

    def start(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = self.index
            if vis and all(x.strip().isdigit() for x in vis.split(",")):
                idx = int(vis.split(",")[self.index])
            h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            mx = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
            reasons = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons", None) or pynvml.nvmlDeviceGetCurrentClocksThrottleReasons
            pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM), reasons(h)

            def poll():
                while self.run:
                    try:
                        m = int(reasons(h))
                        self.rows.append((time.perf_counter(), float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)), mx,
                                          [n for n, b in zip(self.NVML_NAMES, self.NVML_BITS) if m & b]))
                    except Exception:
                        pass
                    time.sleep(0.005)
            self.run, self.source = True, "nvml"
            self.th = threading.Thread(target=poll, daemon=True)
            self.th.start()
            return
        except Exception:
            pass
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.source = "nvidia-smi"
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            r = [x.strip() for x in line.split(",")]
            try:
                self.rows.append((time.perf_counter(), float(r[0]), float(r[1]),
                                  [n for n, v in zip(self.NAMES, r[2:6]) if v.lower().startswith("active")]))
            except Exception:
                pass

    def mark(self):
        """the timed region opens here"""
        self.t0 = time.perf_counter()

    def stop(self):
        t1 = time.perf_counter()
        if self.source is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "samples": 0, "reasons": ["nvml and nvidia-smi unavailable"]}
        self.run = False
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        rows = [r for r in list(self.rows) if self.t0 is None or self.t0 <= r[0] <= t1]
        window = "timed region"
        if not rows:                      # region shorter than one sampling period: the warm-up ran the same kernels
            rows, window = list(self.rows), "warm-up + timed region"
        sm, mx = [r[1] for r in rows], [r[2] for r in rows]
        reasons = sorted(set(n for r in rows for n in r[3]))
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": reasons, "source": self.source, "window": window}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], d["bf16_tflops"], d.get("bf16_tflops_sustained", d["bf16_tflops"]), "measured"
    return 6650.0, 1590.0, 1400.0, "fallback"


def flops_per_frame(Ch, G=7, Cin=512, Cfeat=1024, lstm=False):
    """ALGORITHMIC flops (in-bounds taps only), SURVEY.md 8(d)."""
    P = G * G
    inb = (3 * G - 2) ** 2
    ng_x = 4 if lstm else 3
    return {"proj": 2.0 * P * Cfeat * Cin, "i2h": 2.0 * ng_x * inb * Cin * Ch, "h2h": 2.0 * 3 * inb * Ch * Ch}


# ----------------------------------------------------------------------------------------------
def cpu_reference_step(model, p, c3d, gt, train=False):
    """The oracle's torch-CPU fp32 restatement with the reference's op structure (T-unrolled, per-gate convs)."""
    from oracle import model_oracle as O
    pt = O.to_torch(p, torch.float32, requires_grad=train)
    x = torch.from_numpy(c3d)
    g = torch.from_numpy(gt)
    with torch.set_grad_enabled(train):
        logits = O.recurrent_network(model, pt, c3d=x)
        probs = O.softmax_2d(logits)
        loss = O.loss_fn(logits, g, "xentropy")
        if train:
            loss.backward()
    return float(loss), probs


def use_all_host_threads():
    """torchrun exports OMP_NUM_THREADS=1 to every rank; the CPU legs are meant to use every host core this process may
    run on (the other ranks do no CPU work while rank 0 times the reference)."""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    if torch.get_num_threads() < n:
        torch.set_num_threads(n)
    return torch.get_num_threads()


def cpu_sample_clips(args, cores):
    """clips per CPU pass: enough frames per conv call to occupy every core (one clip per two threads, at least 8)"""
    return args.cpu_clips if args.cpu_clips > 0 else max(8, min(32, cores // 2))


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU implementation of the path.  TensorFlow 1 is not installable here
    (DESIGN.md section 0), so this is the oracle's torch-CPU fp32 restatement of the reference graph (kind "port"), on all
    host threads, each step a bounded sample of the workload."""
    if rank != 0:
        return
    cores = use_all_host_threads()
    from oracle import model_oracle as O
    from recurrent_gaze_prediction_b200 import synthetic as SY
    B = cpu_sample_clips(args, cores)
    p = O.init_params(args.model, hidden=args.hidden, T=args.T, grid=args.grid, seed=0, regime="R")
    rng = np.random.RandomState(0)
    c3d = SY.make_c3d(rng, B, args.T, 1024, args.grid)
    GH = args.grid if args.model == "gaze_grcn77" else 49
    gt, _ = SY.make_gazemaps(rng, B, args.T, GH, GH, normalize=True)
    train = args.mode == "train"
    for _ in range(args.warmup):
        cpu_reference_step(args.model, p, c3d, gt, train)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_reference_step(args.model, p, c3d, gt, train)
    dt = (time.perf_counter() - t0) / args.steps
    fps = B * args.T / dt
    sample = "each step = %d clips x T=%d (a bounded sample of the %d-clip workload named in config; frames/s normalises " \
             "the size), torch-CPU fp32 restatement of the reference graph (oracle/model_oracle.py; TF1 not installable), " \
             "%d threads of %d cores" % (B, args.T, args.batch, cores, os.cpu_count())
    out = {"impl": "reference", "metric": metric_name(args), "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": workload_config(args), "sample_clips_per_step": B,
           "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": cores, "kind": "port", "sample": sample},
           "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out))


def metric_name(args):
    return "GRCN frames/sec (%s)" % ("fwd" if args.mode == "inference" else "train: fwd+BPTT+Adam")


def workload_config(args):
    feat = "C3D conv5b 1024x%dx%d -> proj 512" % (args.grid, args.grid) if args.entry == "c3d" else \
        "embedded features 512x%dx%d" % (args.grid, args.grid)
    maps = "49x49" if (args.head or ("fc" if args.model == "gaze_grcn77" else "deconv")) == "deconv" else "%dx%d" % (args.grid, args.grid)
    return {"workload": "%s %s, %d clips/GPU x T=%d, %dx%d grid, %s, hidden %d, %s gaze maps"
                        % (args.model, args.mode, args.batch, args.T, args.grid, args.grid, feat, args.hidden, maps),
            "clips_per_gpu": args.batch, "T": args.T, "grid": args.grid, "hidden": args.hidden,
            "l2_policy": "inputs larger than L2 (features 1.8 GB/step at the default size)",
            "parallelism": ("dp%d (clips sharded across ranks; one NCCL all-reduce of the flat fp32 gradient per step)"
                            if args.mode == "train" else
                            "dp%d (clips sharded across ranks, no data-path collective in inference)") % args.gpus}


# ----------------------------------------------------------------------------------------------
class Harness(object):
    """timing plumbing shared by the records: barrier + synchronize on both sides, CUDA events on the launching stream,
    max over ranks"""

    def __init__(self, dev, local, rank, world):
        self.dev, self.local, self.rank, self.world = dev, local, rank, world

    def barrier(self):
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()

    def timed(self, fn, steps, warmup, lib=None, per_replay=0, sample_clocks=False):
        cs = ClockSampler(self.local) if sample_clocks else None
        if cs:
            cs.start()
        for _ in range(warmup):
            fn()
        self.barrier()
        if cs:
            cs.mark()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = lib.launches if lib is not None else 0
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        self.barrier()
        ms = e0.elapsed_time(e1)
        clocks = cs.stop() if cs else None
        t = torch.tensor([ms], device=self.dev)
        if self.world > 1:
            import torch.distributed as dist
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        launches = (lib.launches - l0 if lib is not None else 0) + per_replay * steps
        return float(t[0]) / steps, clocks, launches


def make_inputs(rank, B, T, G, GH, feature_dtype=torch.float32, emb=False):
    """synthetic inputs: a small pool tiled to the batch (host RNG cost), distinct per rank; pinned host tensors"""
    from recurrent_gaze_prediction_b200 import synthetic as SY
    rng = np.random.RandomState(1234 + rank)
    pool = min(B, 8)
    feat_pool = SY.make_embedded(rng, pool, T, G, 512) if emb else SY.make_c3d(rng, pool, T, 1024, G)
    gt_pool, _ = SY.make_gazemaps(rng, pool, T, GH, GH, normalize=True)
    reps = (B + pool - 1) // pool
    feat = torch.from_numpy(np.tile(feat_pool, (reps,) + (1,) * (feat_pool.ndim - 1))[:B])
    gt = torch.from_numpy(np.tile(gt_pool, (reps, 1, 1, 1))[:B])
    return feat.to(feature_dtype).pin_memory(), gt.pin_memory(), feat_pool, gt_pool


def build_model(args, session, B_local, world, train, feature_dtype, dev):
    """the reference-facing model class (models/gaze_grcn.py etc.) at the bench size, random-init (reference initialisers)"""
    from recurrent_gaze_prediction_b200.models import train_gaze
    TheModel, _, GRUModelConfig = train_gaze.get_model_class(args.model)
    c = GRUModelConfig()
    c.batch_size, c.n_lstm_steps, c.rnn_state_size, c.grid = B_local * world, args.T, args.hidden, args.grid
    c.dtype, c.device, c.trainable = args.dtype, str(dev), train
    c.feature_dtype = feature_dtype
    c.cuda_graph = not args.no_graph
    c.use_flip_batch = False
    return TheModel(session, None, c)


def bench_model_record(args, H, session, mode, B, feature_dtype="bf16", with_roofline=False, sample_clocks=False,
                       also_fp32_e2e=False):
    """One record (value / ms_per_step / e2e / gpu_launches [/ roofline]) of `mode` at B clips per GPU, THROUGH THE MODEL
    CLASS: `value` = model.run() on resident placeholders (one CUDA-graph submission per step), `e2e` = the public
    pipelined feed (model.prefetch -> acquire -> run) from pinned host tensors + device->host read of the result."""
    import torch.distributed as dist
    train = mode == "train"
    dev, world, rank = H.dev, H.world, H.rank
    T, G, Ch = args.T, args.grid, args.hidden
    model = build_model(args, session, B, world, train, feature_dtype, dev)
    eng = model.engine
    fdt = torch.bfloat16 if feature_dtype == "bf16" else torch.float32
    feat_h, gt_h, feat_pool, gt_pool = make_inputs(rank, B, T, G, eng.GH, fdt)
    # global-batch host tensors are what the model API takes; every rank feeds its own shard: build the shard view
    model.c3d_input.copy_(feat_h)
    model.gt_gazemap.copy_(gt_h)
    probs_h = torch.empty(B, T, eng.GH, eng.GH, dtype=torch.float32).pin_memory()

    def step_resident():
        model.run(train=train, sync=False)

    step_resident()                                   # builds the step graph (undone for train: no update is kept)
    torch.cuda.synchronize()
    per_replay = 0
    if model._use_graph():
        per_replay = model._step_graph(train).kernels
    ms_step, clocks, launches = H.timed(step_resident, args.steps, args.warmup, lib=eng.lib, per_replay=per_replay,
                                        sample_clocks=sample_clocks)
    frames = B * T * world
    rec = {"value": frames / (ms_step * 1e-3), "unit": "frames/s", "ms_per_step": ms_step, "clips_per_gpu": B,
           "gpu_launches": launches, "submission": "cuda-graph" if model._use_graph() else "eager",
           "api": "models.%s: run() on resident placeholders" % type(model).__name__}
    if clocks is not None:
        rec["clocks"] = clocks

    # ---- allreduce share of the training step (device time of the NCCL call, events on the launching stream)
    if train and world > 1:
        sg = model._step_graph(True)
        evs = []

        def step_timed_ar():
            sg.g1.replay()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            sg.allreduce(eng.grads)
            b.record()
            sg.g2.replay()
            eng.note_replayed_steps(1)
            evs.append((a, b))
        for _ in range(2):
            step_timed_ar()
        H.barrier()
        evs.clear()
        for _ in range(max(4, args.steps)):
            step_timed_ar()
        H.barrier()
        ar = torch.tensor([float(np.mean([a.elapsed_time(b) for a, b in evs]))], device=dev)
        dist.all_reduce(ar, op=dist.ReduceOp.MAX)
        rec["allreduce_ms"] = float(ar[0])
        rec["allreduce_bytes"] = int(eng.nparams) * 4
        rec["allreduce_share_of_step"] = float(ar[0]) / ms_step

    # ---- end to end through the public pipelined feed
    def e2e_of(m, fg, gg):
        fh, gh = fg, gg                        # every rank feeds its own shard of the global batch (local=True)

        def step_e2e():
            m.prefetch(fg, gg, local=True)     # upload of the next batch starts on the copy stream
            m.acquire()                        # the oldest uploaded batch becomes the fed batch
            m.run(train=train, sync=False)     # fwd (+BPTT+clip/Adam); the loss read is queued
            if not train:
                probs_h.copy_(m.predicted_gazemaps, non_blocking=True)
        m.prefetch(fg, gg, local=True)         # pipeline prologue
        ms, _, _ = H.timed(step_e2e, max(4, args.steps), 2)
        m._get_feeder().drain()
        _ = m.last_loss
        h2d = fh.numel() * fh.element_size() + gh.numel() * 4
        d2h = (0 if train else probs_h.numel() * 4) + 4
        return {"value": frames / (ms * 1e-3), "unit": "frames/s", "ms_per_step": ms,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "feeding": "models.%s.prefetch/acquire/run: double-buffered H2D of pinned host %s features + fp32 gaze maps "
                           "on a copy stream, result read back to pinned host memory" % (type(m).__name__, str(fh.dtype).split(".")[-1])}

    if not args.no_e2e:
        rec["e2e"] = e2e_of(model, feat_h, gt_h)
        if also_fp32_e2e and feature_dtype != "fp32":
            m32 = build_model(args, session, B, world, train, "fp32", dev)
            f32_h = feat_h.float().pin_memory()
            rec["e2e_fp32_features"] = e2e_of(m32, f32_h, gt_h)
            # the synchronous single call a reference user would write: predict(host batch) -> host maps
            if not train:
                sync_ms, _, _ = H.timed(lambda: m32.predict(f32_h, out=probs_h, local=True), max(4, args.steps), 2)
                rec["e2e_sync_predict_fp32"] = {"value": frames / (sync_ms * 1e-3), "unit": "frames/s", "ms_per_step": sync_ms,
                                                "api": "model.predict(pinned fp32 batch, out=pinned maps), no pipelining"}
            del m32
            torch.cuda.empty_cache()

    # ---- roofline of the dominant kernel + stage breakdown
    if with_roofline and rank == 0:
        hbm, tf_burst, tf_sus, which = peaks()
        fl = flops_per_frame(Ch, G, lstm=eng.lstm)
        evs = eng.time_stages(model.c3d_input, model.gt_gazemap, iters=3)
        i2h_ms = evs["i2h"]
        ach = fl["i2h"] * B * T / (i2h_ms * 1e-3) / 1e12
        default_cfg = (args.model, B, T, Ch, G, args.dtype) == ("gaze_grcn", 256, 35, 256, 7, "bf16")
        # the stage timings are isolated launches (3 iterations at boost clocks): the BURST peak is their denominator;
        # the fraction against the sustained peak is printed beside it
        roof = {"kernel": "tapgemm i2h (3x3 SAME conv, all T frames, M=%d N=%d K=9x512)" % (T * B * eng.S, eng.ngx * Ch),
                "bound": "tensor", "achieved": ach, "peak": tf_burst, "unit": "TFLOP/s", "frac": ach / tf_burst,
                "frac_of_sustained_peak": ach / tf_sus,
                "peak_source": which + " (burst bf16 %.1f: kernel timed alone; sustained %.1f)" % (tf_burst, tf_sus),
                "traffic": TRAFFIC.get("i2h") if default_cfg else None,
                "traffic_source": TRAFFIC.get("i2h_source") if default_cfg else None,
                "ms_per_launch": i2h_ms, "stage_ms": evs}
        rec_ms = evs["recurrence"]
        rec_bytes = 4.0 * Ch * G * G * 4 * B * T
        roof["recurrence"] = {
            "kernel": ("gru_seq (persistent ConvGRU recurrence, all T steps in one launch)" if eng.use_seq else
                       "per-step tap-GEMMs (%d launches)" % (2 * T)),
            "ms": rec_ms, "tensor_TFLOPs": fl["h2h"] * B * T / (rec_ms * 1e-3) / 1e12,
            "tensor_frac": fl["h2h"] * B * T / (rec_ms * 1e-3) / 1e12 / tf_burst,
            "tensor_frac_of_sustained_peak": fl["h2h"] * B * T / (rec_ms * 1e-3) / 1e12 / tf_sus,
            "hbm_GBps": rec_bytes / (rec_ms * 1e-3) / 1e9, "hbm_frac": rec_bytes / (rec_ms * 1e-3) / 1e9 / hbm,
            "traffic": TRAFFIC.get("gru_seq") if (default_cfg and eng.use_seq) else None}
        if train:
            eng.forward(c3d=model.c3d_input, gt=model.gt_gazemap, global_batch=B * world)
            snap = eng.snapshot_train_state()
            roof["stage_ms_bwd"] = eng.time_backward(iters=3)
            eng.restore_train_state(snap)
        rec["roofline"] = roof
        rec["step_tflops"] = sum(fl.values()) * B * T / (ms_step * 1e-3) / 1e12 * (3.0 if train else 1.0)
    rec["_pools"] = (feat_pool, gt_pool)
    rec["_params"] = None
    del model
    torch.cuda.empty_cache()
    return rec


# DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) from the committed `ncu --set full` captures of
# exactly the default configuration (profiles/); None for any other configuration
TRAFFIC = {"i2h": 1.7787e9,
           "i2h_source": "profiles/r02_ncu_full4_fwd256_position_major.txt (position-major tiles: 0.479 GB read + 1.300 GB "
                         "written per launch)",
           "gru_seq": 3.6568e9}


def bench_config4(args, H):
    """BASELINE config 4: gaze_lstm, 14x14 grid, T = 128, hidden 128, embedded-feature entry [B,T,14,14,512], per-position
    FC head, inference at 8 and 64 clips (one GPU).  Engine-level: the model classes take the reference's C3D 7x7 entry
    only; 14x14 is reachable through the cell's spatial_shape."""
    from recurrent_gaze_prediction_b200.engine import GazeEngine, StepGraph
    from recurrent_gaze_prediction_b200.models.model_util import init_parameters
    out = {"workload": "gaze_lstm inference, T=128, 14x14 grid, embedded features 512x14x14, hidden 128, per-position FC head",
           "api": "GazeEngine.forward(emb=...) as one CUDA graph", "n_gpus": 1}
    T, G, Ch = 128, 14, 128
    H = Harness(H.dev, H.local, 0, 1)      # rank 0 runs this record alone: no barrier / max over ranks with the others
    p = dict(init_parameters("gaze_lstm", hidden=Ch, T=T, grid=G, seed=0))
    p["out_W"] = np.random.RandomState(7).uniform(-0.1, 0.1, (Ch, 1)).astype(np.float32)
    hbm, tf_burst, tf_sus, _ = peaks()
    fl = flops_per_frame(Ch, G, lstm=True)
    for B in (8, 64):
        eng = GazeEngine("gaze_lstm", B=B, T=T, G=G, Ch=Ch, dtype=1, device=H.dev, head="fc")
        eng.load_tf_params(p)
        feat_h, gt_h, _, _ = make_inputs(0, B, T, G, G, emb=True)
        emb_d, gt_d = feat_h.to(H.dev), gt_h.to(H.dev)
        sg = StepGraph(eng, lambda: eng.forward(emb=emb_d, gt=gt_d))
        ms, _, _ = H.timed(sg.replay, max(3, args.steps // 2), 3)
        st = eng.time_stages(emb_d, gt_d, iters=2, emb=True)
        out["B%d" % B] = {"value": B * T / (ms * 1e-3), "unit": "frames/s", "ms_per_step": ms, "gpu_launches_per_step": sg.kernels,
                          "stage_ms": st, "persistent_recurrence": bool(getattr(eng, "use_seq_lstm", False)),
                          "recurrence_tensor_frac": fl["h2h"] * B * T / (st["recurrence"] * 1e-3) / 1e12 / tf_burst,
                          "i2h_tensor_frac": fl["i2h"] * B * T / (st["i2h"] * 1e-3) / 1e12 / tf_burst}
        del eng, sg, emb_d, gt_d
        torch.cuda.empty_cache()
    return out


def bench_reference_batch(args, H):
    """The reference's own training configuration (train_gaze.py:74-75 batch 28; gaze_rnn.py:50 T = 42; hidden 128,
    gaze_grcn.py:211) through the model class on one GPU: one CUDA-graph submission per step against eager launches (the
    path is launch-bound at this size), and inference."""
    H1 = Harness(H.dev, H.local, 0, 1)
    out = {"workload": "gaze_grcn, batch 28 x T=42, hidden 128 (the reference's train_gaze.py configuration), bf16 operands",
           "api": "models.GazePredictionGRCN.run() on resident placeholders", "n_gpus": 1}
    for mode in ("train", "inference"):
        for graph in (True, False):
            a = argparse.Namespace(**vars(args))
            a.model, a.batch, a.T, a.hidden, a.grid, a.mode, a.no_graph = "gaze_grcn", 28, 42, 128, 7, mode, not graph
            model = build_model(a, {"rank": 0, "world_size": 1, "allreduce": None}, 28, 1, mode == "train", "bf16", H.dev)
            feat_h, gt_h, _, _ = make_inputs(0, 28, 42, 7, 49, torch.bfloat16)
            model.c3d_input.copy_(feat_h)
            model.gt_gazemap.copy_(gt_h)
            step = lambda: model.run(train=(mode == "train"), sync=False)
            step()
            ms, _, _ = H1.timed(step, max(4, args.steps), 3)
            out["%s_%s" % (mode, "graph" if graph else "eager")] = {"ms_per_step": ms, "frames_per_s": 28 * 42 / (ms * 1e-3)}
            del model
            torch.cuda.empty_cache()
    return out


def bench_config5(args, H):
    """BASELINE config 5: the cascade (gaze_grcn_cascade.py: bottom ConvGRU 256 -> 11x11/s7 transposed conv -> 5x5
    ConvGRU on 49x49 -> two maxout FCs) through its model class, inference at 64 clips x T=35, bf16; then the saliency
    suite (sim, cc, NSS, AUC-Judd, AUC-Borji, shuffled AUC) of 1024 of ITS predicted maps against 360x480 fixation maps
    through MetricsEngine.evaluate_batch (the product's evaluation call)."""
    from recurrent_gaze_prediction_b200 import evaluation_metrics as EM
    from recurrent_gaze_prediction_b200 import synthetic as SY
    H = Harness(H.dev, H.local, 0, 1)
    B, T = 64, 35
    cargs = argparse.Namespace(**vars(args))
    cargs.model, cargs.batch, cargs.T, cargs.grid, cargs.mode = "gaze_grcn_cascade", B, T, 7, "inference"
    model = build_model(cargs, {"rank": 0, "world_size": 1, "allreduce": None}, B, 1, False, "bf16", H.dev)
    eng = model.engine
    feat_h, gt_h, _, gt_pool = make_inputs(0, B, T, 7, 49, torch.bfloat16)
    model.c3d_input.copy_(feat_h)
    model.gt_gazemap.copy_(gt_h)
    model.run(train=False, sync=True)
    ms, _, _ = H.timed(lambda: model.run(train=False, sync=False), max(4, args.steps), 3)
    st = eng.time_stages(model.c3d_input, model.gt_gazemap, iters=3)
    out = {"workload": "gaze_grcn_cascade inference, %d clips x T=%d, C3D 1024x7x7 -> bottom ConvGRU 256 -> 49x49x64 -> 5x5 "
                       "ConvGRU (3 units) -> maxout FCs 7203-4802-2401-4802-2401, bf16 operands" % (B, T),
           "api": "models.gaze_grcn_cascade.GazePredictionGRCN: run() on resident placeholders (one CUDA graph)",
           "value": B * T / (ms * 1e-3), "unit": "frames/s", "ms_per_step": ms, "clips_per_gpu": B, "n_gpus": 1,
           "gpu_launches": model._step_graph(False).kernels, "stage_ms": st,
           "note": "the reference defines the cascade on the 7x7 C3D grid (its 11x11 / stride-7 transposed conv maps 7 -> 49); "
                   "a 14x14 cascade does not exist in it"}
    # ---- full saliency evaluation of the cascade's own maps at VAS scale
    N = 1024
    pred = model.predicted_gazemaps.reshape(B * T, 49, 49)[:N].float().cpu().numpy()
    rng = np.random.RandomState(3)
    reps = (N + gt_pool.shape[0] * T - 1) // (gt_pool.shape[0] * T)
    gtm = np.tile(gt_pool.reshape(-1, 49, 49), (reps, 1, 1))[:N]
    fix64 = SY.make_fixationmaps(rng, gtm[:64, None], 360, 480).reshape(64, 360, 480)
    fix = np.tile(fix64, (N // 64, 1, 1))
    union = (fix64[:10].sum(axis=0) > 0).astype(np.uint8)
    me = EM.MetricsEngine(device=H.dev, chunk=N)
    mets = ("sim", "cc", "NSS", "AUC_Judd", "AUC_Borji", "AUC_shuffled")
    np.random.seed(0)
    r = me.evaluate_batch(pred, gtm, fix, union=union, metrics=mets)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        np.random.seed(0)
        r = me.evaluate_batch(pred, gtm, fix, union=union, metrics=mets)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 3
    out["evaluation"] = {"call": "MetricsEngine.evaluate_batch(pred[1024,49,49], gt, fix[1024,360,480], union, metrics=%s), "
                                 "host arrays in, host scores out (H2D of the fixation maps and the host-drawn random index "
                                 "tensors included)" % (mets,),
                         "frames_per_s": N / dt, "ms_per_1024_frames": dt * 1e3,
                         "mean_scores": {k: float(np.nanmean(v)) for k, v in r.items()}}
    # the same call with the opt-in device generator (jitter / Borji negatives drawn on the device, not NumPy's global stream)
    r2 = me.evaluate_batch(pred, gtm, fix, union=union, metrics=mets, rng="device", seed=1)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        r2 = me.evaluate_batch(pred, gtm, fix, union=union, metrics=mets, rng="device", seed=1)
    torch.cuda.synchronize()
    dt2 = (time.perf_counter() - t0) / 3
    out["evaluation_device_rng"] = {"call": "the same with rng='device' (statistically equivalent draws, not NumPy's stream)",
                                    "frames_per_s": N / dt2, "ms_per_1024_frames": dt2 * 1e3,
                                    "mean_scores": {k: float(np.nanmean(v)) for k, v in r2.items()}}
    del model
    torch.cuda.empty_cache()
    return out


# ----------------------------------------------------------------------------------------------
def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    import torch.distributed as dist
    from recurrent_gaze_prediction_b200 import numa

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa_info = numa.bind_to_gpu_node(local) if world > 1 else {"node": None, "note": "single rank: not bound"}
    session = {"rank": rank, "world_size": world, "allreduce": None}
    if world > 1:
        import datetime
        # a rank that dies must fail the others within minutes, not after NCCL's default 10-minute watchdog
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=240))
        session["allreduce"] = lambda t: dist.all_reduce(t, op=dist.ReduceOp.SUM)
    H = Harness(dev, local, rank, world)
    if args.entry == "emb" or args.head is not None:
        return main_engine_level(args, H, session)

    # ---- headline record: the mode asked for (default: inference = BASELINE config 2)
    head = bench_model_record(args, H, session, args.mode, args.batch, feature_dtype=args.features, with_roofline=True,
                              sample_clocks=(rank == 0), also_fp32_e2e=True)
    feat_pool, gt_pool = head.pop("_pools")
    head.pop("_params")

    sub = {}
    if args.mode == "inference" and not args.no_sub:
        # BASELINE.json's metric is "fwd & train": the training half (config 3's per-GPU share, 128 clips/GPU, with the
        # NCCL all-reduce of the flat gradient when N > 1) rides in the same line, then config 4 and the metric suite
        targs = argparse.Namespace(**vars(args))
        targs.mode, targs.batch = "train", args.train_batch
        tr = bench_model_record(targs, H, session, "train", args.train_batch, feature_dtype=args.features, with_roofline=True)
        tr.pop("_pools"), tr.pop("_params")
        tr["metric"] = metric_name(targs)
        tr["config"] = workload_config(targs)
        sub["train"] = tr
        if rank == 0:
            try:
                sub["reference_batch"] = bench_reference_batch(args, H)
            except Exception as e:
                sub["reference_batch"] = {"error": "%s: %s" % (type(e).__name__, e)}
            try:
                sub["config4"] = bench_config4(args, H)
            except Exception as e:       # a sub-record must not take the headline down
                sub["config4"] = {"error": "%s: %s" % (type(e).__name__, e)}
            try:
                sub["config5"] = bench_config5(args, H)
            except Exception as e:
                sub["config5"] = {"error": "%s: %s" % (type(e).__name__, e)}
            try:
                sys.path.insert(0, os.path.join(ROOT, "tools"))
                import bench_metrics
                sub["metrics"] = bench_metrics.run(1024, cpu_frames=2)
            except Exception as e:
                sub["metrics"] = {"error": "%s: %s" % (type(e).__name__, e)}
        H.barrier()

    cpu = None
    if rank == 0 and not args.no_cpu_baseline:
        from recurrent_gaze_prediction_b200.models.model_util import init_parameters
        cores = use_all_host_threads()
        nb = cpu_sample_clips(args, cores)
        reps = (nb + len(feat_pool) - 1) // len(feat_pool)
        c3 = np.tile(feat_pool, (reps, 1, 1, 1, 1))[:nb]
        g = np.tile(gt_pool, (reps, 1, 1, 1))[:nb]
        pc = init_parameters(args.model, hidden=args.hidden, T=args.T, grid=args.grid, seed=0)
        train = args.mode == "train"
        cpu_reference_step(args.model, pc, c3, g, train)
        t0 = time.perf_counter()
        n = 0
        while time.perf_counter() - t0 < 10.0 or n < 2:
            cpu_reference_step(args.model, pc, c3, g, train)
            n += 1
        dt = (time.perf_counter() - t0) / n
        cpu = {"value": nb * args.T / dt, "unit": "frames/s", "cores": torch.get_num_threads(), "kind": "port",
               "sample": "%d clips x T=%d per pass, %d passes, torch-CPU fp32 restatement of the reference graph "
                         "(oracle/model_oracle.py; TF1 not installable)" % (nb, args.T, n)}
    if rank == 0:
        out = {"metric": metric_name(args), "value": head["value"], "unit": "frames/s", "n_gpus": world, "steps": args.steps,
               "warmup": args.warmup, "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": "weak",
               "vs_baseline": None, "dtype": args.dtype, "data": "synthetic", "config": workload_config(args),
               "submission": head["submission"], "api": head["api"],
               "clocks": head.get("clocks"), "e2e": head.get("e2e"), "e2e_fp32_features": head.get("e2e_fp32_features"),
               "e2e_sync_predict_fp32": head.get("e2e_sync_predict_fp32"),
               "gpu_launches": head["gpu_launches"], "roofline": head.get("roofline"), "cpu_baseline": cpu,
               "step_tflops": head.get("step_tflops"), "numa": numa_info}
        if "allreduce_ms" in head:
            out.update({k: head[k] for k in ("allreduce_ms", "allreduce_bytes", "allreduce_share_of_step")})
        out.update(sub)
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def main_engine_level(args, H, session):
    """--entry emb / --head: shapes the model classes do not expose (embedded-feature entry, 14x14 grids, FC head on a
    deconv model) timed on the engine directly — the round-1 harness, kept for the config-4 family and A/B runs."""
    import torch.distributed as dist
    from recurrent_gaze_prediction_b200.engine import GazeEngine, StepGraph
    from recurrent_gaze_prediction_b200.models.model_util import init_parameters
    rank, world, dev = H.rank, H.world, H.dev
    B, T, G, Ch = args.batch, args.T, args.grid, args.hidden
    dtype = 1 if args.dtype == "bf16" else 0
    train = args.mode == "train"
    p = init_parameters(args.model, hidden=Ch, T=T, grid=G, seed=0)
    eng = GazeEngine(args.model, B=B, T=T, G=G, Ch=Ch, dtype=dtype, device=dev, train=train, head=args.head)
    if eng.head == "fc" and np.asarray(p["out_W"]).size != Ch:
        p = dict(p)
        p["out_W"] = np.random.RandomState(7).uniform(-0.1, 0.1, (Ch, 1)).astype(np.float32)
    eng.load_tf_params(p)
    emb_entry = args.entry == "emb"
    feat_h, gt_h, _, _ = make_inputs(rank, B, T, G, eng.GH, emb=emb_entry)
    feat_d, gt_d = feat_h.to(dev), gt_h.to(dev)
    fwd = (lambda: eng.forward(emb=feat_d, gt=gt_d, global_batch=B * world)) if emb_entry else \
        (lambda: eng.forward(c3d=feat_d, gt=gt_d, global_batch=B * world))
    if args.no_graph:
        def step():
            fwd()
            if train:
                eng.backward()
                eng.optimizer_step(allreduce=session["allreduce"])
        per_replay = 0
    else:
        sg = StepGraph(eng, fwd, train=train, allreduce=session["allreduce"] if train else None, opt={})
        step, per_replay = sg.replay, sg.kernels
    ms_step, clocks, launches = H.timed(step, args.steps, args.warmup, lib=eng.lib, per_replay=per_replay,
                                        sample_clocks=(rank == 0))
    frames = B * T * world
    if rank == 0:
        st = eng.time_stages(feat_d, gt_d, iters=3, emb=emb_entry)
        hbm, tf_burst, tf_sus, which = peaks()
        fl = flops_per_frame(Ch, G, lstm=eng.lstm)
        roof = {"kernel": "tapgemm i2h", "bound": "tensor", "achieved": fl["i2h"] * B * T / (st["i2h"] * 1e-3) / 1e12,
                "peak": tf_burst, "unit": "TFLOP/s", "traffic": None, "stage_ms": st}
        roof["frac"] = roof["achieved"] / tf_burst
        if train:
            fwd()
            snap = eng.snapshot_train_state()
            roof["stage_ms_bwd"] = eng.time_backward(iters=3)
            eng.restore_train_state(snap)
        print(json.dumps({"metric": metric_name(args), "value": frames / (ms_step * 1e-3), "unit": "frames/s", "n_gpus": world,
                          "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
                          "scaling": "weak", "vs_baseline": None, "dtype": args.dtype, "data": "synthetic",
                          "config": workload_config(args), "submission": "eager" if args.no_graph else "cuda-graph",
                          "api": "GazeEngine (engine level)", "clocks": clocks, "e2e": None, "gpu_launches": launches,
                          "roofline": roof, "cpu_baseline": None}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
