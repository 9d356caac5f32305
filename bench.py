#!/usr/bin/env python
"""bench.py — GRCN frames/sec on B200 (BASELINE.json metric), one process per GPU.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--mode inference|train] [--impl ours|reference]

Default workload = BASELINE.json configs[1]: gaze_grcn inference, 256 clips x T=35 per GPU, 7x7 grid,
hidden 256, bf16 operands / fp32 accumulate, synthetic C3D conv5b features [B,T,1024,7,7] and
random-init weights (reference init).  A "step" is one forward pass of the whole hot path
(pack -> projection -> input convs -> T recurrent steps -> composed head -> softmax+loss) over one batch.
`value`: inputs resident in HBM.  `e2e`: through the model-class API with pinned HOST buffers, H2D of the
features + gt and D2H of the gaze maps + loss inside the timed region.
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
    ap.add_argument("--cpu-clips", type=int, default=2)
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


def run_reference(args, rank, world):
    if rank != 0:
        return
    use_all_host_threads()
    from oracle import model_oracle as O
    from recurrent_gaze_prediction_b200 import synthetic as SY
    B = args.cpu_clips
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
    cores = torch.get_num_threads()
    sample = "%d clips x T=%d per step (bounded sample of the %d-clip workload), torch-CPU fp32 restatement of the " \
             "reference graph (oracle/model_oracle.py), %d threads of %d cores" % (B, args.T, args.batch, cores,
                                                                                 os.cpu_count())
    out = {"impl": "reference", "metric": metric_name(args), "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True,
           "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
           "config": workload_config(args),
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
def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    import torch.distributed as dist
    from recurrent_gaze_prediction_b200 import synthetic as SY
    from recurrent_gaze_prediction_b200.models.model_util import init_parameters
    from recurrent_gaze_prediction_b200.engine import GazeEngine
    from recurrent_gaze_prediction_b200 import lib as L

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, T, G, Ch = args.batch, args.T, args.grid, args.hidden
    dtype = 1 if args.dtype == "bf16" else 0
    train = args.mode == "train"
    p = init_parameters(args.model, hidden=Ch, T=T, grid=G, seed=0)      # the reference's initialisers
    eng = GazeEngine(args.model, B=B, T=T, G=G, Ch=Ch, dtype=dtype, device=dev, train=train, head=args.head)
    if eng.head == "fc" and np.asarray(p["out_W"]).size != Ch:   # per-position FC head on a model whose own head is the deconv chain
        p = dict(p)
        p["out_W"] = np.random.RandomState(7).uniform(-0.1, 0.1, (Ch, 1)).astype(np.float32)
    eng.load_tf_params(p)
    GH = eng.GH
    # synthetic inputs: a small pool tiled to the batch (host RNG cost), distinct per rank
    rng = np.random.RandomState(1234 + rank)
    pool = min(B, 8)
    emb_entry = args.entry == "emb"
    c3d_pool = SY.make_embedded(rng, pool, T, G, 512) if emb_entry else SY.make_c3d(rng, pool, T, 1024, G)
    gt_pool, _ = SY.make_gazemaps(rng, pool, T, GH, GH, normalize=True)
    reps = (B + pool - 1) // pool
    c3d_h = torch.from_numpy(np.tile(c3d_pool, (reps, 1, 1, 1, 1))[:B]).pin_memory()
    gt_h = torch.from_numpy(np.tile(gt_pool, (reps, 1, 1, 1))[:B]).pin_memory()
    c3d_d, gt_d = c3d_h.to(dev), gt_h.to(dev)
    probs_h = torch.empty(B, T, GH, GH, dtype=torch.float32).pin_memory()
    loss_h = torch.empty(1, dtype=torch.float32).pin_memory()

    allreduce = (lambda t: dist.all_reduce(t, op=dist.ReduceOp.SUM)) if world > 1 else None

    def fwd():
        if emb_entry:
            eng.forward(emb=c3d_d, gt=gt_d, global_batch=B * world)
        else:
            eng.forward(c3d=c3d_d, gt=gt_d, global_batch=B * world)

    def fwd_bwd():
        fwd()
        if train:
            eng.backward()

    def opt():
        # data parallel: ONE NCCL all-reduce (sum) of the flat fp32 gradient, then the global-norm clip + TF-Adam
        # identically on every rank (base.py:287-297: the norm is over the GLOBAL gradient)
        eng.optimizer_step(allreduce=allreduce)

    def step_eager():
        fwd_bwd()
        if train:
            opt()

    graph, per_replay, graph2 = None, 0, None
    if not args.no_graph:
        if train and world > 1:      # the collective stays outside the captured region: fwd+BPTT graph, all-reduce, Adam graph
            graph, per_replay = eng.make_graph(fwd_bwd)
            allreduce, keep = None, allreduce
            graph2, n2 = eng.make_graph(opt)
            allreduce = keep
            per_replay += n2
        else:
            graph, per_replay = eng.make_graph(step_eager)

    def step_resident():
        if graph is None:
            step_eager()
        elif graph2 is not None:
            graph.replay()
            dist.all_reduce(eng.grads, op=dist.ReduceOp.SUM)
            graph2.replay()
        else:
            graph.replay()

    from recurrent_gaze_prediction_b200.pipeline import PrefetchFeeder
    feeder = PrefetchFeeder([c3d_d, gt_d], device=dev)

    def step_e2e():
        # every step uploads one batch of pinned HOST features + gt (copy stream, overlapping the previous step's
        # kernels), runs the pass, and reads the gaze maps / loss back to pinned host memory
        feeder.prefetch([c3d_h, gt_h])
        feeder.acquire()
        step_resident()
        if not train:
            probs_h.copy_(eng.probs, non_blocking=True)
        loss_h.copy_(eng.loss, non_blocking=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warmup, sample_clocks=False):
        cs = ClockSampler(local) if sample_clocks else None
        if cs:
            cs.start()
        for _ in range(warmup):
            fn()
        barrier()
        if cs:
            cs.mark()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = eng.lib.launches
        if graph is not None:
            l0 -= per_replay * steps
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        clocks = cs.stop() if cs else None
        t = torch.tensor([ms], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]) / steps, clocks, (eng.lib.launches - l0)

    ms_step, clocks, launches = timed(step_resident, args.steps, args.warmup, sample_clocks=(rank == 0))
    frames = B * T * world
    value = frames / (ms_step * 1e-3)
    e2e = None
    if not args.no_e2e:
        feeder.prefetch([c3d_h, gt_h])        # pipeline prologue: step i consumes the batch uploaded during step i-1
        ms_e2e, _, _ = timed(step_e2e, max(4, args.steps), 2)
        feeder.drain()
        h2d = c3d_h.numel() * 4 + gt_h.numel() * 4
        d2h = (0 if train else probs_h.numel() * 4) + 4
        e2e = {"value": frames / (ms_e2e * 1e-3), "unit": "frames/s", "ms_per_step": ms_e2e, "feeding": "double-buffered H2D on a copy stream (PrefetchFeeder)",
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h}

    # roofline of the dominant kernel: the input-half conv GEMM (all T frames at once), timed live with CUDA events
    hbm, tf_burst, tf_sus, which = peaks()
    fl = flops_per_frame(Ch, G, lstm=eng.lstm)
    roof = None
    if rank == 0:
        evs = eng.time_stages(c3d_d, gt_d, iters=3, emb=emb_entry)
        i2h_ms = evs["i2h"]
        ach = fl["i2h"] * B * T / (i2h_ms * 1e-3) / 1e12
        # DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) from the committed ncu --set full captures
        # of exactly this configuration; None for any other configuration
        default_cfg = (args.model, B, T, Ch, G, args.entry, dtype) == ("gaze_grcn", 256, 35, 256, 7, "c3d", 1)
        roof = {"kernel": "tapgemm i2h (3x3 SAME conv, all T frames, M=%d N=%d K=9x512)" % (T * B * eng.S, eng.ngx * Ch),
                "bound": "tensor", "achieved": ach, "peak": tf_sus, "unit": "TFLOP/s", "frac": ach / tf_sus,
                "peak_source": which + " (sustained bf16; burst %.1f)" % tf_burst,
                "traffic": 2.3048e9 if default_cfg else None,
                "traffic_source": "profiles/r01_ncu_full_fwd_pass17.txt (0.595 GB read + 1.710 GB written per launch)" if default_cfg else None,
                "ms_per_launch": i2h_ms, "stage_ms": evs}
        # second view: the recurrence (persistent kernel when the engine picked it).  Algorithmic work per frame
        # (SURVEY 8d): 2*3*(3G-2)^2*Ch^2 flop; bytes = read 3 gate pre-activations + write h, fp32
        rec_ms = evs["recurrence"]
        rec_bytes = 4.0 * Ch * G * G * 4 * B * T
        rec = {"kernel": ("gru_seq (persistent ConvGRU recurrence, all T steps in one launch)" if eng.use_seq else
                          "per-step tap-GEMMs (%d launches)" % (2 * T)),
               "ms": rec_ms, "tensor_TFLOPs": fl["h2h"] * B * T / (rec_ms * 1e-3) / 1e12,
               "tensor_frac": fl["h2h"] * B * T / (rec_ms * 1e-3) / 1e12 / tf_sus,
               "hbm_GBps": rec_bytes / (rec_ms * 1e-3) / 1e9, "hbm_frac": rec_bytes / (rec_ms * 1e-3) / 1e9 / hbm,
               "traffic": 3.5349e9 if (default_cfg and eng.use_seq) else None}
        roof["recurrence"] = rec
        if train:
            fwd()
            roof["stage_ms_bwd"] = eng.time_backward(iters=3)

    cpu = None
    if rank == 0 and not args.no_cpu_baseline:
        nb = args.cpu_clips
        pc = p
        if emb_entry:
            raise SystemExit("--entry emb has no CPU reference leg: pass --no-cpu-baseline")
        c3 = c3d_pool[:nb]
        g = gt_pool[:nb]
        use_all_host_threads()
        cpu_reference_step(args.model, pc, c3, g, train)
        t0 = time.perf_counter()
        n = 0
        while time.perf_counter() - t0 < 10.0 or n < 2:
            cpu_reference_step(args.model, pc, c3, g, train)
            n += 1
        dt = (time.perf_counter() - t0) / n
        cpu = {"value": nb * T / dt, "unit": "frames/s", "cores": torch.get_num_threads(), "kind": "port",
               "sample": "%d clips x T=%d per pass, %d passes, torch-CPU fp32 restatement of the reference graph "
                         "(oracle/model_oracle.py; TF1 not installable)" % (nb, T, n)}
    if rank == 0:
        out = {"metric": metric_name(args), "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
               "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
               "vs_baseline": None, "dtype": args.dtype, "data": "synthetic", "config": dict(workload_config(args), submission="eager" if graph is None else "cuda-graph"),
               "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roof, "cpu_baseline": cpu,
               "step_tflops": sum(fl.values()) * B * T / (ms_step * 1e-3) / 1e12}
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
