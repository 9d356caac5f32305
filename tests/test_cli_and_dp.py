"""train_gaze / evaluate_gaze entry points (reference flag names) and data-parallel training over 2 gloo ranks,
on the NumPy emulation of the C-ABI.  CPU only."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from recurrent_gaze_prediction_b200.models import evaluate_gaze, train_gaze
from recurrent_gaze_prediction_b200 import synthetic as SY
from emu_lib import EmuLib

TINY = ["--hidden", "16", "--n_lstm_steps", "2", "--dtype", "fp32", "--device", "cpu", "--dataset", "synthetic"]


def test_train_flags_match_reference():
    p = train_gaze.make_parser()
    a = p.parse_args(["--model", "gaze_grcn", "--tag", "x", "--max_grad_norm", "5", "--learning_rate", "0.01",
                      "--learning_rate_decay", "0.5", "--loss_type", "l2", "--batch_size", "6", "--max_steps", "3",
                      "--batch_norm", "--gpu_fraction", "0.3", "--train_dir", "/tmp/x", "--dataset", "crc"])
    _, _, Cfg = train_gaze.get_model_class(a.model)
    c = train_gaze.build_config(a, Cfg)
    assert (c.batch_size, c.max_grad_norm, c.initial_learning_rate, c.learning_rate_decay, c.loss_type, c.max_steps,
            c.train_tag) == (6, 5.0, 0.01, 0.5, "l2", 3, "x")
    assert (c.steps_per_evaluation, c.steps_per_validation, c.steps_per_checkpoint) == (100, 20, 100)
    d = train_gaze.build_config(p.parse_args(["--model", "gaze_grcn"]), Cfg)
    assert d.batch_size == 28 and d.initial_learning_rate == 0.0001 and d.use_flip_batch       # train_gaze.py:75-81
    with pytest.raises(NotImplementedError):
        train_gaze.get_model_class("gaze_c3d_conv")
    with pytest.raises(NotImplementedError):
        train_gaze.load_data_sets(a, c, None)


def test_train_then_evaluate_end_to_end(tmp_path):
    args = train_gaze.make_parser().parse_args(["--model", "gaze_grcn", "--max_steps", "3", "--batch_size", "4",
                                                "--train_dir", str(tmp_path), "--learning_rate", "0.001"] + TINY)
    # shrink the feature width through the config hook (the CLI keeps the reference's 1024)
    orig = train_gaze.build_config

    def small(a, Cfg):
        c = orig(a, Cfg)
        c.dim_feature, c.dim_cnn_proj, c.steps_per_checkpoint, c.steps_per_logprint = 64, 32, 2, 1000
        c.steps_per_validation, c.steps_per_evaluation = 2, 1000
        return c
    train_gaze.build_config = small
    try:
        np.random.seed(0)
        model, scores = train_gaze.train(args, session={"lib": EmuLib()})
    finally:
        train_gaze.build_config = orig
    assert model.current_step == 4 and sorted(scores) == ['AUC_Borji', 'AUC_shuffled', 'cc', 'sim']
    ckpt = os.path.join(str(tmp_path), "model", "GazePredictionGRCN-4.npz")
    assert os.path.isfile(ckpt)
    # evaluate_gaze pieces on the trained model
    eargs = evaluate_gaze.make_parser().parse_args(["--model", "gaze_grcn", "--checkpoint_path", ckpt, "--num_frames",
                                                    "20", "-j", "3", "--batch_size", "4"] + TINY)
    assert evaluate_gaze.get_out_dir("synthetic", ckpt).endswith("generated-synthetic-4")
    ds = SY.read_synthetic_data_sets(4, 16, 4, T=2, dim_feature=64, fix_hw=(60, 80))
    evaluate_gaze.reload_checkpoint(model, ckpt)
    np.random.seed(1)
    out = evaluate_gaze.run_evaluation(model, eargs, ds, str(tmp_path / "gen"))
    assert list(out) == ['sim', 'cc', 'AUC_Borji', 'AUC_Judd', 'AUC_shuffled']
    n = len(out['sim'])
    assert n == 24 and all(len(v) == n for v in out.values())
    txt = open(str(tmp_path / "gen" / "overall.txt")).read()
    assert txt.startswith("Average sim : ") and txt.count("Average ") == 5
    with pytest.raises(ValueError):
        evaluate_gaze.main(["--model", "gaze_grcn"] + TINY)


def _dp_worker(rank, world, initfile, out):
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    dist.init_process_group("gloo", init_method="file://" + initfile, rank=rank, world_size=world)
    from recurrent_gaze_prediction_b200.models import gaze_grcn
    from test_models_emu import small_config, data
    c = small_config(T=2, B=4)
    sess = {"rank": rank, "world_size": world, "lib": EmuLib(),
            "allreduce": lambda t: dist.all_reduce(t, op=dist.ReduceOp.SUM)}
    m = gaze_grcn.GazePredictionGRCN(sess, data(T=2), c)
    assert m.local_batch == 2 and m.engine.B == 2
    np.random.seed(7)
    m.single_step()
    m.single_step()
    ret = m.generate(m.data_sets.valid, max_instances=4)
    if rank == 0:
        np.savez(out, params=m.engine.params.numpy(), loss=m.last_loss, pred=ret['pred_gazemap_list'])
    dist.barrier()
    dist.destroy_process_group()


def test_data_parallel_two_ranks_equals_single_process(tmp_path):
    """2 gloo ranks x 2 clips == 1 process x 4 clips: same loss, same parameters after two clip+Adam steps, and
    generate() returns the gathered maps of the whole batch on every rank."""
    from recurrent_gaze_prediction_b200.models import gaze_grcn
    from test_models_emu import small_config, data
    out = str(tmp_path / "dp.npz")
    mp.spawn(_dp_worker, args=(2, str(tmp_path / "init"), out), nprocs=2, join=True)
    z = np.load(out)
    m = gaze_grcn.GazePredictionGRCN({"lib": EmuLib()}, data(T=2), small_config(T=2, B=4))
    np.random.seed(7)
    m.single_step()
    m.single_step()
    ret = m.generate(m.data_sets.valid, max_instances=4)
    assert abs(float(z["loss"]) - m.last_loss) <= 1e-6 * abs(m.last_loss)
    ref = m.engine.params.numpy()
    off, n = m.engine.offsets["out_b"]
    mask = np.ones(ref.size, bool)
    mask[off:off + n] = False                 # zero-gradient parameter: Adam amplifies summation-order noise
    assert np.abs(z["params"][mask] - ref[mask]).max() <= 2e-5
    assert np.abs(z["pred"] - ret['pred_gazemap_list']).max() <= 1e-6


def test_bench_clock_sampler_reports_only_rows_inside_the_timed_region():
    """bench.py's ClockSampler: rows are time-stamped, mark() opens the timed region, stop() reports the median clock and
    the reasons of the rows inside it; a region shorter than one sampling period falls back to warm-up + region rows."""
    import importlib.util
    import sys
    import time
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    argv, sys.argv = sys.argv, ["bench.py"]
    try:
        spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(root, "bench.py"))
        bench = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(bench)
    finally:
        sys.argv = argv
    cs = bench.ClockSampler(0)
    cs.source = "nvml"
    now = time.perf_counter()
    cs.rows = [(now - 2.0, 1200.0, 1965.0, ["sw_power_cap"]), (now - 1.5, 1300.0, 1965.0, [])]   # warm-up rows
    cs.t0 = now - 1.0
    cs.rows += [(now - 0.8, 1965.0, 1965.0, []), (now - 0.5, 1950.0, 1965.0, ["sw_thermal_slowdown"]), (now - 0.2, 1965.0, 1965.0, [])]
    out = cs.stop()
    assert out["samples"] == 3 and out["sm_mhz"] == 1965.0 and out["sm_max_mhz"] == 1965.0
    assert out["reasons"] == ["sw_thermal_slowdown"] and out["window"] == "timed region"
    cs2 = bench.ClockSampler(0)
    cs2.source = "nvml"
    cs2.rows = [(now - 2.0, 1200.0, 1965.0, ["sw_power_cap"])]
    cs2.t0 = time.perf_counter()
    out2 = cs2.stop()
    assert out2["samples"] == 1 and out2["window"] == "warm-up + timed region" and out2["reasons"] == ["sw_power_cap"]
    assert bench.ClockSampler(0).stop()["samples"] == 0
