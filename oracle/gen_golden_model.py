"""Mint tests/golden/model_ref_<model>.npz by RUNNING the reference's own graph builders.

Run in the build container only:   python oracle/gen_golden_model.py

``models/gaze_grcn.py``, ``gaze_grcn77.py``, ``gaze_lstm.py`` (network builders and
cells), ``gaze_rnn.py`` (``create_loss_and_summary``) and ``model_util.py`` are imported
UNMODIFIED from /root/reference and executed over ``oracle/tf1_eager_shim.py`` (TensorFlow
is not installable here).  Variables are created by the reference code in its own order
from ``RandomState(seed)``; ``oracle.model_oracle.init_params`` draws in the same order,
so the fixture only stores the seed, per-variable fingerprints, the reference-run
logits/probabilities/losses, and gradient fingerprints (sum, L2 norm, 64 sampled entries).
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import tf1_eager_shim as SH  # noqa: E402
from recurrent_gaze_prediction_b200 import synthetic as SY  # noqa: E402

B, T = 2, 4
SEED = 11
STDDEV_SCALE = 200.0

# reference variable name (scope-qualified by the shim) -> oracle parameter name
GRU_MAP = {"GRU_Conv_Wz": "GRU_Conv_Wz", "GRU_Conv_Uz": "GRU_Conv_Uz", "GRU_Conv_Wr": "GRU_Conv_Wr",
           "GRU_Conv_Ur": "GRU_Conv_Ur", "GRU_Conv_W": "GRU_Conv_W", "GRU_Conv_U": "GRU_Conv_U"}
# LSTM_RCN_Cell gives several variables the same TF name (gaze_lstm.py:64-88); creation order disambiguates
LSTM_ORDER = ["ConvLSTM_Wxi", "ConvLSTM_Whi", "ConvLSTM_Wci", "ConvLSTM_Wxf", "ConvLSTM_Whf", "ConvLSTM_Wcf",
              "ConvLSTM_Wxc", "ConvLSTM_Whc", "ConvLSTM_Wxo", "ConvLSTM_Who", "ConvLSTM_Wco"]


def fingerprint(a, rs):
    a = np.asarray(a, dtype=np.float64).ravel()
    idx = rs.randint(0, a.size, size=min(64, a.size))
    return np.array([a.sum(), np.sqrt((a * a).sum())]), idx.astype(np.int64), a[idx]


def oracle_names(model, variables):
    names = []
    lstm_i = 0
    bn_t = 0
    for full, _ in variables:
        leaf = full.split("/")[-1]
        if leaf in ("proj_c3d_W", "proj_c3d_b", "out_W", "out_b"):
            names.append(leaf)
        elif leaf.startswith("GRU_Conv_"):
            names.append(GRU_MAP[leaf])
        elif leaf.startswith("ConvLSTM_"):
            names.append(LSTM_ORDER[lstm_i])
            lstm_i += 1
        elif leaf.startswith("weight") and "Upsampling" in full:
            names.append("Upsampling/" + leaf)
        elif leaf == "gamma":
            names.append("bn_gamma[%d]" % bn_t)
        elif leaf == "beta":
            names.append("bn_beta[%d]" % bn_t)
            bn_t += 1
        else:
            raise KeyError(full)
    return names


def run(model):
    SH.install()
    import importlib
    mod = importlib.import_module("models." + model)
    rnn = importlib.import_module("models.gaze_rnn")
    mu = importlib.import_module("models.model_util")
    cls = {"gaze_grcn": "GazePredictionGRCN", "gaze_grcn77": "GazePredictionGRCN", "gaze_lstm": "GazePredictionLSTM"}[model]
    TheModel = getattr(mod, cls)
    GH, GW = mod.CONSTANTS.gazemap_height, mod.CONSTANTS.gazemap_width

    rng = np.random.RandomState(SEED + 1000)
    c3d = SY.make_c3d(rng, B, T).astype(np.float64)
    gt, _ = SY.make_gazemaps(rng, B, T, GH, GW, normalize=True)
    gt = gt.astype(np.float64)
    gt /= gt.sum(axis=(2, 3), keepdims=True)

    SH.reset(seed=SEED, stddev_scale=STDDEV_SCALE, bn_randomize=True)
    frame_images = torch.zeros(B, T, mod.CONSTANTS.image_height, mod.CONSTANTS.image_width, 3, dtype=SH.DT)
    net = {}
    logits = TheModel.create_gazeprediction_network(frame_images, torch.tensor(c3d), 1.0, net)
    gt_t = torch.tensor(gt)
    loss_x = rnn.GazePredictionGRU.create_loss_and_summary(logits, gt_t, "xentropy")
    loss_l2 = rnn.GazePredictionGRU.create_loss_and_summary(logits, gt_t, "l2")
    probs = torch.stack([mu.tf_softmax_2d(logits[:, i, :, :]) for i in range(T)]).permute(1, 0, 2, 3)  # gaze_rnn.py:149-159
    variables = SH.STATE["variables"]
    names = oracle_names(model, variables)
    gx = torch.autograd.grad(loss_x, [v for _, v in variables], allow_unused=True, retain_graph=True)
    gl = torch.autograd.grad(loss_l2, [v for _, v in variables], allow_unused=True)

    out = {"seed": SEED, "stddev_scale": STDDEV_SCALE, "B": B, "T": T, "input_seed": SEED + 1000,
           "logits": logits.detach().numpy(), "probs": probs.detach().numpy(),
           "loss_xentropy": float(loss_x), "loss_l2": float(loss_l2),
           "rcn_last": net["rcn_outputs"][-1].detach().numpy(),
           "emb_fp": fingerprint(net["c3d_embedded"].detach().numpy(), np.random.RandomState(5))[0],
           "var_names": np.array(names)}
    rs = np.random.RandomState(99)
    for nm, (full, v), g1, g2 in zip(names, variables, gx, gl):
        s, idx, samp = fingerprint(v.detach().numpy(), rs)
        out["var/" + nm + "/sum"] = s
        out["var/" + nm + "/idx"] = idx
        out["var/" + nm + "/samp"] = samp
        for tag, g in (("gx", g1), ("gl2", g2)):
            gn = np.zeros(tuple(v.shape)) if g is None else g.detach().numpy()
            a = gn.ravel()
            out["%s/%s/sum" % (tag, nm)] = np.array([a.sum(), np.sqrt((a * a).sum())])
            out["%s/%s/samp" % (tag, nm)] = a[idx]
            out["%s/%s/none" % (tag, nm)] = np.array(g is None)
    path = os.path.join(ROOT, "tests", "golden", "model_ref_%s.npz" % model)
    np.savez_compressed(path, **out)
    print(model, "vars:", len(variables), "loss", float(loss_x), float(loss_l2), "->", path, os.path.getsize(path), "bytes")


def run_cascade():
    """models/gaze_grcn_cascade.py: the file ends in Python-2 `print` statements (self_test / main, :541-), so it is not
    importable; its text is read from /root/reference AT RUN TIME, cut before `def self_test` (nothing on the path
    lives below) and executed unmodified over the shim.  One binding is replaced in the executed namespace, nothing in
    the text: GRU_RCN_Cell is wrapped so that the top cell's declared 65 input channels become the 64 it is fed
    (reading A of the shape bug at :346-351,379, see oracle/model_oracle.py) - with 65 the reference's own conv2d
    raises."""
    SH.install()
    import importlib
    importlib.import_module("models.gaze_rnn")            # the cascade imports GazePredictionGRU from it
    src = open("/root/reference/models/gaze_grcn_cascade.py").read()
    cut = src.index("def self_test(")
    ns = {"__name__": "models.gaze_grcn_cascade"}
    exec(compile(src[:cut], "/root/reference/models/gaze_grcn_cascade.py", "exec"), ns)
    Orig = ns["GRU_RCN_Cell"]

    class FedChannelsCell(Orig):
        def __init__(self, num_units, dim_feature, *a, **kw):
            Orig.__init__(self, num_units, 64 if dim_feature == 65 else dim_feature, *a, **kw)
    ns["GRU_RCN_Cell"] = FedChannelsCell
    TheModel = ns["GazePredictionGRCN"]
    Bc, Tc = 2, 3
    rng = np.random.RandomState(SEED + 2000)
    c3d = SY.make_c3d(rng, Bc, Tc).astype(np.float64)
    gt, _ = SY.make_gazemaps(rng, Bc, Tc, 49, 49, normalize=False)
    gt = gt.astype(np.float64)
    SH.reset(seed=SEED, stddev_scale=STDDEV_SCALE)
    frame_images = torch.zeros(Bc, Tc, 98, 98, 3, dtype=SH.DT)
    net = {}
    pred, loss, _ = TheModel.create_gazeprediction_network(frame_images, torch.tensor(c3d), torch.tensor(gt), None, net)
    variables = SH.STATE["variables"]
    names = []
    for full, _ in variables:
        parts = full.split("/")
        leaf = parts[-1]
        if leaf in ("proj_c3d_W", "proj_c3d_b"):
            names.append(leaf)
        elif leaf.startswith("GRU_Conv_"):
            names.append(parts[0] + "/" + leaf)                    # RCNBottom/..., RCNGaze/...
        elif "Upsampling" in full:
            names.append("Upsampling/weight")
        elif leaf in ("weights", "bias"):
            names.append("LastProjection/%s/%s" % (parts[-2], leaf))
        else:
            raise KeyError(full)
    out = {"seed": SEED, "stddev_scale": STDDEV_SCALE, "B": Bc, "T": Tc, "input_seed": SEED + 2000,
           "pred": pred.detach().numpy(), "loss_l2": float(loss),
           "rcn_last": net["rcn_outputs"][Tc - 1].detach().numpy(),
           "up_fp": fingerprint(net["rcn_upsampled_outputs"][-1].detach().numpy(), np.random.RandomState(5))[0],
           "var_names": np.array(names)}
    rs = np.random.RandomState(99)
    for nm, (full, v) in zip(names, variables):
        s_, idx, samp = fingerprint(v.detach().numpy(), rs)
        out["var/" + nm + "/sum"] = s_
        out["var/" + nm + "/idx"] = idx
        out["var/" + nm + "/samp"] = samp
        out["var/" + nm + "/shape"] = np.array(v.shape)
    path = os.path.join(ROOT, "tests", "golden", "model_ref_gaze_grcn_cascade.npz")
    np.savez_compressed(path, **out)
    print("gaze_grcn_cascade vars:", len(variables), "loss", float(loss), "->", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    for m in sys.argv[1:] or ["gaze_grcn", "gaze_grcn77", "gaze_lstm", "gaze_grcn_cascade"]:
        if m == "gaze_grcn_cascade":
            run_cascade()
        else:
            run(m)
