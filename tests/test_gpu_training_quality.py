"""Do the reduced-precision engines actually train?  native tf32, native bf16 and the fp32 torch engine (TF32 off)
start from the same weights and see the same 320 batches of a synthetic task that is hard enough not to saturate
(class prototypes buried in unit noise: fp32 ends far from 100 % validation accuracy), under the reference's
optimiser (Adam 1e-3, Balanced All-Reduce/main.py:53).  Final training loss and validation accuracy must agree."""
import json
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

SIGNAL = float(os.environ.get("LDNN_QUALITY_SIGNAL", "0.06"))
EVAL_AT = (200, 240, 280, 320)      # validation accuracy is averaged over these steps (one snapshot of a net that is
                                    # memorising noisy data swings by several points from step to step)
STEPS, BATCH = 320, 64


def _validate(eng, bundle):
    vx, vy = bundle.valset.images.cuda(), torch.from_numpy(bundle.valset.labels).cuda()
    correct, vloss, nb = 0.0, 0.0, vx.shape[0] // BATCH
    for b in range(nb):
        l, c, _ = eng.eval_step(vx[b * BATCH:(b + 1) * BATCH], vy[b * BATCH:(b + 1) * BATCH])
        correct += float(c); vloss += float(l)
    return 100.0 * correct / (nb * BATCH), vloss / nb


def _engine(kind, sd):
    from gpu_util import fp32_mode
    from ldnn_b200.engine import OptimConfig, TorchEngine
    from ldnn_b200.models import enhanced_cnn_spec
    from ldnn_b200.models.native import NativeEngine
    fp32_mode()
    opt = OptimConfig(kind="adam", lr=1e-3)
    if kind == "fp32":
        eng = TorchEngine(enhanced_cnn_spec(), "cuda", opt, init=False)
    else:
        eng = NativeEngine(enhanced_cnn_spec(), "cuda", opt, batch_size=BATCH, init=False, dtype=kind)
    eng.load_state_dict(sd)
    return eng


def _run(kind, bundle, sd, steps=STEPS):
    eng = _engine(kind, sd)
    xs, ys = bundle.trainset.images.cuda(), torch.from_numpy(bundle.trainset.labels).cuda()
    n = xs.shape[0] // BATCH
    losses, accs, vlosses = [], [], []
    for it in range(steps):
        b = it % n
        loss, _ = eng.train_step(xs[b * BATCH:(b + 1) * BATCH], ys[b * BATCH:(b + 1) * BATCH])
        losses.append(loss.detach().float().reshape(()).clone())
        if it + 1 in EVAL_AT:
            a_, l_ = _validate(eng, bundle)
            accs.append(a_); vlosses.append(l_)
    losses = torch.stack(losses).cpu().tolist()
    # the same final weights evaluated by the fp32 torch engine: separates "trained differently" from "evaluates
    # differently" (BN running statistics and the eval kernels of the native engine are part of what is compared)
    cross_acc, _ = _validate(_engine("fp32", eng.state_dict()), bundle)
    return {"losses": losses, "final_train_loss": sum(losses[-40:]) / 40, "val_acc": sum(accs) / len(accs),
            "val_acc_last": accs[-1], "val_loss": sum(vlosses) / len(vlosses), "val_acc_last_in_fp32_engine": cross_acc}


def test_tf32_and_bf16_engines_train_like_fp32():
    from ldnn_b200.data import make_synthetic
    from ldnn_b200.models import enhanced_cnn_spec
    from ldnn_b200.models.torch_net import SpecNet, xavier_init_
    torch.manual_seed(0)
    bundle = make_synthetic(n_train=BATCH * 160, n_val=BATCH * 48, n_test=64, seed=11, signal=SIGNAL)
    ref = SpecNet(enhanced_cnn_spec())
    xavier_init_(ref)
    sd = {k: v.clone() for k, v in ref.state_dict().items()}
    res = {k: _run(k, bundle, sd) for k in ("fp32", "tf32", "bf16")}
    summary = {k: {kk: round(v[kk], 4) for kk in ("final_train_loss", "val_acc", "val_loss", "val_acc_last",
                                                   "val_acc_last_in_fp32_engine")} for k, v in res.items()}
    print("[training quality] signal", SIGNAL, json.dumps(summary))
    out = os.environ.get("LDNN_QUALITY_OUT")
    if out:
        with open(out, "w") as f:
            json.dump({"signal": SIGNAL, "steps": STEPS, "batch": BATCH, "optimizer": "adam lr 1e-3", "summary": summary,
                       "loss_curves_every_8_steps": {k: [round(x, 4) for x in v["losses"][::8]] for k, v in res.items()}}, f)
    f32 = res["fp32"]
    assert f32["losses"][0] > 2.0 and f32["final_train_loss"] < 0.8 * f32["losses"][0], "the oracle itself did not train"
    assert 25.0 < f32["val_acc"] < 97.0, f"task saturated or unlearnable for the oracle: {f32['val_acc']}"
    for k in ("tf32", "bf16"):
        r = res[k]
        # measured gaps 0.02-0.03 at a loss of ~0.8 (3-4 %); two runs of one engine differ by about as much
        assert abs(r["final_train_loss"] - f32["final_train_loss"]) <= 0.07 * f32["final_train_loss"] + 0.02, (k, summary)
        # measured spread of this metric between two runs of the SAME engine (atomics / cuDNN algorithm choice make the
        # trajectories non-reproducible): ~1.5 points; engines differ by 0.3-2.6 points (profiles/training_quality.json)
        assert abs(r["val_acc"] - f32["val_acc"]) <= 3.5, (k, summary)
        # the native eval path (running statistics, eval-mode BN kernel, head) agrees with torch on the same weights
        assert abs(r["val_acc_last"] - r["val_acc_last_in_fp32_engine"]) <= 1.0, (k, summary)
