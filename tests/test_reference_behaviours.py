"""SURVEY §2.3 behaviours reproduced on the reference schedule (2 ranks, gloo, tiny widths):
B1 — with --aggregation_by gradients the exchange never reaches the weights (no optimizer step follows it);
     --aggregation_by weights mixes the models;  the product schedule (--sync step) keeps ranks identical;
B6 — BN running statistics are broadcast once and then drift per rank (the exchange walks parameters only)."""
import os

import pytest
import torch


def _worker(rank, world, mode, out_dir):
    os.environ["LDNN_TINY_WIDTHS"] = "1"
    from ldnn_b200 import data as D
    from ldnn_b200.config import Config, resolve
    from ldnn_b200.engine import make_engine
    from ldnn_b200.launch import init_runtime, shutdown
    from ldnn_b200.parallel import make_exchanger
    from ldnn_b200.trainer import train_global
    kw = dict(backend="gloo", synthetic=True, synthetic_train=128, synthetic_val=32, synthetic_test=32,
              epochs_global=2, epochs_local=1, batch_size=32, probe_batches=1, seed=11, quiet=True,
              topology="allreduce", engine="torch", exchange_impl="dist", output_folder=out_dir)
    if mode == "gradients":
        kw.update(aggregation_by="gradients")
    elif mode == "weights":
        kw.update(aggregation_by="weights")
    elif mode == "step_weights":
        kw.update(sync="step", aggregation_by="weights")
    else:
        kw.update(sync="step")
    cfg = resolve(Config(**kw))
    rt = init_runtime("gloo", seed=cfg.seed)
    engine = make_engine(cfg, rt)
    bundle = D.get_data(cfg, rt.rank)
    ex = make_exchanger(cfg, rt, engine)
    ex.broadcast_state(engine, src=0)
    p0 = engine.flat.params.clone()
    shards = D.initial_shards(bundle, cfg, rt, engine)
    train_global(engine, bundle, shards, cfg, rt, ex)
    sd = engine.state_dict()
    torch.save({"params": engine.flat.params.clone(), "p0": p0,
                "bn_mean": torch.cat([v.flatten().float() for k, v in sd.items() if k.endswith("running_mean")])},
               os.path.join(out_dir, f"{mode}_{rank}.pt"))
    ex.close()
    shutdown(rt)


@pytest.mark.parametrize("mode", ["gradients", "weights", "step", "step_weights"])
def test_b1_b6_what_the_exchange_reaches(dist_runner, tmp_path, mode):
    dist_runner(_worker, 2, mode, str(tmp_path))
    a, b = (torch.load(tmp_path / f"{mode}_{r}.pt") for r in range(2))
    assert torch.equal(a["p0"], b["p0"])                              # rank-0 broadcast at start (C1)
    assert not torch.equal(a["params"], a["p0"])                      # training moved the weights
    if mode == "gradients":
        # B1: the exchanged gradient is never applied -> the ranks' models only ever saw their own shard
        assert not torch.allclose(a["params"], b["params"], atol=1e-6)
    else:
        # weights are averaged after the last global epoch / every step's update uses the same mean gradient
        assert torch.allclose(a["params"], b["params"], atol=1e-6)
    # B6: running statistics are per rank in every mode (different shards, never exchanged after the broadcast)
    assert not torch.allclose(a["bn_mean"], b["bn_mean"], atol=1e-7)


def _shard_weight_worker(rank, world, out_dir):
    os.environ["LDNN_TINY_WIDTHS"] = "1"
    import json
    from ldnn_b200 import data as D
    from ldnn_b200.config import Config, resolve
    from ldnn_b200.engine import make_engine
    from ldnn_b200.launch import init_runtime, shutdown
    from ldnn_b200.parallel import make_exchanger
    from ldnn_b200.trainer import refresh_shard_weights, train_global
    # no explicit --shard_ratios: shard sizes follow the timing-derived ratios and move at every re-shard
    cfg = resolve(Config(backend="gloo", synthetic=True, synthetic_train=160, synthetic_val=32, synthetic_test=32,
                         epochs_global=3, epochs_local=1, batch_size=16, probe_batches=1, seed=5, quiet=True,
                         topology="allreduce", engine="torch", exchange_impl="dist", aggregation_type="shard",
                         aggregation_by="weights", output_folder=out_dir))
    rt = init_runtime("gloo", seed=cfg.seed)
    engine = make_engine(cfg, rt)
    bundle = D.get_data(cfg, rt.rank)
    ex = make_exchanger(cfg, rt, engine)
    ex.broadcast_state(engine, src=0)
    shards = D.initial_shards(bundle, cfg, rt, engine)
    refresh_shard_weights(cfg, rt, shards, ex)
    seen = []

    def on_epoch_end(ge, h, sh):
        sizes = D.all_gather_floats(float(len(sh.train_indices)), rt)
        seen.append({"sizes": sizes, "weights": list(ex.spec.shard_weights)})

    train_global(engine, bundle, shards, cfg, rt, ex, on_epoch_end=on_epoch_end)
    if rank == 0:
        with open(os.path.join(out_dir, "shard_weights.json"), "w") as f:
            json.dump(seen, f)
    ex.close()
    shutdown(rt)


def test_shard_weights_follow_the_shards_across_reshards(dist_runner, tmp_path):
    """--aggregation_type shard: w_r = n_r / sum(n) of the shards the ranks hold NOW — re-computed after every re-shard
    (the sizes follow ratio x total and move with the measured epoch durations), not frozen at start-up."""
    import json
    dist_runner(_shard_weight_worker, 2, str(tmp_path))
    seen = json.load(open(tmp_path / "shard_weights.json"))
    assert len(seen) == 3
    for e in seen:
        tot = sum(e["sizes"])
        assert tot > 0 and abs(sum(e["weights"]) - 1.0) < 1e-9
        for w, n in zip(e["weights"], e["sizes"]):
            assert abs(w - n / tot) < 1e-9, e


def _empty_shard_worker(rank, world, out_dir):
    os.environ["LDNN_TINY_WIDTHS"] = "1"
    from ldnn_b200 import data as D
    from ldnn_b200.config import Config, resolve
    from ldnn_b200.engine import make_engine
    from ldnn_b200.launch import init_runtime, shutdown
    from ldnn_b200.parallel import make_exchanger
    from ldnn_b200.trainer import train_global
    cfg = resolve(Config(backend="gloo", synthetic=True, synthetic_train=64, synthetic_val=32, synthetic_test=32,
                         epochs_global=1, epochs_local=1, batch_size=16, probe_batches=1, seed=5, quiet=True, sync="step",
                         topology="allreduce", engine="torch", exchange_impl="dist", shard_ratios=[1000.0, 1.0],
                         output_folder=out_dir))
    rt = init_runtime("gloo", seed=cfg.seed)
    engine = make_engine(cfg, rt)
    bundle = D.get_data(cfg, rt.rank)
    ex = make_exchanger(cfg, rt, engine)
    shards = D.initial_shards(bundle, cfg, rt, engine)
    msg = "no error"
    try:
        train_global(engine, bundle, shards, cfg, rt, ex)
    except RuntimeError as e:
        msg = str(e)
    with open(os.path.join(out_dir, f"empty_{rank}.txt"), "w") as f:
        f.write(f"{len(shards.train_indices)}|{msg}")
    shutdown(rt)


def test_per_step_schedule_reports_an_empty_shard_instead_of_hanging(dist_runner, tmp_path):
    """--sync step: every step is a collective, so a rank with nothing to feed it would leave its peers waiting.  All
    ranks see the same minimum shard length and stop together with a clear error."""
    dist_runner(_empty_shard_worker, 2, str(tmp_path))
    out = [open(tmp_path / f"empty_{r}.txt").read().split("|", 1) for r in range(2)]
    assert int(out[1][0]) == 0 and int(out[0][0]) > 0            # ratios 1000:1 of 64 samples: rank 1 holds nothing
    for n, msg in out:
        assert "empty training shard" in msg, msg
