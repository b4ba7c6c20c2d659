"""Flag surface, schedule resolution, checkpoint round trip and the PNG rasteriser (CPU)."""
import json
import struct
import zlib

import numpy as np
import torch

from ldnn_b200.config import Config, build_parser, parse_args, resolve


def test_reference_flags_and_defaults():
    cfg = parse_args([], topology="ring", balance="balanced")
    assert (cfg.backend, cfg.epochs_local, cfg.epochs_global, cfg.batch_size, cfg.lr) == ("gloo", 5, 20, 64, 1e-3)
    assert (cfg.time_limit, cfg.prev_fraction, cfg.next_fraction, cfg.local_weight, cfg.fixed_ratio) == (60, 0.5, 0.5, 0.5, 0.5)
    assert (cfg.aggregation_type, cfg.aggregation_by) == ("equal", "gradients")
    # dead upstream flags still parse
    cfg = parse_args(["--local-rank", "3", "--gpu_weight", "7", "--dist-url", "tcp://x:1", "--backend", "nccl"])
    assert cfg.local_rank == 3 and cfg.gpu_weight == 7 and cfg.dist_url == "tcp://x:1"
    names = {a.dest for a in build_parser()._actions}
    for flag in ["epochs_local", "epochs_global", "batch_size", "lr", "time_limit", "prev_fraction", "next_fraction",
                 "aggregation_type", "aggregation_by", "local_weight", "fixed_ratio", "gpu_weight", "dist_url", "backend"]:
        assert flag in names, flag


def test_schedule_resolution():
    c = resolve(Config(topology="ring", sync="global_epoch"))
    assert c.semantics == "gossip" and c.apply_exchanged_grads is False          # reference behaviour
    c = resolve(Config(topology="ring", sync="step"))
    assert c.semantics == "allreduce" and c.apply_exchanged_grads is True        # product path
    c = resolve(Config(topology="allreduce", sync="global_epoch"))
    assert c.semantics == "allreduce"
    # sampling with replacement: BalAR/BalR without, BalDR and every Dis* with
    assert resolve(Config(topology="ring", balance="balanced")).replace is False
    assert resolve(Config(topology="double_ring", balance="balanced")).replace is True
    assert resolve(Config(topology="allreduce", balance="disbalanced")).replace is True


def test_checkpoint_roundtrip_torch_engine(tmp_path):
    from ldnn_b200.engine import OptimConfig, TorchEngine
    from ldnn_b200.utils.checkpoint import load_checkpoint, save_checkpoint
    from test_flat_optimizer import tiny_spec
    torch.manual_seed(0)
    a = TorchEngine(tiny_spec(), "cpu", OptimConfig(lr=1e-2))
    x, y = torch.randn(8, 3, 8, 8), torch.randint(0, 10, (8,))
    for _ in range(3):
        a.train_step(x, y)
    a.scheduler_step()
    path = str(tmp_path / "ck.pt")
    save_checkpoint(path, a, epoch=5)
    b = TorchEngine(tiny_spec(), "cpu", OptimConfig(lr=1e-2))
    assert load_checkpoint(path, b) == 5
    assert torch.equal(a.flat.params, b.flat.params) and torch.equal(a.flat.m, b.flat.m) and torch.equal(a.flat.v, b.flat.v)
    assert (a.flat.step, a.flat.epoch) == (b.flat.step, b.flat.epoch)
    la, _ = a.train_step(x, y)
    lb, _ = b.train_step(x, y)
    assert torch.equal(la, lb)
    # a bare upstream-style state_dict file loads too
    torch.save(a.state_dict(), str(tmp_path / "sd.pt"))
    c = TorchEngine(tiny_spec(), "cpu", OptimConfig())
    assert load_checkpoint(str(tmp_path / "sd.pt"), c) == 0
    assert torch.equal(c.flat.params, a.flat.params)


def test_rasteriser_writes_valid_png(tmp_path):
    from ldnn_b200.utils import viz
    viz.plot_metrics_global(3, [1.0, 0.5, 0.2], [10, 50, 80], [1.2, 0.9, 0.7], [9, 40, 60], output_folder=str(tmp_path))
    viz.plot_loss_distribution_per_epoch([[0.1, 0.2, 0.9], [0.05, 0.1]], output_folder=str(tmp_path))
    viz.plot_time_throughput([1.0, 0.9], [1000, 1100], [5.0, 4.0], output_folder=str(tmp_path))
    for name in ("training_metrics.png", "loss_distribution_per_epoch.png", "time_throughput.png"):
        raw = (tmp_path / name).read_bytes()
        assert raw[:8] == b"\x89PNG\r\n\x1a\n"
        w, h = struct.unpack(">II", raw[16:24])
        assert (w, h) == (1600, 1000) or w > 100          # builtin canvas (matplotlib, if present, picks its own)
        # IDAT decompresses to h rows of 1 + 3*w bytes for the builtin writer
        if (w, h) == (1600, 1000):
            i = raw.index(b"IDAT")
            n = struct.unpack(">I", raw[i - 4:i])[0]
            assert len(zlib.decompress(raw[i + 4:i + 4 + n])) == h * (1 + 3 * w)


def test_tune_cache_roundtrip_and_replay(tmp_path, monkeypatch):
    """Persistent conv-tuning file: written atomically, merged on load, replayed when timing is disabled."""
    from ldnn_b200.ops import native_ops as K
    path = tmp_path / "tune.json"
    monkeypatch.setenv("LDNN_TUNE_CACHE", str(path))
    monkeypatch.setattr(K, "_TUNE_CACHE", {})
    monkeypatch.setattr(K, "_TUNE_FILE_LOADED", False)
    geo = (64, 32, 32, 64, 128, 3, 1, 1)
    # keys are (dtype, mode, geometry): dtype 0 = bf16 kernels, 1 = tf32 kernels (tuned separately)
    K._TUNE_CACHE[(0, K.FPROP) + geo] = (128, 2, True)
    K._TUNE_CACHE[(1, K.FPROP) + geo] = (64, 3, False)
    K._TUNE_CACHE[(0, K.WGRAD) + geo] = (0, 0, False)
    assert K.save_tune_cache()
    monkeypatch.setattr(K, "_TUNE_CACHE", {})
    assert K.load_tune_cache() == 3
    assert K._TUNE_CACHE[(0, K.FPROP) + geo] == (128, 2, True)
    # timing disabled: known choices are replayed per dtype, unknown shapes fall back to heuristics;
    # make_plan must never be called on this path (no GPU here)
    monkeypatch.setenv("LDNN_NO_TUNE_DEFAULTS", "1")
    boom = lambda *a: (_ for _ in ()).throw(AssertionError("timed"))
    assert K.tuned_conv_config(K.FPROP, geo, boom, enabled=False) == (128, 2, True)
    assert K.tuned_conv_config(K.FPROP, geo, boom, enabled=False, dtype=1) == (64, 3, False)
    assert K.tuned_conv_config(K.DGRAD, geo, boom, enabled=False) == (0, 0, False)
    # a round-1 file (9-int keys, bf16 kernels only) still loads, as dtype 0
    old = tmp_path / "old.json"
    old.write_text('{"version": 1, "configs": {"1,64,32,32,64,128,3,1,1": [64, 4, false]}}')
    assert K.load_tune_cache(str(old)) == 1
    assert K._TUNE_CACHE[(0, K.DGRAD) + geo] == (64, 4, False)
    # the shipped defaults cover the repo model's shapes at batch 64, so a fresh process does not re-time them
    monkeypatch.setattr(K, "_TUNE_CACHE", {})
    assert K.load_tune_cache(K.DEFAULT_TUNE_FILE) > 0
    # a corrupt file is ignored
    path.write_text("{not json")
    assert K.load_tune_cache(str(path)) == 0


def test_exchange_impl_follows_the_node_topology(monkeypatch):
    """auto = fused P2P kernels inside one host, torch.distributed when the job spans hosts; an explicit `fused`
    across hosts is an error, not a hang."""
    import types
    import torch
    from ldnn_b200.launch import Runtime
    from ldnn_b200.parallel import make_exchanger
    from ldnn_b200.parallel.exchange import DistExchanger
    from ldnn_b200.config import Config, resolve
    rt = Runtime(rank=0, world_size=16, local_rank=0, device=torch.device("cpu"), backend="gloo", owns_group=False)
    monkeypatch.setenv("LOCAL_WORLD_SIZE", "8")
    assert rt.single_node() is False
    rt2 = Runtime(rank=0, world_size=8, local_rank=0, device=torch.device("cpu"), backend="gloo", owns_group=False)
    assert rt2.single_node() is True
    assert Runtime(0, 1, 0, torch.device("cpu"), "gloo", False).single_node() is True
    # a CUDA runtime that spans two hosts: auto -> dist, explicit fused -> clear error
    fake = types.SimpleNamespace(rank=0, world_size=16, is_cuda=True, device=torch.device("cpu"), single_node=lambda: False)
    ex = make_exchanger(resolve(Config(exchange_impl="auto")), fake, engine=None)
    assert isinstance(ex, DistExchanger)
    import pytest
    with pytest.raises(RuntimeError, match="one host"):
        make_exchanger(resolve(Config(exchange_impl="fused")), fake, engine=None)


def test_checkpoint_restores_each_ranks_own_shard(tmp_path):
    import types
    import numpy as np
    from ldnn_b200.data import ShardState
    from ldnn_b200.engine import OptimConfig, TorchEngine
    from ldnn_b200.utils.checkpoint import load_checkpoint, save_checkpoint
    from test_flat_optimizer import tiny_spec
    eng = TorchEngine(tiny_spec(), "cpu", OptimConfig())
    mk = lambda lo, hi: ShardState(np.arange(lo, hi), np.arange(lo, lo + 2), [0.75, 0.25], None)
    blob = lambda st: {"train": st.train_indices, "val": st.val_indices, "ratios": st.ratios, "fixed_classes": None}
    path = str(tmp_path / "c.pt")
    save_checkpoint(path, eng, shards=mk(0, 30), epoch=1, all_shards=[blob(mk(0, 30)), blob(mk(30, 40))])
    for rank, want in ((0, (0, 30)), (1, (30, 40))):
        st = mk(100, 110)
        assert load_checkpoint(path, eng, shards=st, rank=rank, world_size=2) == 1
        assert st.train_indices[0] == want[0] and st.train_indices[-1] == want[1] - 1
    # the job now has 4 ranks: nobody inherits rank 0's shard, the fresh split is kept
    st = mk(100, 110)
    load_checkpoint(path, eng, shards=st, rank=3, world_size=4)
    assert st.train_indices[0] == 100
    # single-process checkpoints restore into a single process
    save_checkpoint(path, eng, shards=mk(5, 9), epoch=2)
    st = mk(100, 110)
    load_checkpoint(path, eng, shards=st)
    assert st.train_indices[0] == 5
    st = mk(100, 110)
    load_checkpoint(path, eng, shards=st, rank=1, world_size=2)
    assert st.train_indices[0] == 100


def test_dtype_flag_selects_the_engine():
    import types
    import pytest
    import torch
    from ldnn_b200.config import Config
    from ldnn_b200.engine import TorchEngine, make_engine
    cpu = types.SimpleNamespace(is_cuda=False, device=torch.device("cpu"), world_size=1, rank=0)
    import os
    os.environ["LDNN_TINY_WIDTHS"] = "1"
    try:
        assert isinstance(make_engine(Config(dtype="fp32"), cpu), TorchEngine)
        assert isinstance(make_engine(Config(dtype="bf16"), cpu), TorchEngine)          # no GPU: torch engine
        fake_gpu = types.SimpleNamespace(is_cuda=True, device=torch.device("cpu"), world_size=1, rank=0)
        assert isinstance(make_engine(Config(dtype="fp32"), fake_gpu), TorchEngine)     # fp32 on a GPU: torch engine
        with pytest.raises(RuntimeError, match="bf16"):
            make_engine(Config(dtype="fp32", engine="native"), fake_gpu)
    finally:
        del os.environ["LDNN_TINY_WIDTHS"]


def test_launcher_env_understands_mpirun_and_srun():
    """`mpirun -n N python main.py` (how upstream starts the ring variants) must not start N one-rank jobs."""
    from ldnn_b200.launch import launcher_env
    assert launcher_env({}) == (0, 1, 0)
    assert launcher_env({"RANK": "3", "WORLD_SIZE": "8", "LOCAL_RANK": "3"}) == (3, 8, 3)
    assert launcher_env({"OMPI_COMM_WORLD_RANK": "2", "OMPI_COMM_WORLD_SIZE": "4", "OMPI_COMM_WORLD_LOCAL_RANK": "0"}) == (2, 4, 0)
    assert launcher_env({"PMI_RANK": "1", "PMI_SIZE": "2"}) == (1, 2, 1)
    assert launcher_env({"SLURM_PROCID": "5", "SLURM_NTASKS": "16", "SLURM_LOCALID": "1"}) == (5, 16, 1)
    # torchrun's variables win when several launch systems are visible (srun -> torchrun)
    assert launcher_env({"RANK": "1", "WORLD_SIZE": "2", "SLURM_PROCID": "0", "SLURM_NTASKS": "1"}) == (1, 2, 1)


def test_checkpoint_keeps_every_ranks_own_replica(tmp_path):
    """Replicas diverge under the reference's schedule; a resume must not collapse them onto rank 0."""
    import warnings
    from ldnn_b200.engine import OptimConfig, TorchEngine
    from ldnn_b200.utils.checkpoint import load_checkpoint, rank_path, save_checkpoint, save_rank_state
    from test_flat_optimizer import tiny_spec
    engs = [TorchEngine(tiny_spec(), "cpu", OptimConfig()) for _ in range(2)]
    for r, e in enumerate(engs):
        e.flat.params.add_(0.1 * (r + 1))
        e.flat.m.fill_(float(r + 1))
        e.flat.step = 7
    path = str(tmp_path / "c.pt")
    save_checkpoint(path, engs[0], epoch=4)
    save_rank_state(path, engs[1], rank=1, world_size=2, epoch=4)
    assert rank_path(path, 0) == path and rank_path(path, 1) == path + ".rank1"
    for r in range(2):
        fresh = TorchEngine(tiny_spec(), "cpu", OptimConfig())
        assert load_checkpoint(path, fresh, rank=r, world_size=2) == 4
        for sl in fresh.flat.slots:           # per tensor: the alignment padding between tensors is not state
            assert torch.equal(fresh.flat.view(fresh.flat.params, sl.key), engs[r].flat.view(engs[r].flat.params, sl.key))
            assert torch.equal(fresh.flat.view(fresh.flat.m, sl.key), engs[r].flat.view(engs[r].flat.m, sl.key))
    # a side file from another epoch is not trusted: rank 0's state, with a warning
    save_rank_state(path, engs[1], rank=1, world_size=2, epoch=3)
    fresh = TorchEngine(tiny_spec(), "cpu", OptimConfig())
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        load_checkpoint(path, fresh, rank=1, world_size=2)
    k0 = fresh.flat.slots[0].key
    assert torch.equal(fresh.flat.view(fresh.flat.params, k0), engs[0].flat.view(engs[0].flat.params, k0))
    assert any("replica" in str(x.message) for x in w)


def test_bench_arms_emit_the_same_config_keys():
    """The driver compares the two arms' `config` objects: both come from one function, arm-specific facts go to `extra`."""
    import importlib.util
    import os
    import types
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("bench_main", os.path.join(root, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    a = types.SimpleNamespace(model="enhanced_cnn", batch=64, aggregation_type="equal", shard_ratios=None)
    ours = bench.config_block(a, 8, 512, "allreduce", "global_epoch")
    ref = bench.config_block(a, 8, 512, "allreduce", "global_epoch")
    assert ours == ref and {"model", "global_batch", "parallelism", "topology", "sync", "l2"} <= set(ours)
    assert ours["parallelism"] == "dp8" and ours["global_batch"] == 512
    assert bench.DTYPE_HEADLINE == "tf32"
    assert bench.ratios_for(types.SimpleNamespace(shard_ratios=None), 4) == [1.0] * 4
    assert bench.ratios_for(types.SimpleNamespace(shard_ratios=[4, 2, 1, 1]), 4) == [4.0, 2.0, 1.0, 1.0]
