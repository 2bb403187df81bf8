"""`Trainer`: epochs over per-rank batch slices, versioned checkpoints in the reference's directory layout, exact resume.

Reference behaviour: `model.save()` writes `<saveDir>/<ms timestamp>/model.json + weights.bin` and moves the `current`
link, `setup()` resumes from the newest version (`/root/reference/src/server/models.ts:98-150`); the server loop applies
one micro-batch gradient per version (`/root/reference/src/server/asynchronousSGD_server.ts:95-108`).
"""
import json
import os
import socket

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from distriflow_b200 import ConfigError, load_layers_model
from distriflow_b200.parallel.engine import DataParallelEngine
from distriflow_b200.trainer import Trainer

CFG = {"optimizer": "adam", "learningRate": 0.01, "dtype": "fp32", "gradDtype": "fp32"}


def _net(seed=0):
    torch.manual_seed(seed)
    return nn.Sequential(nn.Linear(10, 12), nn.Tanh(), nn.Linear(12, 3))


def _loss(m, x, y):
    return F.cross_entropy(m(x), y)


def _data(n=70):
    g = torch.Generator().manual_seed(7)
    return torch.randn(n, 10, generator=g), torch.randint(0, 3, (n,), generator=g)


def _weights(eng):
    return torch.cat([t.reshape(-1) for t in eng.full_precision_weights()])


def test_trainer_runs_epochs_and_writes_reference_layout(tmp_path):
    x, y = _data()
    eng = DataParallelEngine(_net(), _loss, CFG, device=torch.device("cpu"))
    seen, ckpts = [], []
    tr = Trainer(eng, x, y, {"batchSize": 8, "epochs": 2, "saveDir": str(tmp_path / "saved-models"), "checkpointEvery": 5,
                             "logPath": str(tmp_path / "log.jsonl")})
    tr.on_step(lambda step, epoch, b, loss: seen.append((step, epoch, b)))
    tr.on_checkpoint(ckpts.append)
    out = tr.fit()
    assert tr.batches == 8                                   # 70 // 8, ragged tail dropped (smallLastBatch False)
    assert out["steps"] == 16 and out["finished"] and eng.version == 16
    assert seen[0] == (1, 0, 0) and seen[8] == (9, 1, 0) and seen[-1] == (16, 1, 7)
    assert len(ckpts) == 4                                   # steps 5, 10, 15 and the end
    save_dir = str(tmp_path / "saved-models")
    cur = os.path.join(save_dir, "current")
    assert os.path.islink(cur) and os.path.realpath(cur) == os.path.realpath(ckpts[-1])
    for f in ("model.json", "weights.bin", "engine_state.pt"):
        assert os.path.exists(os.path.join(cur, f))
    loaded = load_layers_model(cur, builder=lambda topo: _net(5))      # tf.js artifact holds the fp32 masters
    for a, b in zip(loaded.parameters(), eng.full_precision_weights()):
        assert torch.allclose(a.detach(), b, atol=1e-7)
    rows = [json.loads(line) for line in open(tmp_path / "log.jsonl")]
    assert [r["step"] for r in rows] == list(range(1, 17)) and rows[0]["lr"] == 0.01


def test_trainer_interrupted_run_resumes_bit_for_bit(tmp_path):
    x, y = _data()
    conf = {"batchSize": 8, "epochs": 3, "checkpointEvery": 4}
    ref_eng = DataParallelEngine(_net(), _loss, CFG, device=torch.device("cpu"))
    Trainer(ref_eng, x, y, conf).fit()
    d = str(tmp_path / "ck")
    eng1 = DataParallelEngine(_net(), _loss, CFG, device=torch.device("cpu"))
    out1 = Trainer(eng1, x, y, {**conf, "saveDir": d}).fit(max_steps=10)          # "crash" after step 10
    assert out1["steps"] == 10 and not out1["finished"]
    eng2 = DataParallelEngine(_net(123), _loss, CFG, device=torch.device("cpu"))   # fresh process, different init
    out2 = Trainer(eng2, x, y, {**conf, "saveDir": d}).fit()
    assert out2["resumed_from"] == 10 and out2["steps"] == 24 and out2["finished"]
    assert torch.equal(_weights(eng2), _weights(ref_eng))
    # resume=False starts over from the engine's own weights
    eng3 = DataParallelEngine(_net(), _loss, CFG, device=torch.device("cpu"))
    out3 = Trainer(eng3, x, y, {**conf, "saveDir": d, "resume": False, "epochs": 1}).fit()
    assert out3["resumed_from"] == 0 and out3["steps"] == 8


def test_trainer_keep_versions_schedule_and_config_errors(tmp_path):
    x, y = _data()
    eng = DataParallelEngine(_net(), _loss, {**CFG, "optimizer": "sgd"}, device=torch.device("cpu"))
    lrs = []
    tr = Trainer(eng, x, y, {"batchSize": 8, "epochs": 1, "saveDir": str(tmp_path), "checkpointEvery": 2, "keepVersions": 2},
                 lr_schedule=lambda step: 0.1 / (1 + step))
    tr.on_step(lambda *a: lrs.append(eng.lr))
    tr.fit()
    assert lrs[0] == 0.1 and abs(lrs[-1] - 0.1 / 8) < 1e-12
    names = [n for n in os.listdir(tmp_path) if n != "current"]
    assert len(names) == 2                                   # older version directories were pruned
    with pytest.raises(ConfigError, match="Unrecognized key"):
        Trainer(eng, x, y, {"batchsize": 8})
    with pytest.raises(ValueError):
        Trainer(eng, x[:3], y[:3], {"batchSize": 8})


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_dir):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    x, y = _data(64)
    eng = DataParallelEngine(_net(rank), _loss, {**CFG, "optimizer": "momentum", "backend": "gloo"}, rank=rank, world=world,
                             device=torch.device("cpu"))
    out = Trainer(eng, x, y, {"batchSize": 4, "epochs": 2, "saveDir": os.path.join(out_dir, "ck"), "checkpointEvery": 5}).fit()
    if rank == 0:
        torch.save({"w": _weights(eng), "out": out}, os.path.join(out_dir, "res.pt"))
    else:
        _weights(eng)                                        # collective
    dist.destroy_process_group()


def test_trainer_two_ranks_slice_the_global_batch(tmp_path):
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    res = torch.load(os.path.join(str(tmp_path), "res.pt"), weights_only=False)
    assert res["out"]["steps"] == 16 and res["out"]["finished"]            # 64 // (4 * 2) = 8 global batches per epoch
    solo = DataParallelEngine(_net(0), _loss, {**CFG, "optimizer": "momentum"}, device=torch.device("cpu"))
    Trainer(solo, *_data(64), {"batchSize": 8, "epochs": 2}).fit()           # one rank on the same global batches
    assert torch.allclose(_weights(solo), res["w"], atol=1e-5)
    assert os.path.exists(os.path.join(str(tmp_path), "ck", "current", "engine_state.pt"))


def test_command_line_trainer_on_cpu(tmp_path, capsys):
    from distriflow_b200 import train
    rc = train.main(["--model", "mnist-mlp", "--device", "cpu", "--epochs", "2", "--batch", "16", "--examples", "64",
                     "--optimizer", "momentum", "--lr", "0.0005", "--warmup-steps", "3", "--save-dir", str(tmp_path / "ck"),
                     "--checkpoint-every", "3", "--log", str(tmp_path / "log.jsonl"), "--trace"])
    assert rc == 0
    out = json.loads(capsys.readouterr().out.strip().splitlines()[-1])
    assert out["steps"] == 8 and out["finished"] and out["version"] == 8 and out["device"] == "cpu"
    assert set(out["phases_ms"]) == {"forward", "backward", "collective_tail"}
    assert os.path.exists(tmp_path / "ck" / "current" / "model.json")
    rows = [json.loads(line) for line in open(tmp_path / "log.jsonl")]
    assert abs(rows[0]["lr"] - 0.0005 / 3) < 1e-12 and rows[-1]["lr"] == 0.0005
    # a second invocation finds the finished run and has nothing left to do
    rc = train.main(["--model", "mnist-mlp", "--device", "cpu", "--epochs", "2", "--batch", "16", "--examples", "64",
                     "--optimizer", "momentum", "--save-dir", str(tmp_path / "ck")])
    out = json.loads(capsys.readouterr().out.strip().splitlines()[-1])
    assert rc == 0 and out["resumed_from"] == 8 and out["steps_this_call"] == 0


def _shrink_worker(rank, world, port, out_dir):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    x, y = _data(64)
    eng = DataParallelEngine(_net(0), _loss, {**CFG, "backend": "gloo"}, rank=rank, world=world, device=torch.device("cpu"))
    Trainer(eng, x, y, {"batchSize": 4, "epochs": 2, "saveDir": os.path.join(out_dir, "ck"), "checkpointEvery": 3}).fit(max_steps=6)
    dist.destroy_process_group()


def test_run_started_on_two_ranks_resumes_on_one(tmp_path):
    """Restart-based recovery after a rank is lost: the checkpoint (fp32 masters, Adam moments, step) does not depend on the
    world size, so the job continues on the surviving rank with the global batch kept (4 x 2 ranks -> 8 x 1) and ends where
    an uninterrupted single-rank run ends."""
    import torch.multiprocessing as mp
    mp.spawn(_shrink_worker, args=(2, _free_port(), str(tmp_path)), nprocs=2, join=True)
    x, y = _data(64)
    eng = DataParallelEngine(_net(5), _loss, CFG, device=torch.device("cpu"))           # different init: the state replaces it
    out = Trainer(eng, x, y, {"batchSize": 8, "epochs": 2, "saveDir": os.path.join(str(tmp_path), "ck"), "checkpointEvery": 3}).fit()
    assert out["resumed_from"] == 6 and out["finished"] and out["steps"] == 16
    solo = DataParallelEngine(_net(0), _loss, CFG, device=torch.device("cpu"))
    Trainer(solo, x, y, {"batchSize": 8, "epochs": 2}).fit()
    assert torch.allclose(_weights(solo), _weights(eng), atol=1e-5)


def test_no_checkpoint_is_written_after_a_peer_was_lost(tmp_path):
    """`Trainer.save` asks the engine first: a step taken after an in-kernel barrier gave up must not become a resume point,
    the previous version stays `current` and the error reaches the caller (the CLI turns it into exit code 75)."""
    from distriflow_b200.parallel.engine import PeerLostError
    x, y = _data(64)
    eng = DataParallelEngine(_net(0), _loss, CFG, device=torch.device("cpu"))
    tr = Trainer(eng, x, y, {"batchSize": 8, "epochs": 2, "saveDir": str(tmp_path / "ck"), "checkpointEvery": 4})
    tr.fit(max_steps=4)
    good = os.path.realpath(os.path.join(str(tmp_path / "ck"), "current"))
    eng.peer_failure = lambda _side=False: (1, 42)                       # what a timed-out barrier leaves in the signal pad
    with pytest.raises(PeerLostError, match="rank 1 never reached barrier generation 42"):
        tr.fit(max_steps=4)
    assert os.path.realpath(os.path.join(str(tmp_path / "ck"), "current")) == good
    assert len([d for d in os.listdir(str(tmp_path / "ck")) if d != "current"]) == 1
