"""End-to-end rounds through the public API (``Federation``): every mode, every attack, failure handling,
checkpoint / resume.  CPU twin by default, GPU variants marked ``gpu``."""
import os

import pytest
import torch

from attackfl_b200 import from_dict, models
from attackfl_b200.engine import ClientSpec, Federation

DEVICES = ["cpu", pytest.param("cuda", marks=pytest.mark.gpu)]


def _cfg(mode, model, n, rounds=2, **extra):
    d = {"server": {"num-round": rounds, "clients": n, "mode": mode, "model": model, "data-name": "ICU",
                    "data-distribution": {"num-data-range": [300, 400]}, "random-seed": 1},
         "learning": {"epoch": 2, "batch-size": 32, "learning-rate": 0.004, "hyper-lr": 0.001, "clip-grad-norm": 1.0},
         "runtime": {"synthetic_train_rows": 1024, "synthetic_test_rows": 512}}
    for k, v in extra.items():
        d.setdefault(k, {}).update(v)
    return from_dict(d)


def _specs(n, **attackers):
    sp = [ClientSpec(i) for i in range(n)]
    for i, (mode, rnd, args) in attackers.items():
        sp[int(i[1:])] = ClientSpec(int(i[1:]), True, mode, rnd, args)
    return sp


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("model", ["TransformerModel", "RNNModel", "CNNModel"])
def test_fedavg_learns(tmp_path, device, model):
    fed = Federation(_cfg("fedavg", model, 3, rounds=3), _specs(3), device=device, workdir=str(tmp_path), verbose=False)
    res = fed.run()
    assert [r.ok for r in res] == [True] * 3
    assert res[-1].metric > 0.7
    sd = torch.load(tmp_path / f"{model}.pth", weights_only=True)
    models.build(model).load_state_dict(sd)                     # reference-compatible .pth
    assert os.path.exists(tmp_path / "metrics.jsonl")


@pytest.mark.parametrize("device", DEVICES)
def test_attacks_activate_from_their_round(tmp_path, device):
    sp = _specs(6, c3=("Random", 2, [0.5]), c4=("LIE", 2, [0.74]), c5=("Min-Max", 3, None))
    fed = Federation(_cfg("fedavg", "RNNModel", 6, rounds=3), sp, device=device, workdir=str(tmp_path), verbose=False)
    res = fed.run()
    assert res[0].attackers == [] and res[1].attackers == [3, 4] and res[2].attackers == [3, 4, 5]
    assert all(r.ok for r in res)


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("mode", ["Min-Sum", "Opt-Fang"])
def test_gram_attacks_run(tmp_path, device, mode):
    sp = _specs(5, c4=(mode, 2, None))
    fed = Federation(_cfg("fedavg", "TransformerModel", 5, rounds=2), sp, device=device, workdir=str(tmp_path), verbose=False)
    res = fed.run()
    assert res[1].attackers == [4] and all(r.ok for r in res)


@pytest.mark.parametrize("device", DEVICES)
def test_hyper_mode_and_resume(tmp_path, device):
    cfg = _cfg("hyper", "TransformerModel", 3, rounds=2)
    fed = Federation(cfg, _specs(3), device=device, workdir=str(tmp_path), verbose=False)
    res = fed.run()
    assert all(r.ok for r in res)
    p = tmp_path / "TransformerModel_hyper_3.pth"
    sd = torch.load(p, weights_only=True)
    assert len(sd) == 121 and "hyper_layers.fc1__weight.weight" in sd
    h = models.HyperNetwork(models.TransformerModel(), 3, 8, 100, False, 2)
    h.load_state_dict(sd)
    cfg2 = _cfg("hyper", "TransformerModel", 3, rounds=1, server={"parameters": {"load": True}})
    fed2 = Federation(cfg2, _specs(3), device=device, workdir=str(tmp_path), verbose=False)
    got = fed2.hyper.export_state_dict()
    assert all(torch.allclose(got[k], sd[k]) for k in sd)       # real resume


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("mode", ["trimmed_mean", "median", "krum", "shieldfl", "scionfl", "FLTrust", "fltracer",
                                  "byzantine", "gmm"])
def test_defences_run(tmp_path, device, mode):
    fed = Federation(_cfg(mode, "TransformerModel", 5), _specs(5), device=device, workdir=str(tmp_path), verbose=False)
    res = fed.run()
    assert all(r.ok for r in res) and res[-1].metric > 0.6


def test_nan_update_fails_round_and_retries(tmp_path):
    """A NaN model => validation fails => round discarded and repeated (server.py:546-563)."""
    fed = Federation(_cfg("fedavg", "TransformerModel", 3, rounds=2), _specs(3), device="cpu", workdir=str(tmp_path),
                     verbose=False)
    r1 = fed.run_round()
    assert r1.ok and fed.remaining == 1
    good = fed.global_row.clone()
    orig = fed._aggregate

    def poisoned(rows_all, reporting):
        info = orig(rows_all, reporting)
        fed.global_row.fill_(float("nan"))
        return info
    fed._aggregate = poisoned
    r2 = fed.run_round()
    assert not r2.ok and fed.remaining == 1 and fed.failures_in_a_row == 1
    assert torch.equal(fed.global_row, good)                    # poisoned global is not propagated (Q10)
    fed._aggregate = orig
    r3 = fed.run_round()
    assert r3.ok and fed.remaining == 0


def test_torch_backend_matches_native_without_dropout(tmp_path):
    """The pure-PyTorch fall-back trainer and the native kernel implement the same round."""
    out = []
    for backend in ("native", "torch"):
        cfg = _cfg("fedavg", "TransformerModel", 2, rounds=1, learning={"dropout": False}, runtime={"backend": backend})
        fed = Federation(cfg, _specs(2), device="cpu", workdir=str(tmp_path), verbose=False, save=False)
        fed.run()
        out.append(fed.global_row.clone())
    assert torch.allclose(out[0], out[1], atol=5e-4, rtol=1e-2)


def test_har_round(tmp_path):
    d = {"server": {"num-round": 1, "clients": 2, "mode": "fedavg", "model": "TransformerClassifier", "data-name": "HAR",
                    "data-distribution": {"num-data-range": [16, 24]}, "random-seed": 1},
         "learning": {"epoch": 1, "batch-size": 8, "learning-rate": 0.001},
         "runtime": {"synthetic_train_rows": 32, "synthetic_test_rows": 16}}
    fed = Federation(from_dict(d), _specs(2), device="cpu", workdir=str(tmp_path), verbose=False)
    res = fed.run()
    assert res[0].ok and 0.0 <= res[0].metric <= 1.0


def test_fault_injection_retries_the_round(tmp_path):
    """a NaN loss (client.py:100-102) or a diverged update discards the round and re-issues it;
    ``round`` only advances on success (server.py:555)."""
    from attackfl_b200.utils.timers import Watchdog
    cfg = _cfg("fedavg", "TransformerModel", 3, rounds=3)
    cfg.runtime.fault_inject = [{"attempt": 2, "client": 1, "kind": "nan-loss"},
                                {"attempt": 4, "client": 0, "kind": "nan-weights"}]
    fed = Federation(cfg, [ClientSpec(i) for i in range(3)], device="cpu", workdir=str(tmp_path), verbose=False)
    res = fed.run()
    assert [r.ok for r in res] == [True, False, True, False, True]
    assert [r.round for r in res] == [1, 2, 2, 3, 3]
    assert fed.remaining == 0
    fired = []
    wd = Watchdog(0.05, on_expire=fired.append)
    wd.arm("round attempt 1")
    import time
    time.sleep(0.2)
    assert fired == ["round attempt 1"] and wd.expired
    wd2 = Watchdog(0.05, on_expire=fired.append)
    wd2.arm("x"); wd2.disarm(); time.sleep(0.1)
    assert len(fired) == 1


def test_hyper_failed_round_keeps_the_last_successful_hypernet(tmp_path):
    """A round that fails BEFORE aggregation (client NaN loss) must leave the hyper-network exactly as the previous,
    successful round left it: the rollback snapshot belongs to the attempt that took it."""
    cfg = _cfg("hyper", "TransformerModel", 3, rounds=2)
    cfg.runtime.fault_inject = [{"attempt": 2, "client": 1, "kind": "nan-loss"}]
    fed = Federation(cfg, [ClientSpec(i) for i in range(3)], device="cpu", workdir=str(tmp_path), verbose=False)
    r1 = fed.run_round()
    assert r1.ok
    after_first = {k: v.clone() for k, v in fed.state_dict().items()}
    r2 = fed.run_round()
    assert not r2.ok and r2.round == 2
    after_failed = fed.state_dict()
    for k, v in after_first.items():
        assert torch.equal(v, after_failed[k]), k
    r3 = fed.run_round()
    assert r3.ok and r3.round == 2


def test_client_device_cpu_is_accepted_on_a_cpu_server(tmp_path):
    """``client.py --device cpu`` (reference client.py:50-61): on a CPU server it is the server's own path."""
    cfg = _cfg("fedavg", "TransformerModel", 2, rounds=1)
    fed = Federation(cfg, [ClientSpec(0, device="cpu"), ClientSpec(1)], device="cpu", workdir=str(tmp_path), verbose=False)
    assert fed._cpu_slots() == []
    assert fed.run()[0].ok


@pytest.mark.gpu
def test_cpu_client_slot_on_a_gpu_server(tmp_path):
    """A client registered with ``--device cpu`` on a CUDA server trains its slot through the CPU twin and its row is
    copied into the device arena (reference client.py:50-61: every client picks its own device)."""
    cfg = _cfg("fedavg", "TransformerModel", 3, rounds=2)
    fed = Federation(cfg, [ClientSpec(0, device="cpu"), ClientSpec(1), ClientSpec(2)], device="cuda", workdir=str(tmp_path),
                     verbose=False, save=False)
    assert fed._cpu_slots() == [0]
    res = fed.run()
    assert [r.ok for r in res] == [True, True] and res[-1].metric > 0.6
    assert all(x == x and x > 0 for x in res[-1].losses)          # the CPU slot reported a loss like the GPU slots


@pytest.mark.gpu
@pytest.mark.parametrize("trainer,clients,want", [("auto", 3, "ffma"), ("tcgen05", 3, "tcgen05"), ("ffma", 12, "ffma"),
                                                  ("auto", 12, "tcgen05")])
def test_trainer_selection_and_both_trainers_learn(tmp_path, trainer, clients, want):
    """``runtime.trainer``: the interpreter while it can give every client 16 blocks, the tensor-core trainer beyond
    (or when forced); both must learn the task through the public API."""
    cfg = _cfg("fedavg", "TransformerModel", clients, rounds=2, runtime={"trainer": trainer},
               learning={"batch-size": 128})
    fed = Federation(cfg, _specs(clients), device="cuda", workdir=str(tmp_path), verbose=False, save=False)
    assert fed.trainer == want
    res = fed.run()
    assert all(r.ok for r in res) and res[-1].metric > 0.6


def test_trainer_key_is_validated():
    with pytest.raises(Exception):
        _cfg("fedavg", "TransformerModel", 2, runtime={"trainer": "wgmma"})


@pytest.mark.parametrize("rng", ["perm", "torch"])
def test_both_index_generators_train(tmp_path, rng):
    """``runtime.index-rng``: keyed permutations (default, one native launch) or argsort of torch.rand keys."""
    cfg = _cfg("fedavg", "TransformerModel", 3, rounds=2, runtime={"index-rng": rng})
    fed = Federation(cfg, _specs(3), device="cpu", workdir=str(tmp_path), verbose=False, save=False)
    idx = fed._round_indices([300, 0, 350])
    assert idx.shape[0] == 3 and idx.dtype == torch.int32 and int(idx[0, 0, :300].min()) >= 0
    assert len(set(idx[2, 1, :350].tolist())) == 350 and int(idx[0, 0, 300:].max()) == -1
    res = fed.run()
    assert all(r.ok for r in res) and res[-1].metric > 0.6
