"""CLI shims + coordinator + control socket: BASELINE config 1 (server.py + 3 x client.py --device cpu)."""
import os
import subprocess
import sys
import time

import pytest
import torch
import yaml

from attackfl_b200 import from_dict, load_config
from attackfl_b200.cli import client_parser, parse_bool, validate_client_args

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_attack_flag_parsing_q11():
    p = client_parser()
    assert p.parse_args(["--attack", "True"]).attack is True
    assert p.parse_args(["--attack", "False"]).attack is False       # the reference would say True here
    assert p.parse_args(["--attack", "banana"]).attack is True
    assert p.parse_args([]).attack is False
    a = p.parse_args(["--attack", "True", "--attack_mode", "LIE", "--attack_round", "1", "--attack_args", "0.74"])
    validate_client_args(a)
    assert a.attack_args == [0.74]
    with pytest.raises(SystemExit):
        validate_client_args(p.parse_args(["--attack", "True", "--attack_mode", "LIE"]))
    with pytest.raises(SystemExit):
        p.parse_args(["--attack_mode", "Nope"])
    assert parse_bool("0") is False and parse_bool(1) is True


def test_config_accepts_shipped_and_readme_variants(tmp_path):
    shipped = yaml.safe_load(open(os.path.join(ROOT, "baseline", "_ref", "config.yaml"))) if os.path.exists(
        os.path.join(ROOT, "baseline", "_ref", "config.yaml")) else yaml.safe_load(open(os.path.join(ROOT, "config.yaml")))
    cfg = from_dict(shipped)
    assert cfg.server.mode == "hyper" and cfg.learning.clip_grad_norm == 1.0 and cfg.server.num_data_range == [12000, 15000]
    readme = dict(shipped)
    readme["server"] = {k: v for k, v in shipped["server"].items() if k != "hyper-detection"}
    readme["learning"] = {k: v for k, v in shipped["learning"].items() if k != "clip-grad-norm"}
    cfg2 = from_dict(readme)                                          # the reference KeyErrors on this one
    assert cfg2.server.hyper_detection.enable is False and cfg2.learning.clip_grad_norm == 0.0
    with pytest.raises(ValueError, match="Server mode 'nope' is not valid."):
        from_dict({"server": {"mode": "nope"}})
    assert load_config(os.path.join(ROOT, "config.yaml")).server.clients == 3


@pytest.mark.parametrize("attack", [False, True])
def test_baseline_config1_server_plus_three_clients_cpu(tmp_path, ref, attack):
    port = 56000 + (os.getpid() % 500) * 2 + int(attack)
    cfg = {"name": "t", "server": {"num-round": 2, "clients": 3, "mode": "fedavg", "model": "RNNModel", "data-name": "ICU",
                                   "parameters": {"load": False}, "validation": True, "genuine-rate": 0.5, "random-seed": 1,
                                   "data-distribution": {"num-data-range": [200, 260]},
                                   "hyper-detection": {"enable": False, "cosine-search": 10, "n_components": 3,
                                                       "eps": 0.007, "min_samples": 3}},
           "rabbit": {"address": "127.0.0.1", "username": "admin", "password": "admin"}, "log_path": ".",
           "learning": {"epoch": 1, "learning-rate": 0.004, "hyper-lr": 0.001, "momentum": 0.5, "batch-size": 64,
                        "clip-grad-norm": 1.0},
           "runtime": {"control_port": port, "synthetic_train_rows": 512, "synthetic_test_rows": 256}}
    with open(tmp_path / "config.yaml", "w") as fh:
        yaml.safe_dump(cfg, fh)
    env = dict(os.environ, PYTHONPATH=ROOT, ATTACKFL_QUIET="0")
    srv = subprocess.Popen([sys.executable, os.path.join(ROOT, "server.py"), "--device", "cpu", "--register-timeout", "120"],
                           cwd=tmp_path, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    time.sleep(1.0)
    clients = []
    for i in range(3):
        cmd = [sys.executable, os.path.join(ROOT, "client.py"), "--device", "cpu"]
        if attack and i == 2:
            cmd += ["--attack", "True", "--attack_mode", "LIE", "--attack_round", "2", "--attack_args", "0.74"]
        clients.append(subprocess.Popen(cmd, cwd=tmp_path, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    out, _ = srv.communicate(timeout=300)
    assert srv.returncode == 0, out
    couts = [c.communicate(timeout=60)[0] for c in clients]
    assert all(c.returncode == 0 for c in clients), couts
    assert "Start training round 2" in out and "Ok, ready!" in out
    assert all("Stop training!" in c for c in couts)
    if attack:
        assert any("start attacking LIE" in c for c in couts)
    log = open(tmp_path / "app.log").read()
    assert "### Application start ###" in log and "Active with 3 client: [0, 1, 2]" in log and log.count("ROC_AUC:") == 2
    sd = torch.load(tmp_path / "RNNModel.pth", weights_only=True)
    ref.model.RNNModel().load_state_dict(sd)                          # the reference's module loads our checkpoint
