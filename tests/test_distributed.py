"""SPMD engine under torch.distributed: gloo / world_size 2 on CPU here, NCCL + fused P2P kernels on GPUs."""
import json
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _launch(tmp_path, world, mode, attack="", env_extra=None, timeout=600):
    out = str(tmp_path / "res")
    port = 29600 + (os.getpid() % 300)
    env = dict(os.environ, PYTHONPATH=ROOT, ATTACKFL_QUIET="1")
    env.update(env_extra or {})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "dist_worker.py"), out, mode, attack]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=timeout)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    return [json.load(open(f"{out}.rank{k}.json")) for k in range(world)]


@pytest.mark.parametrize("mode,attack", [("fedavg", "LIE"), ("fedavg", "Min-Max"), ("hyper", ""), ("median", "Random")])
def test_two_ranks_gloo_agree(tmp_path, mode, attack):
    recs = _launch(tmp_path, 2, mode, attack)
    a, b = recs
    assert a["world"] == 2 and a["local"] == [0, 1, 2] and b["local"] == [3, 4, 5]
    assert a["ok"] == b["ok"] == [True, True]
    assert a["checksum"] == pytest.approx(b["checksum"], rel=1e-6)          # replicated state stays in sync
    assert a["metric"] == pytest.approx(b["metric"], rel=1e-6)
    if attack:
        assert a["attackers"][1] == [5]
    assert a["metric"][-1] > (0.6 if mode != "hyper" else 0.45)      # the hyper-network needs more rounds to learn


@pytest.mark.gpu
@pytest.mark.parametrize("attack", ["LIE", "Min-Max", "Opt-Fang"])
def test_two_gpus_fused_matches_nccl_path(tmp_path, attack):
    """K1-fused (+LIE moments) and the K-slice Gram attack over peer memory vs the NCCL gather path."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    out = {}
    for collectives in ("nccl", "fused"):
        d = tmp_path / collectives
        d.mkdir()
        recs = _launch(d, 2, "fedavg", attack, {"AFL_TEST_DEVICE": "cuda", "AFL_TEST_COLLECTIVES": collectives})
        a, b = recs
        assert a["ok"] == b["ok"] == [True, True]
        assert a["checksum"] == pytest.approx(b["checksum"], rel=1e-5)
        assert a["fused"] == (collectives == "fused")
        assert a["attackers"][1] == [5]
        out[collectives] = a
    # same seeds, same algorithm: the two data paths agree up to fp summation order / tf32 Gram rounding
    assert out["fused"]["checksum"] == pytest.approx(out["nccl"]["checksum"], rel=2e-3, abs=2e-2)
    assert out["fused"]["metric"][-1] == pytest.approx(out["nccl"]["metric"][-1], abs=0.02)


@pytest.mark.gpu
def test_two_gpus_fused_with_random_rank_delays(tmp_path):
    """Cross-GPU flag protocol under skew: every fused collective is preceded by a random per-rank device delay; the
    result must equal the undelayed run bit for bit (same kernels, same order of operations)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    out = {}
    for tag, env in (("plain", {}), ("jitter", {"AFL_TEST_JITTER": "1"})):
        d = tmp_path / tag
        d.mkdir()
        recs = _launch(d, 2, "fedavg", "Min-Max", {"AFL_TEST_DEVICE": "cuda", "AFL_TEST_COLLECTIVES": "fused",
                                                    "AFL_TEST_ROUNDS": "4", **env})
        a, b = recs
        assert a["ok"] == b["ok"] == [True] * 4 and a["fused"]
        assert a["checksum"] == pytest.approx(b["checksum"], rel=1e-6)
        out[tag] = a
    assert out["jitter"]["checksum"] == pytest.approx(out["plain"]["checksum"], rel=1e-6)
    assert out["jitter"]["metric"] == pytest.approx(out["plain"]["metric"], abs=1e-6)
