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
    recs = [json.load(open(f"{out}.rank{k}.json")) for k in range(world)]
    for k, r in enumerate(recs):
        r["vector"] = torch.load(f"{out}.rank{k}.pt")
    return recs


@pytest.mark.parametrize("mode,attack", [("fedavg", "LIE"), ("fedavg", "Min-Max"), ("hyper", ""), ("median", "Random")])
def test_two_ranks_gloo_agree(tmp_path, mode, attack):
    recs = _launch(tmp_path, 2, mode, attack)
    a, b = recs
    assert a["world"] == 2 and a["local"] == [0, 1, 2] and b["local"] == [3, 4, 5]
    assert a["ok"] == b["ok"] == [True, True]
    assert float((a["vector"] - b["vector"]).abs().max()) <= 1e-6 * max(1.0, float(a["vector"].abs().max()))   # replicated state stays in sync
    assert a["metric"] == pytest.approx(b["metric"], rel=1e-6)
    if attack:
        assert a["attackers"][1] == [5]
    assert a["metric"][-1] > (0.6 if mode != "hyper" else 0.45)      # the hyper-network needs more rounds to learn


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("attack", ["LIE", "Min-Max", "Opt-Fang"])
def test_gpus_fused_matches_nccl_path(tmp_path, attack, world):
    """K1-fused (+LIE moments) and the K-slice Gram attack over peer memory vs the NCCL gather path, at 2, 4 and 8 ranks
    (from 4 ranks on the reduction is the NVLS multimem path).  The FULL global vectors are compared."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    out = {}
    for collectives in ("nccl", "fused"):
        d = tmp_path / collectives
        d.mkdir()
        recs = _launch(d, world, "fedavg", attack, {"AFL_TEST_DEVICE": "cuda", "AFL_TEST_COLLECTIVES": collectives, "ATTACKFL_DEBUG_OWNERSHIP": "1",
                                                    "AFL_TEST_CLIENTS": "8"})
        a = recs[0]
        for b in recs[1:]:
            assert a["ok"] == b["ok"] == [True, True]
            assert torch.equal(a["vector"], b["vector"]) or float((a["vector"] - b["vector"]).abs().max()) < 1e-6
        assert a["fused"] == (collectives == "fused")
        assert a["attackers"][1] == [7]
        out[collectives] = a
    # same seeds, same algorithm: the two data paths agree up to fp summation order / tf32 Gram rounding
    diff = float((out["fused"]["vector"] - out["nccl"]["vector"]).abs().max())
    scale = float(out["nccl"]["vector"].abs().max())
    assert diff <= 2e-3 * scale, (diff, scale)
    assert out["fused"]["metric"][-1] == pytest.approx(out["nccl"]["metric"][-1], abs=0.02)


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("mode,attack", [("median", "Random"), ("trimmed_mean", ""), ("krum", "Min-Max"), ("shieldfl", ""),
                                         ("FLTrust", "Random"), ("scionfl", "Random")])
def test_gpus_slice_owned_defences_match_gathered_path(tmp_path, mode, attack, world):
    """Defences without gathering rows (K6 over peer rows, K-slice Gram + K1-fused weights, ScionFL on 1-bit codes read
    in place) vs all-gather + replicated aggregation; full global vectors compared."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    out = {}
    for collectives in ("nccl", "fused"):
        d = tmp_path / collectives
        d.mkdir()
        recs = _launch(d, world, mode, attack, {"AFL_TEST_DEVICE": "cuda", "AFL_TEST_COLLECTIVES": collectives, "ATTACKFL_DEBUG_OWNERSHIP": "1",
                                                "AFL_TEST_CLIENTS": "8", "AFL_TEST_ROUNDS": "3"})
        a = recs[0]
        assert a["ok"] == [True] * 3
        for b in recs[1:]:
            assert float((a["vector"] - b["vector"]).abs().max()) < 1e-6
        out[collectives] = a
    diff = float((out["fused"]["vector"] - out["nccl"]["vector"]).abs().max())
    scale = float(out["nccl"]["vector"].abs().max())
    if mode == "scionfl":
        # the filter decision is discrete: both paths must keep the same clients in every round; the kept rows are then
        # averaged in a different summation order and three rounds of Adam training amplify that rounding difference
        assert out["fused"]["kept"] == out["nccl"]["kept"] and all(k for k in out["fused"]["kept"])
        assert diff <= 5e-3 * scale, (mode, diff, scale)
    else:
        assert diff <= 2e-3 * scale, (mode, diff, scale)


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 8])
def test_gpus_sharded_hyper_matches_gathered_path(tmp_path, world):
    """Hyper mode: row-sharded head + persistent per-rank kernel reading client rows in place (collectives: fused) vs the
    all-gather + replicated per-client steps (collectives: nccl).  Full head-bias vectors compared."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    out = {}
    for collectives in ("nccl", "fused"):
        d = tmp_path / collectives
        d.mkdir()
        recs = _launch(d, world, "hyper", "", {"AFL_TEST_DEVICE": "cuda", "AFL_TEST_COLLECTIVES": collectives, "ATTACKFL_DEBUG_OWNERSHIP": "1",
                                               "AFL_TEST_CLIENTS": "8", "AFL_TEST_ROUNDS": "3"})
        a = recs[0]
        assert a["ok"] == [True] * 3
        out[collectives] = a
        if collectives == "fused":       # W_h / b_h / both moments are stored as 1 / world row shards on every rank
            assert all(r["head_rows"] == r["d_pad"] // world for r in recs), [(r["head_rows"], r["d_pad"]) for r in recs]
        else:
            assert all(r["head_rows"] == r["d_pad"] for r in recs)
    # Adam normalises gradients: a coordinate whose gradient is rounding noise can move by +-lr per step in either run, so
    # require that almost all coordinates agree tightly and that none is off by more than a few steps of hyper-lr = 1e-3
    d = (out["fused"]["vector"] - out["nccl"]["vector"]).abs()
    assert float((d > 1e-5).float().mean()) < 0.05 and float(d.max()) < 24 * 1e-3, (float(d.max()), float((d > 1e-5).float().mean()))
    assert out["fused"]["metric"] == pytest.approx(out["nccl"]["metric"], abs=5e-3)


@pytest.mark.gpu
@pytest.mark.parametrize("world", [2, 4, 8])
def test_fused_kernel_against_torch_and_flag_stress(tmp_path, world):
    """afl_fedavg_fused itself: one-chunk and multi-chunk pipelines, P2P and NVLS, 1 / 3 rows per rank, LIE moment sets,
    full-vector comparison with torch on gathered rows; then 1,000 epochs with random per-rank delays."""
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    out = str(tmp_path / "kernel.json")
    port = 29900 + (os.getpid() % 90)
    env = dict(os.environ, PYTHONPATH=ROOT)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "dist_kernel_worker.py"), out, "1000"]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    rec = json.load(open(out))
    assert rec["world"] == world and len(rec["cases"]) >= 9
    if world >= 4:
        assert any(c["nvls"] for c in rec["cases"]), "NVLS path not exercised"
    for c in rec["cases"]:
        # the LIE rows contain a standard deviation from shifted sums: same value as torch.std up to fp32 summation order
        assert c["max_abs_err"] < (2e-4 if c["A"] else 2e-5), c
    assert rec["stress"]["max_abs_err"] < 2e-4 and rec["stress"]["multi_chunk_max_abs_err"] < 2e-5, rec["stress"]


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
