"""bench.py bookkeeping that can be checked without a GPU: the headline configuration is valid for the engine, both arms
print the same ``config`` object, and the bench runs the reference's init semantics."""
import importlib
import inspect
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def test_headline_config_is_shared_and_valid():
    bench = importlib.import_module("bench")
    from attackfl_b200 import from_dict
    cfg = from_dict(bench.bench_config(8 * 2))
    assert cfg.server.model == "TransformerModel" and cfg.server.mode == "fedavg" and cfg.server.clients == 16
    assert cfg.server.common_init is False                      # every client starts from its own init, like the reference
    assert cfg.learning.epoch == 5 and cfg.learning.batch_size == 128
    assert tuple(cfg.server.num_data_range) == (12000, 15000)
    hc = bench.headline_config(16, 2)
    for key in ("model", "global_batch", "seq_len", "parallelism", "clients", "clients_per_gpu"):
        assert key in hc, key
    assert hc["clients"] == 16 and hc["clients_per_gpu"] == 8 and hc["global_batch"] == 16 * 128
    # the reference arm receives the very same function, so the two JSON lines carry an identical `config`
    src = inspect.getsource(bench.run_reference)
    assert "headline_config" in src
    from baseline import harness_ref
    assert "headline_config(C, world)" in inspect.getsource(harness_ref.run_bench)


def test_bench_defaults_finish_quickly():
    bench = importlib.import_module("bench")
    src = inspect.getsource(bench.main)
    assert 'default=1' in src and 'default=5' in src and 'default=3' in src      # --gpus 1, --steps 5, --warmup 3
