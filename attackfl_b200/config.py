"""Config loader: the reference's ``config.yaml`` schema, validated, with defaults and additive keys.

Reference parity: ``/root/reference/config.yaml:1-37`` read into module globals at
``server.py:55-89`` and ``client.py:42-48`` (no validation, KeyError on the README's own sample
config because ``hyper-detection`` / ``clip-grad-norm`` are missing there — SURVEY §2.4).  This loader
accepts both the shipped file and the README sample, tolerates and ignores ``rabbit.*`` (there is
no broker; the address is reused as the control-socket host when it is a loopback address) and adds
the switches of the quirk ledger under the same top-level sections.
"""
from __future__ import annotations

import dataclasses
import hashlib
import json
import warnings
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional

import yaml

SERVER_MODES = (
    "fedavg", "hyper", "FLTrust", "trimmed_mean", "shieldfl", "gmm", "krum", "median",
    "scionfl", "fltracer", "byzantine",
)
MODEL_NAMES = ("CNNModel", "RNNModel", "TransformerModel", "TransformerClassifier")
DATA_NAMES = ("ICU", "HAR")
ATTACK_MODES = ("Random", "Min-Max", "Min-Sum", "Opt-Fang", "LIE")


@dataclass
class HyperDetection:
    enable: bool = False
    cosine_search: int = 10          # reference reads it and then hard-codes 5 (server.py:133) - Q12
    n_components: int = 3
    eps: float = 0.007
    min_samples: int = 3
    start_round: int = 18            # hard-coded at server.py:513,524
    honour_cosine_search: bool = False  # False => deque(maxlen=5) like the reference


@dataclass
class ServerCfg:
    num_round: int = 30
    clients: int = 3
    mode: str = "hyper"
    hyper_detection: HyperDetection = field(default_factory=HyperDetection)
    model: str = "TransformerModel"
    data_name: str = "ICU"
    load_parameters: bool = False
    validation: bool = True
    num_data_range: List[int] = field(default_factory=lambda: [12000, 15000])
    genuine_rate: float = 0.5
    random_seed: int = 1
    # --- additive keys ---
    common_init: bool = True         # Q4: round-1 START carries one common init (False = reference)
    krum_f: int = 0                  # Q13
    krum_reference_f: bool = False   # f = int(n * genuine-rate) with the reference's negative-slice semantics (server.py:384)
    max_retries: int = 0             # Q10: 0 = retry forever like the reference
    trim_ratio: float = 0.1          # src/Utils.py:267
    hyper_update: str = "sequential"  # sequential (parity) | batched
    scionfl_keep: str = "reference"  # Q14: reference keeps the most dissimilar half | "similar"
    fltrust_root_size: int = 200     # server.py:292
    fltrust_root_batch: int = 100    # server.py:293


@dataclass
class LearningCfg:
    epoch: int = 5
    learning_rate: float = 0.004
    hyper_lr: float = 0.001
    momentum: float = 0.5
    batch_size: int = 128
    clip_grad_norm: float = 0.0
    # --- additive keys ---
    optimizer: str = "adam"          # Q3: adam (re-created per round) | sgd (honours momentum)
    clip_mode: str = "reference"     # Q2: reference = client clip is a no-op | proper
    dropout: bool = True             # False disables dropout in local training (for exact parity tests)


@dataclass
class AttackCfg:
    metric: str = "flat"             # flat | layer-fro | reference   
    search: str = "reference"        # reference (last tried gamma) | paper (gamma_succ)
    strict_reference: bool = False   # Q9: True reproduces LIE NaN at k=1
    gram_precision: str = "tf32"     # tf32 (fp32 rows as TF32 operands, no cast pass) | bf16 | bf16x3 | fp32


@dataclass
class RuntimeCfg:
    gpus: int = 1
    clients_per_gpu: int = 0         # 0 = ceil(clients / gpus)
    control_port: int = 56720
    control_host: str = "127.0.0.1"
    backend: str = "auto"            # auto | native | torch  (torch = pure-PyTorch fallback/oracle)
    cta_per_client: int = 0          # 0 = auto
    trainer: str = "auto"            # auto | ffma (op-program interpreter) | tcgen05 (tensor-core trainer, TransformerModel)
    collectives: str = "fused"       # fused (P2P/multimem kernels) | nccl (debug fall-back)
    index_rng: str = "perm"          # per-round sample indices: perm (keyed permutations, one launch, device-independent)
                                     # | torch (argsort of torch.rand keys on the run's device generator)
    synthetic_train_rows: int = 16384
    synthetic_test_rows: int = 4096
    metrics_file: str = "metrics.jsonl"
    nvtx: bool = True                # NVTX range per phase (train / craft / aggregate / validate ...)
    profile_round: int = 0           # >0: bracket that attempt with cudaProfilerStart/Stop (ncu/nsys --capture-range)
    round_timeout_s: float = 0.0     # >0: rank-health watchdog, aborts the process if one round exceeds it
    fault_inject: List[Dict[str, Any]] = field(default_factory=list)   # test-only: [{attempt, client, kind}]


@dataclass
class Config:
    name: str = "Federated Learning poisoning attack testbed"
    server: ServerCfg = field(default_factory=ServerCfg)
    learning: LearningCfg = field(default_factory=LearningCfg)
    attack: AttackCfg = field(default_factory=AttackCfg)
    runtime: RuntimeCfg = field(default_factory=RuntimeCfg)
    rabbit: Dict[str, Any] = field(default_factory=dict)
    log_path: str = "."

    def digest(self) -> str:
        blob = json.dumps(dataclasses.asdict(self), sort_keys=True, default=str).encode()
        return hashlib.sha256(blob).hexdigest()[:16]

    def to_dict(self) -> Dict[str, Any]:
        return dataclasses.asdict(self)


def _norm(key: str) -> str:
    return key.replace("-", "_")


def _fill(obj, raw: Dict[str, Any], section: str, renames: Optional[Dict[str, str]] = None):
    renames = renames or {}
    names = {f.name for f in dataclasses.fields(obj)}
    for k, v in (raw or {}).items():
        nk = renames.get(k, _norm(k))
        if nk not in names:
            warnings.warn(f"config: unknown key '{section}.{k}' ignored")
            continue
        cur = getattr(obj, nk)
        if dataclasses.is_dataclass(cur):
            _fill(cur, v, f"{section}.{k}")
        else:
            setattr(obj, nk, v)


def from_dict(raw: Dict[str, Any]) -> Config:
    cfg = Config()
    raw = dict(raw or {})
    cfg.name = raw.pop("name", cfg.name)
    cfg.log_path = raw.pop("log_path", cfg.log_path)
    cfg.rabbit = raw.pop("rabbit", {}) or {}
    srv = dict(raw.pop("server", {}) or {})
    # nested reference keys that do not map 1:1 on a dataclass field
    params = srv.pop("parameters", None)
    if params is not None:
        cfg.server.load_parameters = bool(params.get("load", False))
    dist = srv.pop("data-distribution", srv.pop("data_distribution", None))
    if dist is not None:
        rng = dist.get("num-data-range", dist.get("num_data_range"))
        if rng is not None:
            cfg.server.num_data_range = [int(rng[0]), int(rng[1])]
    _fill(cfg.server, srv, "server")
    _fill(cfg.learning, raw.pop("learning", {}), "learning")
    _fill(cfg.attack, raw.pop("attack", {}), "attack")
    _fill(cfg.runtime, raw.pop("runtime", {}), "runtime")
    for k in raw:
        warnings.warn(f"config: unknown top-level key '{k}' ignored")
    validate(cfg)
    return cfg


def validate(cfg: Config) -> None:
    s, l = cfg.server, cfg.learning
    if s.mode not in SERVER_MODES:
        # same wording as server.py:494 so scrapers of the reference's error keep working
        raise ValueError(f"Server mode '{s.mode}' is not valid.")
    if s.model not in MODEL_NAMES:
        raise ValueError(f"Model name '{s.model}' is not valid.")
    if s.data_name not in DATA_NAMES:
        raise ValueError(f"Data name '{s.data_name}' is not valid.")
    if s.clients < 1 or s.num_round < 1:
        raise ValueError("server.clients and server.num-round must be >= 1")
    lo, hi = s.num_data_range
    if lo < 1 or hi < lo:
        raise ValueError(f"invalid num-data-range {s.num_data_range}")
    if l.optimizer not in ("adam", "sgd"):
        raise ValueError(f"learning.optimizer '{l.optimizer}' is not valid (adam|sgd)")
    if l.clip_mode not in ("reference", "proper"):
        raise ValueError(f"learning.clip-mode '{l.clip_mode}' is not valid (reference|proper)")
    if cfg.attack.metric not in ("flat", "layer-fro", "reference"):
        raise ValueError(f"attack.metric '{cfg.attack.metric}' is not valid")
    if cfg.attack.search not in ("reference", "paper"):
        raise ValueError(f"attack.search '{cfg.attack.search}' is not valid")
    if s.hyper_update not in ("sequential", "batched"):
        raise ValueError(f"server.hyper-update '{s.hyper_update}' is not valid")
    if cfg.runtime.index_rng not in ("perm", "torch"):
        raise ValueError(f"runtime.index-rng '{cfg.runtime.index_rng}' is not valid (perm|torch)")
    if cfg.runtime.trainer not in ("auto", "ffma", "tcgen05"):
        raise ValueError(f"runtime.trainer '{cfg.runtime.trainer}' is not valid (auto|ffma|tcgen05)")
    for f in cfg.runtime.fault_inject:
        if f.get("kind") not in ("nan-loss", "nan-weights", "stall"):
            raise ValueError(f"runtime.fault-inject kind '{f.get('kind')}' is not valid (nan-loss|nan-weights|stall)")
    if l.batch_size < 1 or l.epoch < 1:
        raise ValueError("learning.batch-size and learning.epoch must be >= 1")


def load_config(path: str = "config.yaml") -> Config:
    """Reads ``config.yaml`` from the CWD like the reference (server.py:55, client.py:42)."""
    with open(path, "r") as fh:
        raw = yaml.safe_load(fh) or {}
    return from_dict(raw)
