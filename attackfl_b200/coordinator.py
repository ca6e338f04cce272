"""Host round coordinator: REGISTER -> (START -> UPDATE)* -> STOP, on top of the ``Federation`` engine.

Reference parity: ``Server.__init__`` / ``on_request`` / ``client_selection`` / ``notify_clients`` /
``process_consumer`` (``/root/reference/server.py:101-182,205-275,626-635,569-624,277-567``) and
``delete_old_queues`` (``:803-835``, n/a without a broker).  Semantics kept: the server waits for exactly
``server.clients`` registrations, registration order is the client index, attackers are identified at REGISTER,
all clients are selected every round, ``round`` only advances on success (failed rounds are repeated), STOP goes to
every registered client, ``app.log`` carries ``### Application start ###``, ``Active with N client: [...]`` and one
``ROC_AUC:`` line per validated round.
"""
from __future__ import annotations

import os
import uuid
from typing import List, Optional, Sequence

from .cli import parse_bool
from .config import ATTACK_MODES, Config
from .control import ControlServer
from .engine import ClientSpec, Federation
from .log import Logger, print_with_color


def spec_from_register(index: int, msg: dict) -> ClientSpec:
    attack = parse_bool(msg.get("attack", False))
    mode = msg.get("attack_mode")
    if attack and mode not in ATTACK_MODES:
        raise ValueError(f"Attack client not contain '{mode}' algorithm.")
    return ClientSpec(index=index, attack=attack, attack_mode=mode, attack_round=int(msg.get("attack_round") or 0),
                      attack_args=[float(x) for x in (msg.get("attack_args") or [])] or None,
                      device=msg.get("device"), client_id=str(msg.get("client_id")))


class Coordinator:
    def __init__(self, cfg: Config, device: Optional[str] = None, workdir: str = ".", ctx=None):
        self.cfg, self.device, self.workdir, self.ctx = cfg, device, workdir, ctx
        os.makedirs(cfg.log_path or ".", exist_ok=True)
        self.rank0 = ctx is None or ctx.rank == 0
        if self.rank0:
            self.logger = Logger(os.path.join(cfg.log_path or ".", "app.log"))
        else:                                   # only rank 0 talks to the user / the log file
            from .log import NullLogger, set_quiet
            self.logger = NullLogger()
            set_quiet(True)
        self.server: Optional[ControlServer] = None
        self.fed: Optional[Federation] = None

    # ------------------------------------------------------------------ registration
    def listen(self) -> ControlServer:
        host = self.cfg.runtime.control_host
        self.server = ControlServer(host, self.cfg.runtime.control_port)
        self.logger.log_info("### Application start ###\n")
        print_with_color(f"Server is waiting for {self.cfg.server.clients} clients.", "green")
        return self.server

    def wait_for_clients(self, timeout: Optional[float] = None) -> List[ClientSpec]:
        assert self.server is not None
        msgs = self.server.wait_for(self.cfg.server.clients,
                                    lambda m: print_with_color(f"[<<<] Received message from client: {m}", "blue"),
                                    timeout)
        print_with_color("All clients are connected. Sending notifications.", "green")
        return [spec_from_register(i, m) for i, m in enumerate(msgs)]

    @staticmethod
    def local_specs(n: int, attackers: Sequence[str] = ()) -> List[ClientSpec]:
        """Bulk registration (extension for 64-128 simulated clients): ``attackers`` items are
        ``index:mode:round[:arg,arg]``."""
        specs = [ClientSpec(i, client_id=str(uuid.uuid4())) for i in range(n)]
        for item in attackers:
            parts = item.split(":")
            idx, mode, rnd = int(parts[0]), parts[1], int(parts[2])
            args = [float(x) for x in parts[3].split(",")] if len(parts) > 3 and parts[3] else None
            if mode not in ATTACK_MODES:
                raise ValueError(f"Attack client not contain '{mode}' algorithm.")
            specs[idx] = ClientSpec(idx, True, mode, rnd, args, client_id=specs[idx].client_id)
        return specs

    # ------------------------------------------------------------------ rounds
    def start(self, specs: Sequence[ClientSpec]) -> Federation:
        self.fed = Federation(self.cfg, specs, device=self.device, ctx=self.ctx, logger=self.logger,
                              workdir=self.workdir)
        # client_selection(): every registered client, logged like server.py:635
        self.logger.log_info(f"Active with {len(self.fed.selected)} client: {self.fed.selected}")
        return self.fed

    def run(self) -> int:
        """Runs until ``num-round`` successful rounds; returns the number of attempted rounds."""
        fed = self.fed
        assert fed is not None
        attempts = 0
        while fed.remaining > 0:
            rno = fed.num_round - fed.remaining + 1
            print_with_color(f"Start training round {rno}", "yellow")
            if self.server is not None:
                for c in fed.selected:
                    print_with_color(f"[>>>] Sent start training request to client {fed.specs[c].client_id}", "red")
                    self.server.notify(c, {"action": "START", "message": "Server accept the connection!", "round": rno,
                                           "model_name": fed.model_name, "data_name": fed.data_name})
            res = fed.run_round()
            attempts += 1
            print_with_color("Collected all parameters.", "yellow")
            if self.server is not None:
                for c in fed.selected:
                    self.server.notify(c, {"action": "UPDATE", "round": rno, "result": res.ok, "size": res.sizes[c],
                                           "loss": res.losses[c], "attacking": c in res.attackers,
                                           "metric": res.metric})
            mr = self.cfg.server.max_retries
            if not res.ok and mr and fed.failures_in_a_row > mr:
                print_with_color(f"Round {rno} failed {fed.failures_in_a_row} times; giving up "
                                 f"(server.max-retries={mr})", "red")
                break
        self.stop()
        return attempts

    def stop(self) -> None:
        if self.server is not None:
            for _, m in self.server.clients:
                print_with_color(f"[>>>] Sent stop training request to client {m.get('client_id')}", "red")
            self.server.broadcast({"action": "STOP", "message": "Stop training!", "parameters": None})
            self.server.close()
            self.server = None
        if self.fed is not None:
            self.fed.close()
