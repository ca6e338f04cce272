"""Imports the UNMODIFIED reference (``baseline/_ref``) in-process.

* puts the in-memory ``pika`` stand-in (``baseline/stubs``) and ``baseline/_ref`` on ``sys.path``;
* ``load_utils()`` / ``load_model()`` import ``src.Utils`` / ``src.Model`` (pure functions/classes);
* ``load_server_client(workdir, config, device)`` writes ``config.yaml`` into ``workdir``, chdirs
  there, sets ``sys.argv`` and imports ``server`` (which star-imports ``client``, server.py:31);
  module bodies run exactly as shipped.  The only edit ever applied is the documented one-token
  ``fedavg`` fix (``if`` -> ``elif`` at server.py:289), done on the *source text in
  memory* when ``patch_fedavg=True`` — the files on disk stay byte-identical to the reference.
"""
from __future__ import annotations

import importlib
import importlib.util
import os
import subprocess
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.path.join(HERE, "_ref")
STUBS = os.path.join(HERE, "stubs")


def available() -> bool:
    return os.path.exists(os.path.join(REF, "server.py"))


def ensure_installed() -> bool:
    if available():
        return True
    if os.path.isdir("/root/reference"):
        subprocess.run([os.path.join(HERE, "install_ref.sh")], check=False,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return available()


def _paths():
    for p in (REF, STUBS):
        if p not in sys.path:
            sys.path.insert(0, p)


def load_utils():
    _paths()
    return importlib.import_module("src.Utils")


def load_model():
    _paths()
    return importlib.import_module("src.Model")


def load_validation():
    _paths()
    return importlib.import_module("src.Validation")


def load_server_client(workdir: str, config: dict, device: str = "cpu", patch_fedavg: bool = True):
    """Returns ``(server_module, client_module)`` freshly executed inside ``workdir``."""
    import yaml
    _paths()
    import pika
    pika.reset()
    os.makedirs(workdir, exist_ok=True)
    with open(os.path.join(workdir, "config.yaml"), "w") as fh:
        yaml.safe_dump(config, fh)
    os.chdir(workdir)
    old_argv = sys.argv
    sys.argv = ["ref", "--device", device]
    try:
        for name in ("server", "client"):
            sys.modules.pop(name, None)
        # load the reference's client.py by path: this repo has its own client.py / server.py CLI shims at its root,
        # which must not shadow the reference's modules (server.py:31 does `from client import *`)
        spec = importlib.util.spec_from_file_location("client", os.path.join(REF, "client.py"))
        client = importlib.util.module_from_spec(spec)
        sys.modules["client"] = client
        spec.loader.exec_module(client)
        src = open(os.path.join(REF, "server.py")).read()
        if patch_fedavg:
            a = '                self.avg_all_parameters()\n            if server_mode == "FLTrust":'
            b = '                self.avg_all_parameters()\n            elif server_mode == "FLTrust":'
            assert src.count(a) == 1, "reference server.py changed; fedavg patch site not found"
            src = src.replace(a, b)
        server = types.ModuleType("server")
        server.__file__ = os.path.join(REF, "server.py")
        sys.modules["server"] = server
        exec(compile(src, server.__file__, "exec"), server.__dict__)
    finally:
        sys.argv = old_argv
    return server, client


def pump(server_obj, clients, max_messages: int = 10**9) -> int:
    """Delivers queued messages until the server calls ``sys.exit()`` (server.py:567).

    ``clients``: list of ``RpcClient``.  Returns the number of messages delivered."""
    import pika
    n = 0
    by_queue = {f"reply_{c.client_id}": c for c in clients}
    try:
        while n < max_messages:
            progressed = False
            q = pika.QUEUES["rpc_queue"]
            while q:
                body = q.popleft()
                server_obj.on_request(pika.Channel(), pika._Method(0), pika._Props(), body)
                n += 1
                progressed = True
            for name, c in by_queue.items():
                cq = pika.QUEUES[name]
                while cq:
                    c.response_message(cq.popleft())
                    n += 1
                    progressed = True
            if not progressed:
                break
    except SystemExit:
        pass
    return n
