"""Local control plane: newline-delimited JSON over a loopback TCP socket.

Replaces the reference's RabbitMQ/AMQP transport (``/root/reference/server.py:102-108,187-203``;
``src/RpcClient.py:42-56,174-188``) for the *control* messages only — REGISTER / START / UPDATE / STOP
(SURVEY §2.5).  Tensors never travel with them: client models live in the device arena of the server process, so
a START is ~100 bytes instead of a pickled ``state_dict`` (0.2-0.8 MB, plus k whole models for attackers).
No ``pickle.loads`` of network input (the reference unpickles whatever arrives on ``rpc_queue``, server.py:214).
"""
from __future__ import annotations

import json
import socket
import time
from typing import Callable, List, Optional, Tuple


def _send(sock: socket.socket, obj: dict) -> None:
    sock.sendall((json.dumps(obj) + "\n").encode())


class _LineReader:
    def __init__(self, sock: socket.socket):
        self.sock, self.buf = sock, b""

    def read(self, timeout: Optional[float] = None) -> Optional[dict]:
        self.sock.settimeout(timeout)
        while b"\n" not in self.buf:
            try:
                chunk = self.sock.recv(65536)
            except socket.timeout:
                return None
            if not chunk:
                raise ConnectionError("peer closed the control connection")
            self.buf += chunk
        line, self.buf = self.buf.split(b"\n", 1)
        return json.loads(line.decode())


class ControlServer:
    """Accepts client registrations; ``clients[i]`` is the i-th REGISTER to arrive (server.py:223-227)."""

    def __init__(self, host: str = "127.0.0.1", port: int = 56720):
        self.sock = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        self.sock.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        self.sock.bind((host, port))
        self.sock.listen(256)
        self.host, self.port = self.sock.getsockname()
        self.clients: List[Tuple[socket.socket, dict]] = []
        self._seen = set()

    def wait_for(self, n: int, on_register: Optional[Callable[[dict], None]] = None,
                 timeout: Optional[float] = None) -> List[dict]:
        """Blocks until ``n`` distinct clients registered (the reference triggers at exact equality, :231)."""
        deadline = None if timeout is None else time.time() + timeout
        while len(self.clients) < n:
            self.sock.settimeout(None if deadline is None else max(0.05, deadline - time.time()))
            try:
                conn, _ = self.sock.accept()
            except socket.timeout:
                raise TimeoutError(f"only {len(self.clients)} of {n} clients registered")
            try:
                msg = _LineReader(conn).read(timeout=10.0)
            except Exception:
                conn.close()
                continue
            if not msg or msg.get("action") != "REGISTER":
                conn.close()
                continue
            cid = str(msg.get("client_id"))
            if cid in self._seen:
                conn.close()
                continue
            self._seen.add(cid)
            self.clients.append((conn, msg))
            if on_register:
                on_register(msg)
        return [m for _, m in self.clients]

    def notify(self, index: int, obj: dict) -> None:
        try:
            _send(self.clients[index][0], obj)
        except OSError:
            pass                      # a vanished client process does not stop the federation (its slot lives here)

    def broadcast(self, obj: dict) -> None:
        for i in range(len(self.clients)):
            self.notify(i, obj)

    def close(self) -> None:
        for conn, _ in self.clients:
            try:
                conn.close()
            except OSError:
                pass
        self.sock.close()


class ControlClient:
    """``client.py`` side: register, then print what the coordinator reports until STOP."""

    def __init__(self, host: str, port: int, retry_s: float = 1.0, max_wait_s: float = 120.0):
        t0 = time.time()
        last: Optional[Exception] = None
        while time.time() - t0 < max_wait_s:
            try:
                self.sock = socket.create_connection((host, port), timeout=5.0)
                self.reader = _LineReader(self.sock)
                return
            except OSError as e:       # same idea as the reference's reconnect loop (RpcClient.py:53-56)
                last = e
                print(f"Connection failed, retrying in {retry_s:.0f} seconds: {e}")
                time.sleep(retry_s)
        raise ConnectionError(f"cannot reach the coordinator at {host}:{port}: {last}")

    def register(self, msg: dict) -> None:
        _send(self.sock, msg)

    def messages(self):
        while True:
            try:
                m = self.reader.read(timeout=None)
            except ConnectionError:
                return
            if m is None:
                continue
            yield m
            if m.get("action") == "STOP":
                return

    def close(self) -> None:
        try:
            self.sock.close()
        except OSError:
            pass
