"""TCP transport for state_dicts (multi-host client/server deployments).

Reference: ``send_saved_model`` streams ``model.pt`` in 1024-byte chunks to a hard-coded
address (client.py:191-210); the server accepts ``world_size-1`` connections, writes
``received_model_{n}.pt`` and averages them (server.py:25-55, 80-98). Kept as a
*transport option* for clients that are not on the same NVLink domain; on one B200 box the
upload + mean + broadcast is a single kernel (``parallel/aggregate.py``). Differences:
address/port are configuration, frames are length-prefixed, the listening socket is
created once with SO_REUSEADDR and closed, chunks are 4 MiB.
"""
from __future__ import annotations

import io
import os
import socket
import struct
import threading
from typing import Dict, List, Optional

import torch

_CHUNK = 4 << 20


def send_state_dict(state: Dict[str, torch.Tensor], host: str, port: int, client_id: int = 0,
                    timeout_s: float = 600.0) -> int:
    buf = io.BytesIO()
    torch.save(state, buf)
    payload = buf.getbuffer()
    with socket.create_connection((host, port), timeout=timeout_s) as s:
        s.sendall(struct.pack("!IQ", client_id, len(payload)))
        for off in range(0, len(payload), _CHUNK):
            s.sendall(payload[off:off + _CHUNK])
    return len(payload)


def send_saved_model(path: str = "model.pt", host: str = "127.0.0.1", port: int = 12345, client_id: int = 0) -> int:
    return send_state_dict(torch.load(path, map_location="cpu", weights_only=True), host, port, client_id)


class StateDictServer:
    """Receives ``expected`` uploads per round into ``received_model_{i}.pt`` files."""

    def __init__(self, host: str = "0.0.0.0", port: int = 12345, out_dir: str = "."):
        self.out_dir = out_dir
        self.sock = socket.socket(socket.AF_INET, socket.SOCK_STREAM)
        self.sock.setsockopt(socket.SOL_SOCKET, socket.SO_REUSEADDR, 1)
        self.sock.bind((host, port))
        self.sock.listen(64)
        self.port = self.sock.getsockname()[1]
        self.last_client_ids = {}       # slot → client id announced in the upload header (for weighted means)

    def _handle(self, conn: socket.socket, slot: int, paths: List[Optional[str]]) -> None:
        with conn:
            hdr = _recv_exact(conn, 12)
            cid, n = struct.unpack("!IQ", hdr)
            self.last_client_ids[slot] = cid
            path = os.path.join(self.out_dir, f"received_model_{slot}.pt")
            with open(path, "wb") as f:
                left = n
                while left:
                    chunk = conn.recv(min(_CHUNK, left))
                    if not chunk:
                        raise ConnectionError("client closed the connection mid-upload")
                    f.write(chunk)
                    left -= len(chunk)
            paths[slot - 1] = path

    def receive_round(self, expected: int, timeout_s: float = 600.0) -> List[str]:
        paths: List[Optional[str]] = [None] * expected
        self.last_client_ids = {}
        threads = []
        self.sock.settimeout(timeout_s)
        for slot in range(1, expected + 1):
            conn, _ = self.sock.accept()
            t = threading.Thread(target=self._handle, args=(conn, slot, paths), daemon=True)
            t.start()
            threads.append(t)
        for t in threads:
            t.join(timeout_s)
        if any(p is None for p in paths):
            raise TimeoutError("not every client upload completed")
        return [p for p in paths if p]

    def close(self) -> None:
        self.sock.close()


def _recv_exact(conn: socket.socket, n: int) -> bytes:
    out = bytearray()
    while len(out) < n:
        chunk = conn.recv(n - len(out))
        if not chunk:
            raise ConnectionError("short read")
        out += chunk
    return bytes(out)


def average_state_dicts(paths: List[str], weights: Optional[List[float]] = None) -> Dict[str, torch.Tensor]:
    """Per-key weighted mean of bare state_dict files (server.py:37-55 uses uniform weights)."""
    sds = [torch.load(p, map_location="cpu", weights_only=True) for p in paths]
    w = [1.0] * len(sds) if weights is None else list(weights)
    tot = float(sum(w))
    out = {}
    for k in sds[0]:
        if sds[0][k].is_floating_point():
            acc = torch.zeros_like(sds[0][k], dtype=torch.float32)
            for sd, wi in zip(sds, w):
                acc.add_(sd[k].float(), alpha=wi / tot)
            out[k] = acc.to(sds[0][k].dtype)
        else:
            out[k] = sds[0][k].clone()
    return out
