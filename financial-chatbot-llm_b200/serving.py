"""One resident engine for many worker processes.

The reference runs 3–6 gunicorn workers (gunicorn.conf.py:8, Dockerfile:39), each with its own backend
client built at import (tools/qdrant_tool.py:73) and ONE query in flight at a time (§3.2). A 61 GB corpus
cannot be loaded once per worker, and single queries against small tenants are launch-latency-bound. So:
one ``EngineServer`` process owns the GPU-resident collection; workers talk to it over a Unix-domain
socket with ``EngineClient``, which duck-types ``QdrantClient.query_points``; the server MICRO-BATCHES the
requests that arrive within a short window into one grouped ``fcb_search`` launch (the path that does
270k q/s at 64 tenants per call instead of 36k q/s one by one) and returns scored points with payloads.

Transport: length-prefixed frames over an AF_UNIX stream socket. A request is a FIXED BINARY header (struct, below) +
the collection name + the qdrant filter as JSON + the query vector as raw float64; a reply is JSON (scored points with
their payload dicts; Python's float repr round-trips float64 scores bit for bit). Nothing is unpickled on either side,
so a peer can at worst send a malformed request, which is answered with an error frame. The socket file is created with
mode 0600: the trust boundary is "processes of the same user on this host" (INTEGRATION.md §3).
"""
from __future__ import annotations

import json
import math
import os
import queue
import socket
import struct
import threading
import time
from typing import Any, Dict, List, Optional

import numpy as np

from .client import Collection, merge_conjunct_lists
from .qdrant_models import Filter, QueryResponse, ScoredPoint

# request header: magic, version, flags (bit 0: with_payload), limit, offset, dim, len(collection name), len(filter JSON),
# score_threshold (NaN: none). Then: name (utf-8), filter JSON (utf-8, may be empty), dim x float64 (little endian).
_REQ = struct.Struct("<4sHHiiIIId")
_MAGIC = b"FCB1"
_LEN = struct.Struct("<I")
_MAX_FRAME = 64 << 20                  # a request is ~12 KB, a reply with 10 000 payloads a few MB


def _recv_exact(sock: socket.socket, n: int) -> bytes:
    buf = bytearray(n)
    view, got = memoryview(buf), 0
    while got < n:
        r = sock.recv_into(view[got:], n - got)
        if r == 0:
            raise EOFError("peer closed the connection")
        got += r
    return bytes(buf)


def _send_frame(sock: socket.socket, body: bytes) -> None:
    sock.sendall(_LEN.pack(len(body)) + body)


def _recv_frame(sock: socket.socket) -> bytes:
    (n,) = _LEN.unpack(_recv_exact(sock, _LEN.size))
    if n > _MAX_FRAME:
        raise ValueError(f"frame of {n} bytes refused")
    return _recv_exact(sock, n)


def filter_to_json(flt: Any) -> str:
    """A qdrant filter (our models or qdrant-client's: both pydantic) as JSON; '' for no filter."""
    if flt is None:
        return ""
    if hasattr(flt, "model_dump"):
        return json.dumps(flt.model_dump(exclude_none=True))
    raise NotImplementedError(f"cannot serialise a filter of type {type(flt).__name__}")


def encode_request(collection_name: str, query, query_filter: Any, limit: int, offset: Optional[int], with_payload: bool,
                   score_threshold: Optional[float]) -> bytes:
    q = np.ascontiguousarray(query, dtype="<f8").reshape(-1)
    name, fj = collection_name.encode("utf-8"), filter_to_json(query_filter).encode("utf-8")
    head = _REQ.pack(_MAGIC, 1, 1 if with_payload else 0, int(limit), int(offset or 0), q.shape[0], len(name), len(fj),
                     math.nan if score_threshold is None else float(score_threshold))
    return head + name + fj + q.tobytes()


def decode_request(body: bytes) -> Dict[str, Any]:
    if len(body) < _REQ.size:
        raise ValueError("short request")
    magic, ver, flags, limit, offset, dim, n_name, n_flt, thr = _REQ.unpack_from(body, 0)
    if magic != _MAGIC or ver != 1:
        raise ValueError("not an fcb200 request frame")
    if len(body) != _REQ.size + n_name + n_flt + 8 * dim:
        raise ValueError("request length does not match its header")
    at = _REQ.size
    name = body[at:at + n_name].decode("utf-8"); at += n_name
    fj = body[at:at + n_flt].decode("utf-8"); at += n_flt
    return {"collection_name": name, "query": np.frombuffer(body, dtype="<f8", count=dim, offset=at),
            "query_filter": Filter.model_validate(json.loads(fj)) if fj else None, "limit": limit, "offset": offset,
            "with_payload": bool(flags & 1), "score_threshold": None if math.isnan(thr) else thr}


class _Request:
    __slots__ = ("conn", "query", "filts", "k", "skip", "with_payload", "score_threshold", "lock")

    def __init__(self, conn, lock, query, filts, k, skip, with_payload, score_threshold):
        self.conn, self.lock, self.query, self.filts, self.k = conn, lock, query, filts, k
        self.skip, self.with_payload, self.score_threshold = skip, with_payload, score_threshold


class EngineServer:
    """Owns a loaded ``Collection``; serves ``query_points`` requests from ``EngineClient``s with micro-batching."""

    def __init__(self, collection: Collection, address: str, *, max_batch: Optional[int] = None,
                 window_us: float = 200.0):
        self.col, self.address = collection, address
        self.max_batch = min(max_batch or collection.engine.max_batch, collection.engine.max_batch)
        self.window_s = window_us * 1e-6
        self._q: "queue.Queue[_Request]" = queue.Queue()
        self._stop = threading.Event()
        self._threads: List[threading.Thread] = []
        self.stats = {"requests": 0, "launches": 0, "max_batch": 0, "batch_hist": {}}
        if os.path.exists(address):
            os.unlink(address)
        self._listener = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        old_umask = os.umask(0o177)                     # the socket file is born 0600: same-user processes only
        try:
            self._listener.bind(address)
        finally:
            os.umask(old_umask)
        os.chmod(address, 0o600)
        self._listener.listen(64)

    # ---- lifecycle
    def start(self) -> "EngineServer":
        for target in (self._accept_loop, self._batch_loop):
            t = threading.Thread(target=target, daemon=True)
            t.start()
            self._threads.append(t)
        return self

    def stop(self) -> None:
        self._stop.set()
        try:
            c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)    # unblock accept()
            c.connect(self.address)
            c.close()
        except OSError:
            pass
        self._listener.close()
        for t in self._threads:
            t.join(timeout=2.0)
        if os.path.exists(self.address):
            os.unlink(self.address)

    # ---- network side: one reader thread per connection
    def _accept_loop(self) -> None:
        while not self._stop.is_set():
            try:
                conn, _ = self._listener.accept()
            except OSError:
                return
            if self._stop.is_set():
                conn.close()
                return
            t = threading.Thread(target=self._reader, args=(conn,), daemon=True)
            t.start()

    @staticmethod
    def _reply(conn, lock, obj) -> None:
        with lock:
            _send_frame(conn, json.dumps(obj, separators=(",", ":")).encode("utf-8"))

    def _reader(self, conn) -> None:
        lock = threading.Lock()
        try:
            while not self._stop.is_set():
                body = _recv_frame(conn)
                try:
                    msg = decode_request(body)
                    if msg["collection_name"] != self.col.name:
                        raise ValueError(f"Collection {msg['collection_name']} not found")
                    q = msg["query"]
                    if q.shape != (self.col.engine.dim,):
                        raise ValueError(f"query must be a vector of length {self.col.engine.dim}")
                    skip = int(msg.get("offset") or 0)
                    k = int(msg["limit"]) + skip
                    if k < 1:
                        self._reply(conn, lock, {"points": []})
                        continue
                    filts = self.col.resolve_dnf(msg.get("query_filter"))   # should / must_not: several conjuncts
                    if not filts:                                            # nothing can match
                        self._reply(conn, lock, {"points": []})
                        continue
                    self._q.put(_Request(conn, lock, q, filts, k, skip, bool(msg.get("with_payload", True)),
                                         msg.get("score_threshold")))
                except Exception as e:                             # the caller's blanket except turns it into []
                    self._reply(conn, lock, {"error": f"{type(e).__name__}: {e}"})
        except (EOFError, OSError, ValueError):
            pass
        finally:
            conn.close()

    # ---- GPU side: collect what arrives within the window, one grouped launch per distinct k
    def _batch_loop(self) -> None:
        while not self._stop.is_set():
            try:
                first = self._q.get(timeout=0.05)
            except queue.Empty:
                continue
            batch = [first]
            deadline = time.perf_counter() + self.window_s
            while len(batch) < self.max_batch:
                left = deadline - time.perf_counter()
                if left <= 0:
                    break
                try:
                    batch.append(self._q.get(timeout=left))
                except queue.Empty:
                    break
            by_k: Dict[int, List[_Request]] = {}
            for r in batch:
                by_k.setdefault(r.k, []).append(r)
            for k, reqs in by_k.items():
                self._serve(k, reqs)

    def _serve(self, k: int, reqs: List[_Request]) -> None:
        try:
            # one row per (request, conjunct); chunks of at most max_batch rows per launch
            rows_q = [r.query for r in reqs for _ in r.filts]
            rows_f = [f for r in reqs for f in r.filts]
            parts = []
            for lo in range(0, len(rows_f), self.max_batch):
                parts.append(self.col.engine.search_host(np.stack(rows_q[lo:lo + self.max_batch]),
                                                         rows_f[lo:lo + self.max_batch], k))
                nb = len(rows_f[lo:lo + self.max_batch])
                self.stats["launches"] += 1
                self.stats["max_batch"] = max(self.stats["max_batch"], nb)
                self.stats["batch_hist"][nb] = self.stats["batch_hist"].get(nb, 0) + 1
            ids = np.concatenate([p[0] for p in parts]); scores = np.concatenate([p[1] for p in parts])
            counts = np.concatenate([p[2] for p in parts])
            self.stats["requests"] += len(reqs)
            at = 0
            for r in reqs:
                n = len(r.filts)
                if n == 1:
                    r_ids, r_sc = ids[at, :int(counts[at])], scores[at, :int(counts[at])]
                else:
                    r_ids, r_sc = merge_conjunct_lists(ids[at:at + n], scores[at:at + n], counts[at:at + n], k)
                at += n
                pts = []
                for j in range(r.skip, r_ids.shape[0]):
                    s = float(r_sc[j])
                    if r.score_threshold is not None and s < r.score_threshold:
                        break
                    pid = int(r_ids[j])
                    row = self.col.row_of_id(pid)
                    pts.append({"id": self.col.point_id_fn(row) if self.col.point_id_fn else pid, "version": 0, "score": s,
                                "payload": self.col.payload_fn(row) if r.with_payload else None, "vector": None})
                self._reply(r.conn, r.lock, {"points": pts})
        except Exception as e:
            for r in reqs:
                try:
                    self._reply(r.conn, r.lock, {"error": f"{type(e).__name__}: {e}"})
                except OSError:
                    pass


class EngineClient:
    """``QdrantClient`` look-alike for a worker process: ``query_points`` is forwarded to the ``EngineServer``.
    Constructor keywords of the reference (url, api_key) are accepted and ignored; ``address`` is the socket path
    (default: environment variable FCB_ENGINE_SOCKET)."""

    def __init__(self, location: Optional[str] = None, url: Optional[str] = None, api_key: Optional[str] = None,
                 address: Optional[str] = None, **_ignored: Any):
        self.address = address or os.environ.get("FCB_ENGINE_SOCKET")
        if not self.address:
            raise ValueError("EngineClient needs address= or FCB_ENGINE_SOCKET")
        self._conn = None
        self._lock = threading.Lock()

    def _connection(self):
        if self._conn is None:
            c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            c.connect(self.address)
            self._conn = c
        return self._conn

    def close(self) -> None:
        if self._conn is not None:
            self._conn.close()
            self._conn = None

    def query_points(self, collection_name: str, query: Any = None, using: Optional[str] = None, prefetch: Any = None,
                     query_filter: Any = None, search_params: Any = None, limit: int = 10, offset: Optional[int] = None,
                     with_payload: Any = True, with_vectors: Any = False, score_threshold: Optional[float] = None,
                     **kwargs: Any) -> QueryResponse:
        if prefetch is not None or using is not None or with_vectors or query is None:
            raise NotImplementedError("unsupported query_points arguments")
        frame = encode_request(collection_name, query, query_filter, limit, offset, bool(with_payload), score_threshold)
        with self._lock:
            conn = self._connection()
            _send_frame(conn, frame)
            reply = json.loads(_recv_frame(conn))
        if "error" in reply:
            raise RuntimeError(reply["error"])
        return QueryResponse(points=[ScoredPoint(**p) for p in reply["points"]])
