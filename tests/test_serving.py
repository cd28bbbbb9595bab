"""Resident engine + micro-batching (serving.py): many clients, one grouped launch per window.
CPU: host logic with a fake engine. GPU: the real engine, results against the oracle."""
import os
import threading

import numpy as np
import pytest

from financial_chatbot_llm_b200 import client, serving
from financial_chatbot_llm_b200 import qdrant_models as m


class FakeEngine:
    dim, row_base, max_batch = 4, 0, 16

    def __init__(self):
        self.batches = []

    def search_host(self, q, filters, k):
        B = q.shape[0]
        self.batches.append(B)
        ids = np.full((B, k), np.iinfo(np.uint64).max, dtype=np.uint64)
        sc = np.full((B, k), -np.inf)
        cnt = np.zeros(B, dtype=np.int32)
        for b, f in enumerate(filters):
            rows = np.arange(f.row_begin, f.row_end, dtype=np.uint64)[:k]
            ids[b, :rows.size] = rows
            sc[b, :rows.size] = q[b, 0] - 0.01 * np.arange(rows.size)        # the query leaves a fingerprint in the score
            cnt[b] = rows.size
        return ids, sc, cnt


def user_filter(u):
    return m.Filter(must=[m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value=u))])


def test_micro_batching_host_logic(tmp_path):
    eng = FakeEngine()
    users = ["alice", "bob", "carol"]
    col = client.Collection("transactions", eng, users, np.array([0, 3, 8, 20]),
                            lambda row: {"page_content": f"txn {row}", "metadata": {"user_id": users[0 if row < 3 else 1 if row < 8 else 2]}})
    addr = str(tmp_path / "engine.sock")
    srv = serving.EngineServer(col, addr, window_us=20000).start()
    errors, results = [], {}

    def worker(i):
        try:
            cl = serving.EngineClient(url="http://ignored", api_key="x", address=addr)
            for j in range(10):
                u = users[(i + j) % 3]
                pts = cl.query_points("transactions", query=[float(i * 100 + j), 0, 0, 0], limit=4, query_filter=user_filter(u)).points
                results[(i, j)] = (u, [p.id for p in pts], pts[0].score, pts[0].payload["metadata"]["user_id"])
            with pytest.raises(RuntimeError, match="not found"):
                cl.query_points("missing", query=[0, 0, 0, 0])
            cl.close()
        except Exception as e:                                              # surfaced below
            errors.append(repr(e))
    threads = [threading.Thread(target=worker, args=(i,)) for i in range(8)]
    [t.start() for t in threads]
    [t.join(timeout=30) for t in threads]
    srv.stop()
    assert not errors, errors
    assert len(results) == 80
    for (i, j), (u, ids, s0, pu) in results.items():
        lo = {"alice": 0, "bob": 3, "carol": 8}[u]
        assert ids[0] == lo and len(ids) == (3 if u == "alice" else 4) and pu == u
        assert s0 == float(i * 100 + j)                                    # each reply went back to ITS requester
    assert srv.stats["requests"] == 80 and srv.stats["max_batch"] >= 2 and srv.stats["launches"] < 80
    assert max(eng.batches) <= eng.max_batch


def test_or_and_not_filters_through_the_server(tmp_path):
    """should / must_not requests become several query rows of the same grouped launch; the server merges them per request."""
    eng = FakeEngine()
    users = ["alice", "bob", "carol"]
    col = client.Collection("transactions", eng, users, np.array([0, 3, 8, 20]),
                            lambda row: {"page_content": f"txn {row}", "metadata": {"user_id": users[0 if row < 3 else 1 if row < 8 else 2]}})
    addr = str(tmp_path / "engine2.sock")
    srv = serving.EngineServer(col, addr, window_us=1000).start()
    try:
        cl = serving.EngineClient(address=addr)
        U = lambda u: m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value=u))
        # alice OR carol: two conjuncts; each list is scored q0 - 0.01 * position, so the merge interleaves them by score, id
        pts = cl.query_points("transactions", query=[5.0, 0, 0, 0], limit=5, query_filter=m.Filter(should=[U("alice"), U("carol")])).points
        assert [p.id for p in pts] == [0, 8, 1, 9, 2] and [p.payload["metadata"]["user_id"] for p in pts] == ["alice", "carol"] * 2 + ["alice"]
        assert eng.batches[-1] == 2
        # NOT bob: the rows before and after bob's segment
        pts = cl.query_points("transactions", query=[5.0, 0, 0, 0], limit=3, query_filter=m.Filter(must_not=[U("bob")])).points
        assert [p.id for p in pts] == [0, 8, 1]
        # nothing can match: answered without a launch
        n = len(eng.batches)
        assert cl.query_points("transactions", query=[5.0, 0, 0, 0], limit=3, query_filter=m.Filter(must=[U("bob")], must_not=[U("bob")])).points == []
        assert len(eng.batches) == n
        with pytest.raises(RuntimeError, match="NotImplementedError"):
            cl.query_points("transactions", query=[5.0, 0, 0, 0], query_filter=m.Filter(min_should=2, should=[U("alice")]))
        cl.close()
    finally:
        srv.stop()


@pytest.mark.gpu
def test_server_with_real_engine_matches_oracle(tmp_path, built):
    from financial_chatbot_llm_b200 import loader, synth
    from financial_chatbot_llm_b200.engine import Engine
    from oracle import canon
    spec = synth.SynthSpec(n_rows=20000, dim=256, n_users=30, seed=61)
    hc = synth.make_host_corpus(spec)
    eng = Engine(device=0, max_batch=32, max_k=16)
    col = loader.collection_from_host(hc, eng)
    addr = str(tmp_path / "engine.sock")
    srv = serving.EngineServer(col, addr, window_us=2000).start()
    bad = []

    def worker(i):
        cl = serving.EngineClient(address=addr)
        for j in range(15):
            u = (3 * i + j) % 30
            q = synth.query_vector(spec, 100 * i + j)
            pts = cl.query_points("transactions", query=q.tolist(), limit=10, query_filter=user_filter(hc.users[u])).points
            oi, os_ = canon.search(hc.vectors, q, 10, row_begin=int(hc.offsets[u]), row_end=int(hc.offsets[u + 1]))
            if [p.id for p in pts] != oi.tolist() or [p.score for p in pts] != os_.tolist() or \
                    any(p.payload != hc.payload(p.id) for p in pts):
                bad.append((i, j))
        cl.close()
    threads = [threading.Thread(target=worker, args=(i,)) for i in range(12)]
    [t.start() for t in threads]
    [t.join(timeout=120) for t in threads]
    srv.stop()
    eng.close()
    assert not bad and srv.stats["requests"] == 180 and srv.stats["max_batch"] >= 2


def test_frames_carry_no_pickle_and_the_socket_is_private(tmp_path):
    """The wire format is a fixed binary header + JSON + raw float64: requests round-trip exactly, a malformed frame is
    answered with an error (never executed), and the socket file is mode 0600 (same-user trust boundary)."""
    import inspect
    import socket
    import stat
    src = inspect.getsource(serving)
    assert "import pickle" not in src and "multiprocessing.connection" not in src.split('"""', 2)[2]
    flt = m.Filter(must=[m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value="bob")),
                         m.FieldCondition(key="metadata.date", range=m.Range(gte=1700000000))],
                   must_not=[m.FieldCondition(key="metadata.category", match=m.MatchAny(any=["Rent", "Fuel"]))])
    q = np.array([0.1, -2.5e-7, 3.0, 1e300])
    msg = serving.decode_request(serving.encode_request("transactions", q.tolist(), flt, 7, 2, False, 0.25))
    assert msg["collection_name"] == "transactions" and msg["limit"] == 7 and msg["offset"] == 2
    assert msg["with_payload"] is False and msg["score_threshold"] == 0.25 and msg["query"].tolist() == q.tolist()
    assert msg["query_filter"] == flt
    assert serving.decode_request(serving.encode_request("t", q, None, 1, None, True, None))["query_filter"] is None
    with pytest.raises(ValueError):
        serving.decode_request(b"not a frame at all, but long enough to hold a header ...............")

    eng = FakeEngine()
    col = client.Collection("transactions", eng, ["alice"], np.array([0, 3]), lambda row: {"page_content": "x", "metadata": {"user_id": "alice"}})
    addr = str(tmp_path / "p.sock")
    srv = serving.EngineServer(col, addr).start()
    try:
        assert stat.S_IMODE(os.stat(addr).st_mode) == 0o600
        raw = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        raw.connect(addr)
        serving._send_frame(raw, b"\\x80\\x04garbage that a pickle-based server would have tried to load")
        import json as _json
        assert "error" in _json.loads(serving._recv_frame(raw))
        raw.close()
        cl = serving.EngineClient(address=addr)                       # the server is still serving
        assert [p.id for p in cl.query_points("transactions", query=[1.0, 0, 0, 0], limit=2, query_filter=user_filter("alice")).points] == [0, 1]
        cl.close()
    finally:
        srv.stop()
