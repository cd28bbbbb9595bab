"""On-disk corpus format + loader.

The reference has no ingestion side: nothing in it writes to Qdrant, the payload layout
``{"page_content": str, "metadata": {"user_id": str, "date": int, ...}}`` is only visible from the
reads at tools/qdrant_tool.py:107,120,160-164. This module is the missing half for the B200 engine:
it turns points in that shape (e.g. the result of a Qdrant ``scroll``) into memory-mappable column
files laid out exactly as the kernels want them (rows sorted by user, CSR offsets, int32 dates,
category codes, integer cents), and streams them into HBM in bounded chunks — host memory use is
one chunk, not one corpus.

Directory layout (all little-endian, no headers; ``manifest.json`` carries shapes and dtypes):
  manifest.json     version, n_rows, dim, dtype ("float32" | "bfloat16"), categories, has_* flags
  vectors.bin       [n_rows, dim] row-major
  ids.u64           [n_rows] point ids (absent: id = row)        id_rank.u32  (only if ids do not ascend)
  dates.i32  categories.u8  cents.i32  seg_offsets.i64 ([n_users + 1])
                    (category code 255 = the row has no metadata.category: matches no category filter)
  users.json        user_id strings in segment order (ascending)
  point_ids.json    original point ids when they are not unsigned 64-bit integers (UUID strings)
  payload.blob / payload.off   UTF-8 JSON of each row's payload, and u64 offsets ([n_rows + 1])
"""
from __future__ import annotations

import json
import os
from typing import Any, Dict, Iterable, List, Optional, Sequence

import numpy as np

from . import synth

FORMAT_VERSION = 1
NO_CATEGORY = 255                       # >= 64: passes when there is no category filter, never matches one
NO_DATE = -(2 ** 31)                    # row without metadata.date: Qdrant's Range never matches a missing key. The adapter
                                        # keeps every lower bound above this value and refuses upper bounds / must_not on the
                                        # date while such rows exist (manifest "missing_date"), so the sentinel never passes
_COLUMNS = {"ids.u64": np.uint64, "id_rank.u32": np.uint32, "dates.i32": np.int32, "categories.u8": np.uint8,
            "cents.i32": np.int32, "seg_offsets.i64": np.int64, "payload.off": np.uint64}


def _amount_to_cents(md: Dict[str, Any]) -> int:
    if "amount_cents" in md:
        return int(md["amount_cents"])
    if "amount" in md and md["amount"] is not None:
        c = round(float(md["amount"]) * 100.0)
        if abs(c / 100.0 - float(md["amount"])) > 1e-9:
            raise ValueError(f"amount {md['amount']!r} is not exact cents")
        return int(c)
    return 0


def write_corpus(path: str, points: Iterable[Dict[str, Any]], *, dim: int, dtype: str = "float32",
                 categories: Optional[Sequence[Any]] = None) -> Dict[str, Any]:
    """Write points ``{"id": int | str, "vector": sequence[float], "payload": {...}}`` (the shape a Qdrant
    scroll yields, payload as the reference reads it) as a corpus directory. Rows are sorted by
    ``metadata.user_id`` (stable, so the input order survives within a user). Returns the manifest.
    Everything is materialised in host memory first: meant for ingestion-sized inputs, not for 60 GB."""
    if dtype not in ("float32", "bfloat16"):
        raise ValueError("dtype must be float32 or bfloat16")
    pts = list(points)
    n = len(pts)
    if n == 0:
        raise ValueError("no points")
    users_per_row = [p["payload"]["metadata"]["user_id"] for p in pts]
    order = sorted(range(n), key=lambda i: users_per_row[i])            # stable
    users = sorted(set(users_per_row))
    uidx = {u: i for i, u in enumerate(users)}
    counts = np.bincount([uidx[u] for u in users_per_row], minlength=len(users))
    offsets = np.zeros(len(users) + 1, dtype=np.int64)
    np.cumsum(counts, out=offsets[1:])

    cat_values = list(categories) if categories is not None else []
    cat_code = {c: i for i, c in enumerate(cat_values)}
    vec = np.empty((n, dim), dtype=np.float32)
    dates = np.zeros(n, dtype=np.int32)
    cats = np.full(n, NO_CATEGORY, dtype=np.uint8)
    cents = np.zeros(n, dtype=np.int32)
    raw_ids: List[Any] = []
    blobs: List[bytes] = []
    missing_date = False
    for r, i in enumerate(order):
        p = pts[i]
        v = np.asarray(p["vector"], dtype=np.float32)
        if v.shape != (dim,):
            raise ValueError(f"point {p.get('id')!r}: vector length {v.shape} != {dim}")
        vec[r] = v
        md = p["payload"]["metadata"]
        if md.get("date") is None:
            dates[r] = NO_DATE
            missing_date = True
        else:
            d = int(md["date"])
            if not NO_DATE < d < 2 ** 31:
                raise ValueError(f"metadata.date {d} does not fit int32 unix seconds")
            dates[r] = d
        c = md.get("category")
        if c is not None:
            if c not in cat_code:
                if categories is not None or len(cat_code) >= 64:
                    raise ValueError(f"category {c!r} not in the declared list / more than 64 categories")
                cat_code[c] = len(cat_values)
                cat_values.append(c)
            cats[r] = cat_code[c]
        cents[r] = _amount_to_cents(md)
        raw_ids.append(p.get("id", r))
        blobs.append(json.dumps(p["payload"], separators=(",", ":")).encode("utf-8"))

    os.makedirs(path, exist_ok=True)
    int_ids = all(isinstance(x, (int, np.integer)) and 0 <= int(x) < 2 ** 64 for x in raw_ids)
    ids = np.array([int(x) for x in raw_ids], dtype=np.uint64) if int_ids else np.arange(n, dtype=np.uint64)
    if len(set(ids.tolist())) != n:
        raise ValueError("point ids are not unique")
    ascending = bool(np.all(ids[1:] > ids[:-1])) if n > 1 else True
    if dtype == "float32":
        vec.tofile(os.path.join(path, "vectors.bin"))
    else:
        synth.f32_to_bf16_bits(vec).reshape(n, dim).tofile(os.path.join(path, "vectors.bin"))
    ids.tofile(os.path.join(path, "ids.u64"))
    if not ascending:
        rank = np.empty(n, dtype=np.uint32)
        rank[np.argsort(ids, kind="stable")] = np.arange(n, dtype=np.uint32)
        rank.tofile(os.path.join(path, "id_rank.u32"))
    dates.tofile(os.path.join(path, "dates.i32"))
    cats.tofile(os.path.join(path, "categories.u8"))
    cents.tofile(os.path.join(path, "cents.i32"))
    offsets.tofile(os.path.join(path, "seg_offsets.i64"))
    off = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum([len(b) for b in blobs], out=off[1:])
    off.tofile(os.path.join(path, "payload.off"))
    with open(os.path.join(path, "payload.blob"), "wb") as f:
        for b in blobs:
            f.write(b)
    with open(os.path.join(path, "users.json"), "w") as f:
        json.dump(users, f)
    if not int_ids:
        with open(os.path.join(path, "point_ids.json"), "w") as f:
            json.dump(raw_ids, f)
    manifest = {"version": FORMAT_VERSION, "n_rows": n, "dim": dim, "dtype": dtype, "n_users": len(users),
                "categories": cat_values, "ids_ascending": ascending, "original_ids": not int_ids,
                "missing_category": bool(np.any(cats == NO_CATEGORY)), "missing_date": missing_date}
    with open(os.path.join(path, "manifest.json"), "w") as f:
        json.dump(manifest, f, indent=1)
    return manifest


class CorpusFiles:
    """Memory-mapped view of a corpus directory (host side; opens nothing on the GPU)."""

    def __init__(self, path: str):
        self.path = path
        with open(os.path.join(path, "manifest.json")) as f:
            self.manifest = json.load(f)
        if self.manifest.get("version") != FORMAT_VERSION:
            raise ValueError(f"unsupported corpus format version {self.manifest.get('version')!r}")
        m = self.manifest
        self.n_rows, self.dim, self.dtype = int(m["n_rows"]), int(m["dim"]), m["dtype"]
        vdt = np.float32 if self.dtype == "float32" else np.uint16
        self.vectors = np.memmap(os.path.join(path, "vectors.bin"), dtype=vdt, mode="r", shape=(self.n_rows, self.dim))
        self.columns: Dict[str, np.ndarray] = {}
        for name, dt in _COLUMNS.items():
            fp = os.path.join(path, name)
            if os.path.exists(fp):
                self.columns[name] = np.memmap(fp, dtype=dt, mode="r")
        with open(os.path.join(path, "users.json")) as f:
            self.users: List[str] = json.load(f)
        self.offsets = np.array(self.columns["seg_offsets.i64"])
        self.point_ids: Optional[List[Any]] = None
        if m.get("original_ids"):
            with open(os.path.join(path, "point_ids.json")) as f:
                self.point_ids = json.load(f)
        self._blob = np.memmap(os.path.join(path, "payload.blob"), dtype=np.uint8, mode="r") \
            if os.path.getsize(os.path.join(path, "payload.blob")) else np.zeros(0, np.uint8)

    def payload(self, row: int) -> Dict[str, Any]:
        off = self.columns["payload.off"]
        return json.loads(bytes(self._blob[int(off[row]):int(off[row + 1])]).decode("utf-8"))


def load_collection(path: str, engine, *, name: str = "transactions", row_lo: int = 0, row_hi: Optional[int] = None,
                    chunk_rows: int = 1 << 18):
    """Stream rows [row_lo, row_hi) of a corpus directory into HBM (pinned staging, ``chunk_rows`` at a time) and
    attach them to ``engine`` as its shard. Returns the ``client.Collection`` to hand to ``B200QdrantClient``."""
    import torch
    from .client import Collection
    cf = CorpusFiles(path)
    row_hi = cf.n_rows if row_hi is None else row_hi
    n = row_hi - row_lo
    dev = torch.device("cuda", engine.device)
    tdt = torch.float32 if cf.dtype == "float32" else torch.bfloat16
    vec = torch.empty((n, cf.dim), dtype=tdt, device=dev)
    stage = torch.empty((min(chunk_rows, n), cf.dim), dtype=torch.float32 if cf.dtype == "float32" else torch.int16).pin_memory()
    for lo in range(0, n, chunk_rows):
        hi = min(n, lo + chunk_rows)
        src = np.asarray(cf.vectors[row_lo + lo:row_lo + hi])
        stage_np = stage.numpy()[:hi - lo]
        stage_np[...] = src if cf.dtype == "float32" else src.view(np.int16)
        dst = vec[lo:hi] if cf.dtype == "float32" else vec[lo:hi].view(torch.int16)
        dst.copy_(stage[:hi - lo], non_blocking=True)
        torch.cuda.synchronize(dev)                                   # the staging buffer is reused

    def col(fname, view=None):
        if fname not in cf.columns:
            return None
        a = np.array(cf.columns[fname][row_lo:row_hi])                 # copy: the memory map is read-only
        return torch.from_numpy(a.view(view) if view is not None else a).to(dev)
    ids = col("ids.u64", np.int64)
    id_rank = col("id_rank.u32", np.int32)
    if id_rank is not None and (row_lo != 0 or row_hi != cf.n_rows):
        raise ValueError("id_rank is global to the corpus: load the whole corpus or re-export the shard")
    dates = col("dates.i32")
    cats = col("categories.u8")
    cents = col("cents.i32")
    offs = torch.from_numpy(cf.offsets).to(dev)
    engine.load_shard(vec, ids=ids, id_rank=id_rank, dates=dates, categories=cats, cents=cents, seg_offsets=offs,
                      row_base=row_lo)
    # id -> row map over the WHOLE corpus (8 host bytes per row): a sharded search returns points of other shards too
    ids_host = np.array(cf.columns["ids.u64"]) if "ids.u64" in cf.columns else None
    codes = {c: i for i, c in enumerate(cf.manifest["categories"])}
    collection = Collection(name, engine, cf.users, cf.offsets, cf.payload, n_total_rows=cf.n_rows,
                            category_codes=codes, ids_host=ids_host, ids_row_base=0)
    collection.files = cf
    collection.has_missing_category = bool(cf.manifest.get("missing_category", False))
    collection.has_missing_date = bool(cf.manifest.get("missing_date", False))
    if cf.point_ids is not None:                                      # UUID-style ids: hand the original id back to the caller
        collection.point_id_fn = lambda row: cf.point_ids[row]
    return collection
