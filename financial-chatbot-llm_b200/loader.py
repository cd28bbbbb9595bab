"""Put a corpus into HBM and describe it to the engine.

Two sources: a materialised host corpus (tests, BASELINE config 1) and the on-device
synthetic generator (10M+ rows never touch host memory). PyTorch only allocates and holds
the device buffers; the library reads them through raw pointers.
"""
from __future__ import annotations

from typing import Optional

import numpy as np

from . import synth
from .client import Collection
from .engine import Engine, synth_columns, synth_vectors

CATEGORY_CODES = {name: i for i, name in enumerate(synth.CATEGORIES)}


def _u64_tensor(a: np.ndarray, device):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a, dtype=np.uint64).view(np.int64)).to(device)


def collection_from_host(hc: "synth.HostCorpus", engine: Engine, *, name: str = "transactions",
                         bf16: bool = False, with_ids: bool = True,
                         row_lo: int = 0, row_hi: Optional[int] = None) -> Collection:
    """Upload rows [row_lo, row_hi) of a host corpus as this engine's shard."""
    import torch
    dev = torch.device("cuda", engine.device)
    row_hi = hc.spec.n_rows if row_hi is None else row_hi
    sl = slice(row_lo, row_hi)
    if bf16:
        bits = hc.bf16_bits if hc.bf16_bits is not None else \
            synth.f32_to_bf16_bits(hc.vectors).reshape(hc.vectors.shape)
        vec = torch.from_numpy(np.ascontiguousarray(bits[sl]).view(np.int16)).to(dev).view(torch.bfloat16)
    else:
        vec = torch.from_numpy(np.ascontiguousarray(hc.vectors[sl])).to(dev)
    ids = id_rank = None
    ids_host = None
    if with_ids:
        ids_host = hc.ids[sl]
        ids = _u64_tensor(ids_host, dev)
        if ids_host.size > 1 and not np.all(ids_host[1:] > ids_host[:-1]):
            rank = np.empty(ids_host.size, dtype=np.int64)
            rank[np.argsort(ids_host, kind="stable")] = np.arange(ids_host.size)
            id_rank = torch.from_numpy(rank.astype(np.int32)).to(dev)
    dates = torch.from_numpy(np.ascontiguousarray(hc.dates[sl])).to(dev)
    cats = torch.from_numpy(np.ascontiguousarray(hc.categories[sl])).to(dev)
    cents = torch.from_numpy(np.ascontiguousarray(hc.cents[sl])).to(dev)
    offs = torch.from_numpy(hc.offsets).to(dev)
    engine.load_shard(vec, ids=ids, id_rank=id_rank, dates=dates, categories=cats, cents=cents,
                      seg_offsets=offs, row_base=row_lo)
    # the host-side id -> row map covers the WHOLE corpus (a sharded search returns points of other shards too);
    # only this shard's slice of the ids went to the device
    return Collection(name, engine, hc.users, hc.offsets, hc.payload, n_total_rows=hc.spec.n_rows,
                      category_codes=CATEGORY_CODES, ids_host=hc.ids if with_ids else None, ids_row_base=0)


def collection_synthetic(spec: "synth.SynthSpec", engine: Engine, *, name: str = "transactions",
                         bf16: bool = False, row_lo: int = 0, row_hi: Optional[int] = None,
                         with_cents: bool = True, chunk_rows: int = 1 << 20) -> Collection:
    """Generate rows [row_lo, row_hi) of the synthetic corpus directly in HBM (ids = row)."""
    import torch
    dev = torch.device("cuda", engine.device)
    row_hi = spec.n_rows if row_hi is None else row_hi
    n = row_hi - row_lo
    with torch.cuda.device(dev):
        vec = torch.empty((n, spec.dim), dtype=torch.bfloat16 if bf16 else torch.float32, device=dev)
        for lo in range(0, n, chunk_rows):
            hi = min(n, lo + chunk_rows)
            synth_vectors(vec[lo:hi], row_lo + lo, spec.seed, spec.n_clusters)
        dates = torch.empty(n, dtype=torch.int32, device=dev)
        cats = torch.empty(n, dtype=torch.uint8, device=dev)
        cents = torch.empty(n, dtype=torch.int32, device=dev) if with_cents else None
        synth_columns(row_lo, n, spec.seed, spec.n_categories, spec.date_span_seconds,
                      dates=dates, categories=cats, cents=cents)
        offsets = synth.segment_offsets(spec)
        offs = torch.from_numpy(offsets).to(dev)
        torch.cuda.synchronize(dev)
    engine.load_shard(vec, dates=dates, categories=cats, cents=cents, seg_offsets=offs, row_base=row_lo)
    users = synth.user_ids(spec.n_users)
    return Collection(name, engine, users, offsets,
                      lambda row: synth.payload_for_row(spec, offsets, users, row),
                      n_total_rows=spec.n_rows, category_codes=CATEGORY_CODES)
