"""Host-side mirror of the aggregation in reference tools/plot_tool.py.

The reference's ``create_financial_plot(transactions_json, plot_config)`` (tools/plot_tool.py:16-78)
parses a JSON string with pandas (:29), computes
``transactions.groupby(plot_config.group_by)[plot_config.y_axis].sum()`` for ``bar`` (:45) and
``pie`` (:52), and draws the result with matplotlib (:32-75). Only the group-by sum is on the hot
path; the drawing is serial CPU work over a handful of groups and is out of
scope: the tool-shaped entries here return the grouped series as a JSON data URI instead of a PNG.

Three entry levels:
  * ``groupby_user_category_month``  — the bulk path: (user, category, UTC month) sums over the
    device-resident columns of a loaded shard (BASELINE config 5), in pandas' output order.
  * ``grouped_sum_json`` / ``create_financial_plot`` — the tool-shaped path: same arguments as the
    reference tool; the key column becomes the CSR segment, the value column integer cents, and
    the same CUDA kernel does the reduction. Values must be integers or exact cents (x.xx): the
    reference sums float64, this engine sums integers exactly — a deliberate strengthening.
  * ``grouped_sum_retrieved`` / ``plot_retrieved`` — retrieval and group-by fused on the device (SURVEY.md
    §8f.4): the agent's round trip retrieve_transactions -> JSON -> create_financial_plot collapses into one
    search launch plus one tiny reduction launch over the returned points; nothing but the grouped sums
    comes back to the host.
"""
from __future__ import annotations

import base64
import io
import json
from typing import Literal, Optional

import numpy as np
from pydantic import BaseModel, Field

from .engine import Engine, aggregate_columns


class PlotConfig(BaseModel):
    """Same fields as reference tools/plot_tool.py:9-14."""
    plot_type: Literal["line", "bar", "pie", "scatter", "histogram"] = Field(description="Type of plot to create")
    x_axis: str = Field(description="Column for x-axis")
    y_axis: Optional[str] = Field(description="Column for y-axis", default=None)
    title: str = Field(description="Plot title")
    group_by: Optional[str] = Field(description="Column to group by", default=None)


def groupby_user_category_month(engine: Engine, seg_begin: int, seg_end: int, month_lo: int, n_months: int,
                                n_categories: int):
    """Sparse (user, category, month, sum) arrays in ascending key order — exactly the index order
    of ``DataFrame.groupby([user, category, month], sort=True)[cents].sum()`` (observed groups only).
    ``month`` is months since 1970-01 (UTC). Raises if any row fell outside the dense window."""
    import torch
    sums, present, dropped = engine.aggregate(seg_begin, seg_end, month_lo, n_months, n_categories)
    torch.cuda.synchronize(engine.device)
    if int(dropped.item()) != 0:
        raise ValueError(f"{int(dropped.item())} rows fall outside the category/month window")
    idx = torch.nonzero(present, as_tuple=False)               # row-major => ascending (user, category, month)
    vals = sums[present.bool()]
    idx = idx.cpu().numpy()
    return (idx[:, 0].astype(np.int64) + seg_begin, idx[:, 1].astype(np.int32),
            idx[:, 2].astype(np.int32) + month_lo, vals.cpu().numpy())


def _to_integer_units(y: np.ndarray):
    """Value column -> (int32 units, scale): integers as they are; floats must be exact cents."""
    if np.issubdtype(y.dtype, np.integer) or y.dtype == bool:
        v, scale = y.astype(np.int64), 1
    else:
        c = np.rint(y.astype(np.float64) * 100.0)
        if not np.allclose(c / 100.0, y, rtol=0, atol=1e-9):
            raise ValueError("y_axis values are neither integers nor exact cents; the engine sums integers")
        v, scale = c.astype(np.int64), 100
    if v.size and (v.min() < -(2 ** 31) or v.max() >= 2 ** 31):
        raise ValueError("y_axis values do not fit int32 units")
    return v.astype(np.int32), scale


def grouped_sum_json(transactions_json: str, group_by: str, y_axis: str, device: int = 0):
    """``pd.read_json(transactions_json).groupby(group_by)[y_axis].sum()`` with the reduction on the GPU.
    Returns a pandas Series (index = ascending keys, null keys dropped) like the reference's."""
    import pandas as pd
    import torch
    transactions = pd.read_json(io.StringIO(transactions_json))          # tools/plot_tool.py:29
    keys = transactions[group_by]
    keep = keys.notna().to_numpy()                                        # groupby drops null keys
    y_all = transactions[y_axis].to_numpy()
    inv, uniq = pd.factorize(keys.to_numpy()[keep], sort=True)            # ascending keys, like groupby's sort=True (hash-based:
    uniq = np.asarray(uniq)                                               #   several times cheaper than np.unique on strings)
    v, scale = _to_integer_units(y_all[keep])
    order = np.argsort(inv, kind="stable")
    offsets = np.zeros(len(uniq) + 1, dtype=np.int64)
    np.cumsum(np.bincount(inv, minlength=len(uniq)), out=offsets[1:])
    n = int(v.size)
    if n == 0:
        return pd.Series([], index=pd.Index([], name=group_by), name=y_axis, dtype=np.int64)
    # the key column becomes the CSR segments, the value column the integer units: one engine-less call, no vectors
    dev = torch.device("cuda", device)
    zeros_i = torch.zeros(n, dtype=torch.int32, device=dev)               # one month, one category: the window is 1 x 1
    zeros_b = torch.zeros(n, dtype=torch.uint8, device=dev)
    sums, present, dropped = aggregate_columns(device, zeros_i, zeros_b, torch.from_numpy(v[order]).to(dev),
                                               torch.from_numpy(offsets).to(dev), 0, len(uniq), 0, 1, 1)
    torch.cuda.synchronize(dev)
    if int(dropped.item()) != 0 or not bool(present.all()):
        raise RuntimeError("fcb_aggregate_columns dropped rows of a 1 x 1 window")
    out = sums.view(-1).cpu().numpy()
    data = out if scale == 1 else out.astype(np.float64) / scale
    return pd.Series(data, index=pd.Index(uniq, name=group_by), name=y_axis)


def month_index(ts: int) -> int:
    """Months since 1970-01 of a unix timestamp (UTC) — the device kernels' month key."""
    import datetime as _dt
    d = _dt.datetime.fromtimestamp(int(ts), _dt.timezone.utc)
    return (d.year - 1970) * 12 + d.month - 1


def month_label(mi: int) -> str:
    return "%04d-%02d" % (1970 + mi // 12, mi % 12 + 1)


def grouped_sum_retrieved(collection, query, query_filter, limit: int, group_by, y_axis: str = "amount_cents", *,
                          month_lo: Optional[int] = None, n_months: Optional[int] = None):
    """``groupby(group_by)[y_axis].sum()`` over the ``limit`` best points of a filtered search, without the
    points leaving the GPU. ``group_by``: "category", "month" ("YYYY-MM", UTC) or ["category", "month"];
    ``y_axis``: "amount_cents" (int64) or "amount" (cents / 100). The month window defaults to the filter's
    date range (at most 240 months back from its upper bound, or from now). Returns the pandas Series the
    reference's group-by would give for the retrieved transactions (ascending keys, observed groups only)."""
    import time
    import pandas as pd
    import torch
    from .client import parse_filter
    keys = [group_by] if isinstance(group_by, str) else list(group_by)
    if not keys or any(key not in ("category", "month") for key in keys) or len(set(keys)) != len(keys):
        raise NotImplementedError(f"group_by {group_by!r}: the device columns offer 'category' and 'month'")
    if y_axis not in ("amount_cents", "amount"):
        raise NotImplementedError(f"y_axis {y_axis!r}: the device value column is the amount in integer cents")
    eng = collection.engine
    pf = parse_filter(query_filter)                         # one conjunction, like the reference's own filter
    filt = collection.resolve(pf)
    names = sorted(collection.category_codes, key=collection.category_codes.get)
    n_cat = max(collection.category_codes.values()) + 1 if names else 1
    if n_months is None or month_lo is None:
        hi = month_index(pf.date_hi if pf.date_hi is not None else time.time())
        lo = month_index(pf.date_lo) if pf.date_lo is not None else hi - 239
        lo = max(lo, hi - min(239, 4096 // n_cat - 1))
        month_lo, n_months = lo, hi - lo + 1
    dev = torch.device("cuda", eng.device)
    k = int(limit)
    q = torch.from_numpy(np.ascontiguousarray(np.asarray(query, dtype=np.float64)[None, :])).to(dev)
    ids = torch.empty((1, k), dtype=torch.int64, device=dev)
    scores = torch.empty((1, k), dtype=torch.float64, device=dev)
    counts = torch.empty(1, dtype=torch.int32, device=dev)
    eng.search_device(q, [filt], k, ids, scores, counts)
    sums, present, stats = eng.aggregate_hits(ids, counts, month_lo, n_months, n_cat)
    sums, present, stats = sums[0].cpu().numpy(), present[0].cpu().numpy().astype(bool), stats.cpu().numpy()
    if int(stats[0]):
        raise ValueError(f"{int(stats[0])} retrieved rows fall outside the category/month window")
    # fold the observed (category, month) cells along the key that is not grouped on — exact integer arithmetic in numpy,
    # observed groups only, keys ascending: what pandas' groupby(...).sum() yields for the retrieved payloads
    code_name = {c: n for n, c in collection.category_codes.items()}
    cc, mm = np.nonzero(present)
    cat_labels = np.array([code_name.get(int(c), int(c)) for c in range(sums.shape[0])], dtype=object)
    month_labels = np.array([month_label(month_lo + m) for m in range(sums.shape[1])], dtype=object)
    if len(keys) == 2:
        tuples = [(cat_labels[c], month_labels[m]) if keys[0] == "category" else (month_labels[m], cat_labels[c]) for c, m in zip(cc, mm)]
        order = sorted(range(len(tuples)), key=lambda i: tuples[i])
        index = pd.MultiIndex.from_tuples([tuples[i] for i in order], names=keys)
        data = np.array([int(sums[cc[i], mm[i]]) for i in order], dtype=np.int64)
    else:
        axis = 1 if keys[0] == "category" else 0                          # sum over the OTHER key
        seen = present.any(axis=axis)
        labels = (cat_labels if keys[0] == "category" else month_labels)[seen]
        totals = np.where(present, sums, 0).sum(axis=axis, dtype=np.int64)[seen]
        order = np.argsort(labels.astype(str), kind="stable")
        index = pd.Index(labels[order], name=keys[0])
        data = totals[order]
    out = pd.Series(data, index=index, name=y_axis, dtype=np.int64)
    return out if y_axis == "amount_cents" else out.astype(np.float64) / 100.0


def _render(grouped, plot_config: PlotConfig) -> str:
    """What stands in for the drawing half of tools/plot_tool.py:43-75: the grouped series as a JSON data URI (title,
    plot type, labels, values). Rendering a figure is out of scope."""
    body = json.dumps({"title": plot_config.title, "plot_type": plot_config.plot_type,
                       "labels": [str(x) for x in grouped.index.tolist()],
                       "values": [float(x) for x in grouped.tolist()]})
    return "data:application/json;base64," + base64.b64encode(body.encode()).decode()


def create_financial_plot(transactions_json: str, plot_config: PlotConfig) -> str:
    """Tool-shaped entry (reference tools/plot_tool.py:16-78) for the aggregating plot types.
    The group-by runs on the GPU; the grouped series comes back as a JSON data URI (no figure is drawn: out of
    scope). Errors come back as the
    reference's ``"Error creating plot: ..."`` string (:77-78)."""
    try:
        if plot_config.plot_type in ("bar", "pie") and plot_config.group_by and plot_config.y_axis:
            return _render(grouped_sum_json(transactions_json, plot_config.group_by, plot_config.y_axis), plot_config)
        raise NotImplementedError(f"plot_type {plot_config.plot_type!r} without group_by/y_axis has no "
                                  "reduction and is not part of the GPU path")
    except Exception as e:
        return f"Error creating plot: {str(e)}"


def plot_retrieved(collection, query, query_filter, limit: int, plot_config: PlotConfig) -> str:
    """``create_financial_plot`` over the transactions a retrieval selects, with both steps on the device:
    same ``PlotConfig`` and same return convention, but the first argument is the retrieval (collection, query
    vector, qdrant filter, limit) instead of its JSON."""
    try:
        if plot_config.plot_type in ("bar", "pie") and plot_config.group_by and plot_config.y_axis:
            return _render(grouped_sum_retrieved(collection, query, query_filter, limit, plot_config.group_by,
                                                 plot_config.y_axis), plot_config)
        raise NotImplementedError(f"plot_type {plot_config.plot_type!r} without group_by/y_axis has no "
                                  "reduction and is not part of the GPU path")
    except Exception as e:
        return f"Error creating plot: {str(e)}"
