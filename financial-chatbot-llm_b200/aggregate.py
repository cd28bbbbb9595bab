"""Host-side mirror of the aggregation in reference tools/plot_tool.py.

The reference's ``create_financial_plot(transactions_json, plot_config)`` (tools/plot_tool.py:16-78)
parses a JSON string with pandas (:29), computes
``transactions.groupby(plot_config.group_by)[plot_config.y_axis].sum()`` for ``bar`` (:45) and
``pie`` (:52), and draws the result with matplotlib (:32-75). Only the group-by sum is on the hot
path; the drawing is serial CPU work over a handful of groups and is out of
scope (matplotlib is not installed here).

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

from .engine import Engine


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
    uniq, inv = np.unique(keys.to_numpy()[keep], return_inverse=True)     # ascending, like sort=True
    v, scale = _to_integer_units(y_all[keep])
    order = np.argsort(inv, kind="stable")
    offsets = np.zeros(len(uniq) + 1, dtype=np.int64)
    np.cumsum(np.bincount(inv, minlength=len(uniq)), out=offsets[1:])
    n = int(v.size)
    if n == 0:
        return pd.Series([], index=pd.Index([], name=group_by), name=y_axis, dtype=np.int64)
    dev = torch.device("cuda", device)
    eng = Engine(device=device, max_batch=1, max_k=1)
    try:
        eng.load_shard(torch.zeros((n, 4), dtype=torch.float32, device=dev),
                       dates=torch.zeros(n, dtype=torch.int32, device=dev),
                       categories=torch.zeros(n, dtype=torch.uint8, device=dev),
                       cents=torch.from_numpy(v[order]).to(dev), seg_offsets=torch.from_numpy(offsets).to(dev))
        sums, present, dropped = eng.aggregate(0, len(uniq), 0, 1, 1)
        torch.cuda.synchronize(dev)
        assert int(dropped.item()) == 0 and bool(present.all())
        out = sums.view(-1).cpu().numpy()
    finally:
        eng.close()
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
    code_name = {c: n for n, c in collection.category_codes.items()}
    rows = []                                                # (key tuple, cents) of the observed cells
    for c, m in zip(*np.nonzero(present)):
        rows.append(({"category": code_name.get(int(c), int(c)), "month": month_label(month_lo + int(m))}, int(sums[c, m])))
    frame = pd.DataFrame([{**kv, y_axis: v} for kv, v in rows], columns=["category", "month", y_axis])
    frame[y_axis] = frame[y_axis].astype(np.int64)
    out = frame.groupby(keys if len(keys) > 1 else keys[0])[y_axis].sum()    # folds the cells of the key not grouped on
    return out if y_axis == "amount_cents" else out.astype(np.float64) / 100.0


def _render(grouped, plot_config: PlotConfig) -> str:
    """The drawing half of tools/plot_tool.py:43-75 for an already grouped series (a JSON data URI when
    matplotlib is not importable)."""
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
    except ImportError:
        body = json.dumps({"title": plot_config.title, "plot_type": plot_config.plot_type,
                           "labels": [str(x) for x in grouped.index.tolist()],
                           "values": [float(x) for x in grouped.tolist()]})
        return "data:application/json;base64," + base64.b64encode(body.encode()).decode()
    plt.figure(figsize=(10, 6))
    if plot_config.plot_type == "bar":
        grouped.plot(kind="bar")
    else:
        plt.pie(grouped, labels=grouped.index.astype(str).tolist(), autopct="%1.1f%%")
    plt.title(plot_config.title)
    plt.tight_layout()
    buf = io.BytesIO()
    plt.savefig(buf, format="png")
    plt.close()
    return "data:image/png;base64," + base64.b64encode(buf.getvalue()).decode("utf-8")


def create_financial_plot(transactions_json: str, plot_config: PlotConfig) -> str:
    """Tool-shaped entry (reference tools/plot_tool.py:16-78) for the aggregating plot types.
    The group-by runs on the GPU; the figure itself is drawn only when matplotlib is importable,
    otherwise the grouped series is returned as a JSON data URI. Errors come back as the
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
