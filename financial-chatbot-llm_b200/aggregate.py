"""Host-side mirror of the aggregation in reference tools/plot_tool.py.

The reference's ``create_financial_plot(transactions_json, plot_config)`` (tools/plot_tool.py:16-78)
parses a JSON string with pandas (:29), computes
``transactions.groupby(plot_config.group_by)[plot_config.y_axis].sum()`` for ``bar`` (:45) and
``pie`` (:52), and draws the result with matplotlib (:32-75). Only the group-by sum is on the hot
path; the drawing is serial CPU work over a handful of groups and is out of
scope (matplotlib is not installed here).

Two entry levels:
  * ``groupby_user_category_month``  — the bulk path: (user, category, UTC month) sums over the
    device-resident columns of a loaded shard (BASELINE config 5), in pandas' output order.
  * ``grouped_sum_json`` / ``create_financial_plot`` — the tool-shaped path: same arguments as the
    reference tool; the key column becomes the CSR segment, the value column integer cents, and
    the same CUDA kernel does the reduction. Values must be integers or exact cents (x.xx): the
    reference sums float64, this engine sums integers exactly — a deliberate strengthening.
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


def create_financial_plot(transactions_json: str, plot_config: PlotConfig) -> str:
    """Tool-shaped entry (reference tools/plot_tool.py:16-78) for the aggregating plot types.
    The group-by runs on the GPU; the figure itself is drawn only when matplotlib is importable,
    otherwise the grouped series is returned as a JSON data URI. Errors come back as the
    reference's ``"Error creating plot: ..."`` string (:77-78)."""
    try:
        if plot_config.plot_type in ("bar", "pie") and plot_config.group_by and plot_config.y_axis:
            grouped = grouped_sum_json(transactions_json, plot_config.group_by, plot_config.y_axis)
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
        raise NotImplementedError(f"plot_type {plot_config.plot_type!r} without group_by/y_axis has no "
                                  "reduction and is not part of the GPU path")
    except Exception as e:
        return f"Error creating plot: {str(e)}"
