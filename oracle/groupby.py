"""ORACLE — test infrastructure only.

The reference's aggregation IS a pandas call:
    transactions.groupby(plot_config.group_by)[plot_config.y_axis].sum()
(reference tools/plot_tool.py:45 for ``bar`` and :52 for ``pie``), applied to a
DataFrame parsed by ``pd.read_json(transactions_json)`` (tools/plot_tool.py:29).
pandas is installed here (3.0.2; the reference pins 2.3.0), so this part of the
oracle is the reference's own library routine, not a restatement, and the golden
fixtures under tests/golden/ are its outputs (tests/golden/make_golden.py).

pandas semantics relied on (probed): keys ascending, rows with a
null key dropped, int64 in -> int64 out, only observed groups present.
"""
from __future__ import annotations

import io

import numpy as np
import pandas as pd


def read_transactions_json(transactions_json: str) -> pd.DataFrame:
    """pd.read_json on a literal JSON string (tools/plot_tool.py:29). pandas >= 3
    no longer accepts a bare string, so it is wrapped in StringIO; the parse is the same."""
    return pd.read_json(io.StringIO(transactions_json))


def groupby_sum_json(transactions_json: str, group_by: str, y_axis: str) -> pd.Series:
    """The exact expression of tools/plot_tool.py:45 / :52."""
    transactions = read_transactions_json(transactions_json)
    return transactions.groupby(group_by)[y_axis].sum()


def month_index(dates_s: np.ndarray) -> np.ndarray:
    """UTC calendar month as months since 1970-01 (int32)."""
    m = np.asarray(dates_s, dtype="int64").astype("datetime64[s]").astype("datetime64[M]")
    return m.astype(np.int64).astype(np.int32)


def groupby_user_category_month(users: np.ndarray, categories: np.ndarray, dates_s: np.ndarray,
                                cents: np.ndarray) -> pd.Series:
    """DataFrame.groupby([user, category, month], sort=True)[cents].sum() with int64 cents
    (BASELINE.json config 5 / SURVEY.md §8c)."""
    df = pd.DataFrame({"user": np.asarray(users, dtype=np.int64),
                       "category": np.asarray(categories, dtype=np.int32),
                       "month": month_index(dates_s),
                       "cents": np.asarray(cents, dtype=np.int64)})
    return df.groupby(["user", "category", "month"], sort=True)["cents"].sum()
