"""Aggregation kernel (fcb_aggregate) against pandas (the reference's own library call,
tools/plot_tool.py:45,52), the C restatement and the committed golden fixtures. Integer sums:
bit-exact."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def make_engine(spec, cents=None, dates=None, cats=None, offsets=None):
    import torch
    from financial_chatbot_llm_b200 import synth
    from financial_chatbot_llm_b200.engine import Engine
    n = spec.n_rows
    dev = torch.device("cuda", 0)
    offsets = synth.segment_offsets(spec) if offsets is None else offsets
    d = synth.dates(spec, 0, n) if dates is None else dates
    c = synth.categories(spec, 0, n) if cats is None else cats
    v = synth.cents(spec, 0, n) if cents is None else cents
    eng = Engine(device=0, max_batch=1, max_k=1)
    eng.load_shard(torch.zeros((n, 4), dtype=torch.float32, device=dev), dates=torch.from_numpy(d).to(dev),
                   categories=torch.from_numpy(c).to(dev), cents=torch.from_numpy(v).to(dev),
                   seg_offsets=torch.from_numpy(offsets).to(dev))
    return eng, offsets, d, c, v


def month_window(d):
    from oracle import groupby
    m = groupby.month_index(d)
    return int(m.min()), int(m.max() - m.min() + 1)


def test_user_category_month_equals_pandas_golden(built):
    from financial_chatbot_llm_b200 import aggregate, synth
    g = json.load(open(os.path.join(HERE, "golden", "groupby_golden.json")))["columns"]
    spec = synth.SynthSpec(**g["spec"])
    eng, off, d, c, v = make_engine(spec)
    mlo, nm = month_window(d)
    u, cat, mo, s = aggregate.groupby_user_category_month(eng, 0, spec.n_users, mlo, nm, 16)
    assert u.tolist() == g["user"] and cat.tolist() == g["category"] and mo.tolist() == g["month"]
    assert s.tolist() == g["sum"]
    # a sub-range of users
    u2, _, _, s2 = aggregate.groupby_user_category_month(eng, 5, 9, mlo, nm, 16)
    sel = (np.array(g["user"]) >= 5) & (np.array(g["user"]) < 9)
    assert u2.tolist() == np.array(g["user"])[sel].tolist() and s2.tolist() == np.array(g["sum"])[sel].tolist()
    eng.close()


def test_ragged_users_extreme_values_and_zero_sums(built):
    """Tiny users (global-atomic path), a whale (shared-memory path, many tiles), full-range int32
    values (lo/hi split), groups whose sum is exactly 0, dates before 1970."""
    from financial_chatbot_llm_b200 import aggregate, synth
    from oracle import canon, groupby
    rng = np.random.default_rng(5)
    sizes = np.array([1, 2, 3, 50000, 95, 96, 97, 4096, 4097, 1, 20000, 7], dtype=np.int64)
    off = np.zeros(len(sizes) + 1, dtype=np.int64)
    np.cumsum(sizes, out=off[1:])
    n = int(off[-1])
    spec = synth.SynthSpec(n_rows=n, dim=4, n_users=len(sizes))
    v = rng.integers(-(2 ** 31), 2 ** 31, size=n, dtype=np.int64).astype(np.int32)
    v[::7] = 2 ** 31 - 1
    v[1::7] = -(2 ** 31)
    v[int(off[4]):int(off[5])] = np.where(np.arange(95) % 2 == 0, 777, -777)[:95]
    v[int(off[4])] = 0                                          # 47 x +777, 47 x -777, one 0 -> group sums can be 0
    d = rng.integers(-86400 * 400, 86400 * 400, size=n).astype(np.int32)     # 1968-11 .. 1971-02
    c = rng.integers(0, 5, size=n).astype(np.uint8)
    c[int(off[4]):int(off[5])] = 2
    d[int(off[4]):int(off[5])] = 86400 * 45
    eng, *_ = make_engine(spec, cents=v, dates=d, cats=c, offsets=off)
    mlo, nm = month_window(d)
    u, cat, mo, s = aggregate.groupby_user_category_month(eng, 0, len(sizes), mlo, nm, 5)
    ser = groupby.groupby_user_category_month(synth.row_users(off, np.arange(n)), c, d, v)
    idx = ser.index.to_frame(index=False)
    assert u.tolist() == idx["user"].tolist() and cat.tolist() == idx["category"].tolist()
    assert mo.tolist() == idx["month"].tolist() and s.tolist() == ser.tolist()
    cu, cc, cm, cs, _ = canon.groupby(off, d, c, v)
    assert s.tolist() == cs.tolist() and mo.tolist() == cm.tolist()
    assert 0 in s[u == 4].tolist()                              # the zero-sum group is present
    # window too small -> rows are reported as dropped, never silently lost
    with pytest.raises(ValueError, match="outside"):
        aggregate.groupby_user_category_month(eng, 0, len(sizes), mlo + 1, nm - 1, 5)
    with pytest.raises(ValueError, match="outside"):
        aggregate.groupby_user_category_month(eng, 0, len(sizes), mlo, nm, 4)
    eng.close()


def test_size_independent_properties_at_scale(built):
    """20M rows generated on the device: total of all cells == independent torch reduction of the
    column; per-user totals == segment sums; a random sample of users == host pandas on regenerated rows."""
    import torch
    from financial_chatbot_llm_b200 import synth
    from financial_chatbot_llm_b200.engine import Engine, synth_columns
    from oracle import groupby
    n, users = 20_000_000, 2000
    spec = synth.SynthSpec(n_rows=n, dim=4, n_users=users, seed=4242)
    dev = torch.device("cuda", 0)
    d = torch.empty(n, dtype=torch.int32, device=dev)
    c = torch.empty(n, dtype=torch.uint8, device=dev)
    v = torch.empty(n, dtype=torch.int32, device=dev)
    synth_columns(0, n, spec.seed, spec.n_categories, spec.date_span_seconds, dates=d, categories=c, cents=v)
    off = synth.segment_offsets(spec)
    eng = Engine(device=0, max_batch=1, max_k=1)
    eng.load_shard(torch.zeros((n, 4), dtype=torch.float32, device=dev), dates=d, categories=c, cents=v,
                   seg_offsets=torch.from_numpy(off).to(dev))
    mlo = int(groupby.month_index(np.array([synth.DATE_BASE]))[0])
    sums, present, dropped = eng.aggregate(0, users, mlo, 24, 16)
    torch.cuda.synchronize()
    assert int(dropped.item()) == 0
    assert int(sums.sum().item()) == int(v.to(torch.int64).sum().item())
    csum = torch.cat([torch.zeros(1, dtype=torch.int64, device=dev), v.to(torch.int64).cumsum(0)])
    offs_d = torch.from_numpy(off).to(dev)
    assert torch.equal(sums.sum(dim=(1, 2)), csum[offs_d[1:]] - csum[offs_d[:-1]])
    for u in (0, 1, 17, 999, 1999):
        lo, hi = int(off[u]), int(off[u + 1])
        ser = groupby.groupby_user_category_month(np.zeros(hi - lo, dtype=np.int64), synth.categories(spec, lo, hi),
                                                  synth.dates(spec, lo, hi), synth.cents(spec, lo, hi))
        idx = ser.index.to_frame(index=False)
        got = sums[u].cpu().numpy()
        pres = present[u].cpu().numpy().astype(bool)
        assert int(pres.sum()) == len(ser)
        assert got[idx["category"].to_numpy(), idx["month"].to_numpy() - mlo].tolist() == ser.tolist()
    eng.close()


def test_tool_shaped_json_groupby_equals_pandas_golden(built):
    from financial_chatbot_llm_b200 import aggregate
    g = json.load(open(os.path.join(HERE, "golden", "groupby_golden.json")))
    for key, want in g["json_groupby"].items():
        gb, y = key.split("|")
        s = aggregate.grouped_sum_json(g["transactions_json"], gb, y)
        assert [x if isinstance(x, str) else int(x) for x in s.index.tolist()] == want["index"]
        if want["dtype"] == "int64":
            assert s.dtype == np.int64 and s.tolist() == want["values"]
        else:
            assert np.allclose(s.to_numpy(), want["values"], rtol=0, atol=1e-9)   # float cents: exact to 1e-9
    out = aggregate.create_financial_plot(g["transactions_json"], aggregate.PlotConfig(
        plot_type="bar", x_axis="category", y_axis="amount_cents", title="t", group_by="category"))
    assert out.startswith("data:")
    err = aggregate.create_financial_plot("not json", aggregate.PlotConfig(plot_type="bar", x_axis="a", y_axis="b", title="t", group_by="a"))
    assert err.startswith("Error creating plot:")


@pytest.mark.parametrize("kernel", ["default", "sorted"])
def test_time_ordered_log_uses_the_atomic_free_kernel_and_matches_pandas(built, monkeypatch, kernel):
    """Dates non-decreasing within every user (the load-time check sees it). An engine shard is aggregated from its group
    keys by default; the date columns alone (FCB_AGG_KERNEL=sorted here, fcb_aggregate_columns in production) take
    aggregate_sorted_kernel, the atomic-free kernel for ordered logs. Same bit-exact answer as pandas either way; the
    device generator of ordered dates equals the host recipe."""
    if kernel == "sorted":
        monkeypatch.setenv("FCB_AGG_KERNEL", "sorted")
    import torch
    from financial_chatbot_llm_b200 import aggregate, synth
    from financial_chatbot_llm_b200.engine import synth_dates_sorted
    from oracle import groupby
    sizes = np.array([1, 5, 40000, 3, 129, 4096, 4100, 2, 30000, 127, 128, 7000], dtype=np.int64)
    off = np.zeros(len(sizes) + 1, dtype=np.int64)
    np.cumsum(sizes, out=off[1:])
    n = int(off[-1])
    spec = synth.SynthSpec(n_rows=n, dim=4, n_users=len(sizes), seed=77)
    d = synth.dates_sorted(spec, off, 0, n)
    dd = torch.empty(n - 100, dtype=torch.int32, device="cuda")
    synth_dates_sorted(dd, torch.from_numpy(off).cuda(), 100, spec.date_span_seconds)
    assert np.array_equal(dd.cpu().numpy(), d[100:])
    rng = np.random.default_rng(8)
    v = rng.integers(-(2 ** 31), 2 ** 31, size=n, dtype=np.int64).astype(np.int32)
    c = rng.integers(0, 16, size=n).astype(np.uint8)
    eng, *_ = make_engine(spec, cents=v, dates=d, cats=c, offsets=off)
    assert eng.dates_sorted
    mlo, nm = month_window(d)
    u, cat, mo, s = aggregate.groupby_user_category_month(eng, 0, len(sizes), mlo, nm, 16)
    ser = groupby.groupby_user_category_month(synth.row_users(off, np.arange(n)), c, d, v)
    idx = ser.index.to_frame(index=False)
    assert u.tolist() == idx["user"].tolist() and cat.tolist() == idx["category"].tolist()
    assert mo.tolist() == idx["month"].tolist() and s.tolist() == ser.tolist()
    u2, _, _, s2 = aggregate.groupby_user_category_month(eng, 2, 9, mlo, nm, 16)
    sel = (u >= 2) & (u < 9)
    assert u2.tolist() == u[sel].tolist() and s2.tolist() == s[sel].tolist()
    eng.close()


def test_atomic_free_kernel_is_correct_for_any_row_order(built, monkeypatch):
    """Forced onto shuffled dates the ordered-log kernel is slower, never wrong."""
    from financial_chatbot_llm_b200 import aggregate, synth
    from oracle import groupby
    spec = synth.SynthSpec(n_rows=60000, dim=4, n_users=11, seed=78)
    eng, off, d, c, v = make_engine(spec)
    assert not eng.dates_sorted
    mlo, nm = month_window(d)
    monkeypatch.setenv("FCB_AGG_KERNEL", "sorted")
    u, cat, mo, s = aggregate.groupby_user_category_month(eng, 0, spec.n_users, mlo, nm, 16)
    monkeypatch.delenv("FCB_AGG_KERNEL")
    ser = groupby.groupby_user_category_month(synth.row_users(off, np.arange(spec.n_rows)), c, d, v)
    idx = ser.index.to_frame(index=False)
    assert u.tolist() == idx["user"].tolist() and cat.tolist() == idx["category"].tolist()
    assert mo.tolist() == idx["month"].tolist() and s.tolist() == ser.tolist()
    eng.close()


@pytest.mark.parametrize("id_layout", ["rows", "ascending", "permuted"])
def test_group_by_over_retrieved_points_equals_pandas_on_their_payloads(built, id_layout):
    """Retrieval + group-by fused on the device (fcb_aggregate_hits) vs the reference's round trip: take the
    payloads query_points returns, build the DataFrame, groupby(...)[...].sum() (tools/plot_tool.py:29,45)."""
    import dataclasses
    import pandas as pd
    from financial_chatbot_llm_b200 import aggregate, loader, synth
    from financial_chatbot_llm_b200 import qdrant_models as m
    from financial_chatbot_llm_b200.client import B200QdrantClient
    from financial_chatbot_llm_b200.engine import Engine
    spec = synth.SynthSpec(n_rows=20000, dim=64, n_users=7, seed=31)
    hc = synth.make_host_corpus(spec)
    if id_layout == "ascending":
        hc = dataclasses.replace(hc, ids=(np.arange(spec.n_rows, dtype=np.uint64) * np.uint64(3) + np.uint64(10 ** 9)))
    elif id_layout == "permuted":
        hc = dataclasses.replace(hc, ids=np.random.default_rng(5).permutation(spec.n_rows).astype(np.uint64) + np.uint64(77))
    eng = Engine(device=0, max_batch=1, max_k=2000)
    col = loader.collection_from_host(hc, eng, with_ids=id_layout != "rows")
    cl = B200QdrantClient()
    cl.attach_collection(col)
    t0 = int(np.percentile(hc.dates, 35))
    for u, limit in [(2, 10), (4, 700), (0, 2000)]:
        q = synth.query_vector(spec, 90 + u)
        flt = m.Filter(must=[m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value=hc.users[u])),
                             m.FieldCondition(key="metadata.date", range=m.Range(gte=t0))])
        pts = cl.query_points("transactions", query=q.tolist(), limit=limit, query_filter=flt).points
        assert len(pts) == min(limit, int(np.sum(hc.dates[hc.offsets[u]:hc.offsets[u + 1]] >= t0)))
        frame = pd.DataFrame([p.payload["metadata"] for p in pts])
        frame["month"] = pd.to_datetime(frame["date"], unit="s", utc=True).dt.strftime("%Y-%m")
        frame["amount"] = frame["amount_cents"] / 100.0
        for keys in ("category", "month", ["category", "month"], ["month", "category"]):
            want = frame.groupby(keys)["amount_cents"].sum()
            got = aggregate.grouped_sum_retrieved(col, q, flt, limit, keys, "amount_cents")
            assert got.index.tolist() == want.index.tolist() and got.tolist() == want.tolist(), (id_layout, u, keys)
        got = aggregate.grouped_sum_retrieved(col, q, flt, limit, "category", "amount")
        want = frame.groupby("category")["amount_cents"].sum() / 100.0
        assert got.index.tolist() == want.index.tolist() and got.tolist() == want.tolist()
    # the tool-shaped wrapper: same return convention as create_financial_plot
    out = aggregate.plot_retrieved(col, q, flt, 50, aggregate.PlotConfig(plot_type="bar", x_axis="category",
                                                                        y_axis="amount", title="t", group_by="category"))
    assert out.startswith("data:")
    assert aggregate.plot_retrieved(col, q, flt, 50, aggregate.PlotConfig(plot_type="bar", x_axis="x", y_axis="amount",
                                                                          title="t", group_by="merchant")).startswith("Error creating plot")
    # an unknown user retrieves nothing: empty series, no error
    none = m.Filter(must=[m.FieldCondition(key="metadata.user_id", match=m.MatchValue(value="nobody"))])
    assert len(aggregate.grouped_sum_retrieved(col, q, none, 10, "category")) == 0
    eng.close()


def test_any_order_kernels_on_awkward_shapes_equal_pandas(built, monkeypatch):
    """The two-launch split of any-order data (users under 8192 rows: warp tables; the others: CTA-wide shared cells) on
    awkward shapes: user starts that are not multiples of 4 (ragged quads), sizes around the 8192-row threshold and the
    4096-row tile, a user spanning many tiles, rows outside the category / month window (counted as dropped), full-range
    int32 values; the warp-table kernel (the default for date columns now) and the single-launch kernel (FCB_AGG_SPLIT=0) must give the identical
    dense output; all equal pandas; and
    the engine-less entry point works on a sub-range of users over columns that start in the middle (row_base > 0)."""
    import torch
    from financial_chatbot_llm_b200 import synth
    from financial_chatbot_llm_b200.engine import aggregate_columns
    from oracle import groupby
    rng = np.random.default_rng(11)
    sizes = np.array([3, 8191, 8192, 8193, 1, 4096 * 2 + 1, 200_001, 17, 8192 + 4097, 70_000, 5, 12_345], dtype=np.int64)
    off = np.zeros(len(sizes) + 1, dtype=np.int64)
    np.cumsum(sizes, out=off[1:])
    n = int(off[-1])
    v = rng.integers(-(2 ** 31), 2 ** 31, size=n, dtype=np.int64).astype(np.int32)
    v[::5] = rng.integers(-5000, 5000, size=v[::5].size)
    d = (synth.DATE_BASE + rng.integers(0, 730 * 86400, size=n)).astype(np.int32)
    c = rng.integers(0, 16, size=n).astype(np.uint8)
    c[rng.integers(0, n, size=300)] = 200                       # unknown category -> dropped
    d[rng.integers(0, n, size=300)] = synth.DATE_BASE - 86400   # before the window -> dropped
    dev = torch.device("cuda", 0)
    dd, cc, vv, oo = (torch.from_numpy(a).to(dev) for a in (d, c, v, off))
    mlo = int(groupby.month_index(np.array([synth.DATE_BASE]))[0])
    res = {}
    # "warp": what date columns get by default (the date form of the warp-table kernel); "split" / "single": round 1's kernels
    for mode in ("warp", "split", "single"):
        if mode != "warp":
            monkeypatch.setenv("FCB_AGG_KERNEL", "cta")
        if mode == "single":
            monkeypatch.setenv("FCB_AGG_SPLIT", "0")
        sums, present, dropped = aggregate_columns(0, dd, cc, vv, oo, 0, len(sizes), mlo, 24, 16)
        torch.cuda.synchronize()
        res[mode] = (sums.cpu().numpy(), present.cpu().numpy(), int(dropped.item()))
    monkeypatch.delenv("FCB_AGG_SPLIT")
    monkeypatch.delenv("FCB_AGG_KERNEL")
    assert np.array_equal(res["split"][0], res["single"][0]) and np.array_equal(res["split"][1], res["single"][1])
    assert np.array_equal(res["split"][0], res["warp"][0]) and np.array_equal(res["split"][1], res["warp"][1])
    keep = (c < 16) & (d >= synth.DATE_BASE)
    assert res["split"][2] == res["single"][2] == res["warp"][2] == int((~keep).sum())
    ser = groupby.groupby_user_category_month(synth.row_users(off, np.arange(n))[keep], c[keep], d[keep], v[keep])
    idx = ser.index.to_frame(index=False)
    sums, present, _ = res["split"]
    assert int(present.sum()) == len(ser)
    assert sums[idx["user"].to_numpy(), idx["category"].to_numpy(), idx["month"].to_numpy() - mlo].tolist() == ser.tolist()
    lo_u, hi_u = 5, 10
    a, b = int(off[lo_u]), int(off[hi_u])
    s2, p2, d2 = aggregate_columns(0, dd[a:b].contiguous(), cc[a:b].contiguous(), vv[a:b].contiguous(), oo, lo_u, hi_u, mlo, 24, 16,
                                   row_base=a)
    torch.cuda.synchronize()
    assert np.array_equal(s2.cpu().numpy(), sums[lo_u:hi_u]) and np.array_equal(p2.cpu().numpy(), present[lo_u:hi_u])


def dense_from_engine(eng, lo_u, hi_u, mlo, nm, ncat):
    import torch
    sums, present, dropped = eng.aggregate(lo_u, hi_u, mlo, nm, ncat)
    torch.cuda.synchronize()
    return sums.cpu().numpy(), present.cpu().numpy(), int(dropped.item())


def test_group_key_kernel_equals_the_date_kernels_and_pandas(built, monkeypatch):
    """Engine shards carry a 16-bit group key per row (category, month, "rare amount" flag) built at load; windows inside
    the key range are aggregated from cents + keys by aggregate_key_kernel. Same awkward shapes as above plus what is
    special to this kernel: zero amounts (a cell whose only rows are zero is present with sum 0), amounts that cancel,
    |v| around the 2^18 split, full-range int32, categories >= n_categories and >= 64, dates before / after the window,
    windows narrower than the data. The dense output must be IDENTICAL to the date kernels' (FCB_AGG_KERNEL=atomic) and
    equal pandas."""
    import torch
    from financial_chatbot_llm_b200 import synth
    from financial_chatbot_llm_b200.engine import Engine
    from oracle import groupby
    rng = np.random.default_rng(21)
    sizes = np.array([3, 8191, 8192, 8193, 1, 0, 4096 * 2 + 1, 150_001, 17, 8192 + 4097, 0, 0, 40_000, 5, 12_345, 2], dtype=np.int64)
    off = np.zeros(len(sizes) + 1, dtype=np.int64)
    np.cumsum(sizes, out=off[1:])
    n = int(off[-1])
    v = rng.integers(-(2 ** 31), 2 ** 31, size=n, dtype=np.int64).astype(np.int32)
    v[::3] = rng.integers(-5000, 5000, size=v[::3].size)
    v[1::7] = 0
    edge = np.array([2 ** 18 - 1, 2 ** 18, -(2 ** 18) + 1, -(2 ** 18), 2 ** 31 - 1, -(2 ** 31), 65535, 65536, -65536, 1, -1], dtype=np.int64)
    v[rng.integers(0, n, size=4000)] = edge[rng.integers(0, len(edge), size=4000)].astype(np.int32)
    d = (synth.DATE_BASE + rng.integers(0, 730 * 86400, size=n)).astype(np.int32)
    c = rng.integers(0, 16, size=n).astype(np.uint8)
    c[rng.integers(0, n, size=300)] = 200
    c[rng.integers(0, n, size=300)] = 40
    d[rng.integers(0, n, size=300)] = synth.DATE_BASE - 86400
    d[rng.integers(0, n, size=50)] = synth.DATE_BASE - 600 * 30 * 86400    # ~49 years back: older than the 511 months keys can express
    # user 13 (5 rows): one cell made of zeros only, one of amounts that cancel
    a13 = int(off[13])
    v[a13:a13 + 5] = [0, 0, 77, -77, 5]
    c[a13:a13 + 5] = [3, 3, 4, 4, 5]
    d[a13:a13 + 5] = synth.DATE_BASE + 40 * 86400
    dev = torch.device("cuda", 0)
    mlo = int(groupby.month_index(np.array([synth.DATE_BASE]))[0])
    eng = Engine(device=0, max_batch=1, max_k=1)
    eng.load_shard(torch.zeros((n, 4), dtype=torch.float32, device=dev), dates=torch.from_numpy(d).to(dev),
                   categories=torch.from_numpy(c).to(dev), cents=torch.from_numpy(v).to(dev), seg_offsets=torch.from_numpy(off).to(dev))
    users = synth.row_users(off, np.arange(n))
    month = groupby.month_index(d)
    # (64 x 64 = 4096 cells: the largest table, three warps per CTA; 1 month x 1 category: the smallest)
    for (w_lo, w_n, ncat) in ((mlo, 24, 16), (mlo + 5, 7, 9), (mlo - 1, 26, 64), (mlo - 1, 64, 64), (mlo + 3, 1, 1), (mlo - 3, 30, 64),
                              (mlo - 593, 4, 16)):
        got = dense_from_engine(eng, 0, len(sizes), w_lo, w_n, ncat)
        for other in ("atomic", "warp"):                                  # round 1's kernel, and the date form of the new one
            monkeypatch.setenv("FCB_AGG_KERNEL", other)
            ref = dense_from_engine(eng, 0, len(sizes), w_lo, w_n, ncat)
            monkeypatch.delenv("FCB_AGG_KERNEL")
            assert np.array_equal(got[0], ref[0]) and np.array_equal(got[1], ref[1]) and got[2] == ref[2], (other, w_lo, w_n, ncat)
        keep = (c < ncat) & (month >= w_lo) & (month < w_lo + w_n)
        assert got[2] == int((~keep).sum())
        if keep.any():
            ser = groupby.groupby_user_category_month(users[keep], c[keep], d[keep], v[keep])
            idx = ser.index.to_frame(index=False)
            assert int(got[1].sum()) == len(ser)
            assert got[0][idx["user"].to_numpy(), idx["category"].to_numpy(), idx["month"].to_numpy() - w_lo].tolist() == ser.tolist()
        else:
            assert int(got[1].sum()) == 0
    # the two special cells of user 13
    s, p_, _ = dense_from_engine(eng, 13, 14, mlo, 24, 16)
    m13 = int(groupby.month_index(np.array([synth.DATE_BASE + 40 * 86400]))[0]) - mlo
    assert p_[0, 3, m13] == 1 and s[0, 3, m13] == 0 and p_[0, 4, m13] == 1 and s[0, 4, m13] == 0 and s[0, 5, m13] == 5
    # a sub-range of users
    s2, p2, _ = dense_from_engine(eng, 5, 11, mlo, 24, 16)
    full = dense_from_engine(eng, 0, len(sizes), mlo, 24, 16)
    assert np.array_equal(s2, full[0][5:11]) and np.array_equal(p2, full[1][5:11])
    eng.close()


def test_group_keys_for_caller_owned_columns(built):
    """fcb_group_keys_build + fcb_aggregate_columns_keyed: the engine-less form of the same thing. Identical dense output
    with and without keys (row_base > 0, a sub-range of users, time-ordered dates too); with keys the date / category columns
    may be omitted as long as the window stays inside the key range, and a window outside it is refused, not mis-answered."""
    import torch
    from financial_chatbot_llm_b200 import _lib, synth
    from financial_chatbot_llm_b200.engine import aggregate_columns, build_group_keys
    from oracle import groupby
    rng = np.random.default_rng(31)
    sizes = np.array([7, 20_000, 8192, 1, 0, 33_333, 4095, 4097, 9_000], dtype=np.int64)
    off = np.zeros(len(sizes) + 1, dtype=np.int64)
    np.cumsum(sizes, out=off[1:])
    n = int(off[-1])
    spec = synth.SynthSpec(n_rows=n, dim=4, n_users=len(sizes), seed=5)
    v = rng.integers(-(2 ** 20), 2 ** 20, size=n, dtype=np.int64).astype(np.int32)
    v[::9] = 0
    c = rng.integers(0, 20, size=n).astype(np.uint8)
    mlo = int(groupby.month_index(np.array([synth.DATE_BASE]))[0])
    dev = torch.device("cuda", 0)
    for ordered in (False, True):
        d = synth.dates_sorted(spec, off, 0, n) if ordered else (synth.DATE_BASE + rng.integers(0, 730 * 86400, size=n)).astype(np.int32)
        dd, cc, vv, oo = (torch.from_numpy(a).to(dev) for a in (d, c, v, off))
        keys, base = build_group_keys(0, dd, cc, vv)
        assert keys.numel() == n and base <= mlo
        for (lo_u, hi_u, w_lo, w_n, ncat) in ((0, len(sizes), mlo, 24, 16), (1, 6, mlo + 2, 11, 20), (0, len(sizes), base, 30, 7)):
            a, b = int(off[lo_u]), int(off[hi_u])
            cols = [x[a:b].clone() for x in (dd, cc, vv)]             # (a slice of a 1-D tensor is a view: not aligned)
            k = keys[a:b].clone()
            ref = aggregate_columns(0, *cols, oo, lo_u, hi_u, w_lo, w_n, ncat, row_base=a, dates_sorted=ordered)
            got = aggregate_columns(0, *cols, oo, lo_u, hi_u, w_lo, w_n, ncat, row_base=a, dates_sorted=ordered, keys=k, key_base=base)
            bare = aggregate_columns(0, None, None, cols[2], oo, lo_u, hi_u, w_lo, w_n, ncat, row_base=a, keys=k, key_base=base)
            torch.cuda.synchronize()
            for x, y, z in zip(ref, got, bare):
                assert torch.equal(x, y) and torch.equal(x, z), (ordered, lo_u, hi_u, w_lo, w_n, ncat)
        with pytest.raises(_lib.FcbError):
            aggregate_columns(0, None, None, vv, oo, 0, len(sizes), base - 2, 4, 16, keys=keys, key_base=base)
        with pytest.raises(_lib.FcbError):                                   # misaligned columns are refused, not read
            aggregate_columns(0, dd[7:], cc[7:], vv[7:], oo, 1, len(sizes), mlo, 24, 16, row_base=7)
        s_out, _, dr = aggregate_columns(0, dd, cc, vv, oo, 0, len(sizes), base - 2, 4, 16, keys=keys, key_base=base)   # falls back on the dates
        s_ref, _, dr_ref = aggregate_columns(0, dd, cc, vv, oo, 0, len(sizes), base - 2, 4, 16)
        torch.cuda.synchronize()
        assert torch.equal(s_out, s_ref) and int(dr.item()) == int(dr_ref.item())


def test_windows_at_the_ends_of_the_int32_date_range(built, monkeypatch):
    """Dates are int32 unix seconds (1901-12-13 .. 2038-01-19). A window may start before the first or end after the last
    representable second; the date form of the warp-table kernel tests `(uint32)date - base < length` on the window CLIPPED to
    the int32 range (a shifted day index when the window starts before 1901). Rows at both ends, windows hanging over both
    ends: equal to round 1's kernel (64-bit date arithmetic) and to pandas."""
    import torch
    from financial_chatbot_llm_b200 import synth
    from financial_chatbot_llm_b200.engine import aggregate_columns
    from oracle import groupby
    rng = np.random.default_rng(41)
    n = 30_000
    sizes = np.array([9_000, 1, 12_000, 8_999], dtype=np.int64)
    off = np.zeros(len(sizes) + 1, dtype=np.int64)
    np.cumsum(sizes, out=off[1:])
    lo32, hi32 = -(2 ** 31), 2 ** 31 - 1
    d = np.concatenate([lo32 + rng.integers(0, 70 * 86400, size=n // 2), hi32 - rng.integers(0, 70 * 86400, size=n - n // 2)]).astype(np.int64)
    d[:3] = [lo32, lo32 + 1, lo32 + 86399]
    d[-3:] = [hi32, hi32 - 1, hi32 - 86399]
    rng.shuffle(d)
    d = d.astype(np.int32)
    v = rng.integers(-(2 ** 20), 2 ** 20, size=n, dtype=np.int64).astype(np.int32)
    c = rng.integers(0, 5, size=n).astype(np.uint8)
    dev = torch.device("cuda", 0)
    dd, cc, vv, oo = (torch.from_numpy(a).to(dev) for a in (d, c, v, off))
    users = synth.row_users(off, np.arange(n))
    month = groupby.month_index(d)
    m_first, m_last = int(month.min()), int(month.max())
    for (w_lo, w_n) in ((m_first - 3, 8), (m_first, 2), (m_last - 2, 6), (m_first - 1, m_last - m_first + 3 if m_last - m_first + 3 <= 240 else 240)):
        got = aggregate_columns(0, dd, cc, vv, oo, 0, len(sizes), w_lo, w_n, 5)
        monkeypatch.setenv("FCB_AGG_KERNEL", "cta")
        ref = aggregate_columns(0, dd, cc, vv, oo, 0, len(sizes), w_lo, w_n, 5)
        monkeypatch.delenv("FCB_AGG_KERNEL")
        torch.cuda.synchronize()
        for x, y in zip(got, ref):
            assert torch.equal(x, y), (w_lo, w_n)
        keep = (month >= w_lo) & (month < w_lo + w_n)
        assert int(got[2].item()) == int((~keep).sum())
        if keep.any():
            ser = groupby.groupby_user_category_month(users[keep], c[keep], d[keep], v[keep])
            idx = ser.index.to_frame(index=False)
            sums = got[0].cpu().numpy()
            assert int(got[1].sum().item()) == len(ser)
            assert sums[idx["user"].to_numpy(), idx["category"].to_numpy(), idx["month"].to_numpy() - w_lo].tolist() == ser.tolist()
