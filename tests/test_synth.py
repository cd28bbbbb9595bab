import numpy as np

from financial_chatbot_llm_b200 import synth


def test_generator_is_a_pure_function_of_row_range():
    spec = synth.SynthSpec(n_rows=1000, dim=64, n_users=10)
    a = synth.vectors_f32(spec, 0, 1000)
    b = np.vstack([synth.vectors_f32(spec, 0, 400), synth.vectors_f32(spec, 400, 1000)])
    assert a.dtype == np.float32 and np.array_equal(a, b)
    assert np.array_equal(synth.dates(spec, 100, 200), synth.dates(spec, 0, 1000)[100:200])
    assert np.array_equal(synth.cents(spec, 100, 200), synth.cents(spec, 0, 1000)[100:200])


def test_elements_are_exact_integers_scaled_by_power_of_two():
    spec = synth.SynthSpec(n_rows=64, dim=128, n_users=4)
    vi = synth.vectors_int(spec, 0, 64)
    assert np.abs(vi).max() < 2 ** 24
    assert np.array_equal(synth.vectors_f32(spec, 0, 64).astype(np.float64), vi * synth.VALUE_SCALE)


def test_known_answer_hash():
    # splitmix64 finaliser of 1 (published test value of the mixer)
    assert int(synth.mix64(np.array([1], dtype=np.uint64))[0]) == 0x5692161D100B05E5


def test_column_ranges():
    spec = synth.SynthSpec(n_rows=5000, dim=32, n_users=7)
    d = synth.dates(spec, 0, 5000)
    assert d.dtype == np.int32 and d.min() >= synth.DATE_BASE
    assert d.max() < synth.DATE_BASE + spec.date_span_seconds
    c = synth.categories(spec, 0, 5000)
    assert c.dtype == np.uint8 and c.max() < 16 and len(np.unique(c)) == 16
    v = synth.cents(spec, 0, 5000)
    assert v.dtype == np.int32 and (v < 0).any() and (v > 0).any() and np.abs(v).max() < 2 ** 19


def test_segment_offsets_cover_all_rows():
    for spec in (synth.SynthSpec(n_rows=1000, n_users=1000), synth.SynthSpec(n_rows=100000, n_users=1000),
                 synth.SynthSpec(n_rows=5000, n_users=1), synth.SynthSpec(n_rows=5000, n_users=9, whale_rows=4000)):
        off = synth.segment_offsets(spec)
        assert off[0] == 0 and off[-1] == spec.n_rows and (np.diff(off) >= 1).all()
    assert synth.segment_offsets(synth.SynthSpec(n_rows=5000, n_users=9, whale_rows=4000))[1] == 4000


def test_bf16_round_to_nearest_even():
    x = np.array([1.0, 1.00390625, 1.01171875, -2.5, 3.0e-5, 0.0], dtype=np.float32)
    bits = synth.f32_to_bf16_bits(x)
    back = synth.bf16_bits_to_f32(bits)
    assert back[0] == 1.0 and back[3] == -2.5 and back[5] == 0.0
    assert back[1] == 1.0            # 1 + 2^-8 is a tie -> even mantissa (1.0)
    assert back[2] == np.float32(1.015625)   # 1 + 3*2^-8 is a tie -> even mantissa (1 + 2^-6)
    import torch
    t = torch.from_numpy(x).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    assert np.array_equal(t, bits)


def test_payload_layout_matches_what_the_tool_reads():
    spec = synth.SynthSpec(n_rows=500, dim=32, n_users=5)
    hc = synth.make_host_corpus(spec)
    p = hc.payload(123)
    assert set(p) == {"page_content", "metadata"} and isinstance(p["page_content"], str)
    assert p["metadata"]["user_id"] == hc.users[int(synth.row_users(hc.offsets, np.array([123]))[0])]
    assert isinstance(p["metadata"]["date"], int)
    assert p == synth.payload_for_row(spec, hc.offsets, hc.users, 123)
