"""Property tests (Hypothesis) of the host-side building blocks: flat layout, Philox stream
definition, FedAvg oracle, binary wire codec."""

import torch
from hypothesis import given, settings
from hypothesis import strategies as st

from nanofed_b200.communication.http import codec
from nanofed_b200.ops import philox
from nanofed_b200.ops import reference as ref
from nanofed_b200.runtime.flat import FlatLayout

_shapes = st.lists(st.integers(min_value=1, max_value=5), min_size=0, max_size=3).map(tuple)
_dtypes = st.sampled_from([torch.float32, torch.int64, torch.float64])


@st.composite
def _states(draw):
    n = draw(st.integers(min_value=1, max_value=5))
    state = {}
    for i in range(n):
        shape, dtype = draw(_shapes), draw(_dtypes)
        numel = 1
        for d in shape:
            numel *= d
        vals = torch.arange(numel, dtype=torch.float32).reshape(shape) - i
        state[f"t{i}"] = vals.to(dtype)
    return state


@settings(max_examples=40, deadline=None)
@given(_states(), st.sampled_from([1, 4, 64]))
def test_flat_layout_roundtrip_keeps_values_shapes_dtypes(state, align):
    layout = FlatLayout.from_state(state, align=align)
    assert layout.total % align == 0 or align == 1
    flat = layout.flatten(state)
    assert flat.dtype == torch.float32 and flat.numel() == layout.total
    back = layout.unflatten(flat)
    assert list(back) == list(state)
    for k, v in state.items():
        assert back[k].dtype == v.dtype and back[k].shape == v.shape
        assert torch.equal(back[k], v)
    # entries never overlap and respect the alignment
    spans = sorted((layout[k].offset, layout[k].offset + layout[k].numel) for k in state)
    assert all(a1 <= b0 for (_, a1), (b0, _) in zip(spans, spans[1:]))
    assert all(layout[k].offset % align == 0 for k in state)


@settings(max_examples=40, deadline=None)
@given(st.integers(0, 2**63 - 1), st.integers(0, 500).map(lambda v: v * 4),
       st.integers(1, 300), st.integers(0, 200).map(lambda v: v * 4))
def test_philox_stream_is_position_addressed(seed, offset, count, skip):
    """Element i of a draw at (seed, offset) depends only on offset + i."""
    long = philox.uniform(skip + count, seed, offset)
    part = philox.uniform(count, seed, offset + skip)
    assert torch.equal(long[skip:], part)
    z_long = philox.normal(skip + count, seed, offset)
    z_part = philox.normal(count, seed, offset + skip)
    assert torch.equal(z_long[skip:], z_part)
    assert bool(((long > 0) & (long < 1)).all())


@settings(max_examples=30, deadline=None)
@given(st.integers(2, 6), st.integers(1, 40), st.integers(0, 10**6))
def test_fedavg_oracle_invariants(k, n, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(k, n, generator=g)
    counts = torch.randint(1, 100, (k,), generator=g).float()
    out = ref.fedavg_reduce(x, counts)
    want = (counts / counts.sum()).view(-1, 1).mul(x).sum(0)
    torch.testing.assert_close(out, want, rtol=1e-5, atol=1e-6)
    # permutation of the clients and a common scale of the counts change nothing
    perm = torch.randperm(k, generator=g)
    torch.testing.assert_close(ref.fedavg_reduce(x[perm], counts[perm] * 7), out, rtol=1e-5,
                               atol=1e-6)
    # a masked-out client may hold anything
    mask = torch.ones(k)
    mask[0] = 0
    poisoned = x.clone()
    poisoned[0] = float("nan")
    got = ref.fedavg_reduce(poisoned, counts, mask=mask)
    keep = (counts[1:] / counts[1:].sum()).view(-1, 1).mul(x[1:]).sum(0)
    torch.testing.assert_close(got, keep, rtol=1e-5, atol=1e-6)
    # the average stays inside the convex hull
    assert bool((out <= x.max(0).values + 1e-5).all() and (out >= x.min(0).values - 1e-5).all())


@settings(max_examples=30, deadline=None)
@given(_states())
def test_binary_frame_roundtrip(state):
    meta = {"client_id": "c", "round_number": 3, "metrics": {"loss": 0.25}}
    frame = codec.encode_frame(meta, state)
    meta2, state2 = codec.decode_frame(frame)
    assert meta2 == meta and list(state2) == list(state)
    for key, v in state.items():
        assert state2[key].dtype == v.dtype and state2[key].shape == v.shape
        assert torch.equal(state2[key], v)
