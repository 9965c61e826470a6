"""Property tests (hypothesis) of the host-side logic around the path: the frame plans of sample sharding tile a step
exactly for any world size, the noise slots a rank must pre-load are the two frames before its first one, row ownership
partitions the image, and the alias table reproduces any pdf."""
import numpy as np
from hypothesis import given, settings, strategies as st

import mrt_b200 as m
from mrt_b200 import distributed as D


@settings(max_examples=200, deadline=None)
@given(start=st.integers(0, 10_000), per_rank=st.integers(0, 40), world=st.integers(1, 16))
def test_sample_shard_plans_tile_the_step(start, per_rank, world):
    plans = [D.plan_sample_shard(start, per_rank, g, world) for g in range(world)]
    frames = sorted(p.first_frame + k * p.frame_stride for p in plans for k in range(p.n_frames))
    assert frames == list(range(start, start + per_rank * world))          # every frame once, none missing
    for g, p in enumerate(plans):
        assert p.total_frames == per_rank * world and p.frame_stride == world and p.first_frame == start + g
        assert p.span == (0 if per_rank == 0 else (per_rank - 1) * world + 1)   # tiles first_frame .. last frame inclusive
        lead = D.lead_in_frames(p.first_frame)
        assert lead == [f for f in (p.first_frame - 2, p.first_frame - 1) if f >= 0]
        # slot f % 3 holds frame f's tile: the lead-in frames and the first frame occupy three different slots
        assert len({f % 3 for f in lead + [p.first_frame]}) == len(lead) + 1


@settings(max_examples=200, deadline=None)
@given(height=st.integers(0, 300), world=st.integers(1, 16))
def test_row_ownership_partitions_the_image(height, world):
    rows = [list(D.rows_owned(height, g, world)) for g in range(world)]
    assert sorted(y for r in rows for y in r) == list(range(height))
    assert max(len(r) for r in rows) <= (height + world - 1) // world           # what mrt_shard_layout sizes a shard for


@settings(max_examples=100, deadline=None)
@given(weights=st.lists(st.floats(0.0, 1e3, allow_nan=False, width=32), min_size=1, max_size=300))
def test_alias_table_reproduces_any_pdf(weights):
    pdf = np.asarray(weights, dtype=np.float32)
    prob, alias = m.build_alias_table(pdf)
    n = len(pdf)
    assert ((prob >= 0) & (prob <= 1)).all() and (alias < n).all()
    total = float(pdf.astype(np.float64).sum())
    got = prob.astype(np.float64).copy()
    np.add.at(got, alias, 1.0 - prob.astype(np.float64))
    want = pdf.astype(np.float64) / total if total > 0 else np.full(n, 1.0 / n)   # all-zero pdf: uniform, nothing to prefer
    assert np.allclose(got / n, want, atol=3e-7)
