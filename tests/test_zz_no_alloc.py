"""nori_gpu_trace_static — the allocation-free variant of nori_gpu_trace (static device scratch; no counterpart in the
reference). Collected last on purpose (file name).
Runs alone in a fresh process as well as after the other GPU tests (it only needs the session engine)."""
import numpy as np
import pytest

from conftest import random_rays

pytestmark = pytest.mark.gpu


def _equal(a, b):
    for x, y in zip(a, b):
        np.testing.assert_array_equal(x.view(np.uint32), y.view(np.uint32))      # bit-exact: ids and (t, u, v)


def test_trace_static_matches_trace_on_two_calls_in_a_row(gpu_scene, scenes):
    """Same rays through the existing entry point (cudaMalloc'ed scratch) and the static-scratch one, twice in a row:
    the second call reuses the array with FEWER rays of another kind (any-hit), so stale hits, a stale work cursor or
    stale split arrays of the first call would show."""
    e = gpu_scene("cbox_mis"); s = scenes("cbox_mis")
    rng = np.random.default_rng(7)
    primary = e.camera_rays((rng.random((40000, 2)) * [s.width, s.height]).astype(np.float32))
    first = np.concatenate([primary, random_rays(s, 25536, 11)])                # exactly the maximum: 65 536 rays
    assert first.shape[0] == e.TRACE_STATIC_MAX_RAYS
    want = e.trace(first)
    got = e.trace_static(first)
    _equal(got, want)
    assert (want[0] != 0xffffffff).mean() > 0.5                                   # the comparison is not vacuous
    second = random_rays(s, 12345, 13)
    _equal(e.trace_static(second, shadow=True), e.trace(second, shadow=True))
    _equal(e.trace_static(first[:777]), tuple(x[:777] for x in want))            # and closest hits again after the any-hit call


def test_trace_static_rejects_batches_beyond_its_scratch(gpu_scene, scenes):
    e = gpu_scene("cbox_mis"); s = scenes("cbox_mis")
    rays = random_rays(s, e.TRACE_STATIC_MAX_RAYS + 1, 5)
    with pytest.raises(RuntimeError, match="exceed the static scratch array"):
        e.trace_static(rays)
    tri, mesh, tuv = e.trace_static(rays[:0])                                      # empty input: nothing to do
    assert tri.size == 0 and tuv.shape == (0, 3)
