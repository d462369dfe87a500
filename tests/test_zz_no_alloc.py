"""b200_fr_prefix_product_static -- the entry point that allocates nothing (its scratch is a file-scope static
__device__ array inside the library, csrc/noalloc.cu) -- against the existing b200_fr_prefix_product_dev, through
the C ABI, on two calls in a row that share the array.  Field arithmetic is exact, so the tolerance is the suite's:
bit-for-bit equality.  This file is named to be collected last."""
import numpy as np
import pytest

from tests.conftest import rand_field

pytestmark = pytest.mark.gpu

MAX_N = 1 << 22     # the stated limit of the static scratch


def _both(ctx, pkg, x):
    n = x.shape[0]
    src = pkg.DeviceVector(ctx, max(n, 1) * 32).upload(x)
    want_d = pkg.DeviceVector(ctx, max(n, 1) * 32)
    got_d = pkg.DeviceVector(ctx, max(n, 1) * 32)
    ctx.fr_prefix_product_dev(src, want_d, n)           # existing entry point (scratch from cudaMalloc)
    ctx.fr_prefix_product_static(src, got_d, n)         # new entry point: enqueue only
    ctx.sync()
    want, got = want_d.download((n, 4)), got_d.download((n, 4))
    [v.free() for v in (src, want_d, got_d)]
    return want, got


def test_static_scratch_prefix_product_matches_existing_entry_point_twice(ctx, pkg):
    rng = np.random.default_rng(2024)
    # first call: the largest size the static scratch admits, three levels of chunk summaries
    x = rand_field(rng, MAX_N)
    want, got = _both(ctx, pkg, x)
    assert np.array_equal(got, want)
    assert not np.array_equal(got[-1], x[-1])
    # second call, straight after, different data and a ragged smaller size: the array still holds the first call's
    # chunk products and must not leak into this result
    y = rand_field(rng, (1 << 15) + 77)
    want2, got2 = _both(ctx, pkg, y)
    assert np.array_equal(got2, want2)
    # in place, and tiny sizes (single chunk: no scratch level at all)
    for n in (1, 31, 33, 1025):
        z = rand_field(rng, n)
        d = pkg.DeviceVector(ctx, n * 32).upload(z)
        ctx.fr_prefix_product_static(d, d, n)
        ctx.sync()
        w, _ = _both(ctx, pkg, z)
        assert np.array_equal(d.download((n, 4)), w), n
        d.free()


def test_static_scratch_rejects_sizes_beyond_its_maximum(ctx, pkg):
    d = pkg.DeviceVector(ctx, 64)
    with pytest.raises(pkg.B200Error, match="exceeds the static scratch"):
        ctx.fr_prefix_product_static(d, d, MAX_N + 1)
    ctx.fr_prefix_product_static(d, d, 0)       # empty input: nothing to do, no error
    d.free()
