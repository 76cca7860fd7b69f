"""cvx_graph_align_no_alloc (include/curvix_b200.h): the alignment iterations with the intermediate buffer in a file-scope device array instead of
cudaMalloc-ed scratch.  Called through the C-ABI with caller-owned device buffers and compared with the existing entry point (cvx_graph_align) on
the same graph: same kernel, same neighbour order, so the results are equal bit for bit (asserted with the 1e-12 the alignment tests would
allow for libm); two calls in a row -- the second on a SMALLER graph, so that the array still holds the first call's rows beyond it -- and an odd
and an even iteration count (the last iteration must land in the caller's output either way).  Kept in a file that pytest collects
last."""
import numpy as np
import pytest

from conftest import pkg

pytestmark = pytest.mark.gpu


def _graph(rng, n, deg):
    """A random symmetric neighbour structure (CSR, ascending neighbour index) with some isolated nodes."""
    a = rng.integers(0, n, n * deg // 2); b = rng.integers(0, n, n * deg // 2)
    keep = (a != b) & (a % 7 != 0) & (b % 7 != 0)                     # every 7th node stays isolated
    pairs = np.unique(np.stack([np.minimum(a[keep], b[keep]), np.maximum(a[keep], b[keep])], 1), axis=0)
    src = np.concatenate([pairs[:, 0], pairs[:, 1]]); dst = np.concatenate([pairs[:, 1], pairs[:, 0]])
    o = np.lexsort((dst, src))
    off = np.concatenate([[0], np.cumsum(np.bincount(src, minlength=n))]).astype(np.int64)
    return off, dst[o].astype(np.int32)


def _reference_align(api, sk, loc, ori, off, nb, iterations):
    g = api.EdgeGraph(sk, loc, ori)
    g.set_neighbors(off, nb)
    lo, oo, _ = g.align(ori, iterations=iterations)
    g.close()
    return lo, oo


def test_no_alloc_alignment_equals_the_allocating_entry_point():
    import torch
    api = pkg("api")
    dev = torch.device("cuda", 0)
    sk = api.EdgeSketch(api.default_config(n_views=4, img_rows=800, img_cols=800), 0)
    rng = np.random.default_rng(5)
    for n, iterations in ((6000, 7), (2500, 4)):                      # two calls in a row: odd then even count, larger then smaller graph
        loc = rng.uniform(-0.5, 0.5, (n, 3))
        ori = rng.normal(size=(n, 3)); ori /= np.linalg.norm(ori, axis=1, keepdims=True)
        off, nb = _graph(rng, n, 6)
        want_loc, want_ori = _reference_align(api, sk, loc, ori, off, nb, iterations)
        d_off = torch.from_numpy(off).to(dev); d_nb = torch.from_numpy(nb).to(dev)
        d_loc = torch.from_numpy(loc).to(dev); d_ori = torch.from_numpy(ori).to(dev)
        o_loc = torch.full((n, 3), float("nan"), dtype=torch.float64, device=dev); o_ori = torch.full_like(o_loc, float("nan"))
        torch.cuda.synchronize()
        rc, msg = api.graph_align_no_alloc(0, torch.cuda.current_stream(dev).cuda_stream, n, d_off.data_ptr(), d_nb.data_ptr(), d_loc.data_ptr(),
                                           d_ori.data_ptr(), iterations, 0.01, 0.01, o_loc.data_ptr(), o_ori.data_ptr())
        assert rc == 0, msg
        torch.cuda.synchronize()
        got_loc, got_ori = o_loc.cpu().numpy(), o_ori.cpu().numpy()
        assert np.isfinite(got_loc).all() and np.isfinite(got_ori).all()
        assert np.abs(got_loc - loc).max() > 1e-6                     # the iterations moved something
        np.testing.assert_allclose(got_loc, want_loc, rtol=0, atol=1e-12)
        np.testing.assert_allclose(got_ori, want_ori, rtol=0, atol=1e-12)
        assert np.array_equal(d_loc.cpu().numpy(), loc) and np.array_equal(d_ori.cpu().numpy(), ori)   # inputs untouched
    # iterations == 0 copies the inputs; sizes beyond the fixed scratch are rejected with a message, nothing is launched
    rc, msg = api.graph_align_no_alloc(0, 0, n, d_off.data_ptr(), d_nb.data_ptr(), d_loc.data_ptr(), d_ori.data_ptr(), 0, 0.01, 0.01,
                                       o_loc.data_ptr(), o_ori.data_ptr())
    torch.cuda.synchronize()
    assert rc == 0 and np.array_equal(o_loc.cpu().numpy(), loc)
    rc, msg = api.graph_align_no_alloc(0, 0, api.ALIGN_NO_ALLOC_MAX_NODES + 1, d_off.data_ptr(), d_nb.data_ptr(), d_loc.data_ptr(), d_ori.data_ptr(), 3,
                                       0.01, 0.01, o_loc.data_ptr(), o_ori.data_ptr())
    assert rc != 0 and "exceed" in msg and str(api.ALIGN_NO_ALLOC_MAX_NODES) in msg
    sk.close()
