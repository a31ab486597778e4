"""-m gpu, collected last on purpose: gaia_b200_color_graph_device_noalloc - the allocation-free variant of the device colouring,
scratch in a fixed file-scope device array inside the library - against gaia_b200_color_graph_device (flags = 0: the same
deterministic speculative greedy) through the C ABI. Integer work: the colours must be identical. Two calls in a row on
different graphs (the second smaller: the scratch still holds the first call's colours), then the first graph again."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _csr(n, edges):
    """both directions of every unique edge, grouped by node (the order inside a node's list does not matter to the algorithm)"""
    e = np.asarray(edges, np.int64).reshape(-1, 2)
    src = np.concatenate([e[:, 0], e[:, 1]]); dst = np.concatenate([e[:, 1], e[:, 0]])
    order = np.argsort(src, kind="stable")
    off = np.zeros(n + 1, np.int64)
    np.add.at(off, src + 1, 1)
    return np.cumsum(off).astype(np.int32), dst[order].astype(np.int32)


def _graphs():
    from gaia_b200 import capi, meshio
    v1, t1 = meshio.kuhn_beam(20, 9, 7, h=0.1)
    v2, t2 = meshio.kuhn_beam(5, 4, 3, h=0.1)
    rng = np.random.default_rng(7)
    n3 = 3000                                             # an irregular graph: random sparse edges, a few hubs
    e3 = rng.integers(0, n3, size=(9000, 2))
    e3 = np.concatenate([e3, np.stack([np.zeros(40, np.int64), rng.integers(1, n3, 40)], 1)])
    e3 = np.sort(e3[e3[:, 0] != e3[:, 1]], axis=1)
    e3 = np.unique(e3, axis=0).astype(np.int32)
    return [(len(v1), capi.tet_mesh_edges(len(v1), t1)), (len(v2), capi.tet_mesh_edges(len(v2), t2)), (n3, e3)]


def _expected(n, edges):
    from gaia_b200 import capi
    cats = capi.color_graph_device(n, edges, balance=False)
    col = np.full(n, -1, np.int32)
    for c, ids in enumerate(cats):
        col[ids] = c
    return col, len(cats)


def _noalloc(n, edges):
    import torch
    from gaia_b200 import capi
    off, adj = _csr(n, edges)
    d_off = torch.from_numpy(off).cuda(); d_adj = torch.from_numpy(adj).cuda()
    d_col = torch.full((n,), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    nc = capi.color_graph_device_noalloc(n, d_off.data_ptr(), d_adj.data_ptr(), d_col.data_ptr())
    return d_col.cpu().numpy(), nc


def test_noalloc_colouring_equals_the_allocating_entry_point_on_calls_in_a_row():
    graphs = _graphs()
    for n, e in graphs + graphs[:1]:                      # large, small, irregular, large again - back to back on one scratch
        want, nc_want = _expected(n, e)
        got, nc = _noalloc(n, e)
        assert nc == nc_want and np.array_equal(got, want), (n, nc, nc_want)
        assert (got[e[:, 0]] != got[e[:, 1]]).all() and got.min() == 0      # and it IS a colouring


def test_noalloc_colouring_rejects_what_does_not_fit_or_is_malformed():
    import torch
    from gaia_b200 import capi
    with pytest.raises(capi.GaiaError, match="GAIA_NOALLOC_COLOR_MAX_NODES") as ei:
        capi.color_graph_device_noalloc(capi.NOALLOC_COLOR_MAX_NODES + 1, 256, 256, 256)    # rejected before any pointer is touched
    assert ei.value.code == -6
    n, e = _graphs()[1]
    off, adj = _csr(n, e)
    adj = adj.copy(); adj[3] = n + 5                        # neighbour out of range
    d_off = torch.from_numpy(off).cuda(); d_adj = torch.from_numpy(adj).cuda(); d_col = torch.zeros(n, dtype=torch.int32, device="cuda")
    with pytest.raises(capi.GaiaError, match="out of range"):
        capi.color_graph_device_noalloc(n, d_off.data_ptr(), d_adj.data_ptr(), d_col.data_ptr())
    got, nc = _noalloc(n, e)                                # and the next call is unaffected
    want, nc_want = _expected(n, e)
    assert nc == nc_want and np.array_equal(got, want)
