"""Synthetic OGB-featurised generator: layout, determinism, chemistry-like degree bounds."""
import numpy as np

from gnnepcsaft_b200 import synth


def test_layout_and_determinism():
    a = synth.make_graphs(64, 1234)
    b = synth.make_graphs(64, 1234)
    for ga, gb in zip(a, b):
        assert np.array_equal(ga.x, gb.x) and np.array_equal(ga.edge_index, gb.edge_index)
        assert np.array_equal(ga.edge_attr, gb.edge_attr)
    for g in a:
        n, e = g.num_nodes, g.edge_index.shape[1]
        assert g.x.shape == (n, 9) and g.x.dtype == np.int64 and g.edge_attr.shape == (e, 3)
        assert e % 2 == 0
        # bond b -> columns 2b (i,j) and 2b+1 (j,i) with identical features
        assert np.array_equal(g.edge_index[:, 0::2], g.edge_index[::-1, 1::2])
        assert np.array_equal(g.edge_attr[0::2], g.edge_attr[1::2])
        for k, d in enumerate(synth.ATOM_FEATURE_DIMS):
            assert g.x[:, k].min(initial=0) >= 0 and g.x[:, k].max(initial=0) < d
        if e:
            deg = np.bincount(g.edge_index[1], minlength=n)
            assert deg.max() <= 4 and deg.min() >= 1          # connected, valence <= 4
            assert len({(int(s), int(t)) for s, t in g.edge_index.T}) == e     # no duplicate bonds
        else:
            assert n == 1
        assert (g.para > 0).all()


def test_pack_roundtrip():
    gs = synth.make_graphs(20, 7, 1, 12, 0.2)
    p = synth.pack(gs)
    assert p.num_graphs == 20 and p.x.dtype == np.uint8 and p.src.dtype == np.int16
    for i, g in enumerate(gs):
        ns, ne = slice(p.node_ptr[i], p.node_ptr[i + 1]), slice(p.edge_ptr[i], p.edge_ptr[i + 1])
        assert np.array_equal(p.x[ns], g.x) and np.array_equal(p.src[ne], g.edge_index[0])
        assert np.array_equal(p.dst[ne], g.edge_index[1]) and np.array_equal(p.edge_attr[ne], g.edge_attr)
