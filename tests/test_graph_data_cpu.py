"""Graph containers on the input side of the GNN path: `data_to_graph` / `graph_to_data` against the reference's own
methods (golden: tests/make_golden_graph_data.py), `Batch.from_data_list` against the disjoint-union rule (torch_geometric is
not in this image: properties only) and against the batch layout bench.py feeds the GNN kernels."""
import os

import numpy as np
import pytest
import torch

import fake_scene

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "graph_data_seed6.npz"))


@pytest.mark.parametrize("tag,as_numpy", [("tensor", False), ("numpy", True)])
def test_converters_match_the_reference_methods(tag, as_numpy):
    from meshnet.graph_data import data_to_graph, graph_to_data

    args = fake_scene.graph_case(as_numpy)
    g = data_to_graph(*args)
    for k in ("x", "face", "y", "pos", "pos_target", "vel", "particle_actions"):
        v = getattr(g, k)
        assert v.is_contiguous() and np.array_equal(v.numpy(), G[f"{tag}/{k}"]) and v.numpy().dtype == G[f"{tag}/{k}"].dtype, k
    assert g.grasped_particle == 3 and g.num_nodes == 7
    g.edge_index, g.edge_attr = torch.zeros((2, 1), dtype=torch.long), torch.zeros((1, 4))
    for n, p in enumerate(graph_to_data(g, 2)):
        assert np.array_equal(p.numpy(), G[f"{tag}/part{n}"]), n
    # the inputs are not modified
    assert all(np.array_equal(np.asarray(a), np.asarray(b)) for a, b in zip(args[2:], fake_scene.graph_case(as_numpy)[2:]))


def test_batch_is_the_disjoint_union():
    from meshnet.graph_data import Batch, data_to_graph

    graphs = []
    for n, seed in ((7, 6), (5, 8), (9, 10)):
        g = data_to_graph(*fake_scene.graph_case(False, N=n, seed=seed))
        g.edge_index = torch.randint(0, n, (2, 2 * n), generator=torch.Generator().manual_seed(seed))
        g.edge_attr = torch.randn((2 * n, 4), generator=torch.Generator().manual_seed(seed))
        graphs.append(g)
    b = Batch.from_data_list(graphs)
    assert b.num_graphs == 3 and b.ptr.tolist() == [0, 7, 12, 21] and b.num_nodes == 21
    assert b.batch.tolist() == [0] * 7 + [1] * 5 + [2] * 9
    assert b.grasped_particle == [3, 3, 3]
    e0 = 0
    for i, g in enumerate(graphs):
        lo, hi = int(b.ptr[i]), int(b.ptr[i + 1])
        E = g.edge_index.shape[1]
        assert torch.equal(b.x[lo:hi], g.x) and torch.equal(b.y[lo:hi], g.y) and torch.equal(b.pos[lo:hi], g.pos)
        assert torch.equal(b.edge_index[:, e0:e0 + E] - lo, g.edge_index) and torch.equal(b.edge_attr[e0:e0 + E], g.edge_attr)
        assert torch.equal(b.face[:, 5 * i:5 * i + 5] - lo, g.face)
        e0 += E
    with pytest.raises(ValueError):
        Batch.from_data_list([])


def test_batch_layout_is_what_the_bench_feeds_the_kernels():
    import bench
    from meshnet.graph_data import Batch, Data

    x, ei, ef = bench.gnn_batch_inputs(6, 1)
    one = Data(x=torch.from_numpy(x), edge_index=torch.from_numpy(ei), edge_attr=torch.from_numpy(ef))
    b = Batch.from_data_list([one, one, one])
    bx, bei, bef = bench.gnn_batch_inputs(6, 3)
    assert np.array_equal(b.x.numpy(), bx) and np.array_equal(b.edge_index.numpy(), bei) and np.array_equal(b.edge_attr.numpy(), bef)
