"""TEST INFRASTRUCTURE ONLY: minimal stand-in for torch_geometric (not installed here, no network), just
enough to import the reference's meshnet/graph_network.py and run it on the CPU.  Never shipped, never on
the product path.  It encodes PyG's documented MessagePassing semantics (flow='source_to_target'):
x_i = x[edge_index[1]], x_j = x[edge_index[0]], aggregation at edge_index[1], and update() receives the
ORIGINAL keyword arguments of propagate() (MessagePassing._collect) -- which is why the reference's
InteractionNetwork returns its input edge latents from update()."""
