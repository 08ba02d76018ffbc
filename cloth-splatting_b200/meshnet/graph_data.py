"""The graph containers on the input side of the GNN path: what the reference gets from
torch_geometric (`Data`, `Batch.from_data_list`; torch_geometric is not in this image, so its published collation rule is
restated here -- parity of `Batch` is unpinned) and the two converters of its simulation dataloader, which ARE pinned on the
reference's own methods (meshnet/dataloader_sim.py:352-425; tests/make_golden_graph_data.py):

    data_to_graph(...)   one candidate action of the planner -> a graph (node features = velocity history with the grasped
                         particle's last velocity replaced by its commanded one, ++ node type)
    graph_to_data(g)     a (batched) graph -> the tensors `ClothMeshSimulator.predict_*` / `multi_step_loss` take

A batch of `data_to_graph` graphs over candidate actions is the 128-graph forward the bench measures."""
import torch


class Data:
    """Attribute bag with torch_geometric.data.Data's constructor convention (keyword tensors)."""

    def __init__(self, **kwargs):
        for k, v in kwargs.items():
            setattr(self, k, v)

    def keys(self):
        return [k for k in self.__dict__ if not k.startswith("_")]

    @property
    def num_nodes(self):
        for k in ("x", "pos"):
            if isinstance(getattr(self, k, None), torch.Tensor):
                return getattr(self, k).shape[0]
        raise ValueError("Data: neither `x` nor `pos` to take the number of nodes from")

    def to(self, device):
        return Data(**{k: (v.to(device) if isinstance(v, torch.Tensor) else v) for k, v in self.__dict__.items()})


class Batch(Data):
    @classmethod
    def from_data_list(cls, graphs):
        """Disjoint union (torch_geometric.data.Batch.from_data_list): attributes whose name contains "index" or is "face"
        are shifted by the number of nodes of the graphs before them and concatenated along the last dimension, other
        tensors along dimension 0, anything else is gathered in a list; `batch` [N] maps nodes to graphs, `ptr` [G+1] are
        the node offsets."""
        if not graphs:
            raise ValueError("Batch.from_data_list: empty list")
        counts = torch.tensor([g.num_nodes for g in graphs], dtype=torch.long)
        ptr = torch.cat([counts.new_zeros(1), counts.cumsum(0)])
        out = cls()
        for k in graphs[0].keys():
            vals = [getattr(g, k) for g in graphs]
            if not isinstance(vals[0], torch.Tensor):
                setattr(out, k, vals)
            elif "index" in k or k == "face":
                setattr(out, k, torch.cat([v + int(o) for v, o in zip(vals, ptr[:-1])], dim=-1))
            else:
                setattr(out, k, torch.cat(vals, dim=0))
        dev = graphs[0].x.device if isinstance(getattr(graphs[0], "x", None), torch.Tensor) else None
        out.batch = torch.repeat_interleave(torch.arange(len(graphs)), counts).to(dev)
        out.ptr = ptr.to(dev)
        out.num_graphs = len(graphs)
        return out


def _t(v):
    return v if isinstance(v, torch.Tensor) else torch.tensor(v)


def data_to_graph(action, grasped_particle, velocity, node_type, faces, velocity_target, position_target, positions):
    """`_data_to_graph` (meshnet/dataloader_sim.py:352-399).  `action` [F,3] displaces the grasped particle: its position by
    the first action, the last 3 velocity features by the first target velocity; `particle_actions` carries the action on
    that particle only.  The reference's two branches (tensor / numpy inputs) differ in what they store as `vel` (the
    numpy branch stores the velocities WITH the action applied) -- kept."""
    particle_actions = torch.zeros(velocity_target.shape)
    particle_actions[:, grasped_particle] = _t(action)
    tensors = isinstance(velocity, torch.Tensor)
    positions_actions = positions.clone() if isinstance(positions, torch.Tensor) else positions.copy()
    positions_actions[grasped_particle] += action[0]
    velocity_actions = velocity.clone() if tensors else velocity.copy()
    velocity_actions[grasped_particle, -3:] = _t(velocity_target[0, grasped_particle]) if tensors else velocity_target[0, grasped_particle]
    f32 = lambda v: _t(v).to(torch.float32).contiguous()  # noqa: E731
    x = torch.hstack((f32(velocity_actions), _t(node_type).contiguous()))
    if tensors:
        return Data(x=x, face=faces.to(torch.int64).contiguous(), y=f32(velocity_target.permute(1, 0, 2)),
                    pos=f32(positions_actions), pos_target=f32(position_target.permute(1, 0, 2)), vel=f32(velocity),
                    particle_actions=f32(particle_actions.permute(1, 0, 2)), grasped_particle=grasped_particle)
    return Data(x=x, face=faces.to(torch.int64).contiguous(), y=f32(_t(velocity_target).transpose(1, 0)),
                pos=f32(positions_actions), pos_target=f32(_t(position_target).transpose(1, 0)), vel=f32(velocity_actions),
                particle_actions=f32(particle_actions.transpose(1, 0)), grasped_particle=grasped_particle)


def graph_to_data(graph, input_length_sequence=1):
    """`_graph_to_data` (meshnet/dataloader_sim.py:414-425) -> (velocity, node_types, edge_index, edge_features,
    target_velocities, particle_actions, positions)."""
    return (graph.x[:, :3 * input_length_sequence], graph.x[:, -1].unsqueeze(1), graph.edge_index, graph.edge_attr, graph.y,
            graph.particle_actions, graph.pos)
