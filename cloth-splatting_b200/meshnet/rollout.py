"""Multi-step training loss of the cloth simulator: the body of the reference's optimisation step
(train_meshnet_sim.py:491-527) and its state update between predictions (`update_prediction`, :322-359).

    loss = sum_f mean((pred_acc_f - target_acc_f)^2),   f = 0 .. future_sequence_length-1

Between two predictions the mesh is moved by the predicted velocity (grasped particles follow the action), the
edge features are recomputed from the new positions (torch_geometric Cartesian(norm=False) + Distance(norm=False):
[pos[src] - pos[dst], |pos[src] - pos[dst]|]) and the velocity history is shifted.  The gradient reaches the
earlier predictions through those edge features: EncodeProcessDecode's backward (csb_gnn_backward) returns
dL/d edge_features, the few [N,3] / [E,4] element-wise ops here are left to torch autograd.

Reference quirk kept on purpose: the shifted history is filled with a COPY of the previous last velocity (plus
the action on grasped particles), not with the predicted velocity (train_meshnet_sim.py:354-357)."""
import torch


def compute_edge_features(pos, edge_index):
    """[pos[src] - pos[dst], norm] for edge_index = [src; dst] (PyG Cartesian + Distance, norm=False)."""
    rel = pos[edge_index[0]] - pos[edge_index[1]]
    return torch.cat([rel, rel.norm(dim=-1, keepdim=True)], dim=-1)


def update_prediction(velocity_noise, velocity, pred_acc, init_position, edge_index, old_particle_actions,
                      particle_actions, device=None, input_sequence_length=None):
    """-> (velocity history, edge_features, positions) for the next prediction (train_meshnet_sim.py:322-359).
    Masks are element-wise (`actions != 0` per coordinate), exactly as the reference indexes them."""
    if velocity_noise is not None:
        velocity = velocity + velocity_noise
    last = velocity[:, -3:]
    new_vel = last + pred_acc
    held = old_particle_actions != 0
    new_vel = torch.where(held, old_particle_actions.to(new_vel.dtype), new_vel)
    free = particle_actions == 0
    # the reference adds in place into a clone of init_position: every sum is rounded to ITS dtype
    new_pos = (init_position + torch.where(free, new_vel, torch.zeros_like(new_vel))).to(init_position.dtype)
    new_pos = (new_pos + particle_actions).to(init_position.dtype)
    edge_features = compute_edge_features(new_pos, edge_index)
    acted = particle_actions != 0
    tail = torch.where(acted, particle_actions.to(last.dtype), last.detach())
    velocity = torch.cat([velocity[:, 3:], tail], dim=1)
    return velocity, edge_features, new_pos


def multi_step_loss(simulator, velocity, node_types, edge_index, edge_features, target_velocities,
                    particle_actions, positions, future_sequence_length, velocity_noise=None):
    """Sum over the unrolled predictions of the per-step MSE (train_meshnet_sim.py:491-527).
    target_velocities [N, F, 3], particle_actions [N, F(+1), 3]; the noise is applied to the first step only."""
    loss = 0
    pos = positions
    for f in range(future_sequence_length):
        noise = velocity_noise if f == 0 else None
        pred_acc, target_acc = simulator.predict_acceleration(velocity, node_types, edge_index, edge_features,
                                                              target_velocities[:, f], velocity_noise=noise)
        loss = loss + torch.mean((pred_acc - target_acc) ** 2)
        if future_sequence_length > 1 and f < future_sequence_length - 1:
            acc = simulator._output_normalizer.inverse(pred_acc)
            velocity, edge_features, pos = update_prediction(noise, velocity, acc, pos, edge_index,
                                                             particle_actions[:, f], particle_actions[:, f + 1])
    return loss


def rollout(simulator, features, nsteps, device, input_sequence_length, dt, real_world=False):
    """Open-loop rollout of the simulator over one trajectory (train_meshnet_sim.py:92-280, the `rollout` that
    meshnet/generate_mesh_predictions.py turns into `mesh_predictions/mesh_XXX.hdf5`; see mesh_data.write_mesh_predictions).

    features: 'pos' [T,N,3], 'vel' [T,N,3], 'actions' [T,3], 'node_type' [N,1], 'edge_index' [2,E], 'faces' [3,F],
    'grasped_particle'.  Every step builds the planner graph of the current state (graph_data.data_to_graph), derives the
    edges and their features from the faces (FaceToEdge + Cartesian + Distance), predicts the next velocity, pins the grasped
    particle to its action, and integrates.  `real_world` adds the reference's 10 Adam steps on the predicted velocity that
    pull the edge lengths back to their rest values.  Returns the reference's dictionary ('predicted_rollout' [nsteps,N,3]
    are per-step displacements)."""
    from .graph_data import data_to_graph, graph_to_data
    from .mesh_data import cartesian_distance, face_to_edge
    from .model_utils import NodeType

    node_coords, node_vels, actions = features["pos"], features["vel"], features["actions"]
    node_types, edge_index, faces = features["node_type"], features["edge_index"], features["faces"]
    grasped = features["grasped_particle"]
    L = input_sequence_length
    truth = node_vels[L:]
    history = node_vels[:L].to(device).clone()
    rest = node_coords[0][edge_index.T[:, 1]] - node_coords[0][edge_index.T[:, 0]]
    rest_lengths = torch.norm(rest.to(device), dim=1)
    coords = node_coords[0].to(device).clone()
    types = node_types.to(device)
    predictions = []
    for step in range(nsteps):
        action = actions[step:step + 1].to(device)
        graph = data_to_graph(action, grasped, torch.cat([v for v in history], 1), types, faces.to(device),
                              truth[step:step + 1].to(device), node_coords[step + 1:step + 2].to(device), coords)
        graph.edge_index = face_to_edge(graph.face, graph.pos.shape[0])
        graph.edge_attr = cartesian_distance(graph.pos, graph.edge_index)
        vel, ntype, ei, ef, _, _, _ = graph_to_data(graph, L)
        nxt = simulator.predict_velocity(velocities=vel, node_type=ntype, edge_index=ei, edge_features=ef)
        nxt[grasped] = action
        if real_world:
            free = nxt.detach().clone().requires_grad_(True)
            opt = torch.optim.Adam([free], lr=1e-3)
            with torch.enable_grad():
                for _ in range(10):
                    opt.zero_grad()
                    moved = coords + free
                    dev_len = torch.norm(moved[edge_index[0]] - moved[edge_index[1]], dim=1) - rest_lengths
                    dev_len[grasped] *= 0                        # the reference indexes the EDGE list with the particle id
                    torch.sum(dev_len ** 2).backward()
                    opt.step()
            nxt = free.detach()
            nxt[grasped] = action
        predictions.append(nxt)
        coords = coords + nxt
        history = torch.cat([history[1:], nxt.to(device)[None]], 0)
    predictions = torch.stack(predictions)
    loss = (predictions - truth.to(device)) ** 2
    _ = NodeType  # node-type masks are computed but unused by the reference's rollout (train_meshnet_sim.py:168-172)
    return {"initial_pos": node_coords[0].cpu().numpy(), "predicted_rollout": predictions.detach().cpu().numpy(),
            "ground_truth_rollout": truth.cpu().numpy(), "node_coords": node_coords.cpu().numpy(),
            "node_types": node_types.cpu().numpy(), "edge_index": edge_index.cpu().numpy(), "faces": faces.cpu().numpy(),
            "dt": dt, "mean_loss": loss.mean().detach().cpu().numpy()}
