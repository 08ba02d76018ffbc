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
