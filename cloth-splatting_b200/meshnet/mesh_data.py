"""The mesh files between the GNN rollout and the Gaussian refinement: `init_mesh.hdf5` and
`mesh_predictions/mesh_XXX.hdf5` with the datasets `pos`, `face`, `edge_index`, `norm`, as
meshnet/generate_mesh_predictions.py:34-46,94-119 writes them and meshnet/data_utils.py:450-457 reads them.

The reference builds them with torch_geometric transforms (`FaceToEdge(remove_faces=False)`, `GenerateMeshNormals`,
`Cartesian(norm=False)`, `Distance(norm=False)`); torch_geometric (2.x) is not in this image, so their published rules are
restated here in torch -- parity of this file is UNPINNED and tested by properties (tests/test_mesh_data_cpu.py).  h5py is
absent too: the `.hdf5` container is written and read by utils/hdf5_lite.py (HDF5 file-format structures from the
specification, checked against a libhdf5-written file), so the files are real HDF5 either way."""
import os

import numpy as np
import torch

from .graph_data import Data


def face_to_edge(face, num_nodes):
    """FaceToEdge: the three sides of every triangle in both directions, duplicates removed, sorted by (source, target)."""
    e = torch.cat([face[:2], face[1:], face[::2]], dim=1)
    e = torch.cat([e, e.flip(0)], dim=1)
    key = torch.unique(e[0] * num_nodes + e[1])                    # sorted
    return torch.stack([key // num_nodes, key % num_nodes])


def mesh_normals(pos, face):
    """GenerateMeshNormals: unit face normals (v1 - v0) x (v2 - v0) summed at the three corners, normalised."""
    n = torch.nn.functional.normalize(torch.linalg.cross(pos[face[1]] - pos[face[0]], pos[face[2]] - pos[face[0]], dim=1), dim=-1)
    out = torch.zeros_like(pos).index_add_(0, face.reshape(-1), n.repeat(3, 1))
    return torch.nn.functional.normalize(out, dim=-1)


def cartesian_distance(pos, edge_index):
    """Cartesian(norm=False) ++ Distance(norm=False): [pos[src] - pos[dst], |pos[dst] - pos[src]|] per edge, [E,4]."""
    src, dst = edge_index
    return torch.cat([pos[src] - pos[dst], (pos[dst] - pos[src]).norm(dim=-1, keepdim=True)], dim=-1)


def get_mesh_data(points, faces):
    """generate_mesh_predictions.py:40-45: Data(pos, face [3,F], edge_index, norm)."""
    return Data(pos=points, face=faces, edge_index=face_to_edge(faces, points.shape[0]), norm=mesh_normals(points, faces))


def predicted_positions(initial_pos, predicted_rollout):
    """Positions after every rollout step (generate_mesh_predictions.py:106): the initial positions plus the running sum of
    the predicted displacements -> [T+1, N, 3], entry 0 being the start."""
    steps = np.cumsum(np.asarray(predicted_rollout), axis=0)
    return np.concatenate([np.asarray(initial_pos)[None], np.asarray(initial_pos)[None] + steps], axis=0)


def position_losses(prediction_data):
    """Per-step mean squared position error of a rollout against its trajectory, with a leading 0 for the start
    (train_meshnet_sim.py:296-301, generate_mesh_predictions.py:103-109)."""
    pred = predicted_positions(prediction_data["initial_pos"], prediction_data["predicted_rollout"])
    truth = np.asarray(prediction_data["node_coords"])
    return [0] + [float(((truth[i] - pred[i]) ** 2).mean()) for i in range(1, pred.shape[0])]


def write_datasets(path, arrays):
    """`with h5py.File(path, "w") as f: f.create_dataset(key, data=value)` for every item: through h5py when it is
    importable, else through utils/hdf5_lite (same container format)."""
    try:
        import h5py
    except ImportError:
        from utils import hdf5_lite

        return hdf5_lite.write_hdf5(path, arrays)
    with h5py.File(path, "w") as f:
        for k, v in arrays.items():
            f.create_dataset(k, data=v)
    return path


def read_datasets(path, names=None):
    """{name: array} of an HDF5 file's root datasets (h5py when importable, else utils/hdf5_lite); a legacy `.npz` of an
    earlier version of this package next to a missing `.hdf5` is still read."""
    if not os.path.exists(path) and os.path.exists(os.path.splitext(path)[0] + ".npz"):
        with np.load(os.path.splitext(path)[0] + ".npz") as f:
            return {k: f[k] for k in (names or f.files)}
    try:
        import h5py
    except ImportError:
        from utils import hdf5_lite

        return hdf5_lite.read_hdf5(path, names)
    with h5py.File(path, "r") as f:
        return {k: f[k][:] for k in (names or f.keys())}


def save_mesh(mesh, path, name="mesh.hdf5"):
    """generate_mesh_predictions.py:34-38.  Returns the path written."""
    arrays = {k: getattr(mesh, k).detach().cpu().numpy() for k in mesh.keys() if isinstance(getattr(mesh, k), torch.Tensor)}
    return write_datasets(os.path.join(path, name), arrays)


def load_mesh(path, device="cuda"):
    """load_mesh_from_h5py (meshnet/data_utils.py:450-457): Data(pos, norm, face, edge_index) on `device`."""
    arrays = read_datasets(path, ("pos", "norm", "face", "edge_index"))
    return Data(**{k: torch.tensor(v, device=device) for k, v in arrays.items()})


def write_mesh_predictions(root, initial_pos, predicted_rollout, faces):
    """The files one rollout leaves (generate_mesh_predictions.py:94-119): `<root>/init_mesh.*`, and
    `<root>/mesh_predictions/mesh_000.*` (the start) ... `mesh_TTT.*`.  Returns the paths."""
    pred = os.path.join(root, "mesh_predictions")
    os.makedirs(pred, exist_ok=True)
    faces = torch.as_tensor(np.asarray(faces)).long()
    positions = predicted_positions(initial_pos, predicted_rollout)
    start = get_mesh_data(torch.from_numpy(positions[0]), faces)
    out = [save_mesh(start, root, name="init_mesh.hdf5")]
    for i, p in enumerate(positions):
        out.append(save_mesh(get_mesh_data(torch.from_numpy(p), faces), pred, name=f"mesh_{i:03}.hdf5"))
    return out
