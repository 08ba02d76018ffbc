"""Training samples of the cloth simulator: which window of which trajectory sample `idx` is, and the graph built from
it (`TrajectoriesClothSimDataset.__getitem__`, meshnet/dataloader_sim.py:150-170, over the cumulative lengths of
`TrajectoriesClothDataset._compute_cumulative_lengths`, meshnet/dataloader.py:40-48).  A trajectory of T frames gives
T - input_length_sequence - future_sequence_length + 1 samples; sample `idx` reads the velocity history
[t-L, t), the targets [t, t+F) and the actions [t-1, t+F-1) with t = L + offset inside its trajectory.  Loading the
trajectories themselves (h5 files, Delaunay meshing) is data preparation and stays outside."""
import numpy as np

from .graph_data import data_to_graph


class TrajectorySamples:
    def __init__(self, data, input_length_sequence=1, future_sequence_length=1, transform=None):
        """data: list of dicts with 'pos' [T,N,3], 'velocity' [T,N,3], 'node_type' [T,N,1], 'actions' [T,...],
        'edge_faces' (per frame, [3,F] tensors) and 'grasped_particle'."""
        self._data, self._transform = data, transform
        self._input_length_sequence, self._future_sequence_length = input_length_sequence, future_sequence_length
        self._data_lengths = [x["pos"].shape[0] - input_length_sequence - future_sequence_length + 1 for x in data]
        self._length = sum(self._data_lengths)
        self._precompute_cumlengths = np.cumsum(np.array(self._data_lengths, dtype=int))

    def __len__(self):
        return self._length

    def window(self, idx):
        """-> (trajectory, t): sample idx is frame t of that trajectory (history [t-L, t), targets [t, t+F))."""
        trajectory = int(np.searchsorted(self._precompute_cumlengths - 1, idx, side="left"))
        start = self._precompute_cumlengths[trajectory - 1] if trajectory != 0 else 0
        return trajectory, int(self._input_length_sequence + (idx - start))

    def __getitem__(self, idx):
        k, t = self.window(idx)
        d, L, F = self._data[k], self._input_length_sequence, self._future_sequence_length
        if L == 1:
            velocity = d["velocity"][t - 1]
        else:
            velocity = np.concatenate([v for v in d["velocity"][t - L:t]], 1)
        graph = data_to_graph(d["actions"][t - 1:t + F - 1], d["grasped_particle"], velocity, d["node_type"][t - 1],
                              d["edge_faces"][t - 1], d["velocity"][t:t + F], d["pos"][t:t + F], d["pos"][t - 1])
        return self._transform(graph) if self._transform is not None else graph
