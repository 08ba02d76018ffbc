"""`meshnet.cloth_network.ClothMeshSimulator` (meshnet/cloth_network.py:13-254 of KTH-RPL/cloth-splatting):
node-feature assembly + normalisation around EncodeProcessDecode, with the reference's constructor,
method names, `save` / `load` checkpoint format ({'model', '_output_normalizer', '_node_normalizer'}).
The GNN itself is the fused sm_100a path (meshnet/graph_network.py here); forward AND backward run in
csb_gnn_* kernels, the glue below is a handful of [N, <=8] element-wise torch ops."""
import glob
import os
import re

import torch
import torch.nn as nn

from .graph_network import EncodeProcessDecode
from .model_utils import IdentityNormalizer, Normalizer


class ClothMeshSimulator(nn.Module):
    def __init__(self, simulation_dimensions: int, nnode_in: int, nedge_in: int, latent_dim: int,
                 nmessage_passing_steps: int, nmlp_layers: int, mlp_hidden_dim: int, nnode_types: int,
                 node_type_embedding_size: int, normalize: bool = False, device="cpu"):
        super().__init__()
        self._nnode_types = nnode_types
        self._node_type_embedding_size = node_type_embedding_size
        self._encode_process_decode = EncodeProcessDecode(
            nnode_in_features=nnode_in, nnode_out_features=simulation_dimensions, nedge_in_features=nedge_in,
            latent_dim=latent_dim, nmessage_passing_steps=nmessage_passing_steps, nmlp_layers=nmlp_layers,
            mlp_hidden_dim=mlp_hidden_dim).to(device)
        if normalize:
            self._output_normalizer = Normalizer(size=simulation_dimensions, name="output_normalizer", device=device)
            self._node_normalizer = Normalizer(size=nnode_in, name="node_normalizer", device=device)
        else:  # cloth_network.py:64-65 builds both identities with size=simulation_dimensions
            self._output_normalizer = IdentityNormalizer(size=simulation_dimensions, name="output_normalizer",
                                                         device=device)
            self._node_normalizer = IdentityNormalizer(size=simulation_dimensions, name="output_normalizer",
                                                       device=device)
        self._device = device

    def forward(self):
        pass

    def _encoder_preprocessor(self, velocity, node_type, velocity_noise=None):
        """cat[velocity history (+ noise when training), one_hot(node_type)] -> node normaliser
        (cloth_network.py:72-110)."""
        v = velocity if velocity_noise is None else velocity + velocity_noise
        onehot = torch.nn.functional.one_hot(torch.squeeze(node_type.long()), self._node_type_embedding_size)
        return self._node_normalizer(torch.cat([v, onehot], dim=1), self.training)

    def predict_acceleration(self, velocity, node_type, edge_index, edge_features, target_velocities=None,
                             velocity_noise=None):
        """-> (predicted normalised acceleration, normalised target or None) (cloth_network.py:112-163)."""
        feats = self._encoder_preprocessor(velocity, node_type, velocity_noise)
        pred = self._encode_process_decode(feats.to(torch.float32), edge_index, edge_features)
        if target_velocities is None:
            return pred, None
        last = (velocity if velocity_noise is None else velocity + velocity_noise)[:, -3:]
        return pred, self._output_normalizer(target_velocities - last, self.training)

    def predict_velocity(self, velocities, node_type, edge_index, edge_features):
        """One rollout step: v' = v[:, -3:] + denormalised GNN output (cloth_network.py:165-193)."""
        feats = self._encoder_preprocessor(velocities, node_type, velocity_noise=None)
        pred = self._encode_process_decode(feats, edge_index, edge_features)
        return velocities[:, -3:] + self._output_normalizer.inverse(pred)

    def save(self, path=None):
        torch.save({"model": self.state_dict(), "_output_normalizer": self._output_normalizer.get_variable(),
                    "_node_normalizer": self._node_normalizer.get_variable()}, path)

    def load(self, path: str, file="latest"):
        """`file='latest'` picks model-<STEP>.pt with the largest STEP (cloth_network.py:215-254)."""
        if file == "latest":
            names = glob.glob(os.path.join(path, "*model*pt"))
            if not names:
                raise ValueError(f"Did not find any pre-trained weights for the meshnet in: {path}")
            pat = re.compile(r".*model-(\d+).pt")
            best = max([0] + [int(pat.search(n).groups()[0]) for n in names])
            model_file = os.path.join(path, f"model-{best}.pt")
        else:
            model_file = os.path.join(path, file)
        dicts = torch.load(model_file, map_location=self._device, weights_only=False)
        self.load_state_dict(dicts["model"])
        for owner in (k for k in dicts if k != "model"):
            target = getattr(self, owner)
            for name, value in dicts[owner].items():
                setattr(target, name, value)
        print("Simulator model loaded checkpoint %s" % model_file)
