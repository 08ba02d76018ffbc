"""`meshnet.meshnet_network`: the time-conditioned residual on top of the GNN's mesh predictions that the
refinement loop optimises (meshnet/meshnet_network.py:256-380 of KTH-RPL/cloth-splatting):

    vertices(t) = mesh_predictions[round(t / dt)] + Linear_256->3N(relu(Linear(relu(Linear(sinenc_13(t))))))

Same class names, constructor arguments, parameter names (`input`, `hidden`, `output`, buffer `encoder.freqs`) and
initialisation (output weight ~ N(0, 1e-5), zero bias) so that state_dicts are interchangeable.  The MLP is a
batch-1 GEMV chain (13 -> 256 -> 256 -> 3N): library GEMV, not a hot kernel."""
from typing import Optional

import torch
import torch.nn as nn


class SinusoidalEncoder(nn.Module):
    """NeRF-style encoding [x, sin(2^k s x), sin(2^k s x + pi/2)]_k (meshnet_network.py:256-323)."""

    def __init__(self, input_dim: int, num_freqs: int, min_freq_log2: int = 0, max_freq_log2: Optional[int] = None,
                 scale: float = 1.0, use_identity: bool = True, device="cpu", **kwargs):
        super().__init__()
        self.num_freqs = num_freqs
        self.min_freq_log2 = min_freq_log2
        self.max_freq_log2 = max_freq_log2 if max_freq_log2 else min_freq_log2 + num_freqs - 1.0
        self.scale = scale
        self.use_identity = use_identity
        bands = 2.0 ** torch.linspace(self.min_freq_log2, self.max_freq_log2, int(self.num_freqs), device=device)
        self.register_buffer("freqs", bands.reshape(self.num_freqs, 1))
        self.input_dim = input_dim
        self.output_dim = input_dim * self.num_freqs * 2 + (input_dim if use_identity else 0)

    def __call__(self, x, alpha: Optional[float] = None):
        if self.num_freqs == 0:
            return x
        angles = self.scale * x.unsqueeze(-2) * self.freqs                       # (F, C)
        feats = torch.stack((angles, angles + torch.pi / 2), dim=-2)             # (F, 2, C)
        feats = torch.sin(feats.flatten(start_dim=-3, end_dim=-1))
        return torch.cat([x, feats], dim=-1) if self.use_identity else feats


class ResidualMeshSimulator(nn.Module):
    def __init__(self, mesh_predictions: torch.Tensor, n_times: int = -1, device="cuda"):
        super().__init__()
        self.mesh_predictions = mesh_predictions.to(device)
        self.n_times = n_times if n_times > 0 else self.mesh_predictions.shape[0]
        self.time_delta = 1.0 if self.n_times == 1 else 1.0 / (self.n_times - 1)
        n_nodes = self.mesh_predictions.shape[1]
        self.encoder = SinusoidalEncoder(input_dim=1, num_freqs=6, device=device)
        self.input = nn.Linear(self.encoder.output_dim, 256, device=device)
        self.hidden = nn.Linear(256, 256, device=device)
        self.output = nn.Linear(256, n_nodes * 3, device=device)
        nn.init.normal_(self.output.weight, 0.0, 0.00001)
        nn.init.constant_(self.output.bias, 0.0)

    def check_time(self, time: float):
        """The reference's bounds check (meshnet_network.py:367-369) from a host-side time: callers that know the time as a
        Python number (render() does: viewpoint_camera.time) check here and pass `checked=True` to forward(), which then
        needs no device-to-host synchronisation."""
        if round(float(time) / self.time_delta) >= self.n_times:
            raise ValueError(f"Time {time} is out of bounds for the mesh simulator.")

    def forward(self, time_vector, checked=False):
        time = time_vector[0, :]
        h = torch.relu(self.input(self.encoder(time)))
        h = torch.relu(self.hidden(h))
        residual = self.output(h).reshape(-1, 3)
        time_id = torch.round(time / self.time_delta).to(dtype=torch.long)
        if checked:
            # the host check used double arithmetic: a last-ulp disagreement must not become an out-of-bounds read
            time_id = time_id.clamp(max=self.n_times - 1)
        elif time_id >= self.n_times:  # a device-to-host read, as in the reference
            raise ValueError(f"Time {time} is out of bounds for the mesh simulator.")
        return self.mesh_predictions[time_id].squeeze() + residual

    def forward_times(self, times):
        """The deformed meshes of SEVERAL times at once: `times` [V] (a device tensor) -> [V, N, 3], row v equal to
        `forward(time_vector filled with times[v], checked=True)` up to the rounding of a batched instead of a
        matrix-vector product.  The views of a refinement step (three consecutive times) then cost one pass through the
        MLP instead of three.  The caller checks the bounds on the host (`check_time`)."""
        t = times.reshape(-1, 1)
        h = torch.relu(self.input(self.encoder(t)))
        h = torch.relu(self.hidden(h))
        residual = self.output(h).reshape(t.shape[0], -1, 3)
        time_id = torch.round(t[:, 0] / self.time_delta).to(dtype=torch.long).clamp(max=self.n_times - 1)
        return self.mesh_predictions[time_id] + residual

    def save(self, path):
        torch.save(self.state_dict(), path)

    def load(self, path):
        self.load_state_dict(torch.load(path))
