"""TEST INFRASTRUCTURE ONLY: CPU restatement (plain torch fp32) of the reference's meshnet GNN forward.

Follows meshnet/graph_network.py of KTH-RPL/cloth-splatting @ f7acc90:
  build_mlp :7-45, Encoder :48-111, InteractionNetwork :114-222 (message :178-199, update :201-222,
  residuals :176), Processor :224-292, Decoder :295-332, EncodeProcessDecode :335-408,
and torch_geometric's MessagePassing semantics (third party, unpinned, absent from /root/reference):
x_i = x[edge_index[1]], x_j = x[edge_index[0]], sum aggregation at edge_index[1], update() fed from
propagate()'s original kwargs -> the edge latents are NOT replaced by the edge MLP output, each step
just doubles them (e <- e + e).

Parity status: pinned against the reference's own graph_network.py, imported in the build container
under oracle/pyg_stub (tests/make_golden_gnn.py -> tests/golden/gnn_*.npz).  Real torch_geometric is
not installable here; the stub encodes its documented semantics.

Operates on a state_dict with the reference's key names, so it also checks checkpoint compatibility.
`emulate_bf16=True` rounds the operands of every Linear of the processor/encoder/decoder MLPs to
bf16 (fp32 accumulation), at the places the CUDA kernels do, for a tight comparison.
"""
import torch
import torch.nn.functional as F


def _bf(t, on):
    return t.to(torch.bfloat16).to(torch.float32) if on else t


class _RoundGradBf16(torch.autograd.Function):
    """Identity whose backward rounds the gradient to bf16: the CUDA backward feeds dZ_l to its dgrad / wgrad
    tensor-core contractions as bf16 (fp32 accumulation), the bias gradient still sums the fp32 values."""

    @staticmethod
    def forward(ctx, t):
        return t.view_as(t)

    @staticmethod
    def backward(ctx, g):
        return g.to(torch.bfloat16).to(torch.float32)


def _mlp(sd, prefix, h, nlayers, bf16, first_on_cuda_cores=False, last_on_cuda_cores=False, w0_scale=None):
    """prefix.NN-k.{weight,bias}: Linear + ReLU, last layer Identity (build_mlp, graph_network.py:7-45)."""
    for k in range(nlayers):
        W, b = sd[f"{prefix}.NN-{k}.weight"], sd[f"{prefix}.NN-{k}.bias"]
        if k == 0 and w0_scale is not None:
            W = W * w0_scale
        fp32_layer = (k == 0 and first_on_cuda_cores) or (k == nlayers - 1 and last_on_cuda_cores)
        on = bf16 and not fp32_layer
        hin = _bf(h, on)
        h = F.linear(hin, _bf(W, on))
        if on and h.requires_grad:
            h = _RoundGradBf16.apply(h)
        h = h + b
        if k < nlayers - 1:
            h = F.relu(h)
    return h


def _ln(sd, prefix, h):
    return F.layer_norm(h, (h.size(-1),), sd[f"{prefix}.weight"], sd[f"{prefix}.bias"], 1e-5)


def count_layers(sd, prefix):
    n = 0
    while f"{prefix}.NN-{n}.weight" in sd:
        n += 1
    return n


def forward(sd, x, edge_index, edge_features, emulate_bf16=False, return_intermediates=False, device="cpu",
            differentiable=False):
    """EncodeProcessDecode.forward (graph_network.py:387-408) -> f32[N, nnode_out].

    differentiable=True keeps the autograd graph (sd / x / edge_features are used as they are, already on
    `device`): torch autograd over this function is the oracle of csb_gnn_backward, exactly what the
    reference's loss.backward() differentiates (train_meshnet_sim.py:529)."""
    if differentiable:
        ea = edge_features
    else:
        sd = {k: v.detach().to(torch.float32).to(device) for k, v in sd.items()}
        x = x.detach().float().to(device)
        ea = edge_features.detach().float().to(device)
    src, dst = edge_index[0].long().to(device), edge_index[1].long().to(device)
    bf = emulate_bf16
    nl = count_layers(sd, "_encoder.node_fn.0")
    # encoder (the CUDA path runs the tiny first layers, K = 8 / 4, on the CUDA cores in fp32)
    xl = _ln(sd, "_encoder.node_fn.1", _mlp(sd, "_encoder.node_fn.0", x, nl, bf, first_on_cuda_cores=True))
    e0 = _ln(sd, "_encoder.edge_fn.1", _mlp(sd, "_encoder.edge_fn.0", ea, nl, bf, first_on_cuda_cores=True))
    inter = {"x_enc": xl.clone(), "e_enc": e0.clone()}
    e0 = _bf(e0, bf)  # stored once in bf16; step m sees 2^m * e0 (exact in bf16)
    steps = 0
    while f"_processor.gnn_stacks.{steps}.edge_fn.0.NN-0.weight" in sd:
        steps += 1
    N, L = xl.size(0), xl.size(1)
    for m in range(steps):
        p = f"_processor.gnn_stacks.{m}"
        e = e0 * float(2 ** m)
        h = torch.cat([_bf(xl, bf)[dst], _bf(xl, bf)[src], e], dim=-1)  # [x_i, x_j, e] = [x[dst], x[src], e]
        msg = _ln(sd, f"{p}.edge_fn.1", _mlp(sd, f"{p}.edge_fn.0", h, nl, bf))
        agg = torch.zeros((N, L), dtype=torch.float32, device=msg.device).index_add_(0, dst, msg)
        u = _ln(sd, f"{p}.node_fn.1", _mlp(sd, f"{p}.node_fn.0", torch.cat([agg, xl], dim=-1), nl, bf))
        xl = xl + u
        if m == 0:
            inter["x_step1"] = xl.clone()
            inter["msg_step1"] = msg.clone()
    inter["x_final"] = xl.clone()
    inter["e_final_scale"] = float(2 ** steps)  # e_out = 2^steps * e_enc
    y = _mlp(sd, "_decoder.node_fn", xl, count_layers(sd, "_decoder.node_fn"), bf, last_on_cuda_cores=True)
    return (y, inter) if return_intermediates else y
