"""Drop-in for the reference's `meshnet` GNN modules (meshnet/graph_network.py), backed by the sm_100a
kernels of libcloth_splatting_b200.so (csrc/gnn.cu).  No CPU fallback."""
from .graph_network import (Decoder, EncodeProcessDecode, Encoder, InteractionNetwork, Processor,  # noqa: F401
                            build_mlp)
from .cloth_network import ClothMeshSimulator  # noqa: F401,E402
from .model_utils import IdentityNormalizer, NodeType, Normalizer  # noqa: F401,E402
from .rollout import compute_edge_features, multi_step_loss, update_prediction  # noqa: F401,E402
