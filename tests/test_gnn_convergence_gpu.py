"""Training with the B200 GNN kernels converges like training with the fp32 reference op sequence.

The gradient of a freshly initialised 15-step ReLU / LayerNorm network is chaotic at the 10 % level: two bf16
implementations that differ only in the summation order of the scatter-add are as far from each other as either is from
fp32 (tools/diag_gnn_grad.py; profiles/r2_gnn_grad_deviation.md), so a per-tensor gradient tolerance says little about
whether a trainer can use the kernels.  This test answers that directly: the reference's optimisation step
(train_meshnet_sim.py:491-532: multi-step loss, backward, Adam) on one seeded task, from the same initial weights,
  * with the kernels (bf16 tensor-core MLPs, csb_gnn_forward_train / csb_gnn_backward),
  * with the fp32 oracle under torch autograd (oracle/gnn_oracle.py, pinned on the reference module),
  * and, as the control, with the same fp32 oracle on a permuted edge list (equal in exact arithmetic).
Measured on B200 (profiles/r2_gnn_convergence_*.json, 300 steps, cosine-annealed Adam): final loss kernels / oracle =
1.02 (7x7 x 8 graphs) and 1.05 (20x20 x 2), the fp32 control against the fp32 oracle 0.96 / 0.97; on other runs
0.79 / 1.20 and 1.32 / 1.56 (the fp32 arms are not run-to-run deterministic either)."""
import os
import sys

import pytest

import conftest

sys.path.insert(0, os.path.join(conftest.ROOT, "tools"))

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,graphs", [(7, 8), (20, 2)])
def test_training_converges_like_the_fp32_reference(n, graphs):
    import gnn_convergence as G

    r = G.run(n=n, graphs=graphs, steps=300, lr=1e-3, F=2)
    fin, ini = r["final_loss"], r["initial_loss"]
    for arm in ("b200_kernels", "oracle_fp32", "oracle_fp32_permuted"):
        assert fin[arm] < 0.02 * ini[arm], f"{arm} did not converge: {ini[arm]} -> {fin[arm]}"
    # Two correct fp32 runs (the oracle on the original and on a permuted edge list: equal in exact arithmetic) end 3 .. 57 %
    # apart depending on the run -- torch's index_add_ is itself order-nondeterministic and 300 Adam steps amplify it
    # (measured: 0.96x, 0.97x, 1.20x, 1.56x).  That spread is the yardstick, not a property of the kernels, so it is recorded
    # and not asserted; the kernels must end inside the band the two fp32 runs span, widened by half on either side.
    lo = min(fin["oracle_fp32"], fin["oracle_fp32_permuted"])
    hi = max(fin["oracle_fp32"], fin["oracle_fp32_permuted"])
    assert lo / 1.5 < fin["b200_kernels"] < hi * 1.5, fin
    out = os.path.join(conftest.ROOT, "gpurun_out")
    if os.path.isdir(out):
        import json

        json.dump(r, open(os.path.join(out, f"gnn_convergence_{n}x{n}x{graphs}.json"), "w"))
