import inspect

import torch


class MessagePassing(torch.nn.Module):
    def __init__(self, aggr="add", flow="source_to_target", node_dim=0):
        super().__init__()
        self.aggr = aggr
        assert flow == "source_to_target" and node_dim == 0

    def _args(self, fn, edge_index, kw, skip):
        out = {}
        for name in list(inspect.signature(fn).parameters)[skip:]:
            if name.endswith("_i"):
                out[name] = kw[name[:-2]].index_select(0, edge_index[1])
            elif name.endswith("_j"):
                out[name] = kw[name[:-2]].index_select(0, edge_index[0])
            else:
                out[name] = kw[name]
        return out

    def propagate(self, edge_index, size=None, **kw):
        msg = self.message(**self._args(self.message, edge_index, kw, 0))
        n = kw["x"].size(0)
        assert self.aggr == "add"
        agg = torch.zeros((n, msg.size(1)), dtype=msg.dtype, device=msg.device).index_add_(0, edge_index[1], msg)
        return self.update(agg, **self._args(self.update, edge_index, kw, 1))
