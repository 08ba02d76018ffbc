class Data:  # placeholder so that `from torch_geometric.data import Data` resolves
    pass
