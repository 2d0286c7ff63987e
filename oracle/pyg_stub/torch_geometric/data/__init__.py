"""`Data` / `Batch` holders; `Batch.from_data_list` restates the documented collate for x / edge_index
(the C++ twin in the reference is cpp_hex/hex_graph_game/util.cpp:22-41)."""
import torch


class Data:
    def __init__(self, x=None, edge_index=None, **kw):
        self.x, self.edge_index = x, edge_index
        for k, v in kw.items():
            setattr(self, k, v)

    @property
    def num_nodes(self):
        return self.x.size(0)

    def to(self, device):
        for k, v in list(self.__dict__.items()):
            if torch.is_tensor(v):
                setattr(self, k, v.to(device))
        return self


class Batch(Data):
    @classmethod
    def from_data_list(cls, data_list):
        xs, eis, batch, ptr, off = [], [], [], [0], 0
        for i, d in enumerate(data_list):
            n = d.x.size(0)
            xs.append(d.x)
            eis.append(d.edge_index + off)
            batch.append(torch.full((n,), i, dtype=torch.long, device=d.x.device))
            off += n
            ptr.append(off)
        out = cls(x=torch.cat(xs, 0), edge_index=torch.cat(eis, 1))
        out.batch = torch.cat(batch, 0)
        out.ptr = torch.tensor(ptr, dtype=torch.long, device=out.x.device)
        out.num_graphs = len(data_list)
        return out
