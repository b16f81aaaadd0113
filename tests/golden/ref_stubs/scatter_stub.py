"""torch_scatter.scatter_mean by its definition: sum of the rows sharing an index (in row order, as the CPU kernel's
sequential loop does) divided by their count, empty slots = 0 (count clamped to 1)."""
import torch


def scatter_mean(src, index, dim=0, out=None, dim_size=None):
    assert dim == 0 and out is None
    index = index.long()
    size = int(index.max()) + 1 if dim_size is None else dim_size
    total = torch.zeros((size,) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
    total.index_add_(0, index, src)
    count = torch.zeros(size, dtype=src.dtype, device=src.device)
    count.index_add_(0, index, torch.ones(index.shape[0], dtype=src.dtype, device=src.device))
    count = count.clamp(min=1)
    return total / count.view((-1,) + (1,) * (src.dim() - 1))
