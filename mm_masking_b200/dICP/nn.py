"""Nearest-neighbour search with the reference's interface (`external/dICP/dICP/nn.py:4-124`).

`find_nn(x, y)` returns, for every row of x, the full row of y (xyz or xyz+normal) that is
nearest in Euclidean distance -- `torch.cdist` -> `argmin` -> `gather` in the reference
(`nn.py:32-38`).  Here the index comes from the tiled brute-force CUDA kernel
(`mmb_nn_search`): squared differences evaluated exactly, lowest index on ties, no (N,n,m)
distance tensor.  Gradients flow to `y` through the gather, as in the reference; the index is a
constant.  The Gumbel-softmax variant (`nn.py:43-70`, off in the reference config) is not
supported.
"""
import torch

from .. import _capi
from .. import ops as _ops  # noqa: F401  (registers torch.ops.mmb.*)


def nearest_index(x: torch.Tensor, y: torch.Tensor, dim: int = 3, nn_slices: int = 0,
                  nn_backend: int = _capi.NN_AUTO) -> torch.Tensor:
    """(N,n) int64 index of the nearest row of y (N,m,3|6) for every row of x (N,n,3)."""
    _capi.require_cuda()
    if x.dtype != y.dtype or x.dtype not in (torch.float32, torch.float64):
        raise TypeError("x and y must both be float32 or float64")
    dev = x.device if x.is_cuda else torch.device("cuda", torch.cuda.current_device())
    xq = x.detach().to(dev).contiguous()
    yt = y.detach().to(dev).contiguous()
    idx = torch.ops.mmb.nn_search(xq, yt, int(dim), int(nn_slices), int(nn_backend))
    return idx.to(torch.int64).to(x.device)


class nn:
    def __init__(self, differentiable=True, use_gumbel=True, eps=1e-20, tau=0.1):
        self.differentiable = differentiable
        self.use_gumbel = use_gumbel
        self.eps = eps
        self.tau = tau

    def find_nn(self, x, y):
        if self.differentiable and self.use_gumbel:
            raise NotImplementedError("Gumbel-softmax nearest neighbour is not part of the CUDA path")
        x_use, y_use = self._handle_dimensions(x, y)
        index = nearest_index(x_use, y_use)
        gather_idx = index.unsqueeze(2).expand(-1, -1, y_use.shape[-1])
        return torch.gather(y_use, 1, gather_idx)

    @staticmethod
    def _handle_dimensions(x, y):
        """Accept (n,3|6), (3|6,n), (N,n,3|6) or (N,3|6,n) like `nn.py:94-124`."""
        x_use = x.unsqueeze(0) if x.dim() == 2 else x
        if x_use.shape[-2] == 3 or (x_use.shape[-2] == 6 and x_use.shape[-2] < x_use.shape[-1]):
            x_use = x_use[:, :3, :].transpose(1, 2)
        assert x_use.shape[2] == 3, "x must have 3 elements in the second dimension."
        y_use = y.unsqueeze(0) if y.dim() == 2 else y
        if y_use.shape[-2] == 3 or (y_use.shape[-2] == 6 and y_use.shape[-2] < y_use.shape[-1]):
            y_use = y_use.transpose(1, 2)
        assert y_use.shape[2] == 3 or y_use.shape[2] == 6, "y must have 3 or 6 elements in the second dimension."
        return x_use, y_use
