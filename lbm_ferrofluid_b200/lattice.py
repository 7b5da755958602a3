"""D2Q9 / D3Q19 velocity sets in the reference's population order
(LBM_collision_2d.py:44-83, LBM_collision_3d.py:47-103)."""
import torch

E2 = [(0, 0), (1, 0), (0, 1), (-1, 0), (0, -1), (1, 1), (-1, 1), (-1, -1), (1, -1)]
W2 = [4 / 9] + [1 / 9] * 4 + [1 / 36] * 4
E3 = [(0, 0, 0), (1, 0, 0), (0, 1, 0), (-1, 0, 0), (0, -1, 0), (1, 1, 0), (-1, 1, 0), (-1, -1, 0), (1, -1, 0),
      (0, 0, 1), (1, 0, 1), (0, 1, 1), (-1, 0, 1), (0, -1, 1),
      (0, 0, -1), (1, 0, -1), (0, 1, -1), (-1, 0, -1), (0, -1, -1)]
W3 = [1 / 3] + [1 / 18] * 4 + [1 / 36] * 4 + [1 / 18] + [1 / 36] * 4 + [1 / 18] + [1 / 36] * 4


def lattice_tensors(dim, device, dtype):
    """(_weight [1,Q,1..], _e [1,Q,dim,1..] int64) attribute tensors the reference ops expose."""
    e, w = (E2, W2) if dim == 2 else (E3, W3)
    ones = [1] * dim
    weight = torch.tensor(w, dtype=torch.float32).reshape(1, len(w), *ones).to(device).to(dtype)
    et = torch.tensor(e, dtype=torch.float32).reshape(1, len(e), dim, *ones).to(device).to(torch.int64)
    return weight, et
