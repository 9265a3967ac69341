"""Multi-GPU partition of the assembly, host-side logic only (torch as the array library, runs on CPU
tensors too so that it can be tested under gloo).

Owner-computes: the CSR rows (dofs) are cut into `world` contiguous blocks with about the same number of (row, element)
pairs; rank r assembles exactly the rows of block r and lists every element that touches one of them, so elements on a block
boundary are evaluated by both neighbours instead of exchanging interface rows.  The only collective on the path is the
all-reduce of the Integrate() scalars (elements are cut into contiguous blocks for those)."""
from typing import List, Tuple

import torch


def row_blocks(el2dof_active: torch.Tensor, ndof: int, world: int) -> List[Tuple[int, int]]:
    """contiguous dof ranges [lo, hi) that tile [0, ndof) and carry ~equal numbers of (row, element) pairs.
    el2dof_active: [n][nd] dofs of the elements that will be assembled."""
    counts = torch.bincount(el2dof_active.reshape(-1).long(), minlength=ndof)
    csum = torch.cumsum(counts, 0)
    total = int(csum[-1].item()) if ndof else 0
    bounds = [0]
    for r in range(1, world):
        target = (total * r) // world
        # first dof whose inclusive prefix exceeds the target
        b = int(torch.searchsorted(csum, torch.tensor(target, device=csum.device, dtype=csum.dtype), right=True).item())
        bounds.append(max(bounds[-1], min(b, ndof)))
    bounds.append(ndof)
    return [(bounds[r], bounds[r + 1]) for r in range(world)]


def elements_touching(el2dof: torch.Tensor, active: torch.Tensor, lo: int, hi: int) -> torch.Tensor:
    """the elements of `active` (element numbers) with at least one dof in [lo, hi): the rank's list incl. halo"""
    d = el2dof[active.long()]
    hit = ((d >= lo) & (d < hi)).any(dim=1)
    return active[hit]


def element_block(ne: int, rank: int, world: int) -> Tuple[int, int]:
    """contiguous element block of a rank for element-wise reductions (Integrate)"""
    return (ne * rank) // world, (ne * (rank + 1)) // world
