"""Multi-GPU partition of the assembly, host-side logic only (torch as the array library, runs on CPU
tensors too so that it can be tested under gloo).

Owner-computes: the list of elements to assemble (element order = spatial slabs of the mesh) is cut into `world` contiguous
chunks; a CSR row (dof) belongs to the rank of the FIRST listed element that touches it (row_owners); rank r assembles exactly
its rows and lists every element touching one of them (elements_of_owner: its chunk plus a halo layer out of the later chunks),
so elements on a chunk boundary are evaluated by both neighbours instead of exchanging interface rows.  The only collective on
the path is the all-reduce of the Integrate() scalars (elements are cut into contiguous blocks for those).
(row_blocks / elements_touching -- contiguous dof ranges -- remain for spaces whose dof numbering follows the mesh; for H1
numbered "vertices, then edges" every rank's range meets nearly all elements, measured: k_geom / k_cutmat did not scale.)"""
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


def element_blocks_weighted(weights: torch.Tensor, world: int) -> List[Tuple[int, int]]:
    """contiguous element blocks [lo, hi) tiling [0, ne) with ~equal total weight (element-wise sweeps: a cut element costs
    tens of uncut ones -- cut ratio / cut integral against a sign test -- and the cut elements of a sphere crowd the middle slabs)"""
    ne = weights.numel()
    csum = torch.cumsum(weights.to(torch.float64), 0)
    total = float(csum[-1].item()) if ne else 0.0
    bounds = [0]
    for r in range(1, world):
        b = int(torch.searchsorted(csum, torch.tensor(total * r / world, device=csum.device, dtype=csum.dtype)).item())
        bounds.append(max(bounds[-1], min(b, ne)))
    bounds.append(ne)
    return [(bounds[r], bounds[r + 1]) for r in range(world)]


def row_owners(el2dof_active: torch.Tensor, ndof: int, world: int, weights: torch.Tensor = None) -> torch.Tensor:
    """owner[dof] = rank whose chunk of the active list holds the first element touching the dof; -1: untouched.  int32 [ndof].
    The list is cut into `world` contiguous chunks of equal element count, or of equal total weight when per-element cost
    weights are given (cut elements cost ~3x an uncut one: quadrature block + row gather)"""
    n, nd = el2dof_active.shape
    dev = el2dof_active.device
    first = torch.full((ndof,), n, dtype=torch.int64, device=dev)
    idx = torch.arange(n, device=dev).repeat_interleave(nd)
    first.scatter_reduce_(0, el2dof_active.reshape(-1).long(), idx, reduce="amin")
    if weights is None:
        owner = torch.div(first * world, max(n, 1), rounding_mode="floor")
    else:
        w = weights.to(torch.float64)
        before = torch.cumsum(w, 0) - w                                  # weight in front of each list position
        rank_of_pos = torch.clamp((before * (world / float(w.sum()))).floor().long(), max=world - 1)
        owner = rank_of_pos[torch.clamp(first, max=max(n - 1, 0))]
    return torch.where(first < n, owner, torch.full_like(owner, -1)).to(torch.int32)


def elements_of_owner(el2dof: torch.Tensor, active: torch.Tensor, owner: torch.Tensor, me: int) -> torch.Tensor:
    """the elements of `active` with at least one dof owned by rank `me`: the rank's list incl. halo (in list order)"""
    hit = (owner[el2dof[active.long()].long()] == me).any(dim=1)
    return active[hit]
