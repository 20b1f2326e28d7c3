"""Shot-parallel FWI across the GPUs of one box: one process per GPU, shots sharded, one allreduce.

The reference runs all shots of an iteration serially on one hard-coded GPU, in strided
mini-batches (models/networks.py:5454-5461 ``batch_src_amps = source_amplitudes[:, it::num_batches]``).
Shots are independent forward/adjoint problems sharing only the model, so here each rank takes a
contiguous block of shots, computes its local misfit and model gradient with the same
``Propagator`` call, and the gradients are summed with ONE ``all_reduce`` (NCCL over
NVLink/NVSwitch on GPUs; gloo in the CPU tests of the host logic).  No other collective is on
the data path; receiver traces stay sharded.
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n_shots: int, rank: int, world: int) -> Tuple[int, int]:
    """[lo, hi) of the contiguous block of shots owned by `rank` (sizes differ by at most one)."""
    if world < 1 or not (0 <= rank < world):
        raise RuntimeError(f"bad rank/world {rank}/{world}")
    base, extra = divmod(int(n_shots), world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_shots(tensors: Sequence[torch.Tensor], shot_dims: Sequence[int], rank: int, world: int):
    """Slices every tensor along its shot dimension to this rank's block."""
    out = []
    for t, d in zip(tensors, shot_dims):
        lo, hi = shard_bounds(t.shape[d], rank, world)
        out.append(t.narrow(d, lo, hi - lo))
    return out


def allreduce_gradient(grad: torch.Tensor, loss: Optional[torch.Tensor] = None,
                       group=None) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """Sums the model gradient (and the scalar misfit, folded into the same message) over ranks.

    The misfit rides in one extra element of the gradient buffer so an iteration costs exactly one
    collective.  Enqueued on the current stream, right behind the last imaging kernel.
    """
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return grad, loss
    flat = grad.reshape(-1)
    if loss is not None:
        buf = torch.empty(flat.numel() + 1, dtype=flat.dtype, device=flat.device)
        buf[:-1] = flat
        buf[-1] = loss.detach().to(flat.dtype)
    else:
        buf = flat.contiguous()
    dist.all_reduce(buf, op=dist.ReduceOp.SUM, group=group)
    if loss is not None:
        return buf[:-1].reshape(grad.shape), buf[-1]
    return buf.reshape(grad.shape), None
