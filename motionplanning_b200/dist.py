"""Sample sharding across the GPUs of one box, one process per GPU (torchrun).

Samples are independent; GPU g owns the contiguous global ids [g*K/G, (g+1)*K/G).  Philox
counters use the GLOBAL sample id, so results do not depend on G.  The only coupling is the
softmin normaliser and the weighted sum: every rank reduces against its LOCAL minimum, one
all-gather of the small partial (rho_g, eta_g, argmin_g, S_g[N*m]) follows, and every rank
folds the partials in rank order with the rescale exp(-(rho_g - rho)/lambda) — one latency,
deterministic, identical on all ranks.

Two transports for that exchange:
  * in-library NCCL (mppi_comm_init): the all-gather is captured inside the step graph;
    torch.distributed is used only to ship the NCCL unique id (plumbing);
  * torch.distributed.all_gather of the partial between mppi_step_local and
    mppi_step_combine (works with gloo on CPU for tests of the host logic).
"""
import ctypes as C

import numpy as np

from . import _capi as capi


def shard_range(K, rank, world):
    """Contiguous block of global sample ids owned by `rank`: (k_offset, K_local)."""
    base, rem = divmod(K, world)
    off = rank * base + min(rank, rem)
    return off, base + (1 if rank < rem else 0)


def partial_to_array(part, N):
    """capi.Partial -> flat float64 array [rho, eta, argmin, rollouts, feasible, S(2N)]."""
    a = np.empty(5 + 2 * N)
    a[0], a[1], a[2], a[3], a[4] = part.rho, part.eta, part.argmin_global, part.rollouts, part.feasible
    a[5:] = np.ctypeslib.as_array(part.S)[: 2 * N]
    return a


def array_to_partial(a, N):
    p = capi.Partial()
    p.rho, p.eta = float(a[0]), float(a[1])
    p.argmin_global, p.rollouts, p.feasible = int(a[2]), int(a[3]), int(a[4])
    C.memmove(p.S, np.ascontiguousarray(a[5:5 + 2 * N]).ctypes.data, 2 * N * 8)
    return p


def exchange_partials(part, N, group=None):
    """All-gather one rank's partial over torch.distributed; returns the list in rank order."""
    import torch
    import torch.distributed as dist

    mine = torch.from_numpy(partial_to_array(part, N))
    backend = dist.get_backend(group)
    if backend == "nccl":
        mine = mine.cuda()
    out = [torch.empty_like(mine) for _ in range(dist.get_world_size(group))]
    dist.all_gather(out, mine, group=group)
    return [array_to_partial(t.cpu().numpy(), N) for t in out]


def init_library_comm(mppi, group=None):
    """Create the in-library NCCL communicator: rank 0 makes the unique id, torch.distributed
    broadcasts it, every rank calls mppi_comm_init."""
    import torch
    import torch.distributed as dist

    from .mppi import comm_unique_id

    rank, world = dist.get_rank(group), dist.get_world_size(group)
    buf = torch.zeros(capi.MPPI_COMM_ID_BYTES, dtype=torch.uint8)
    if rank == 0:
        buf = torch.tensor(list(comm_unique_id()), dtype=torch.uint8)
    if dist.get_backend(group) == "nccl":
        buf = buf.cuda()
    dist.broadcast(buf, src=0, group=group)
    mppi.comm_init(bytes(buf.cpu().numpy().tobytes()), rank, world)


def init_peer_exchange(mppi, group=None):
    """Peer-memory exchange: all-gather the CUDA-IPC handles of the ranks' exchange windows over
    torch.distributed (plumbing) and map them; the per-step exchange then runs in-kernel."""
    import torch
    import torch.distributed as dist

    rank, world = dist.get_rank(group), dist.get_world_size(group)
    mine = torch.tensor(list(mppi.peer_export()), dtype=torch.uint8)
    if dist.get_backend(group) == "nccl":
        mine = mine.cuda()
    out = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(out, mine, group=group)
    mppi.peer_init([bytes(t.cpu().numpy().tobytes()) for t in out], rank, world)
