"""Multi-GPU plumbing of the tight-binding step: atom-row partition + the exchanges between stages.

One process per GPU (torch.distributed; NCCL over NVLink on the B200 box, gloo in the CPU tests).  The reference's
MPI build replicates everything, splits atom loops round-robin (`do i = 1+rank, N, nranks`) and MPI_Allreduce(SUM)s
whole dense arrays (Source_files/TB_NRL.f90:267-327, MPI_subroutines.f90:3965-3982).  Here every block has exactly
one owner, so the exchanges are all-gathers of contiguous block ranges and no dense array is ever summed:

    stage                          exchanged                               bytes per rank at au4096, G ranks
    S2 blocks (H, S)               all_gather of the owners' block ranges   2 * 8 * 81 * B / G      (0.42 GB / G)
       dense H_non, Sij            expanded LOCALLY from the gathered blocks (xtb_expand_dense), nothing on the wire
    S3 owner blocks (rho, rho_E)   all_gather of the owners' block ranges   same
    S4 forces                      all_gather of 3 * Na / G doubles
    S5 supercell force             all_reduce(SUM) of 9 doubles

The functions below operate on plain torch tensors (CPU or CUDA), so the same code is exercised by
tests/test_multigpu_cpu.py with the gloo backend and world_size 2.
"""
import torch
import torch.distributed as dist


def partition(na, world, rank):
    """contiguous atom range [a0, a1) of `rank` (0-based, a1 exclusive); the last ranks may be short or empty"""
    per = (na + world - 1) // world
    a0 = min(na, rank * per)
    return a0, min(na, a0 + per)


def ranges(na, world):
    return [partition(na, world, r) for r in range(world)]


_scratch = {}


def _scratch_buf(key, n, like):
    t = _scratch.get(key)
    if t is None or t.numel() < n or t.device != like.device or t.dtype != like.dtype:
        t = torch.empty(n, dtype=like.dtype, device=like.device)
        _scratch[key] = t
    return t[:n]


def _direct_gather(bufs, group):
    """XTB_MG_GATHER=direct selects the copy-free exchange on CUDA tensors over NCCL (default: packed slots)"""
    import os
    if os.environ.get("XTB_MG_GATHER", "packed") != "direct" or not bufs[0].is_cuda:
        return False
    try:
        return dist.get_backend(group) == "nccl"
    except Exception:
        return False


def _gather_segments(bufs, segs, rank, group=None):
    """in-place all-gather of variable-length contiguous segments of the 1-D tensors `bufs` (one tensor or a list of
    tensors sharing the same segmentation): rank r owns buf[segs[r][0]:segs[r][1]] of every buffer; afterwards every rank
    holds all segments.  ONE collective for all buffers: equal lengths that tile a single buffer gather in place;
    otherwise the segments are packed into a flat send buffer of max-length slots, gathered with
    all_gather_into_tensor, and the foreign segments copied out."""
    if isinstance(bufs, torch.Tensor):
        bufs = [bufs]
    world = len(segs)
    if world == 1:
        return
    lens = [e - b for b, e in segs]
    contiguous = all(segs[r][1] == segs[r + 1][0] for r in range(world - 1))
    if len(set(lens)) == 1 and lens[0] > 0 and contiguous:
        # equal segments that tile the buffer (forces; block stores laid out with xtb_set_block_layout): in place
        b0, e1 = segs[0][0], segs[-1][1]
        for buf in bufs:
            mine = buf[segs[rank][0]:segs[rank][1]]
            if not (buf.is_cuda and dist.get_backend(group) == "nccl"):
                mine = mine.clone()          # gloo does not take an input that aliases the output
            dist.all_gather_into_tensor(buf[b0:e1], mine, group=group)
        return
    mx = max(lens)
    if mx == 0:
        return
    if _direct_gather(bufs, group) and min(lens) > 0:
        # NCCL: all_gather with the destination views themselves as outputs.  For unequal lengths ProcessGroupNCCL
        # runs it as one coalesced group of broadcasts (root r -> segment r), so nothing is packed or copied out.
        for buf in bufs:
            outs = [buf[b:e] for b, e in segs]
            dist.all_gather(outs, outs[rank], group=group)
        return
    nb = len(bufs)
    send = _scratch_buf(("send", nb), nb * mx, bufs[0])
    recv = _scratch_buf(("recv", nb), world * nb * mx, bufs[0])
    b, e = segs[rank]
    for k, buf in enumerate(bufs):
        send[k * mx:k * mx + (e - b)] = buf[b:e]          # the tail of a short slot is never read back
    dist.all_gather_into_tensor(recv, send, group=group)
    for r, (b, e) in enumerate(segs):
        if r != rank and e > b:
            for k, buf in enumerate(bufs):
                o = (r * nb + k) * mx
                buf[b:e] = recv[o:o + (e - b)]


def allgather_blocks(blocks, rowptr, na, nb2, rank, world, group=None):
    """blocks: 1-D tensor [n_blocks * nb2] (block store) or a list of such tensors (H and S, or rho and rho_E: one
    collective for the pair); rank r owns the blocks of its atoms, i.e. the contiguous range rowptr[a0] .. rowptr[a1].
    rowptr: host sequence / tensor of length na + 1."""
    segs = [(int(rowptr[a0]) * nb2, int(rowptr[a1]) * nb2) for a0, a1 in ranges(na, world)]
    _gather_segments(blocks, segs, rank, group)


def allgather_panels(dense, na, n_orb, rank, world, group=None):
    """dense: 1-D view of a column-major (N, N) matrix; rank r owns the column panels of its atoms.
    Kept for callers that want the reference's replicated dense matrices moved over the wire; the bench uses
    allgather_blocks + a local expansion instead (50x fewer bytes)."""
    N = na * n_orb
    segs = [(a0 * n_orb * N, a1 * n_orb * N) for a0, a1 in ranges(na, world)]
    _gather_segments(dense, segs, rank, group)


def allgather_forces(F, na, rank, world, group=None):
    """F: 1-D tensor [3 * na] (forces (3, na) column-major); rank r owns atoms [a0, a1)."""
    segs = [(3 * a0, 3 * a1) for a0, a1 in ranges(na, world)]
    _gather_segments(F, segs, rank, group)


def allreduce_scforce(P, group=None):
    """P: tensor [9], partial supercell force of the owned atoms -> total on every rank"""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(P, op=dist.ReduceOp.SUM, group=group)
    return P
