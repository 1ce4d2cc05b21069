"""world_size-2 gloo tests (CPU) of the multi-GPU plumbing in xtant-3_b200/multigpu.py: partition, the all-gathers
of block ranges / column panels / forces and the supercell-force reduction.  The buffers stand in for the device
block stores: each rank fills only what it owns, the exchange must reproduce the single-process buffer bit for bit."""
import importlib
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

mg = importlib.import_module("xtant-3_b200.multigpu")


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, na, n_orb, seed, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(seed)
        nn = rng.integers(1, 7, size=na)                       # ragged neighbour counts
        rowptr = np.concatenate([[0], np.cumsum(nn + 1)])
        nb2 = n_orb * n_orb
        full_blocks = rng.standard_normal(int(rowptr[-1]) * nb2)
        N = na * n_orb
        full_dense = rng.standard_normal(N * N)
        full_F = rng.standard_normal(3 * na)
        a0, a1 = mg.partition(na, world, rank)
        # each rank starts with only its own part (the rest is poison)
        blocks = torch.full((full_blocks.size,), np.nan, dtype=torch.float64)
        b0, b1 = int(rowptr[a0]) * nb2, int(rowptr[a1]) * nb2
        blocks[b0:b1] = torch.from_numpy(full_blocks[b0:b1])
        dense = torch.full((N * N,), np.nan, dtype=torch.float64)
        dense[a0 * n_orb * N:a1 * n_orb * N] = torch.from_numpy(full_dense[a0 * n_orb * N:a1 * n_orb * N])
        F = torch.full((3 * na,), np.nan, dtype=torch.float64)
        F[3 * a0:3 * a1] = torch.from_numpy(full_F[3 * a0:3 * a1])
        mg.allgather_blocks(blocks, rowptr, na, nb2, rank, world)
        # two block stores in one collective (H and S / rho and rho_E)
        pair = [torch.full((full_blocks.size,), np.nan, dtype=torch.float64) for _ in range(2)]
        pair[0][b0:b1] = torch.from_numpy(full_blocks[b0:b1]); pair[1][b0:b1] = torch.from_numpy(-2.0 * full_blocks[b0:b1])
        mg.allgather_blocks(pair, rowptr, na, nb2, rank, world)
        assert np.array_equal(pair[0].numpy(), full_blocks) and np.array_equal(pair[1].numpy(), -2.0 * full_blocks)
        # segmented block layout (xtb_set_block_layout): every rank's range starts at a multiple of the largest range's
        # block count; the segments are equal and tile the buffer, so the pair goes through the in-place path
        per = (na + world - 1) // world
        cnt = [int(rowptr[min(na, (r + 1) * per)] - rowptr[min(na, r * per)]) for r in range(world)]
        if min(cnt) > 0:
            mx = max(cnt)
            rp_seg = np.array([(i // per) * mx + int(rowptr[i] - rowptr[(i // per) * per]) for i in range(na)] + [world * mx])
            full_seg = np.zeros(world * mx * nb2)
            for i in range(na):
                n = int(rowptr[i + 1] - rowptr[i]) * nb2
                full_seg[rp_seg[i] * nb2:rp_seg[i] * nb2 + n] = full_blocks[int(rowptr[i]) * nb2:int(rowptr[i]) * nb2 + n]
            s0, s1 = int(rp_seg[a0]) * nb2, int(rp_seg[a1]) * nb2
            seg = [torch.full((full_seg.size,), np.nan, dtype=torch.float64) for _ in range(2)]
            seg[0][s0:s1] = torch.from_numpy(full_seg[s0:s1]); seg[1][s0:s1] = torch.from_numpy(3.0 * full_seg[s0:s1])
            mg.allgather_blocks(seg, rp_seg, na, nb2, rank, world)
            assert np.array_equal(seg[0].numpy(), full_seg) and np.array_equal(seg[1].numpy(), 3.0 * full_seg)
        mg.allgather_panels(dense, na, n_orb, rank, world)
        mg.allgather_forces(F, na, rank, world)
        P = torch.arange(9, dtype=torch.float64) * (rank + 1)
        mg.allreduce_scforce(P)
        ok = (np.array_equal(blocks.numpy(), full_blocks) and np.array_equal(dense.numpy(), full_dense)
              and np.array_equal(F.numpy(), full_F)
              and np.array_equal(P.numpy(), np.arange(9.0) * sum(range(1, world + 1))))
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("na,n_orb", [(8, 9), (7, 4), (1, 9)])          # even split, ragged split, an empty rank
def test_exchanges_world2(na, n_orb):
    world = 2
    port = _free_port()
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, na, n_orb, 123 + na, out), nprocs=world, join=True)
    assert all(out.get(r) for r in range(world)), dict(out)


def test_partition_covers_all_atoms():
    for na in (1, 5, 64, 4096, 4097):
        for world in (1, 2, 3, 4, 8):
            r = mg.ranges(na, world)
            assert r[0][0] == 0 and r[-1][1] == na
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            assert all(b >= a for a, b in r)
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(s for s in sizes if s > 0 or True) <= max(sizes)    # contiguous, last ranks short
