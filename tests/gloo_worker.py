"""One rank of a CPU-only `gloo` run of the HOST side of the distributed path (spawned by tests/test_dist_host_gloo.py):
block-row partition, process grid, receive plan of the library (b200hypre_block_partition / process_grid / plan_recv, no
GPU needed) and the request / send-list handshake of build_partition_and_halo (csrc/parcsr.cu) restated over
torch.distributed, followed by one halo exchange and the distributed SpMV  y = A_diag x_own + A_offd x_halo  in numpy.
Writes <out>/rank<r>.npz."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "hypre-mini-app_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    out, n, stencil = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    import torch
    import torch.distributed as dist
    import b200hypre as bh
    from oracle_lib import Oracle
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    grid = bh.process_grid(world)                                   # laplace_3d_weak_scaling.hpp:98-169
    o = Oracle()
    o.gen_laplace((n, n, n), p=grid, stencil=stencil)              # global matrix, rank-contiguous numbering
    A = o.scipy(0, "A")
    N = A.shape[0]
    starts = bh.block_partition(N, world)                           # HypreSystem.cpp:525-544
    lo, hi = int(starts[rank]), int(starts[rank + 1])
    assert hi - lo == n ** 3                                        # one box per rank
    Aloc = A[lo:hi].tocsr()
    cols = Aloc.indices.astype(np.int64)
    off = (cols < lo) | (cols >= hi)
    col_map = np.unique(cols[off])                                  # sorted global ids of the halo columns
    recv_rank, recv_off = bh.plan_recv(col_map, starts)             # grouped by owner, contiguous segments
    # how many entries every rank requests from every other rank (allgather #2 of build_partition_and_halo)
    req = torch.zeros(world, dtype=torch.int64)
    for k, r in enumerate(recv_rank):
        req[int(r)] = int(recv_off[k + 1] - recv_off[k])
    allreq = [torch.zeros(world, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(allreq, req)
    send_rank = [r for r in range(world) if int(allreq[r][rank]) > 0]
    # ship the requested global ids to their owners, who turn them into local pack indices
    ops, send_ids = [], {}
    for k, r in enumerate(recv_rank):
        ops.append(dist.isend(torch.from_numpy(col_map[recv_off[k]:recv_off[k + 1]].copy()), int(r)))
    for r in send_rank:
        send_ids[r] = torch.zeros(int(allreq[r][rank]), dtype=torch.int64)
        ops.append(dist.irecv(send_ids[r], r))
    for op in ops:
        op.wait()
    for r in send_rank:
        ids = send_ids[r].numpy()
        assert np.all((ids >= lo) & (ids < hi)), "a rank was asked for rows it does not own"
    # one halo exchange of a global test vector and the distributed product
    rng = np.random.default_rng(99)
    x = rng.standard_normal(N)
    x_own = x[lo:hi]
    halo = np.zeros(len(col_map))
    ops, bufs = [], {}
    for r in send_rank:
        ops.append(dist.isend(torch.from_numpy(x_own[send_ids[r].numpy() - lo].copy()), r))
    for k, r in enumerate(recv_rank):
        bufs[k] = torch.zeros(int(recv_off[k + 1] - recv_off[k]), dtype=torch.float64)
        ops.append(dist.irecv(bufs[k], int(r)))
    for op in ops:
        op.wait()
    for k in bufs:
        halo[recv_off[k]:recv_off[k + 1]] = bufs[k].numpy()
    assert np.array_equal(halo, x[col_map])
    rows = np.repeat(np.arange(hi - lo), np.diff(Aloc.indptr))
    y = np.zeros(hi - lo)
    np.add.at(y, rows[~off], Aloc.data[~off] * x_own[cols[~off] - lo])              # diag block
    np.add.at(y, rows[off], Aloc.data[off] * halo[np.searchsorted(col_map, cols[off])])   # offd block through the column map
    # Krylov-style all-reduce of a partial inner product
    t = torch.tensor([float(x_own @ x_own)], dtype=torch.float64)
    dist.all_reduce(t)
    np.savez(os.path.join(out, f"rank{rank}.npz"), y=y, lo=lo, hi=hi, n_halo=len(col_map), recv_rank=np.asarray(recv_rank),
             send_rank=np.asarray(send_rank), dot=t.item(), grid=np.asarray(grid))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
