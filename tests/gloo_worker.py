"""world_size-2 worker (gloo, CPU): each rank partitions the same mesh on its own, builds ONLY its block, and runs the halo
exchange wire protocol of csrc/cuda/halo.cu (message slot dest-1 <- element src-1 on the sender; element dest-1 <- slot src-1 on
the receiver; one message per (peer, layer), k fastest inside an element) with torch.distributed send/recv in place of
ncclSend/ncclRecv.  Launched by tests/test_gloo_ranks.py."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import common  # noqa: E402


def exchange(blk, kind, field, rank):
    sends = blk.exchange_lists(kind, "send")
    recvs = blk.exchange_lists(kind, "recv")
    reqs, bufs = [], []
    for x in sends:
        msg = torch.zeros((len(x["src"]),) + field.shape[1:], dtype=torch.float64)
        msg[torch.from_numpy(x["dest"].astype(np.int64) - 1)] = torch.from_numpy(field[x["src"] - 1])
        reqs.append(dist.isend(msg, dst=x["endPoint"], tag=x["layer"]))
    for y in recvs:
        buf = torch.zeros((len(y["src"]),) + field.shape[1:], dtype=torch.float64)
        bufs.append(buf)
        reqs.append(dist.irecv(buf, src=y["endPoint"], tag=y["layer"]))
    for r in reqs:
        r.wait()
    for y, buf in zip(recvs, bufs):
        field[y["dest"] - 1] = buf.numpy()[y["src"] - 1]


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo")
    host = common.host
    L = 5
    case = host.HostCase(3, lloyd_iters=2, sfc=True)
    case.partition(world, "metis")
    blk = case.build_block(rank, nVertLevels=L, init_case=2, num_scalars=1, full_fields=True)
    cfg = blk.cfg()
    ok = True
    # every rank must have computed the same partition: owned counts sum to the global count, owned ids are disjoint
    own_ids = blk.field("indexToCellID")[: int(cfg["nCellsSolve"])].astype(np.int64)
    counts = [None] * world
    dist.all_gather_object(counts, own_ids.tolist())
    allids = np.concatenate([np.asarray(c) for c in counts])
    ok &= len(allids) == len(np.unique(allids)) == 10 * 4 ** 3 + 2
    for kind, idname, own, n in (("cell", "indexToCellID", "nCellsSolve", "nCells"), ("edge", "indexToEdgeID", "nEdgesSolve", "nEdges"),
                                 ("vertex", "indexToVertexID", "nVerticesSolve", "nVertices")):
        ids = blk.field(idname)
        nown, ntot = int(cfg[own]), int(cfg[n])
        f = np.full((ntot + 1, L), -1.0)
        f[:nown] = ids[:nown, None] * 100.0 + np.arange(L)[None, :]
        exchange(blk, kind, f, rank)
        want = ids[:ntot, None] * 100.0 + np.arange(L)[None, :]
        good = np.array_equal(f[:ntot], want)
        if not good:
            print("rank", rank, kind, "halo mismatch", int((f[:ntot] != want).sum()), flush=True)
        ok &= good
        ok &= bool((f[ntot] == -1.0).all())  # the garbage column is never a destination
    t = torch.tensor([1 if ok else 0])
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print("GLOO_RANKS world=%d -> %s" % (world, "OK" if t.item() == 1 else "MISMATCH"), flush=True)
    dist.destroy_process_group()
    sys.exit(0 if t.item() == 1 else 1)


main()
