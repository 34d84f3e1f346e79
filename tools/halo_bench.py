"""Time single halo exchanges across ranks:  torchrun --nproc-per-node N tools/halo_bench.py [level] [reps]
Prints, per field, the message size of rank 0 and the device time per exchange (CUDA events on the library stream) for the
transport the dycore has switched to after 3 steps (MPAS_B200_TUNE_p2p=0 keeps NCCL)."""
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import numpy as np
import torch
import torch.distributed as dist
import common


def main():
    level = int(sys.argv[1]) if len(sys.argv) > 1 else 8
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 50
    rank, lrank, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(lrank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lrank))
    host = common.host
    host.hostlib().mpas_host_set_num_threads(max(1, (os.cpu_count() or 1) // world))
    case = host.HostCase(level, lloyd_iters=5, sfc=True)
    case.partition(world, "metis")
    blk = case.build_block(rank, nVertLevels=55, init_case=2, num_scalars=1, full_fields="lean")
    dy = host.Dycore([blk], common.level_config(level), rank=rank, nranks=world, device=lrank)
    obj = [host.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(obj, src=0)
    dy.comm_init(obj[0])
    dy.init()
    dy.step(3)
    L = host.lib()
    cfg = blk.cfg()
    if rank == 0:
        print("HALO_BENCH world=%d level=%d transport=%s owned cells %d, local cells %d, edges %d / %d" % (
            world, level, dy.halo_transport(), cfg["nCellsSolve"], cfg["nCells"], cfg["nEdgesSolve"], cfg["nEdges"]), flush=True)
    for name, layer in (("rho_pp", 1), ("rtheta_pp", 1), ("tend_u", 1), ("u@2", 0), ("w@2", 0), ("scalars@2", 0)):
        kind = {1: "cell", 2: "edge", 3: "vertex"}[blk.info(name)[1]]
        n = sum(len(x["dest"]) for x in blk.exchange_lists(kind, "recv") if layer == 0 or x["layer"] == layer)
        inner = int(np.prod(blk.info(name)[2][:-1]))
        dy.exchange(name, layer)
        dy.sync(); dist.barrier()
        L.mpas_b200_timer_start()
        for _ in range(reps):
            dy.exchange(name, layer)
        ms = L.mpas_b200_timer_stop()
        if rank == 0:
            print("  %-10s layer %d: recv %7d elements x %3d = %7.2f MB  %8.1f us per exchange  (%.1f GB/s)" % (
                name, layer, n, inner, n * inner * 8 / 1e6, 1e3 * ms / reps, n * inner * 8 / 1e9 / (ms / reps / 1e3)), flush=True)
    dy.close()
    dist.destroy_process_group()


main()
