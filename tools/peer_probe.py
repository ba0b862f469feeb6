"""What does the NVLink path sustain from a kernel? (torchrun, 2+ ranks). Rank 0 copies between its own memory and
rank 1's, in each direction, with different load flavours / unrolls / grids; also both directions at once (rank 0
and rank 1 active together)."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gke-ray-train_b200"))


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    from b200_train import _lib
    from b200_train.symm import PeerComm
    comm = PeerComm(dev)
    nbytes = 1 << 30
    t = comm.alloc(nbytes // 2, torch.bfloat16, dev)
    t.fill_(1.0)
    sb, _ = t._b2_symm
    local_buf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    peer = sb.ptrs[(rank + 1) % world]
    stream = torch.cuda.current_stream().cuda_stream

    def run(mode, unroll, flavour, blocks, both):
        active = both or rank == 0
        src, dst = (peer, local_buf.data_ptr()) if mode == 0 else (local_buf.data_ptr(), peer)
        ts = []
        for it in range(4):
            dist.barrier(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            if active:
                # (a probes build: B2_BUILD_PROBES=1 python gke-ray-train_b200/build.py; B2_LIB=.../build/ab/libb2train_probes.so)
                rc = _lib.lib.b2_debug_peer_copy(src, dst, nbytes, unroll, flavour, blocks, stream)
                assert rc == 0, _lib.lib.b2_last_error()
            e1.record(); torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        ms = sorted(ts[1:])[1]
        if rank == 0:
            print(f"mode={'read ' if mode == 0 else 'write'} unroll={unroll} flavour={flavour} blocks={blocks} both_ranks={int(both)}: "
                  f"{ms:7.3f} ms  {nbytes / ms / 1e6:7.1f} GB/s per direction", flush=True)

    for both in (False, True):
        for mode in (0, 1):
            for (u, f, b) in ((1, 0, 296), (4, 0, 296), (8, 0, 296), (4, 1, 296), (8, 1, 296), (4, 2, 296), (4, 0, 148), (4, 0, 592), (8, 1, 592)):
                run(mode, u, f, b, both)
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
