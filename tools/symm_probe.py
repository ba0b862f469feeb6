"""Can torch's symmetric-memory allocator give us peer pointers and an NVLS multicast address on this box? (torchrun)"""
import os

import torch
import torch.distributed as dist


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=dev)
    import torch.distributed._symmetric_memory as symm_mem
    try:
        t = symm_mem.empty(1 << 24, dtype=torch.uint8, device=dev)
        hdl = symm_mem.rendezvous(t, dist.group.WORLD)
        print(f"rank {rank}: buffer_ptrs={[hex(p) for p in hdl.buffer_ptrs]} multicast_ptr={hex(hdl.multicast_ptr)} "
              f"signal_pad_size={hdl.signal_pad_size} local={hex(t.data_ptr())} world={hdl.world_size}", flush=True)
        try:
            from torch._C._autograd import DeviceType
            print(f"rank {rank}: has_multicast_support={symm_mem._SymmetricMemory.has_multicast_support(DeviceType.CUDA, local)}", flush=True)
        except Exception as e:  # noqa: BLE001
            print(f"rank {rank}: has_multicast_support query failed: {e}", flush=True)
    except Exception as e:  # noqa: BLE001
        print(f"rank {rank}: symm_mem failed: {type(e).__name__}: {e}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
