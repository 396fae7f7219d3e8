"""Many-to-one NVLink ceiling: every rank r > 0 copies its 120-frame block (1.5 GB) into rank 0's symmetric buffer at the
same time (plain peer copies, no rendering).  torchrun --nproc-per-node N tools/ingress_probe.py"""
import os, sys, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
dist.init_process_group("nccl", device_id=dev)
F, H, W = 120, 1024, 1024
buf = symm.empty((world * F, H, W, 3), dtype=torch.float32, device=dev)
hdl = symm.rendezvous(buf, dist.group.WORLD)
target = hdl.get_buffer(0, (world * F, H, W, 3), torch.float32)
src = torch.rand(F, H, W, 3, device=dev)
for it in range(3):
    if rank: target[rank * F:(rank + 1) * F].copy_(src)
    hdl.barrier()
torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for it in range(5):
    if rank: target[rank * F:(rank + 1) * F].copy_(src)
    hdl.barrier()
e1.record(); torch.cuda.synchronize()
ms = torch.tensor([e0.elapsed_time(e1) / 5], device=dev); dist.all_reduce(ms, op=dist.ReduceOp.MAX)
if rank == 0:
    gb = (world - 1) * src.numel() * 4 / 1e9
    print(f"many-to-one: {world - 1} senders x {src.numel() * 4 / 1e9:.2f} GB in {ms.item():.2f} ms = {gb / ms.item() * 1e3:.0f} GB/s into rank 0")
dist.destroy_process_group()
