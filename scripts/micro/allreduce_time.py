"""All-reduce of the 6.75 MB fp32 gradient arena alone: device time per call and NCCL's transport.
   NCCL_DEBUG=INFO torchrun --nproc-per-node N scripts/micro/allreduce_time.py"""
import os, torch, torch.distributed as dist
lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
t = torch.ones(1687206, device="cuda")
for _ in range(10):
    dist.all_reduce(t)
torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(200):
    dist.all_reduce(t)
e1.record(); torch.cuda.synchronize()
if dist.get_rank() == 0:
    print("all_reduce 6.75 MB x%d ranks: %.1f us per call" % (dist.get_world_size(), e0.elapsed_time(e1) * 1e3 / 200))
dist.destroy_process_group()
