"""Tiny gloo training loop for the elastic-membership launcher test: prints the world size at every step and keeps
its progress in a file, so that each torchrun re-rendezvous continues where the previous world stopped."""
import os
import sys
import time

import torch
import torch.distributed as dist

progress, total = sys.argv[1], int(sys.argv[2])
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
step = int(open(progress).read()) if os.path.exists(progress) else 0
t = torch.tensor([step], dtype=torch.int64)
dist.broadcast(t, src=0)
step = int(t)
while step < total:
    x = torch.ones(4)
    x[1] = float(rank == 0 and os.path.exists(progress + ".stop"))      # rank 0 tells everyone when to finish
    dist.all_reduce(x)
    assert float(x[0]) == world
    if float(x[1]) > 0:
        break
    step += 1
    if rank == 0:
        tmp = progress + ".tmp"
        open(tmp, "w").write(str(step))
        os.replace(tmp, progress)
    print(f"step {step} world {world} rank {rank} host {os.environ.get('HOSTNAME')}", flush=True)
    time.sleep(0.25)
dist.barrier()
dist.destroy_process_group()
