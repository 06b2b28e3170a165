"""GEMM->reduce-scatter timing only (2+ ranks); B2F_RS_DEBUG=1/2 switch off parts of the protocol."""
import json, os, sys, torch, torch.distributed as dist
sys.path.insert(0, "/root/repo")
from tools.multigpu_check import init_distributed, timeit
init_distributed("nccl")
from b200flow.parallel.fused_tp import FusedTP
rank, world = dist.get_rank(), dist.get_world_size()
dev = torch.device("cuda", torch.cuda.current_device())
ftp = FusedTP.create(dist.group.WORLD, rank, world, dev)
T, N, K = 4096, 4096, 4096
x = torch.randn(T, K, device=dev, dtype=torch.bfloat16)
w = torch.randn(N, K, device=dev, dtype=torch.bfloat16)
t_rs = timeit(lambda: ftp.gemm_rs(x, w))
t_plain = timeit(lambda: torch.matmul(x, w.t()))
from b200flow.ops import gemm as G
t_ours = timeit(lambda: G.gemm(x, w))
if rank == 0:
    print(json.dumps({"dbg": os.environ.get("B2F_RS_DEBUG", "0"), "gemm_rs_ms": t_rs, "cublas_ms": t_plain, "gemm2_ms": t_ours}), flush=True)
dist.barrier(); dist.destroy_process_group()
