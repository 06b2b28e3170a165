"""K3 / K4 at the Llama-2 7B TP2 shapes (2+ ranks): fused kernel vs the same GEMM kernel alone vs NCCL collective + GEMM.
usage: torchrun --nproc-per-node 2 tools/tp_perf.py [comm_ctas ...]      (one JSON line per shape, rank 0)"""
import json
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, "/root/repo")
from tools.multigpu_check import init_distributed, timeit  # noqa: E402

init_distributed("nccl")
from b200flow.ops import _lib  # noqa: E402
from b200flow.ops import gemm as G  # noqa: E402
from b200flow.parallel.fused_tp import FusedTP  # noqa: E402

rank, world = dist.get_rank(), dist.get_world_size()
dev = torch.device("cuda", torch.cuda.current_device())
ctas = [int(a) for a in sys.argv[1:]] or [16]
T = 4096
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def t_of(fn):
    def f():
        flush.zero_()
        fn()
    base = timeit(lambda: flush.zero_())
    return timeit(f) - base


# (name, kind, N, K, b_mn_major): AG shapes take a [T/world, K] shard, RS shapes a full [T, K] operand
shapes = [("fwd_qkv", "ag", 6144, 4096, False), ("fwd_gate_up", "ag", 11008, 4096, False),
          ("fwd_attn_out", "rs", 4096, 2048, False), ("fwd_down", "rs", 4096, 5504, False),
          ("bwd_qkv_dgrad", "rs", 4096, 6144, True), ("bwd_gate_up_dgrad", "rs", 4096, 11008, True),
          ("bwd_attn_out_dgrad", "ag", 2048, 4096, True), ("bwd_down_dgrad", "ag", 5504, 4096, True)]
for c in ctas:
    ftp = FusedTP.create(dist.group.WORLD, rank, world, dev)
    ftp.comm_ctas = c
    for name, kind, N, K, mn in shapes:
        w = (torch.randn(K, N, device=dev) if mn else torch.randn(N, K, device=dev)).to(_lib.HALF)
        if kind == "ag":
            xs = torch.randn(T // world, K, device=dev, dtype=_lib.HALF)
            xf = torch.randn(T, K, device=dev, dtype=_lib.HALF)
            t_f = t_of(lambda: ftp.ag_gemm(xs, w, b_mn_major=mn))
            t_g = t_of(lambda: G.gemm(xf, w, b_mn_major=mn))

            def nccl():
                out = torch.empty(T, K, device=dev, dtype=_lib.HALF)
                dist.all_gather_into_tensor(out, xs)
                G.gemm(out, w, b_mn_major=mn)
            t_n = t_of(nccl)
            wire = T // world * K * 2 * (world - 1)
        else:
            x = torch.randn(T, K, device=dev, dtype=_lib.HALF)
            t_f = t_of(lambda: ftp.gemm_rs(x, w, b_mn_major=mn))
            t_g = t_of(lambda: G.gemm(x, w, b_mn_major=mn))

            def nccl():
                y = G.gemm(x, w, b_mn_major=mn)
                out = torch.empty(T // world, N, device=dev, dtype=_lib.HALF)
                dist.reduce_scatter_tensor(out, y)
            t_n = t_of(nccl)
            wire = T // world * N * 2 * (world - 1)
        if rank == 0:
            fl = 2 * T * N * K
            print(json.dumps({"shape": name, "kind": kind, "T": T, "N": N, "K": K, "comm_ctas": c, "fused_ms": round(t_f, 4),
                              "gemm_alone_ms": round(t_g, 4), "nccl_plus_gemm_ms": round(t_n, 4),
                              "fused_over_gemm": round(t_f / t_g, 3), "fused_tflops": round(fl / t_f / 1e9, 1),
                              "wire_mib": wire / 2 ** 20, "wire_floor_ms": round(wire / 770e9 * 1e3, 4)}), flush=True)
dist.barrier()
dist.destroy_process_group()
