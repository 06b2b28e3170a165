"""Communication log (DeepSpeed `comms_logger` equivalent) over gloo."""
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from b200flow.obs.comms import CommsLogger

    orig = dist.all_reduce
    with CommsLogger() as log:
        for _ in range(3):
            dist.all_reduce(torch.ones(1024))
        dist.all_reduce(torch.ones(10))
        outs = [torch.empty(256) for _ in range(world)]
        dist.all_gather(outs, torch.full((256,), float(rank)))
        dist.broadcast(torch.zeros(64), src=0)
        log._on_native("b2f_gemm_ar_bf16")          # what the native binding reports for a fused kernel
        log._on_native("b2f_gemm_bf16")             # not a communication kernel: ignored
    restored = dist.all_reduce is orig
    dist.all_reduce(torch.ones(4))                   # after disable: not recorded
    s = log.summary()
    lines = []
    log.log_summary(lines.append)
    q.put((rank, s, restored, lines))
    dist.barrier()
    dist.destroy_process_group()


def test_comms_logger_counts_sizes_and_bus_bandwidth():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, 29871, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, s, restored, lines in res:
        assert restored
        ops = {(o["op"], o["message_bytes"]): o for o in s["library"]}
        assert ops[("all_reduce", 4096)]["count"] == 3 and ops[("all_reduce", 40)]["count"] == 1
        assert ("all_reduce", 16) not in ops
        assert ops[("all_gather", 2 * 256 * 4)]["count"] == 1 and ops[("broadcast", 256)]["count"] == 1
        ar = ops[("all_reduce", 4096)]
        assert ar["total_ms"] > 0 and abs(ar["busbw_gbps"] / ar["algbw_gbps"] - 1.0) < 1e-9     # 2(n-1)/n = 1 at n = 2
        assert s["fused_kernels"] == {"fused GEMM -> all-reduce (K4')": 1}
        assert any(l.startswith("all_reduce") for l in lines) and any("K4'" in l for l in lines)
