"""Process-group topology: TP fastest, then DP, then PP (Megatron's rank order).

The reference configures this through Megatron flags
(`--tensor-model-parallel-size 2 --pipeline-model-parallel-size 2`,
examples/megatron-deepspeed/gpt2_345m/pretrain-ddp-tp-pp-zero1.yaml:39-40); the
rank order is the one `mpu.initialize_model_parallel` builds.
Expert-parallel groups are carved out of the data-parallel dimension as in
DeepSpeed-MoE.

`Topology` is pure arithmetic (unit-testable without any process group);
`ParallelState.initialize` materialises torch.distributed groups from it.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.distributed as dist


@dataclass(frozen=True)
class Topology:
    world: int
    tp: int = 1
    pp: int = 1
    ep: int = 1
    sp: int = 1      # DeepSpeed-Ulysses sequence-parallel size, carved out of the data-parallel dimension

    def __post_init__(self):
        assert self.world % (self.tp * self.pp) == 0, "world must be divisible by tp*pp"
        assert self.dp % self.ep == 0, "expert parallel size must divide the data-parallel size"
        assert self.dp % self.sp == 0, "sequence parallel size must divide the data-parallel size"

    @property
    def dp(self) -> int:
        return self.world // (self.tp * self.pp)

    # ---- coordinates -----------------------------------------------------
    def coords(self, rank: int) -> tuple[int, int, int]:
        """rank -> (tp_rank, dp_rank, pp_rank)."""
        tp_rank = rank % self.tp
        dp_rank = (rank // self.tp) % self.dp
        pp_rank = rank // (self.tp * self.dp)
        return tp_rank, dp_rank, pp_rank

    def rank_of(self, tp_rank: int, dp_rank: int, pp_rank: int) -> int:
        return tp_rank + self.tp * (dp_rank + self.dp * pp_rank)

    # ---- group membership lists -------------------------------------------
    def tp_groups(self) -> list[list[int]]:
        return [[self.rank_of(t, d, p) for t in range(self.tp)]
                for p in range(self.pp) for d in range(self.dp)]

    def dp_groups(self) -> list[list[int]]:
        return [[self.rank_of(t, d, p) for d in range(self.dp)]
                for p in range(self.pp) for t in range(self.tp)]

    def pp_groups(self) -> list[list[int]]:
        return [[self.rank_of(t, d, p) for p in range(self.pp)]
                for d in range(self.dp) for t in range(self.tp)]

    def ep_groups(self) -> list[list[int]]:
        """Consecutive blocks of `ep` data-parallel ranks (same tp/pp coordinate)."""
        out = []
        for g in self.dp_groups():
            for i in range(0, len(g), self.ep):
                out.append(g[i:i + self.ep])
        return out

    def sp_groups(self) -> list[list[int]]:
        """Ulysses groups: consecutive blocks of `sp` data-parallel ranks; they share every micro-batch and split
        its sequence.  For gradient reduction they stay ordinary data-parallel ranks."""
        out = []
        for g in self.dp_groups():
            for i in range(0, len(g), self.sp):
                out.append(g[i:i + self.sp])
        return out

    def expert_dp_groups(self) -> list[list[int]]:
        """Ranks holding the same expert shard: stride `ep` through each DP group."""
        out = []
        for g in self.dp_groups():
            for i in range(self.ep):
                out.append(g[i::self.ep])
        return out

    def embedding_groups(self) -> list[list[int]]:
        """First and last pipeline stage of every (tp, dp) column (tied-embedding grads)."""
        if self.pp == 1:
            return []
        return [[g[0], g[-1]] for g in self.pp_groups()]

    def group_containing(self, groups: list[list[int]], rank: int) -> list[int]:
        for g in groups:
            if rank in g:
                return g
        raise ValueError(f"rank {rank} is in no group")

    def describe(self) -> str:
        return f"tp{self.tp}pp{self.pp}dp{self.dp}" + (f"ep{self.ep}" if self.ep > 1 else "") + (
            f"sp{self.sp}" if self.sp > 1 else "")


class ParallelState:
    """Materialised groups for the calling rank."""

    def __init__(self, topo: Topology, rank: int):
        self.topo, self.rank = topo, rank
        self.tp_rank, self.dp_rank, self.pp_rank = topo.coords(rank)
        self.tp_group = self.dp_group = self.pp_group = self.ep_group = None
        self.expert_dp_group = self.embedding_group = self.sp_group = None
        self.sp_ranks = topo.group_containing(topo.sp_groups(), rank)
        self.sp_rank = self.sp_ranks.index(rank)
        self.tp_ranks = topo.group_containing(topo.tp_groups(), rank)
        self.dp_ranks = topo.group_containing(topo.dp_groups(), rank)
        self.pp_ranks = topo.group_containing(topo.pp_groups(), rank)
        self.ep_ranks = topo.group_containing(topo.ep_groups(), rank)

    @classmethod
    def initialize(cls, tp: int = 1, pp: int = 1, ep: int = 1, sp: int = 1) -> "ParallelState":
        """Create every group (all ranks must call; new_group is collective)."""
        assert dist.is_initialized()
        topo = Topology(dist.get_world_size(), tp, pp, ep, sp)
        st = cls(topo, dist.get_rank())

        def make(groups, mine_attr):
            for g in groups:
                pg = dist.new_group(g) if len(g) > 1 else None
                if st.rank in g:
                    setattr(st, mine_attr, pg)

        make(topo.tp_groups(), "tp_group")
        make(topo.dp_groups(), "dp_group")
        make(topo.pp_groups(), "pp_group")
        if ep > 1:
            make(topo.ep_groups(), "ep_group")
            make(topo.expert_dp_groups(), "expert_dp_group")
        if sp > 1:
            make(topo.sp_groups(), "sp_group")
        make(topo.embedding_groups(), "embedding_group")
        return st

    # convenience
    @property
    def is_first_stage(self) -> bool:
        return self.pp_rank == 0

    @property
    def is_last_stage(self) -> bool:
        return self.pp_rank == self.topo.pp - 1

    def prev_stage_rank(self) -> int:
        return self.pp_ranks[self.pp_rank - 1]

    def next_stage_rank(self) -> int:
        return self.pp_ranks[self.pp_rank + 1]


def init_distributed(backend: str | None = None) -> tuple[int, int, int]:
    """torchrun-style bootstrap from RANK / LOCAL_RANK / WORLD_SIZE / MASTER_*."""
    import os

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if torch.cuda.is_available():
        torch.cuda.set_device(local_rank % torch.cuda.device_count())
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        kw = {}
        if backend == "nccl":
            kw["device_id"] = torch.device("cuda", torch.cuda.current_device())
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    return rank, local_rank, world
