"""One-process-per-GPU data parallelism (replaces the reference's nn.DataParallel,
src/models/core/functions.py:96-106): every rank holds all networks, processes its share of the
image PAIRS (pairs must stay together, adain_model.py:113-129), and the flat gradient bucket of each
optimizer is summed with one NCCL all-reduce over NVLink per optimizer phase; the 1/world average is
folded into the fused Adam launch (FlatAdam.grad_scale)."""
import os

import torch
import torch.distributed as dist


class DataParallelGroup:
    def __init__(self, backend=None, device=None):
        if not dist.is_initialized():
            backend = backend or ('nccl' if torch.cuda.is_available() else 'gloo')
            dist.init_process_group(backend=backend)
        self.world = dist.get_world_size()
        self.rank = dist.get_rank()
        self.device = device

    @staticmethod
    def from_env():
        """torchrun environment -> group (None when WORLD_SIZE is 1 or unset)."""
        if int(os.environ.get('WORLD_SIZE', '1')) <= 1:
            return None
        if torch.cuda.is_available():
            torch.cuda.set_device(int(os.environ.get('LOCAL_RANK', '0')))
        return DataParallelGroup()

    def allreduce(self, optimizers):
        """sum the gradient buckets over ranks; Adam then scales by 1/world."""
        if torch.cuda.is_available():
            from . import ops
            ops.join_side()
        for opt in optimizers:
            dist.all_reduce(opt.grad_bucket, op=dist.ReduceOp.SUM)
            opt.grad_scale = 1.0 / self.world

    def broadcast_parameters(self, nets):
        """make every rank start from rank 0's weights (DataParallel replicates module 0's)."""
        with torch.no_grad():
            for net in nets:
                for t in list(net.parameters()) + list(net.buffers()):
                    dist.broadcast(t.data, src=0)

    def parameter_spread(self, nets):
        """Data-parallel self check: every rank applies the same all-reduced gradients to the same weights,
        so the parameters must stay identical.  -> {'max_rel_spread': max over networks of
        (max - min over ranks of sum|p|) / mean, 'identical': spread == 0}."""
        dev = 'cuda' if dist.get_backend() == 'nccl' else 'cpu'
        sums = torch.stack([torch.cat([p.detach().double().abs().sum().reshape(1) for p in net.parameters()]).sum()
                            for net in nets]).to(dev)
        hi, lo = sums.clone(), sums.clone()
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        spread = float(((hi - lo) / hi.clamp_min(1e-30)).max().item())
        return {'max_rel_spread_of_param_checksums': spread, 'identical': spread == 0.0, 'ranks': self.world}

    def max_over_ranks(self, value: float) -> float:
        t = torch.tensor([value], dtype=torch.float64,
                         device='cuda' if dist.get_backend() == 'nccl' else 'cpu')
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def barrier(self):
        dist.barrier()
