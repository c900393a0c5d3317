"""TEST INFRASTRUCTURE — data-parallel gradient equality on real GPUs:

    torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tests/tools/dp_equality.py [dim size Bg]

Every rank runs ONE iteration of AdaINModel.optimize_parameters on its slice of a global batch of Bg image
pairs (slices of the same global random draws: the Ec noise, the reparameterisation eps and the random style
codes are per-sample), with the NCCL all-reduce of the gradient buckets; rank 0 then runs the same iteration
on the whole batch on one GPU.  At every optimizer step the averaged gradient bucket of the N-rank run must
equal the single-GPU bucket (cosine >= 0.9999: the per-sample arithmetic is identical, only fp32 summation
order differs), which covers the mean-reduced losses, the SUM-reduced KL term (pre-scaled by the world size,
adain_model.py:313-314) and the 1/world folded into the fused Adam.  Prints one RESULT json line (rank 0).
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from oracle import synth  # noqa: E402


def main(dim=16, size=64, Bg=4):
    from masterthesis_b200 import models
    from masterthesis_b200.arguments import default_args
    from masterthesis_b200.parallel import DataParallelGroup
    dp = DataParallelGroup.from_env()
    world, rank = (dp.world, dp.rank) if dp is not None else (1, 0)
    assert Bg % world == 0
    B = Bg // world
    nd, seed = 4, 11
    x1, x2 = synth.synth_images(Bg, size, seed, 'x1'), synth.synth_images(Bg, size, seed, 'x2')
    y1, y2 = synth.synth_onehot(Bg, nd, seed, 'y'), synth.synth_onehot(Bg, nd, seed, 'y', offset=1)

    def run(B_local, r, group):
        torch.manual_seed(5)
        m = models.AdaINModel(default_args(dim=dim, crop_size=size, batch_size=B_local, num_domains=nd, init_type='normal'))
        m.initialize()
        m.dp = group
        if group is not None:
            group.broadcast_parameters([m.model[n] for n in m.model])
        src = synth.Draws(seed)
        lo, hi = r * B_local, (r + 1) * B_local

        def sliced(shape):
            """this rank's rows of the GLOBAL draw: [2B, ...] tensors are (A-half | B-half) stacked"""
            n0 = shape[0]
            if n0 == 2 * B_local:
                g = src((2 * Bg,) + tuple(shape[1:]))
                return torch.cat((g[lo:hi], g[Bg + lo:Bg + hi]))
            assert n0 == B_local, shape
            return src((Bg,) + tuple(shape[1:]))[lo:hi]
        m.rng = sliced
        recs = []
        for n in m.optimizer:
            opt = m.optimizer[n]

            def rec(orig=opt.step, n=n, opt=opt):
                recs.append((n, (opt.grad_bucket * opt.grad_scale).detach().clone()))
                return orig()
            opt.step = rec
        m.set_inputs({'x1': x1[lo:hi], 'x2': x2[lo:hi], 'y1': y1[lo:hi], 'y2': y2[lo:hi]})
        m.optimize_parameters(0)
        torch.cuda.synchronize()
        return recs, dict(m.loss)

    multi, loss_multi = run(B, rank, dp)
    out = {}
    if dp is not None:
        # losses are per-rank means / sums: average (sum for kl_zs) over ranks for the comparison
        keys = sorted(loss_multi)
        t = torch.tensor([loss_multi[k] for k in keys], dtype=torch.float64, device='cuda')
        dist.all_reduce(t)
        loss_multi = {k: (v if k == 'kl_zs' else v / world) for k, v in zip(keys, t.tolist())}
        dp.barrier()
    if rank == 0:
        single, loss_single = run(Bg, 0, None)
        for i, ((n1, a), (n2, b)) in enumerate(zip(multi, single)):
            assert n1 == n2
            a, b = a.double(), b.double()
            out[f'cos_step{i}_{n1}'] = float(torch.dot(a, b) / (a.norm() * b.norm()).clamp_min(1e-300))
            out[f'relnorm_step{i}_{n1}'] = float((a - b).norm() / b.norm().clamp_min(1e-300))
        for k in loss_single:
            if k in ('total_g',):       # contains the summed KL term: compare the parts
                continue
            out['loss.' + k] = abs(loss_multi[k] - loss_single[k]) / max(abs(loss_single[k]), 1e-6)
        out['world'] = world
        print('RESULT ' + json.dumps(out))
    if dp is not None:
        dp.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main(*[int(v) for v in sys.argv[1:4]])
