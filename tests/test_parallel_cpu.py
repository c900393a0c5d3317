"""world_size-2 gloo test of the data-parallel host logic (bucket all-reduce + 1/world folding)."""
import os
import sys

import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist
    from masterthesis_b200.parallel import DataParallelGroup
    dist.init_process_group('gloo', rank=rank, world_size=world)
    dp = DataParallelGroup()

    class Opt:                       # the interface allreduce() relies on (FlatAdam provides it)
        grad_bucket = torch.full((5,), float(rank + 1))
        grad_scale = 1.0
    o = Opt()
    dp.allreduce([o])
    w = torch.nn.Linear(2, 2)
    with torch.no_grad():
        w.weight.fill_(float(rank))
    dp.broadcast_parameters([w])
    mx = dp.max_over_ranks(float(rank * 10))
    out.put((rank, o.grad_bucket.tolist(), o.grad_scale, w.weight.flatten().tolist(), mx))
    dist.destroy_process_group()


def test_bucket_allreduce_world2():
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = [q.get(timeout=120) for _ in ps]
    for p in ps:
        p.join(60)
    for rank, bucket, scale, w, mx in res:
        assert bucket == [3.0] * 5            # 1 + 2 summed over ranks
        assert scale == 0.5                   # averaged inside the Adam launch
        assert w == [0.0] * 4                 # rank 0's weights everywhere
        assert mx == 10.0
