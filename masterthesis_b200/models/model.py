"""Base class of the translation models: same public surface as the reference's
src/models/model.py:15-122 (dicts `.model/.optimizer/.scheduler/.loss`, initialize / update_lr /
save / load / write_loss / print_losses / save_images / get_current_lr), so that the reference's
train.py and sample.py drive this implementation unchanged.
"""
import os
from abc import ABC, abstractmethod

import torch
import torch.nn as nn

from ..utils import AttributeDict
from .core.functions import get_scheduler, init_net


class _NullWriter:
    """stand-in when tensorboardX is not installed (the reference hard-requires it, model.py:13)"""

    def __init__(self, *a, **k):
        pass

    def add_scalar(self, *a, **k):
        pass


def _summary_writer(logdir):
    try:
        from tensorboardX import SummaryWriter
        return SummaryWriter(log_dir=logdir)
    except Exception:
        return _NullWriter()


class Model(ABC, nn.Module):
    def __init__(self, args):
        super().__init__()
        self.args = args
        self.device = 'cuda' if torch.cuda.is_available() else 'cpu'
        # AttributeDict is deliberately NOT an nn.Module container: like in the reference
        # (model.py:24) model.train()/.eval()/.to() do not reach the nets.
        self.model = AttributeDict()
        self.optimizer = AttributeDict()
        self.scheduler = AttributeDict()
        self.loss = AttributeDict()
        if 'train' in args.mode:
            self.writer = _summary_writer(getattr(args, 'logdir', None))
        self.print_loss = []
        self.rng = None          # None: device Philox; callable(shape)->cpu tensor: replayed draws
        self.dp = None           # masterthesis_b200.parallel.DataParallelGroup or None
        # The whole training iteration (4 phases, ~2800 kernel launches) is captured once into a CUDA
        # graph and replayed: the host then costs one launch per iteration instead of ~47 ms of
        # enqueueing.  args.cuda_graph=False or MT_CUDA_GRAPH=0 keeps the eager path; replayed CPU
        # draws (self.rng, parity tests) always run eagerly.
        self.use_graph = bool(getattr(args, 'cuda_graph', True)) and os.environ.get('MT_CUDA_GRAPH', '1') != '0'
        self.graph_warmup = 2    # eager iterations before capture (lazy init, allocator warm-up)
        self._graphs = {}        # iteration kind ('main' / 'content') -> captured graph record
        self._eager_iters = {}
        self._scalars = {}
        from .core.blocks import HostContext
        self.ctx = HostContext()     # pack epoch / pack stream / dropout replay of THIS model (core/blocks.py)

    @abstractmethod
    def set_inputs(self, inputs):
        pass

    @abstractmethod
    def optimize_parameters(self, global_iter):
        pass

    # ------------------------------------------------------------------ iteration driver
    def reset_graph(self):
        """forget the captured iteration (call after changing loss weights, batch shape, ...)"""
        self._graphs = {}
        self._eager_iters = {}

    @property
    def _graph(self):
        """the captured main iteration (None while running eagerly)"""
        return self._graphs.get('main')

    def _static_inputs(self, x1, y1, x2, y2):
        """Keep the stacked batch in persistent device buffers (a captured iteration reads them in
        place; reference adain_model.py:87-94 builds `img`, `c_org` by torch.cat)."""
        dev, B = self.device, x1.shape[0]
        shape = (2 * B,) + tuple(x1.shape[1:])
        if getattr(self, 'img', None) is None or tuple(self.img.shape) != shape or \
                tuple(self.c_org.shape) != (2 * B, y1.shape[1]):
            self.img = torch.empty(shape, dtype=torch.float32, device=dev)
            self.c_org = torch.empty((2 * B, y1.shape[1]), dtype=torch.float32, device=dev)
            self.reset_graph()
        self.img[:B].copy_(x1, non_blocking=True)
        self.img[B:].copy_(x2, non_blocking=True)
        self.c_org[:B].copy_(y1, non_blocking=True)
        self.c_org[B:].copy_(y2, non_blocking=True)
        self.img_a, self.img_b = self.img[:B], self.img[B:]
        self.cls_a, self.cls_b = self.c_org[:B], self.c_org[B:]

    def _begin_iteration(self):
        self._scalars = {}          # loss name -> 0-dim device tensor (general torch formulation)
        self._raw_used = set()      # raw loss terms written by the loss kernels (AdaINModel.SLOTS)

    def _loss_vector(self):
        keys = list(self._scalars)
        return keys, torch.stack([self._scalars[k].float().reshape(()) for k in keys])

    def _named_losses(self, keys, values):
        """(name, python float) pairs of the iteration's losses from the vector read back from the device"""
        return zip(keys, values)

    def _run_iteration(self, body, kind='main'):
        """Run `body()` (enqueues one full training iteration and fills self._scalars with 0-dim
        device tensors) eagerly or as a CUDA-graph replay, then read all losses with ONE host sync."""
        from .. import ops
        from ..optim import GraphHyper
        from .core import blocks
        if not self.use_graph or self.rng is not None:
            self._begin_iteration()
            body()
            keys, vec = self._loss_vector()
        elif kind not in self._graphs and self._eager_iters.get(kind, 0) < self.graph_warmup:
            self._eager_iters[kind] = self._eager_iters.get(kind, 0) + 1
            self._begin_iteration()
            body()
            keys, vec = self._loss_vector()
        else:
            # a captured graph holds the raw addresses of the optimizers' flat buckets: if one of them was
            # re-allocated since (load() of an optimizer checkpoint, a network moved or re-built), drop the
            # graphs and capture again
            gen = tuple(getattr(opt, 'generation', 0) for opt in self.optimizer.values())
            if any(rec.get('gen') != gen for rec in self._graphs.values()):
                self._graphs = {}
            if kind not in self._graphs:
                for opt in self.optimizer.values():        # buckets exist (and `generation` is final) before capture
                    if hasattr(opt, '_ensure_flat'):
                        opt._ensure_flat()
                gen = tuple(getattr(opt, 'generation', 0) for opt in self.optimizer.values())
                blocks.invalidate_packs(self.ctx)
                ctx = GraphHyper(self.device)
                for opt in self.optimizer.values():
                    opt.graph_ctx = ctx
                graph = torch.cuda.CUDAGraph()
                l0 = ops.launches()
                torch.cuda.synchronize()
                try:
                    with torch.cuda.graph(graph):
                        self._begin_iteration()
                        body()
                        keys, vec = self._loss_vector()
                finally:
                    for opt in self.optimizer.values():
                        opt.graph_ctx = None
                self._graphs[kind] = dict(graph=graph, ctx=ctx, keys=keys, vec=vec, launches=ops.launches() - l0,
                                          gen=gen)
            else:
                self._graphs[kind]['ctx'].advance()
                ops._count(self._graphs[kind]['launches'])
            g = self._graphs[kind]
            g['ctx'].upload()
            g['graph'].replay()
            keys, vec = g['keys'], g['vec']
        for k, v in self._named_losses(keys, vec.tolist()):     # the one host sync of the iteration
            if k == 'loss_dc':                   # reference keeps it outside self.loss (adain_model.py:142)
                self.loss_dc = v
            else:
                self.loss[k] = v

    # ------------------------------------------------------------------ sampling as a graph
    def _run_sampling(self, kind, fn, *inputs):
        """fn(*inputs) -> image for the sampling entry points (forward_reference / forward_random).
        In test mode (fixed weights) the launch sequence is captured per (kind, input shapes) after
        two eager calls and replayed from static input buffers: the ~100 launches of a sampling call
        are host-bound when enqueued one by one.  Training-mode models, CPU-replayed draws
        (self.rng), grad-enabled calls and MT_CUDA_GRAPH=0 stay eager."""
        if (not self.use_graph or self.rng is not None or torch.is_grad_enabled()
                or 'train' in self.args.mode or not all(t.is_cuda for t in inputs)):
            return fn(*inputs)
        key = (kind,) + tuple((tuple(t.shape), t.dtype) for t in inputs)
        rec = self.__dict__.setdefault('_sample_graphs', {}).get(key)
        if rec is None:
            n = self._eager_iters.get(key, 0)
            if n < self.graph_warmup:
                self._eager_iters[key] = n + 1
                return fn(*inputs)
            static = [torch.empty_like(t).copy_(t) for t in inputs]
            graph = torch.cuda.CUDAGraph()
            torch.cuda.synchronize()
            with torch.cuda.graph(graph):
                out = fn(*static)
            rec = self._sample_graphs[key] = dict(graph=graph, static=static, out=out)
        for d, t in zip(rec['static'], inputs):
            d.copy_(t)
        rec['graph'].replay()
        return rec['out'].clone()

    def reset_sampling_graphs(self):
        """drop the captured sampling graphs (after load(): they hold the packed weights)"""
        self.__dict__.pop('_sample_graphs', None)

    # ------------------------------------------------------------------ lifecycle
    def initialize(self):
        if self.device != 'cuda':
            raise RuntimeError("masterthesis_b200 needs a CUDA device (sm_100a); there is no CPU path")
        init_type = None if self.args.resume else self.args.init_type
        for net in self.model:
            self.model[net] = init_net(self.model[net], init_type=init_type,
                                       init_gain=getattr(self.args, 'init_gain', 0.02),
                                       gpu_ids=self.args.gpu_ids, device=self.device)
        from .core import blocks
        blocks.bind_context(self.ctx, *self.model.values())
        if 'train' in self.args.mode:
            self.args.last_iter = -1 if self.args.resume_opt is None else self.args.last_iter
            self.init_scheduler()
            self.load(self.args.resume, self.args.resume_opt)
        else:
            self.load(self.args.resume)

    def init_scheduler(self):
        for opt in self.optimizer:
            if self.args.last_iter != -1:
                for g in self.optimizer[opt].param_groups:
                    g.setdefault('initial_lr', g['lr'])
            self.scheduler[opt] = get_scheduler(self.optimizer[opt], self.args, self.args.last_iter)

    def get_current_lr(self):
        return {opt: self.optimizer[opt].param_groups[0]['lr'] for opt in self.optimizer}

    def update_lr(self):
        for net in self.model:
            self.scheduler[net].step()

    # ------------------------------------------------------------------ checkpoints
    # Wire format of the reference (model.py:70-100): model_{it}.ckpt = {net: state_dict},
    # opt_{it}.ckpt = {net: optimizer.state_dict()}.  DataParallel-trained reference checkpoints
    # carry a 'module.' prefix (functions.py:101); it is accepted on load.
    def save(self, it):
        path = os.path.join(self.args.checkpoint_dir, f"model_{it}.ckpt")
        torch.save({net: self.model[net].state_dict() for net in self.model}, path)
        path = os.path.join(self.args.checkpoint_dir, f"opt_{it}.ckpt")
        torch.save({opt: self.optimizer[opt].state_dict() for opt in self.optimizer}, path)

    def load(self, checkpoint, opt_ckpt=None):
        # captured graphs hold packed weights and the addresses of the optimizers' buckets: start over
        self.reset_sampling_graphs()
        self.reset_graph()
        if checkpoint is not None:
            ckpt = torch.load(checkpoint, map_location='cpu')
            for net in ckpt:
                if net in self.model.keys():
                    print(f"Loading checkpoint for : {net}")
                    sd = {(k[7:] if k.startswith('module.') else k): v for k, v in ckpt[net].items()}
                    self.model[net].load_state_dict(sd)
                else:
                    print(f"Checkpoint for {net} network is not found.")
        if opt_ckpt is not None:
            ckpt = torch.load(opt_ckpt, map_location='cpu')
            for opt in ckpt:
                if opt in self.optimizer.keys():
                    print(f"Loading checkpoint for {opt} optimizer.")
                    self.optimizer[opt].load_state_dict(ckpt[opt])
                else:
                    print(f"Checkpoint for {opt} optimizer is not found.")

    # ------------------------------------------------------------------ logging
    def save_images(self, it):
        import torchvision
        visuals = self.compute_visuals()
        name = os.path.join(self.args.display_dir, f'gen_{it}.jpg')
        torchvision.utils.save_image(visuals / 2 + 0.5, name, nrow=1)

    def write_loss(self, global_iter):
        for k in self.loss:
            self.writer.add_scalar(k, self.loss[k], global_iter)

    def print_losses(self):
        return {k: self.loss[k] for k in self.loss if k in self.print_loss}

    def compute_metrics(self):
        pass
