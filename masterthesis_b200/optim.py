"""FlatAdam: torch.optim.Adam semantics (coupled L2, bias correction; the reference builds
torch.optim.Adam(lr, betas, weight_decay) per net, adain_model.py:58-61) over ONE flat fp32 bucket
per network, stepped by a single CUDA launch (csrc/ew_optim.cu).  Parameters and their .grad are
re-pointed to views of the flat buffers, so zero_grad is one memset and the data-parallel
all-reduce is one NCCL call on the gradient bucket.  state_dict() keeps torch.optim.Adam's layout
(per-parameter step / exp_avg / exp_avg_sq) so opt_{it}.ckpt files stay interchangeable
(reference model.py:77-82)."""
import math

import torch

from . import _lib, ops


class GraphHyper:
    """Device-resident Adam scalars for an iteration captured in a CUDA graph: one 8-float slot per
    FlatAdam.step() call site of the captured iteration, refreshed from the host (learning-rate
    schedule, step counts) by one small H2D copy before every replay."""

    def __init__(self, device, max_calls=32):
        self.host = torch.zeros(max_calls, 8, dtype=torch.float32).pin_memory()
        self.dev = torch.zeros(max_calls, 8, dtype=torch.float32, device=device)
        self.calls = []

    def slot(self, opt, values):
        i = len(self.calls)
        if i >= self.host.shape[0]:
            raise RuntimeError("GraphHyper: too many optimizer steps in one captured iteration")
        self.calls.append(opt)
        self.host[i, :7] = torch.tensor(values, dtype=torch.float32)
        return self.dev[i].data_ptr()

    def advance(self):
        """host bookkeeping of every captured step() for the next replay"""
        for i, opt in enumerate(self.calls):
            self.host[i, :7] = torch.tensor(opt._host_advance(), dtype=torch.float32)

    def upload(self):
        self.dev.copy_(self.host, non_blocking=True)


class FlatAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        params = list(params)
        defaults = dict(lr=lr, betas=(float(betas[0]), float(betas[1])), eps=eps,
                        weight_decay=weight_decay, amsgrad=False, maximize=False, foreach=None,
                        capturable=False, differentiable=False, fused=None, decoupled_weight_decay=False)
        super().__init__(params, defaults)
        self._flat = None
        self.generation = 0        # bumped whenever the flat buckets are (re)allocated: captured graphs that
                                   # hold the old addresses must be dropped (Model._run_iteration checks it)
        self.grad_scale = 1.0      # set to 1/world by the data-parallel wrapper
        self.graph_ctx = None      # GraphHyper while an iteration is being captured

    # ------------------------------------------------------------------ flat buckets
    def _params(self):
        return [p for g in self.param_groups for p in g['params']]

    def _ensure_flat(self):
        ps = self._params()
        dev = ps[0].device
        if self._flat is not None and self._flat['device'] == dev and \
                all(p.data_ptr() == a for p, a in zip(ps, self._flat['ptrs'])):
            return
        if dev.type != 'cuda':
            raise RuntimeError("FlatAdam has no CPU path: move the network to a CUDA device first")
        offs, n = [], 0
        for p in ps:
            offs.append(n)
            n += (p.numel() + 3) // 4 * 4           # keep every view 16-byte aligned
        flat_p = torch.zeros(n, dtype=torch.float32, device=dev)
        flat_g = torch.zeros(n, dtype=torch.float32, device=dev)
        flat_m = torch.zeros(n, dtype=torch.float32, device=dev)
        flat_v = torch.zeros(n, dtype=torch.float32, device=dev)
        for p, o in zip(ps, offs):
            k = p.numel()
            flat_p[o:o + k].copy_(p.data.reshape(-1))
            old_g = p.grad
            p.data = flat_p[o:o + k].view(p.shape)
            p.grad = flat_g[o:o + k].view(p.shape)
            if old_g is not None:
                p.grad.copy_(old_g)
            st = self.state[p]
            if 'exp_avg' in st:                      # state loaded from a checkpoint
                flat_m[o:o + k].copy_(st['exp_avg'].reshape(-1))
                flat_v[o:o + k].copy_(st['exp_avg_sq'].reshape(-1))
            st['exp_avg'] = flat_m[o:o + k].view(p.shape)
            st['exp_avg_sq'] = flat_v[o:o + k].view(p.shape)
            st.setdefault('step', torch.tensor(0.0))
        self._flat = dict(p=flat_p, g=flat_g, m=flat_m, v=flat_v, n=n, device=dev,
                          ptrs=[p.data_ptr() for p in ps])
        self.generation += 1

    @property
    def grad_bucket(self):
        self._ensure_flat()
        return self._flat['g']

    def zero_grad(self, set_to_none: bool = False):
        self._ensure_flat()
        ops.join_side()               # no side-stream weight gradient may still be adding into the bucket
        self._flat['g'].zero_()
        off = 0
        for p in self._params():      # re-attach views that were replaced (p.grad = None, ...)
            k = p.numel()
            if p.grad is None or p.grad.data_ptr() != self._flat['g'].data_ptr() + 4 * off:
                p.grad = self._flat['g'][off:off + k].view(p.shape)
            off += (k + 3) // 4 * 4

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        self._flat = None                            # rebuild views around the loaded state
        self._steps = None                           # re-read the step count from the loaded state

    @torch.no_grad()
    def step(self, closure=None):
        self._ensure_flat()
        ops.join_side()
        g0 = self.param_groups[0]
        ps = self._params()
        off = 0
        for p in ps:                                  # gradients that autograd re-allocated
            k = p.numel()
            want = self._flat['g'].data_ptr() + 4 * off
            if p.grad is not None and p.grad.data_ptr() != want:
                self._flat['g'][off:off + k].copy_(p.grad.reshape(-1))
                p.grad = self._flat['g'][off:off + k].view(p.shape)
            off += (k + 3) // 4 * 4
        f = self._flat
        hyper = self._host_advance()
        L = _lib.lib()
        if self.graph_ctx is not None:
            slot = self.graph_ctx.slot(self, hyper)
            _lib.check(L.mt_adam_flat_dev(f['p'].data_ptr(), f['g'].data_ptr(), f['m'].data_ptr(),
                                          f['v'].data_ptr(), f['n'], slot, ops._stream()), "mt_adam_flat_dev")
        else:
            # same kernel and the same host-computed scalars as the captured form, so that an eager
            # iteration and a graph replay are bit-identical (a pageable H2D copy: ordered on the stream
            # and staged before it returns)
            hyper_dev = torch.tensor(hyper + (0.0,), dtype=torch.float32).to(f['p'].device)
            _lib.check(L.mt_adam_flat_dev(f['p'].data_ptr(), f['g'].data_ptr(), f['m'].data_ptr(),
                                          f['v'].data_ptr(), f['n'], hyper_dev.data_ptr(), ops._stream()),
                       "mt_adam_flat_dev")
            self._hyper_keep = hyper_dev
        ops._count()
        return None

    def _step_count(self):
        """Steps taken so far.  Kept as one python int (all parameters of a bucket step together); the
        per-parameter `step` entries torch.optim.Adam keeps are materialised by state_dict()."""
        if getattr(self, '_steps', None) is None:
            st0 = self.state[self._params()[0]]
            v = st0.get('step', 0.0)
            self._steps = int(v.item() if torch.is_tensor(v) else v)
        return self._steps

    def state_dict(self):
        """torch.optim.Adam's layout: every parameter gets its OWN `step` tensor (a tensor shared between
        parameters would be incremented once per parameter by torch's foreach Adam after load_state_dict)."""
        self._ensure_flat() if self._params()[0].is_cuda else None
        n = float(self._step_count()) if self.state else 0.0
        for p in self._params():
            if p in self.state:
                self.state[p]['step'] = torch.tensor(n)
        return super().state_dict()

    def _host_advance(self):
        """Host half of one step: advance the step count and autograd's version counters (the kernel
        writes through raw pointers; the packed bf16 weight caches of blocks._ConvParams.packed key on
        the version).  Returns the 7 scalars of mt_adam_flat_dev's `hyper` block."""
        ps = self._params()
        g0 = self.param_groups[0]
        step = self._step_count() + 1
        self._steps = step
        torch._C._autograd._unsafe_set_version_counter(ps, [p._version + 1 for p in ps])
        b1, b2 = g0['betas']
        bc1 = 1.0 - b1 ** step
        bc2 = 1.0 - b2 ** step
        return (float(g0['lr']) / bc1, 1.0 / math.sqrt(bc2), b1, b2, g0['eps'], g0['weight_decay'],
                float(self.grad_scale))
