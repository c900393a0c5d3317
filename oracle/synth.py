"""TEST INFRASTRUCTURE — deterministic synthetic weights and inputs shared by the golden
generator (which loads them INTO the unmodified reference modules) and by every parity test
(which loads the same values into the oracle and into the CUDA path).  Values depend only on
(key name, shape, seed) through a CPU torch.Generator, so no weight file has to be committed."""
import zlib

import torch


def _gen(name, seed):
    g = torch.Generator(device='cpu')
    g.manual_seed((zlib.crc32(name.encode()) + 7919 * seed) % (2 ** 31))
    return g


def synth_tensor(name, shape, seed=0):
    g = _gen(name, seed)
    shape = tuple(shape)
    if name.endswith(('weight', 'weight_orig')) and len(shape) >= 2:    # weight_orig: spectral norm (--dis_sn)
        fan_in = 1
        for s in shape[1:]:
            fan_in *= s
        if len(shape) == 4 and '.dec2.' in name:     # ConvTranspose2d weights are [Cin, Cout, k, k]
            fan_in = shape[0] * shape[2] * shape[3] // (4 if shape[2] == 3 else 1)
        # variance-preserving scale keeps activations O(1) through the deep stacks, which makes
        # the parity check sensitive; (the reference's N(0,0.02) init gives tiny, flat signals)
        return torch.randn(shape, generator=g) * (1.0 / fan_in) ** 0.5
    if name.endswith('weight'):                      # LayerNorm weight [C,1,1] etc.
        return 1.0 + 0.2 * torch.randn(shape, generator=g)
    return 0.1 * torch.randn(shape, generator=g)     # biases


def synth_params(shapes, seed=0, prefix=''):
    """shapes: {key: shape}.  Returns {key: fp32 tensor}."""
    return {k: synth_tensor(prefix + k, s, seed) for k, s in shapes.items()}


def synth_images(n, size, seed=0, name='img'):
    g = _gen(name, seed)
    return torch.rand((n, 3, size, size), generator=g) * 2 - 1


def synth_onehot(n, nd, seed=0, name='cls', offset=0):
    g = _gen(name, seed)
    idx = (torch.randint(0, nd, (n,), generator=g) + offset) % nd
    return torch.eye(nd)[idx]


class Draws:
    """Replayable sequence of N(0,1) draws, named by call order (the reference draws all of its
    noise from the CPU generator: misc.py:25, networks.py:132, adain_model.py:84)."""

    def __init__(self, seed=0):
        self.seed, self.i, self.log = seed, 0, []

    def __call__(self, shape):
        t = torch.randn(tuple(shape), generator=_gen(f'draw{self.i}', self.seed))
        self.i += 1
        self.log.append(tuple(shape))
        return t


def grad_sample(named_params, per_tensor=128):
    """Strided sample of the gradients of `named_params` ((key, parameter) pairs in state_dict order):
    up to `per_tensor` elements of every parameter tensor that has a gradient -> ({key: (offset, count)},
    flat fp32 tensor).  The reference-side recorder (oracle/make_golden.py) and the CUDA-side checks
    (tests/model_cases.py) use the same rule, so the committed samples compare element for element."""
    index, parts, off = {}, [], 0
    for k, p in named_params:
        g = getattr(p, 'grad', None)
        if g is None:
            continue
        g = g.detach().float().flatten().cpu()
        g = g[:: max(1, g.numel() // per_tensor)][:per_tensor]
        index[k] = (off, g.numel())
        parts.append(g)
        off += g.numel()
    return index, (torch.cat(parts) if parts else torch.zeros(0))
