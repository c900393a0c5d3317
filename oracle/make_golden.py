"""TEST INFRASTRUCTURE — generate tests/golden/*.npz by running the UNMODIFIED reference modules
(/root/reference/src, imported through oracle/ref_import.py) on synthetic weights and inputs.

    python -m oracle.make_golden            # only works where /root/reference exists

The three RNG call sites of the reference (misc.py:25, networks.py:132, adain_model.py:84) are
redirected to oracle.synth.Draws so that the noise is reproducible from a seed instead of being
stored; the arithmetic is the reference's own.  Weights are synthesised per state_dict key
(oracle/synth.py) and loaded with load_state_dict, so no weight file is committed either.
"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_import, synth  # noqa: E402

OUT = os.path.join(ROOT, 'tests', 'golden')


def shapes_of(net):
    return {k: list(v.shape) for k, v in net.state_dict().items()}


def load_synth(net, seed, prefix):
    sh = shapes_of(net)
    net.load_state_dict(synth.synth_params(sh, seed, prefix))
    return sh


def patch_rng(models, draws):
    from models.core import misc, networks
    from models import adain_model

    def noise_forward(self, x):
        return x + draws(tuple(x.size())).to(x.device) if self.training else x

    def reparameterize(self, mu, logvar):
        std = logvar.mul(0.5).exp_()
        eps = draws(tuple(std.size())).to(mu.device)
        return eps.mul(std).add_(mu)

    def get_z_random(self, bs, latent_dim):
        return draws((bs, latent_dim)).to(self.device)

    misc.GaussianNoiseLayer.forward = noise_forward
    networks.ReparameterizedStyleEncoder.reparameterize = reparameterize
    adain_model.AdaINModel.get_z_random = get_z_random


def nets_case(name, dim, size, N, nd=4, latent=8, seed=1, sub=1):
    """Each AdaINModel network on its own: forward outputs and the gradients of a fixed linear
    functional of them (independent per net, so that a check of one net does not inherit the
    rounding noise of the others)."""
    from models.core import networks
    torch.manual_seed(0)
    draws = synth.Draws(seed)
    patch_rng(None, draws)
    Ec = networks.ContentEncoder(3, dim)
    Es = networks.ReparameterizedStyleEncoder(3, latent, dim, num_domains=nd, norm_layer=None, activation='lrelu')
    G = networks.AdaINDecoder(3, dim * 4, num_domains=nd, latent_dim=latent)
    D = networks.Discriminator(3, dim, num_domains=nd, image_size=size)
    meta = {'dim': dim, 'size': size, 'N': N, 'nd': nd, 'latent': latent, 'seed': seed, 'sub': sub,
            'shapes': {}}
    for n, net in (('content_encoder', Ec), ('style_encoder', Es), ('decoder', G), ('discriminator', D)):
        meta['shapes'][n] = load_synth(net, seed, n + '.')
    s = sub
    out = {}
    c = synth.synth_onehot(N, nd, seed)
    # content encoder (training mode: draw 0 = noise)
    img = synth.synth_images(N, size, seed).requires_grad_(True)
    z_c = Ec(img)
    (z_c * synth.synth_tensor('proj_zc', z_c.shape, seed)).sum().backward()
    out['z_c'], out['Ec.dimg'] = z_c.detach()[:, :, ::s, ::s], img.grad[:, :, ::s, ::s]
    # style encoder (draw 1 = eps)
    img2 = synth.synth_images(N, size, seed, 'img2').requires_grad_(True)
    z, mu, logvar = Es(img2, c)
    (z.sum() + 2 * mu.sum() + 0.5 * logvar.sum()).backward()
    out['z'], out['mu'], out['logvar'] = z.detach(), mu.detach(), logvar.detach()
    out['Es.dimg'] = img2.grad[:, :, ::s, ::s]
    # decoder on a synthetic unit-variance content code
    zc_in = synth.synth_tensor('zc_in', (N, dim * 4, size // 4, size // 4), seed) * (dim * 4 * (size // 4) ** 2) ** 0.5
    zc_in = zc_in.to(torch.bfloat16).float().requires_grad_(True)
    zs_in = synth.synth_tensor('zs_in', (N, latent), seed) * latent ** 0.5
    zs_in.requires_grad_(True)
    fake = G(zc_in, zs_in, c)
    (fake * synth.synth_tensor('proj', fake.shape, seed)).sum().backward()
    out['fake'] = fake.detach()[:, :, ::s, ::s]
    out['G.dzc'], out['G.dzs'] = zc_in.grad[:, :, ::max(1, s // 4), ::max(1, s // 4)], zs_in.grad
    # discriminator
    img3 = synth.synth_images(N, size, seed, 'img3').requires_grad_(True)
    d_out, d_cls = D(img3)
    (d_out * synth.synth_tensor('proj_d', d_out.shape, seed)).sum().add((d_cls * c).sum()).backward()
    out['d_out'], out['d_cls'], out['D.dimg'] = d_out.detach(), d_cls.detach(), img3.grad[:, :, ::s, ::s]
    for n, net, keys in (('content_encoder', Ec, ['model.0.block.1.weight', 'model.2.block.1.weight',
                                                 'model.6.model.1.block.1.weight']),
                         ('style_encoder', Es, ['model.0.block.1.weight', 'model.2.conv.1.block.1.weight',
                                                'model.3.shortcut.1.weight', 'fc.weight']),
                         ('decoder', G, ['dec1.0.conv1.block.1.weight', 'dec1.3.norm.fc.weight',
                                         'dec2.0.block.0.weight', 'dec2.1.block.1.weight', 'dec2.2.block.0.weight',
                                         'linear.0.weight']),
                         ('discriminator', D, ['model.0.block.1.weight', 'model.3.block.1.weight',
                                               'model.5.block.1.bias', 'conv1.weight', 'conv2.weight'])):
        sd = dict(net.named_parameters())
        for k in keys:
            g = sd[k].grad
            out[f'grad.{n}.{k}'] = g if g.numel() <= 20000 else g.flatten()[:: max(1, g.numel() // 20000)]
    np.savez_compressed(os.path.join(OUT, name + '.npz'), meta=json.dumps(meta),
                        **{k: v.numpy().astype(np.float32) for k, v in out.items()})
    print(name, 'draws', draws.log, {k: float(v.abs().max()) for k, v in out.items() if not k.startswith('grad')})


def step_case(name, dim, size, B, iters=2, nd=4, seed=2):
    """AdaINModel.optimize_parameters (default flags) — the 13 losses per iteration and a
    checksum of every parameter tensor after `iters` iterations."""
    models = ref_import.import_reference()
    draws = synth.Draws(seed)
    patch_rng(models, draws)
    args = ref_import.default_args(dim=dim, crop_size=size, batch_size=B, num_domains=nd, init_type=None)
    torch.manual_seed(0)
    m = models.AdaINModel(args)
    m.device = 'cpu'
    meta = {'dim': dim, 'size': size, 'B': B, 'nd': nd, 'seed': seed, 'iters': iters, 'shapes': {}}
    for n in m.model:
        meta['shapes'][n] = load_synth(m.model[n], seed, n + '.')
    batch = {'x1': synth.synth_images(B, size, seed, 'x1'), 'x2': synth.synth_images(B, size, seed, 'x2'),
             'y1': synth.synth_onehot(B, nd, seed, 'y'), 'y2': synth.synth_onehot(B, nd, seed, 'y', offset=1)}
    out = {}
    for it in range(iters):
        m.set_inputs(batch)
        m.optimize_parameters(it)
        for k, v in m.loss.items():
            out[f'loss.{it}.{k}'] = torch.tensor(v)
    for n in m.model:
        for k, v in m.model[n].state_dict().items():
            v = v.detach().float()
            out[f'param.{n}.{k}'] = torch.stack([v.sum(), v.abs().sum(), (v * v).sum()])
    meta['draws'] = [list(s) for s in draws.log]
    np.savez_compressed(os.path.join(OUT, name + '.npz'), meta=json.dumps(meta),
                        **{k: v.numpy().astype(np.float64 if k.startswith('param') else np.float32)
                           for k, v in out.items()})
    print(name, {k: round(float(v), 5) for k, v in out.items() if k.startswith('loss.%d' % (iters - 1))})


if __name__ == '__main__':
    os.makedirs(OUT, exist_ok=True)
    ref_import.import_reference()
    torch.set_num_threads(os.cpu_count())
    nets_case('nets_dim16_64', dim=16, size=64, N=2)
    nets_case('nets_dim64_256', dim=64, size=256, N=1, sub=8)
    step_case('step_dim8_64', dim=8, size=64, B=1)
