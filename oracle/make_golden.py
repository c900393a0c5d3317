"""TEST INFRASTRUCTURE — generate tests/golden/*.npz by running the UNMODIFIED reference modules
(/root/reference/src, imported through oracle/ref_import.py) on synthetic weights and inputs.

    python -m oracle.make_golden            # only works where /root/reference exists

The RNG call sites of the reference (misc.py:25, networks.py:132, adain_model.py:84, and the
nn.Dropout(0.5) of --use_dropout: blocks.py:133-134,152,190) are
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

    def dropout_forward(self, x):
        # nn.Dropout(0.5) (blocks.py:133-134,152,190) with a replayable mask: keep = draw > 0, scale 1/(1-p)
        if not self.training:
            return x
        return x * (draws(tuple(x.size())) > 0).to(x.dtype) / (1.0 - self.p)

    torch.nn.Dropout.forward = dropout_forward
    misc.GaussianNoiseLayer.forward = noise_forward
    networks.ReparameterizedStyleEncoder.reparameterize = reparameterize
    adain_model.AdaINModel.get_z_random = get_z_random
    from models import base_model
    base_model.BaseModel.get_z_random = get_z_random


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


def nets_base_case(name, dim, size, N, nd=4, latent=8, seed=3, sub=1):
    """The networks only BaseModel / the optional flags use, each on its own: StyleEncoder, Decoder,
    DecoderConcat, ContentDiscriminator, MultiScaleDiscriminator (forward outputs, input gradients
    and a few parameter gradients of a fixed linear functional)."""
    from models.core import networks
    torch.manual_seed(0)
    C = dim * 4
    Es = networks.StyleEncoder(3, latent, dim, num_domains=nd, activation='lrelu')
    G1 = networks.Decoder(3, C, num_domains=nd, latent_dim=latent)
    G2 = networks.DecoderConcat(3, C, num_domains=nd, latent_dim=latent)
    Dc = networks.ContentDiscriminator(C, nd)
    Dm = networks.MultiScaleDiscriminator(3, dim, num_domains=nd)
    meta = {'dim': dim, 'size': size, 'N': N, 'nd': nd, 'latent': latent, 'seed': seed, 'sub': sub, 'shapes': {}}
    named = (('style_encoder', Es), ('decoder', G1), ('decoder_concat', G2), ('content_discriminator', Dc),
             ('ms_discriminator', Dm))
    for n, net in named:
        meta['shapes'][n] = load_synth(net, seed, n + '.')
    s = sub
    out = {}
    c = synth.synth_onehot(N, nd, seed)
    img = synth.synth_images(N, size, seed, 'img_se').requires_grad_(True)
    z = Es(img, c)
    (z * synth.synth_tensor('proj_se', z.shape, seed)).sum().backward()
    out['se.z'], out['se.dimg'] = z.detach(), img.grad[:, :, ::s, ::s]
    q = max(1, s // 4)
    for tag, G in (('dec', G1), ('decc', G2)):
        zc_in = synth.synth_tensor('zc_in', (N, C, size // 4, size // 4), seed) * (C * (size // 4) ** 2) ** 0.5
        zc_in = zc_in.to(torch.bfloat16).float().requires_grad_(True)
        zs_in = (synth.synth_tensor('zs_in', (N, latent), seed) * latent ** 0.5).requires_grad_(True)
        if tag == 'dec':
            # forward only: the reference's DecResnetBlock cannot be back-propagated (`out += residual`
            # modifies the output of a ReLU in place, blocks.py:206; torch raises in backward)
            with torch.no_grad():
                out['dec.fake'] = G(zc_in, zs_in, c)[:, :, ::s, ::s]
            continue
        fake = G(zc_in, zs_in, c)
        (fake * synth.synth_tensor('proj', fake.shape, seed)).sum().backward()
        out[f'{tag}.fake'] = fake.detach()[:, :, ::s, ::s]
        out[f'{tag}.dzc'], out[f'{tag}.dzs'] = zc_in.grad[:, :, ::q, ::q], zs_in.grad
    zc2 = synth.synth_tensor('zc_dc', (N, C, 64, 64), seed) * (C * 64 * 64) ** 0.5
    zc2 = zc2.to(torch.bfloat16).float().requires_grad_(True)
    pred = Dc(zc2)
    (pred * synth.synth_tensor('proj_dc', pred.shape, seed)).sum().backward()
    out['dc.pred'], out['dc.dzc'] = pred.detach(), zc2.grad[:, :, ::4, ::4]
    img4 = synth.synth_images(N, 256, seed, 'img_ms').requires_grad_(True)
    tot = 0
    for i, (dis, cls) in enumerate(Dm(img4)):
        out[f'ms.dis{i}'], out[f'ms.cls{i}'] = dis.detach(), cls.detach()
        tot = tot + (dis * synth.synth_tensor(f'proj_ms{i}', dis.shape, seed)).sum() + (cls * c).sum()
    tot.backward()
    out['ms.dimg'] = img4.grad[:, :, ::8, ::8]
    keys = {'style_encoder': ['model.0.block.1.weight', 'model.3.block.1.weight', 'model.6.weight'],
            'decoder': ['dec1.0.conv1.block.1.weight', 'dec1.1.block1.0.weight', 'dec1.3.block2.2.bias',
                        'dec2.0.block.0.weight', 'linear.4.weight'],
            'decoder_concat': ['dec_share.model.0.block.1.weight', 'dec1.1.model.1.block.1.weight',
                               'dec2.block.0.weight', 'dec2.block.1.weight', 'dec3.block.1.bias', 'dec4.block.0.weight'],
            'content_discriminator': ['model.0.block.1.weight', 'model.2.block.1.bias', 'model.3.block.1.weight',
                                      'model.4.weight'],
            'ms_discriminator': ['model.0.block.0.weight', 'model.4.block.0.weight', 'dis.weight', 'cls.bias']}
    for n, net in named:
        sd = dict(net.named_parameters())
        for k in keys[n]:
            g = sd[k].grad
            if g is None:
                continue
            out[f'grad.{n}.{k}'] = g if g.numel() <= 20000 else g.flatten()[:: max(1, g.numel() // 20000)]
    np.savez_compressed(os.path.join(OUT, name + '.npz'), meta=json.dumps(meta),
                        **{k: v.numpy().astype(np.float32) for k, v in out.items()})
    print(name, {k: float(v.abs().max()) for k, v in out.items() if not k.startswith('grad')})


def step_case(name, dim, size, B, iters=2, nd=4, seed=2, model='AdaINModel', gsamp=128, **flags):
    """Model.optimize_parameters — the losses per iteration, a checksum of every parameter tensor
    after `iters` iterations, and a strided sample (`gsamp` elements per parameter tensor,
    oracle.synth.grad_sample) of the gradients the reference hands to every optimizer.step() of the
    FIRST iteration (`gsample.<i>.<net>` in call order; index in meta['gsample']).  model / flags
    select BaseModel, --concat --reparam, --ms_dis, --use_dis_content (losses of a
    content-discriminator iteration: `loss_dc` only)."""
    models = ref_import.import_reference()
    draws = synth.Draws(seed)
    patch_rng(models, draws)
    args = ref_import.default_args(dim=dim, crop_size=size, batch_size=B, num_domains=nd, init_type=None,
                                   **flags)
    torch.manual_seed(0)
    m = getattr(models, model)(args)
    m.device = 'cpu'
    meta = {'dim': dim, 'size': size, 'B': B, 'nd': nd, 'seed': seed, 'iters': iters, 'shapes': {},
            'model': model, 'flags': flags}
    for n in m.model:
        meta['shapes'][n] = load_synth(m.model[n], seed, n + '.')
    batch = {'x1': synth.synth_images(B, size, seed, 'x1'), 'x2': synth.synth_images(B, size, seed, 'x2'),
             'y1': synth.synth_onehot(B, nd, seed, 'y'), 'y2': synth.synth_onehot(B, nd, seed, 'y', offset=1)}
    out = {}
    gs_index, recording = [], [True]
    for n in m.optimizer:
        def rec(orig=m.optimizer[n].step, n=n):
            if recording[0]:
                idx, flat = synth.grad_sample(m.model[n].named_parameters(), gsamp)
                out[f'gsample.{len(gs_index)}.{n}'] = flat
                gs_index.append([n, {k: list(v) for k, v in idx.items()}])
            return orig()
        m.optimizer[n].step = rec
    for it in range(iters):
        m.set_inputs(batch)
        m.optimize_parameters(it)
        recording[0] = False
        for k, v in m.loss.items():
            out[f'loss.{it}.{k}'] = torch.tensor(v)
        if getattr(m, 'loss_dc', None) is not None:
            out[f'loss.{it}.loss_dc'] = torch.tensor(m.loss_dc)
    for n in m.model:
        for k, v in m.model[n].state_dict().items():
            v = v.detach().float()
            out[f'param.{n}.{k}'] = torch.stack([v.sum(), v.abs().sum(), (v * v).sum()])
    meta['draws'] = [list(s) for s in draws.log]
    meta['gsample'], meta['gsamp'] = gs_index, gsamp
    np.savez_compressed(os.path.join(OUT, name + '.npz'), meta=json.dumps(meta),
                        **{k: v.numpy().astype(np.float64 if k.startswith('param') else np.float32)
                           for k, v in out.items()})
    print(name, {k: round(float(v), 5) for k, v in out.items() if k.startswith('loss.%d' % (iters - 1))})


if __name__ == '__main__':
    os.makedirs(OUT, exist_ok=True)
    ref_import.import_reference()
    torch.set_num_threads(os.cpu_count())
    which = sys.argv[1:] or ['adain', 'base', 'gan', 'opt']
    if 'adain' in which:
        nets_case('nets_dim16_64', dim=16, size=64, N=2)
        nets_case('nets_dim64_256', dim=64, size=256, N=1, sub=8)
        step_case('step_dim8_64', dim=8, size=64, B=1)
    if 'bench' in which:    # the benchmarked geometry (BASELINE config 3): dim 64, 256x256, batch 8
        step_case('step_dim64_256_B8', dim=64, size=256, B=8, gsamp=512)
    if 'gan' in which:      # adversarial-loss variants (loss.py:41-64, adain_model.py:206-215,277-302,359-376)
        step_case('step_ragan_dim8_64', dim=8, size=64, B=1, use_ragan=True)
        step_case('step_hinge_dim8_64', dim=8, size=64, B=1, gan_mode='hinge')
        step_case('step_lsgan_dim8_64', dim=8, size=64, B=1, gan_mode='lsgan')
        step_case('step_wgangp_dim8_64', dim=8, size=64, B=1, gan_mode='wgangp')
    if 'norms' in which or 'opt' in which:      # non-default --enc_norm / --dec_norm / --dis_norm
        step_case('step_norms_dim8_128', dim=8, size=128, B=1, enc_norm='layer', dec_norm='instance',
                  dis_norm='instance')
    if 'opt' in which:      # optional generator / discriminator variants
        step_case('step_nearest_dim8_64', dim=8, size=64, B=1, up_type='nearest')
        step_case('step_dropout_dim8_64', dim=8, size=64, B=1, use_dropout=True)
        step_case('step_sn_dim8_64', dim=8, size=64, B=1, dis_sn=True)
        step_case('step_basec_ndsn_dim8_64', dim=8, size=64, B=1, model='BaseModel', concat=True,
                  up_type='nearest', use_dropout=True, dis_sn=True)
    if 'base' in which:
        nets_base_case('nets_base_dim16_64', dim=16, size=64, N=2)
        step_case('step_basec_dim8_64', dim=8, size=64, B=1, model='BaseModel', concat=True)
        step_case('step_basecr_dim8_64', dim=8, size=64, B=1, model='BaseModel', concat=True, reparam=True)
        step_case('step_dcms_dim8_256', dim=8, size=256, B=1, iters=4, use_dis_content=True, ms_dis=True)
