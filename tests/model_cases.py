"""Network- and step-level parity cases: CUDA path vs (a) the reference's own outputs committed in
tests/golden/ and (b) the fp32 oracle run on the same seeded inputs.  GPU only.

Metrics: outputs/losses max|a-b| / max|b|; gradients cosine similarity.  The tolerances are stated
in tests/test_model_gpu.py.  Keys starting with `emu` are the bf16-storage-emulated oracle vs the
fp32 oracle (tests/emu.py): the precision any bf16-storage implementation can reach on that input.
"""
import contextlib
import json
import os

import numpy as np
import torch

from oracle import nets as onets
from oracle import ref_import, step as ostep, synth

GOLD = os.path.join(os.path.dirname(__file__), 'golden')
DEV = 'cuda'


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-12)).item()


def cos(a, b):
    a, b = a.double().cpu().flatten(), b.double().cpu().flatten()
    return (torch.dot(a, b) / (a.norm() * b.norm()).clamp_min(1e-30)).item()


def loss_err(mine, ref):
    """relative error of one loss scalar: |a - b| / |b| (max relative error <= 1e-2 on each
    of the loss scalars); the floor only guards a reference value that is exactly zero"""
    err = abs(mine - ref) / max(abs(ref), 1e-6)
    return err if err == err else float('inf')


def grad_sample_vs_golden(meta, gold, i, net, named_params):
    """Our gradients at optimizer step `i` of the first iteration against the strided sample the
    REFERENCE recorded at the same step (oracle/make_golden.py: gsample.<i>.<net>): cosine over the whole
    sample, and the worst per-tensor cosine over the tensors that carry signal (rms >= 1e-3 of the
    largest tensor's: a conv bias in front of InstanceNorm has an exactly-zero gradient, the reference
    records rounding noise for it)."""
    n_ref, index = meta['gsample'][i]
    assert n_ref == net, (i, n_ref, net)
    ref = gold[f'gsample.{i}.{net}'].double()
    idx, mine = synth.grad_sample(named_params, meta['gsamp'])
    mine = mine.double()
    a, b, per = [], [], {}
    for k, (off, cnt) in index.items():
        if k not in idx:
            continue
        o2, c2 = idx[k]
        assert c2 == cnt, (k, c2, cnt)
        a.append(mine[o2:o2 + c2]); b.append(ref[off:off + cnt])
        per[k] = (a[-1], b[-1])
    rms = {k: float(v[1].pow(2).mean().sqrt()) for k, v in per.items()}
    top = max(rms.values())
    sig = [k for k in per if rms[k] >= 1e-3 * top]
    worst = min(sig, key=lambda k: cos(*per[k]))
    return cos(torch.cat(a), torch.cat(b)), cos(*per[worst]), worst


def load_gold(name):
    z = np.load(os.path.join(GOLD, name + '.npz'))
    return json.loads(str(z['meta'])), {k: torch.from_numpy(z[k]) for k in z.files if k != 'meta'}


def build_nets(meta, size):
    from masterthesis_b200.models.core import networks
    dim, nd, latent, seed = meta['dim'], meta['nd'], meta['latent'], meta['seed']
    ns = {
        'content_encoder': networks.ContentEncoder(3, dim),
        'style_encoder': networks.ReparameterizedStyleEncoder(3, latent, dim, num_domains=nd,
                                                              norm_layer=None, activation='lrelu'),
        'decoder': networks.AdaINDecoder(3, dim * 4, num_domains=nd, latent_dim=latent),
        'discriminator': networks.Discriminator(3, dim, num_domains=nd, image_size=size),
    }
    for n, net in ns.items():
        net.load_state_dict(synth.synth_params(meta['shapes'][n], seed, n + '.'))
        net.to(DEV)
    return ns


def nets_vs_golden(name):
    """Same computations as oracle/make_golden.py:nets_case, on the CUDA path, vs the REFERENCE's
    recorded outputs and gradients (each network on its own)."""
    from masterthesis_b200.models.core.blocks import from_padded, to_padded
    meta, gold = load_gold(name)
    seed, N, nd, size, s, dim, latent = (meta[k] for k in ('seed', 'N', 'nd', 'size', 'sub', 'dim', 'latent'))
    ns = build_nets(meta, size)
    draws = synth.Draws(seed)
    c = synth.synth_onehot(N, nd, seed).to(DEV)
    r = {}
    img = synth.synth_images(N, size, seed).to(DEV).requires_grad_(True)
    z_c = from_padded(ns['content_encoder'].run(img, draws), dim * 4, 1)
    (z_c * synth.synth_tensor('proj_zc', z_c.shape, seed).to(DEV)).sum().backward()
    r['z_c'] = rel(z_c.detach()[:, :, ::s, ::s], gold['z_c'])
    r['cos_Ec.dimg'] = cos(img.grad[:, :, ::s, ::s], gold['Ec.dimg'])
    img2 = synth.synth_images(N, size, seed, 'img2').to(DEV).requires_grad_(True)
    z, mu, logvar = ns['style_encoder'].run(img2, c, draws)
    (z.sum() + 2 * mu.sum() + 0.5 * logvar.sum()).backward()
    r['z'], r['mu'], r['logvar'] = rel(z.detach(), gold['z']), rel(mu.detach(), gold['mu']), rel(logvar.detach(), gold['logvar'])
    r['cos_Es.dimg'] = cos(img2.grad[:, :, ::s, ::s], gold['Es.dimg'])
    zc_in = synth.synth_tensor('zc_in', (N, dim * 4, size // 4, size // 4), seed) * (dim * 4 * (size // 4) ** 2) ** 0.5
    zc_in = zc_in.to(torch.bfloat16).float().to(DEV).requires_grad_(True)
    zs_in = (synth.synth_tensor('zs_in', (N, latent), seed) * latent ** 0.5).to(DEV).requires_grad_(True)
    fake = ns['decoder'].run(to_padded(zc_in, 1), zs_in, c)
    (fake * synth.synth_tensor('proj', fake.shape, seed).to(DEV)).sum().backward()
    r['fake'] = rel(fake.detach()[:, :, ::s, ::s], gold['fake'])
    q = max(1, s // 4)
    r['cos_G.dzc'], r['cos_G.dzs'] = cos(zc_in.grad[:, :, ::q, ::q], gold['G.dzc']), cos(zs_in.grad, gold['G.dzs'])
    img3 = synth.synth_images(N, size, seed, 'img3').to(DEV).requires_grad_(True)
    d_out, d_cls = ns['discriminator'].run(img3)
    (d_out * synth.synth_tensor('proj_d', d_out.shape, seed).to(DEV)).sum().add((d_cls * c).sum()).backward()
    r['d_out'], r['d_cls'] = rel(d_out.detach(), gold['d_out']), rel(d_cls.detach(), gold['d_cls'])
    r['cos_D.dimg'] = cos(img3.grad[:, :, ::s, ::s], gold['D.dimg'])
    for k, g in gold.items():
        if not k.startswith('grad.'):
            continue
        _, n, key = k.split('.', 2)
        mine = dict(ns[n].named_parameters())[key].grad
        if mine.numel() > 20000:
            mine = mine.flatten()[:: max(1, mine.numel() // 20000)]
        r['cos_' + k[5:]] = cos(mine, g.flatten())
    return r


def make_model(dim, size, B, nd=4, seed=2, shapes=None, model='AdaINModel', **flags):
    from masterthesis_b200 import models
    args = ref_import.default_args(dim=dim, crop_size=size, batch_size=B, num_domains=nd, init_type=None,
                                   **flags)
    m = getattr(models, model)(args)
    if shapes is None:
        shapes = {n: {k: list(v.shape) for k, v in m.model[n].state_dict().items()} for n in m.model}
    params = {n: synth.synth_params(shapes[n], seed, n + '.') for n in m.model}
    for n in m.model:
        m.model[n].load_state_dict(params[n])
    m.initialize()
    return m, params


def batch(B, size, nd, seed):
    return {'x1': synth.synth_images(B, size, seed, 'x1'), 'x2': synth.synth_images(B, size, seed, 'x2'),
            'y1': synth.synth_onehot(B, nd, seed, 'y'), 'y2': synth.synth_onehot(B, nd, seed, 'y', offset=1)}


def step_vs_golden(name='step_dim8_64'):
    """AdaINModel.optimize_parameters on the CUDA path vs the REFERENCE's recorded losses, and its
    gradients at every optimizer step vs the oracle's."""
    meta, gold = load_gold(name)
    seed, B, nd, size, dim = meta['seed'], meta['B'], meta['nd'], meta['size'], meta['dim']
    m, params = make_model(dim, size, B, nd, seed, meta['shapes'])
    m.rng = synth.Draws(seed)
    # oracle twin on the GPU in fp32 (TF32 off) for the gradient comparison
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    st = ostep.AdaINStep({n: {k: v.to(DEV) for k, v in p.items()} for n, p in params.items()}, B)
    st.Ec = lambda x, draw: onets.content_encoder(
        st.P['content_encoder'], x, draw((x.size(0), dim * 4, x.size(2) // 4, x.size(3) // 4)).to(DEV))
    odraw_src = synth.Draws(seed)
    odraw = lambda shape: odraw_src(shape).to(DEV)  # noqa: E731

    mine_g, ref_g, vs_ref = [], [], []
    for n in m.optimizer:
        opt = m.optimizer[n]
        orig = opt.step

        def rec(orig=orig, n=n, opt=opt):
            from masterthesis_b200 import ops
            ops.join_side(all_streams=True)
            i = len(mine_g)
            mine_g.append((n, {k: p.grad.detach().clone() for k, p in m.model[n].named_parameters()
                               if p.grad is not None}))
            if 'gsample' in meta and i < len(meta['gsample']):
                vs_ref.append((i, n) + grad_sample_vs_golden(meta, gold, i, n, m.model[n].named_parameters()))
            return orig()
        opt.step = rec
    for n in st.opt:
        o = st.opt[n]
        orig = o.step

        def rec2(orig=orig, n=n, o=o):
            ref_g.append((n, {k: p.grad.detach().clone() for k, p in o.params.items() if p.grad is not None}))
            return orig()
        o.step = rec2

    # third twin: the oracle with bf16 storage emulated (tests/emu.py) — the precision yardstick
    from tests import emu
    st_e = ostep.AdaINStep({n: {k: v.to(DEV) for k, v in p.items()} for n, p in params.items()}, B)
    st_e.Ec = lambda x, draw: onets.content_encoder(
        st_e.P['content_encoder'], x, draw((x.size(0), dim * 4, x.size(2) // 4, x.size(3) // 4)).to(DEV))
    edraw_src = synth.Draws(seed)
    edraw = lambda shape: edraw_src(shape).to(DEV)  # noqa: E731
    emu_g = []
    for n in st_e.opt:
        o = st_e.opt[n]
        orig = o.step

        def rec3(orig=orig, n=n, o=o):
            emu_g.append((n, {k: p.grad.detach().clone() for k, p in o.params.items() if p.grad is not None}))
            return orig()
        o.step = rec3

    b = batch(B, size, nd, seed)
    r = {}
    for it in range(meta['iters']):
        m.set_inputs(b)
        m.optimize_parameters(it)
        st.optimize_parameters(m.img.clone(), m.c_org.clone(), odraw)
        if it == 0:
            with emu.bf16_storage():
                st_e.optimize_parameters(m.img.clone(), m.c_org.clone(), edraw)
        for k in m.loss:
            r[f'loss_it{it}.{k}'] = loss_err(m.loss[k], float(gold[f'loss.{it}.{k}']))
    assert [list(s) for s in m.rng.log] == meta['draws'], "RNG call order differs from the reference"
    # gradients handed to the optimizers in the first iteration vs the REFERENCE's recorded sample
    for i, n, c_all, c_worst, worst in vs_ref:
        r[f'cosref_step{i}_{n}'] = c_all
        r[f'cosrefmin_step{i}_{n}'] = c_worst
    r['_worst_tensors'] = {f'step{i}_{n}': w for i, n, _, _, w in vs_ref}
    # gradient cosine per (step point, net): first iteration only (identical weights there)
    for i, ((n1, g1), (n2, g2)) in enumerate(zip(mine_g[:7], ref_g[:7])):
        assert n1 == n2
        a = torch.cat([g1[k].flatten() for k in g2 if k in g1])
        bb = torch.cat([g2[k].flatten() for k in g2 if k in g1])
        r[f'cos_step{i}_{n1}'] = cos(a, bb)
        ee = torch.cat([emu_g[i][1][k].flatten() for k in g2 if k in g1])
        r[f'emucos_step{i}_{n1}'] = cos(ee, bb)       # bf16-storage oracle vs fp32 oracle
        if not (torch.isfinite(a).all() and torch.isfinite(bb).all()):
            r[f'nonfinite_step{i}_{n1}'] = float((~torch.isfinite(a)).sum() * 1000 + (~torch.isfinite(bb)).sum())
        # discriminator phases in the G steps skip D/Es grads by design; every key of ours must exist in ref
    return r


def step_losses_vs_golden(name='step_dim64_256_B8', iters=1):
    """Light version of step_vs_golden (no oracle twins): `iters` iterations of optimize_parameters on the
    CUDA path with the replayed draws vs the REFERENCE's recorded losses, plus the gradients of the first
    iteration's optimizer steps vs the reference's recorded samples.  Used by __graft_entry__.smoke()."""
    meta, gold = load_gold(name)
    seed, B, nd, size, dim = meta['seed'], meta['B'], meta['nd'], meta['size'], meta['dim']
    m, _ = make_model(dim, size, B, nd, seed, meta['shapes'], model=meta['model'], **meta['flags'])
    m.rng = synth.Draws(seed)
    vs_ref = []
    for n in m.optimizer:
        opt = m.optimizer[n]

        def rec(orig=opt.step, n=n):
            from masterthesis_b200 import ops
            ops.join_side(all_streams=True)
            i = len(vs_ref)
            if i < len(meta.get('gsample', [])):
                vs_ref.append((i, n) + grad_sample_vs_golden(meta, gold, i, n, m.model[n].named_parameters()))
            return orig()
        opt.step = rec
    b = batch(B, size, nd, seed)
    r = {}
    for it in range(iters):
        m.set_inputs(b)
        m.optimize_parameters(it)
        for k in m.loss:
            r[f'loss_it{it}.{k}'] = loss_err(m.loss[k], float(gold[f'loss.{it}.{k}']))
    for i, n, c_all, c_worst, worst in vs_ref:
        r[f'cosref_step{i}_{n}'] = c_all
    return r


def sampling_vs_oracle(dim=64, size=256, N=2, nd=4, seed=5):
    """forward_reference / forward_random (config 2 shape) vs the fp32 oracle on the GPU."""
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    from masterthesis_b200 import models
    args = ref_import.default_args(dim=dim, crop_size=size, batch_size=N, num_domains=nd, init_type=None,
                                   mode='test')
    m = models.AdaINModel(args)
    shapes = {n: {k: list(v.shape) for k, v in m.model[n].state_dict().items()} for n in m.model}
    params = {n: synth.synth_params(shapes[n], seed, n + '.') for n in m.model}
    for n in m.model:
        m.model[n].load_state_dict(params[n])
    m.initialize()
    m.rng = synth.Draws(seed)
    st = ostep.AdaINStep({n: {k: v.to(DEV) for k, v in p.items()} for n, p in params.items()}, N)
    st.Ec = lambda x, draw: onets.content_encoder(
        st.P['content_encoder'], x, draw((x.size(0), dim * 4, x.size(2) // 4, x.size(3) // 4)).to(DEV))
    src = synth.Draws(seed)
    odraw = lambda shape: src(shape).to(DEV)  # noqa: E731
    a = synth.synth_images(N, size, seed, 'src').to(DEV)
    b = synth.synth_images(N, size, seed, 'ref').to(DEV)
    c = synth.synth_onehot(N, nd, seed).to(DEV)
    z = synth.synth_tensor('zr', (N, 8), seed).to(DEV)
    with torch.no_grad():
        y1, secs, gib = m.forward_reference(a, b, c)
        y2, _, _ = m.forward_random(a, z, c)
    r1 = st.forward_reference(a, b, c, odraw)
    r2 = st.forward_random(a, z, c, odraw)
    from tests import emu
    src2 = synth.Draws(seed)
    edraw = lambda shape: src2(shape).to(DEV)  # noqa: E731
    with emu.bf16_storage():
        e1 = st.forward_reference(a, b, c, edraw)
        e2 = st.forward_random(a, z, c, edraw)
    return {'forward_reference': rel(y1, r1), 'forward_random': rel(y2, r2),
            'emu_forward_reference': rel(e1, r1), 'emu_forward_random': rel(e2, r2)}


def base_nets_vs_golden(name='nets_base_dim16_64'):
    """StyleEncoder / Decoder / DecoderConcat / ContentDiscriminator / MultiScaleDiscriminator on the
    CUDA path vs the reference's recorded outputs and gradients (same computations as
    oracle/make_golden.py:nets_base_case).  The reference cannot back-propagate its Decoder
    (DecResnetBlock's in-place residual add), so Decoder gradients are checked against the oracle."""
    from masterthesis_b200.models.core import networks
    from masterthesis_b200.models.core.blocks import to_padded
    meta, gold = load_gold(name)
    seed, N, nd, size, s, dim, latent = (meta[k] for k in ('seed', 'N', 'nd', 'size', 'sub', 'dim', 'latent'))
    C = dim * 4
    ns = {'style_encoder': networks.StyleEncoder(3, latent, dim, num_domains=nd, activation='lrelu'),
          'decoder': networks.Decoder(3, C, num_domains=nd, latent_dim=latent),
          'decoder_concat': networks.DecoderConcat(3, C, num_domains=nd, latent_dim=latent),
          'content_discriminator': networks.ContentDiscriminator(C, nd),
          'ms_discriminator': networks.MultiScaleDiscriminator(3, dim, num_domains=nd)}
    P = {}
    for n, net in ns.items():
        P[n] = synth.synth_params(meta['shapes'][n], seed, n + '.')
        net.load_state_dict(P[n])
        net.to(DEV)
    c = synth.synth_onehot(N, nd, seed).to(DEV)
    r = {}
    img = synth.synth_images(N, size, seed, 'img_se').to(DEV).requires_grad_(True)
    z = ns['style_encoder'].run(img, c)
    (z * synth.synth_tensor('proj_se', z.shape, seed).to(DEV)).sum().backward()
    r['se.z'], r['cos_se.dimg'] = rel(z.detach(), gold['se.z']), cos(img.grad[:, :, ::s, ::s], gold['se.dimg'])
    q = max(1, s // 4)
    for tag, name_ in (('dec', 'decoder'), ('decc', 'decoder_concat')):
        zc_in = synth.synth_tensor('zc_in', (N, C, size // 4, size // 4), seed) * (C * (size // 4) ** 2) ** 0.5
        zc_in = zc_in.to(torch.bfloat16).float().to(DEV).requires_grad_(True)
        zs_in = (synth.synth_tensor('zs_in', (N, latent), seed) * latent ** 0.5).to(DEV).requires_grad_(True)
        fake = ns[name_].run(to_padded(zc_in, 1), zs_in, c)
        proj = synth.synth_tensor('proj', fake.shape, seed).to(DEV)
        (fake * proj).sum().backward()
        r[f'{tag}.fake'] = rel(fake.detach()[:, :, ::s, ::s], gold[f'{tag}.fake'])
        if tag == 'dec':       # gradients vs the fp32 oracle (out-of-place residual add)
            Po = {k: v.to(DEV).requires_grad_(True) for k, v in P['decoder'].items()}
            zc_o, zs_o = zc_in.detach().clone().requires_grad_(True), zs_in.detach().clone().requires_grad_(True)
            (onets.decoder(Po, zc_o, zs_o, c) * proj).sum().backward()
            r['cos_dec.dzc'], r['cos_dec.dzs'] = cos(zc_in.grad, zc_o.grad), cos(zs_in.grad, zs_o.grad)
            mine = dict(ns['decoder'].named_parameters())
            for k in ('dec1.0.conv1.block.1.weight', 'dec1.1.block1.0.weight', 'dec1.3.block2.2.bias',
                      'dec2.0.block.0.weight', 'linear.4.weight'):
                r[f'cos_decoder.{k}'] = cos(mine[k].grad, Po[k].grad)
        else:
            r['cos_decc.dzc'] = cos(zc_in.grad[:, :, ::q, ::q], gold['decc.dzc'])
            r['cos_decc.dzs'] = cos(zs_in.grad, gold['decc.dzs'])
            # yardstick: the oracle with bf16 storage emulated vs the fp32 oracle on the same input
            from tests import emu
            Po = {k: v.to(DEV) for k, v in P['decoder_concat'].items()}
            with torch.no_grad():
                f32 = onets.decoder_concat(Po, zc_in.detach(), zs_in.detach(), c)
                with emu.bf16_storage():
                    e16 = onets.decoder_concat(Po, zc_in.detach(), zs_in.detach(), c)
            r['emu_decc.fake'] = rel(e16, f32)
    zc2 = synth.synth_tensor('zc_dc', (N, C, 64, 64), seed) * (C * 64 * 64) ** 0.5
    zc2 = zc2.to(torch.bfloat16).float().to(DEV).requires_grad_(True)
    pred = ns['content_discriminator'].run(to_padded(zc2, 1))
    (pred * synth.synth_tensor('proj_dc', pred.shape, seed).to(DEV)).sum().backward()
    r['dc.pred'], r['cos_dc.dzc'] = rel(pred.detach(), gold['dc.pred']), cos(zc2.grad[:, :, ::4, ::4], gold['dc.dzc'])
    img4 = synth.synth_images(N, 256, seed, 'img_ms').to(DEV).requires_grad_(True)
    tot = 0
    for i, (dis, cls) in enumerate(ns['ms_discriminator'].run(img4)):
        r[f'ms.dis{i}'], r[f'ms.cls{i}'] = rel(dis.detach(), gold[f'ms.dis{i}']), rel(cls.detach(), gold[f'ms.cls{i}'])
        tot = tot + (dis * synth.synth_tensor(f'proj_ms{i}', dis.shape, seed).to(DEV)).sum() + (cls * c).sum()
    tot.backward()
    r['cos_ms.dimg'] = cos(img4.grad[:, :, ::8, ::8], gold['ms.dimg'])
    for k, g in gold.items():
        if not k.startswith('grad.'):
            continue
        _, n, key = k.split('.', 2)
        if n == 'content_discriminator' and key == 'model.2.block.1.bias':
            continue    # a bias in front of InstanceNorm: its exact gradient is 0, the recorded one is rounding noise
        mine = dict(ns[n].named_parameters())[key].grad
        if mine.numel() > 20000:
            mine = mine.flatten()[:: max(1, mine.numel() // 20000)]
        r['cos_' + k[5:]] = cos(mine, g.flatten())
    return r


def flag_step_vs_golden(name):
    """BaseModel --concat [--reparam] / AdaINModel --use_dis_content --ms_dis on the CUDA path vs the
    REFERENCE's recorded losses, iteration by iteration (replayed CPU draws, eager launches)."""
    meta, gold = load_gold(name)
    seed, B, nd, size, dim = meta['seed'], meta['B'], meta['nd'], meta['size'], meta['dim']
    m, _ = make_model(dim, size, B, nd, seed, meta['shapes'], model=meta['model'], **meta['flags'])
    m.rng = synth.Draws(seed)
    b = batch(B, size, nd, seed)
    r = {}
    vs_ref = []
    for n in m.optimizer:
        opt = m.optimizer[n]

        def rec(orig=opt.step, n=n):
            from masterthesis_b200 import ops
            ops.join_side(all_streams=True)
            i = len(vs_ref)
            if i < len(meta.get('gsample', [])):
                vs_ref.append((i, n) + grad_sample_vs_golden(meta, gold, i, n, m.model[n].named_parameters()))
            return orig()
        opt.step = rec
    for it in range(meta['iters']):
        m.set_inputs(b)
        m.optimize_parameters(it)
        mine = dict(m.loss)
        if getattr(m, 'loss_dc', None) is not None:
            mine['loss_dc'] = m.loss_dc
        for k, v in mine.items():
            r[f'loss_it{it}.{k}'] = loss_err(v, float(gold[f'loss.{it}.{k}']))
        assert set(mine) == {k.split('.', 2)[2] for k in gold if k.startswith(f'loss.{it}.')}, (it, sorted(mine))
    assert [list(s_) for s_ in m.rng.log] == meta['draws'], "RNG call order differs from the reference"
    for i, n, c_all, c_worst, worst in vs_ref:
        r[f'cosref_step{i}_{n}'] = c_all
    return r


def base_default_step_vs_oracle(dim=8, size=64, B=1, nd=4, seed=4, iters=2):
    """BaseModel with default flags (StyleEncoder + Decoder/DecResnetBlock) trains on the CUDA path; the
    reference itself cannot (DecResnetBlock's in-place add breaks autograd), so the oracle (out-of-place
    add, otherwise base_model.py line for line) is the yardstick for the losses of `iters` iterations."""
    m, params = make_model(dim, size, B, nd, seed, model='BaseModel')
    m.rng = synth.Draws(seed)
    st = ostep.AdaINStep({n: {k: v.to(DEV) for k, v in p.items()} for n, p in params.items()}, B,
                         model='base', reparam=False, concat=False)
    src = synth.Draws(seed)
    odraw = lambda shape: src(shape).to(DEV)  # noqa: E731
    b = batch(B, size, nd, seed)
    r = {}
    for it in range(iters):
        m.set_inputs(b)
        m.optimize_parameters(it)
        ref = st.optimize_parameters(m.img.clone(), m.c_org.clone(), odraw)
        assert set(ref) == set(m.loss)
        for k in ref:
            r[f'loss_it{it}.{k}'] = loss_err(m.loss[k], ref[k])
    return r


def base_sampling_vs_oracle(dim=64, size=256, N=1, nd=4, seed=6):
    """BASELINE config 1 on the GPU: BaseModel (default flags) generator forward, batch 1, 256x256 —
    forward_reference (Ec -> StyleEncoder -> Decoder) and forward_random vs the fp32 oracle."""
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    from masterthesis_b200 import models
    from tests import emu
    args = ref_import.default_args(dim=dim, crop_size=size, batch_size=N, num_domains=nd, init_type=None,
                                   mode='test')
    m = models.BaseModel(args)
    shapes = {n: {k: list(v.shape) for k, v in m.model[n].state_dict().items()} for n in m.model}
    params = {n: synth.synth_params(shapes[n], seed, n + '.') for n in m.model}
    for n in m.model:
        m.model[n].load_state_dict(params[n])
    m.initialize()
    m.rng = synth.Draws(seed)
    st = ostep.AdaINStep({n: {k: v.to(DEV) for k, v in p.items()} for n, p in params.items()}, N,
                         model='base', reparam=False, concat=False)
    a = synth.synth_images(N, size, seed, 'src').to(DEV)
    b = synth.synth_images(N, size, seed, 'ref').to(DEV)
    c = synth.synth_onehot(N, nd, seed).to(DEV)
    z = synth.synth_tensor('zr', (N, 8), seed).to(DEV)
    with torch.no_grad():
        y1, _, _ = m.forward_reference(a, b, c)
        y2, _, _ = m.forward_random(a, z, c)
    out = {}
    for tag, wrap in (('', contextlib.nullcontext), ('emu_', emu.bf16_storage)):
        src = synth.Draws(seed)
        draw = lambda shape: src(shape).to(DEV)  # noqa: E731
        with wrap():
            out[tag + 'ref'] = st.forward_reference(a, b, c, draw)
            out[tag + 'rnd'] = st.forward_random(a, z, c, draw)
    return {'forward_reference': rel(y1, out['ref']), 'forward_random': rel(y2, out['rnd']),
            'emu_forward_reference': rel(out['emu_ref'], out['ref']),
            'emu_forward_random': rel(out['emu_rnd'], out['rnd'])}


def graph_vs_eager(dim=8, size=64, B=2, nd=4, seed=3, iters=6, **flags):
    """The CUDA-graph replay of the iteration (Model._run_iteration) vs the eager launch sequence:
    same weights, same batch, same torch seed (device Philox draws), `iters` iterations including
    the 2 eager warm-up iterations, the capture and several replays, with a changing learning rate."""
    out = {}
    # every sum of the iteration runs in a fixed order (no atomics) and the captured iteration launches the
    # same kernels as the eager one, so the three runs must agree exactly; the second eager run separates
    # "not reproducible" from "graph differs from eager" if that ever breaks
    for mode in ('eager', 'eager2', 'graph'):
        m, _ = make_model(dim, size, B, nd, seed, **flags)
        m.use_graph = mode == 'graph'
        b = {k: v.to(DEV) for k, v in batch(B, size, nd, seed).items()}
        torch.manual_seed(77)
        hist = []
        for it in range(iters):
            for opt in m.optimizer.values():          # what update_lr() does through the schedulers
                opt.param_groups[0]['lr'] = 1e-4 * (1.0 - 0.05 * it)
            m.set_inputs(b)
            m.optimize_parameters(it)
            hist.append(dict(m.loss, **({'loss_dc': m.loss_dc} if getattr(m, 'loss_dc', None) is not None else {})))
        torch.cuda.synchronize()
        out[mode] = (hist, {n: torch.cat([p.detach().flatten() for p in m.model[n].parameters()]).clone()
                            for n in m.model}, m._graph is not None)
    assert out['graph'][2] and not out['eager'][2]
    r = {}
    for it, (he, h2, hg) in enumerate(zip(out['eager'][0], out['eager2'][0], out['graph'][0])):
        for k in he:
            scale = max(1e-6, abs(he[k]))
            r[f'loss_it{it}.{k}'] = abs(he[k] - hg[k]) / scale
            r[f'eager_rerun_it{it}.{k}'] = abs(he[k] - h2[k]) / scale
    for n in out['eager'][1]:
        r[f'params.{n}'] = rel(out['graph'][1][n], out['eager'][1][n])
        r[f'eager_rerun_params.{n}'] = rel(out['eager2'][1][n], out['eager'][1][n])
    return r


def determinism(dim=64, size=256, B=8, nd=4, seed=3, iters=3):
    """Reproducibility of the whole training iteration: two eager runs and one CUDA-graph run of the same
    `iters` iterations (same weights, batch and torch seed) must agree BIT FOR BIT in every loss and every
    parameter — the statistics, the norm-backward sums and the split-K weight gradients are fixed-order
    sums (no atomics), and the captured iteration launches the same kernels as the eager one.
    Returns the number of differing values."""
    out = {}
    for mode in ('eager', 'eager2', 'graph'):
        m, _ = make_model(dim, size, B, nd, seed)
        m.use_graph = mode == 'graph'
        m.graph_warmup = 1
        b = {k: v.to(DEV) for k, v in batch(B, size, nd, seed).items()}
        torch.manual_seed(77)
        hist = []
        for it in range(iters):
            for opt in m.optimizer.values():
                opt.param_groups[0]['lr'] = 1e-4 * (1.0 - 0.05 * it)
            m.set_inputs(b)
            m.optimize_parameters(it)
            hist.append(dict(m.loss))
        torch.cuda.synchronize()
        out[mode] = (hist, {n: torch.cat([p.detach().flatten() for p in m.model[n].parameters()]).clone()
                            for n in m.model}, m._graph is not None)
        del m
    assert out['graph'][2] and not out['eager'][2]
    r = {}
    for other in ('eager2', 'graph'):
        r[f'{other}.loss_mismatches'] = float(sum(
            a[k] != b_[k] for a, b_ in zip(out['eager'][0], out[other][0]) for k in a))
        r[f'{other}.param_mismatches'] = float(sum(
            int((out['eager'][1][n] != out[other][1][n]).sum()) for n in out['eager'][1]))
    return r


def sampling_graph_vs_eager(dim=16, size=64, N=2, nd=4, seed=9):
    """forward_random in test mode: the captured sampling graph (third call on) vs eager launches of
    the same model, on inputs that change from call to call (static input buffers must be refreshed)."""
    from masterthesis_b200 import models
    args = ref_import.default_args(dim=dim, crop_size=size, batch_size=N, num_domains=nd, init_type='normal',
                                   mode='test')
    torch.manual_seed(seed)
    m = models.AdaINModel(args)
    m.initialize()
    for n in m.model:
        m.model[n].eval()          # GaussianNoiseLayer is the identity outside training (misc.py:18-26)
    r = {}
    with torch.no_grad():
        for call in range(5):
            a = synth.synth_images(N, size, seed + call, 'src').to(DEV)
            z = synth.synth_tensor('zr', (N, 8), seed + call).to(DEV)
            c = synth.synth_onehot(N, nd, seed + call).to(DEV)
            m.use_graph = True
            y_g, _, _ = m.forward_random(a, z, c)
            m.use_graph = False
            y_e, _, _ = m.forward_random(a, z, c)
            r[f'call{call}'] = rel(y_g, y_e)
    r['captured'] = 0.0 if getattr(m, '_sample_graphs', None) else 1.0
    return r


CASES = {
    'sampling_graph_vs_eager': sampling_graph_vs_eager,
    'graph_vs_eager': graph_vs_eager,
    'nets_base_dim16_64': base_nets_vs_golden,
    'step_basec': lambda: flag_step_vs_golden('step_basec_dim8_64'),
    'step_basecr': lambda: flag_step_vs_golden('step_basecr_dim8_64'),
    'step_dcms': lambda: flag_step_vs_golden('step_dcms_dim8_256'),
    **{f'step_{g}': (lambda g=g: flag_step_vs_golden(f'step_{g}_dim8_64'))
       for g in ('ragan', 'hinge', 'lsgan', 'wgangp', 'nearest', 'dropout', 'sn')},
    'step_base_default': base_default_step_vs_oracle,
    'sampling_base_256': base_sampling_vs_oracle,
    'graph_vs_eager_dc': lambda: graph_vs_eager(size=256, B=1, iters=10, use_dis_content=True),
    'nets_dim16_64': lambda: nets_vs_golden('nets_dim16_64'),
    'nets_dim64_256': lambda: nets_vs_golden('nets_dim64_256'),
    'step_dim8_64': step_vs_golden,
    'sampling_256': sampling_vs_oracle,
    # the benchmarked geometry (BASELINE config 3: dim 64, 256x256, batch 8 -> 16 / 32 images per call)
    'step_dim64_256_B8': lambda: step_vs_golden('step_dim64_256_B8'),
    'step_dim64_256_B8_light': step_losses_vs_golden,
    'graph_vs_eager_bench': lambda: graph_vs_eager(dim=64, size=256, B=8, iters=5),
    'sampling_256_n8': lambda: sampling_vs_oracle(N=8),
    'determinism_bench': determinism,
    'determinism_small': lambda: determinism(dim=8, size=64, B=2, iters=4),
}
