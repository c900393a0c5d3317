"""AdaINModel — same construction, training-step and sampling API as the reference's
src/models/adain_model.py:9-430, driving the CUDA networks of models/core/networks.py.

Differences that do not change results: the generator forwards of the
discriminator phase run without an autograd graph (the reference builds one and never uses it,
:203); weight gradients nobody consumes are skipped (D1 in phase G1, Es in phase G2, item 11);
the real and fake discriminator passes are batched into one call (two calls under --dis_sn,
where each reference forward advances the power iteration); the 13 loss scalars are read
back with ONE host sync per iteration instead of 13 `.item()` calls.
"""
import contextlib
import os
import time
from contextlib import contextmanager

import torch
import torch.nn.functional as F

from .. import functional as Fn
from .. import ops
from ..optim import FlatAdam
from .core import blocks, networks
from .core.networks import CONTENT_PAD
from .model import Model


def _bce(x, target):
    t = torch.full_like(x, float(target)) if not torch.is_tensor(target) else target
    return F.binary_cross_entropy_with_logits(x, t)


class AdaINModel(Model):
    reparam = True      # ReparameterizedStyleEncoder: Es returns (z, mu, logvar)

    def __init__(self, args):
        super().__init__(args)
        self.latent_dim = args.latent_dim
        if getattr(args, 'vgg_loss', None) is not None:
            raise NotImplementedError("VGG perceptual loss needs downloaded weights; not on the accelerated path")
        m = self.model
        m.content_encoder = networks.ContentEncoder(args.input_dim, dim=args.dim, norm_layer=args.enc_norm)
        self._build_generator(args)
        if 'train' in args.mode:
            for name in ('discriminator1', 'discriminator2'):
                if args.ms_dis:       # dim is NOT forwarded by the reference (adain_model.py:33-42)
                    m[name] = networks.MultiScaleDiscriminator(args.input_dim, norm_layer=args.dis_norm,
                                                               sn=args.dis_sn, num_domains=args.num_domains,
                                                               num_scales=args.num_scales)
                else:
                    m[name] = networks.Discriminator(args.input_dim, dim=args.dim, norm_layer=args.dis_norm,
                                                     sn=args.dis_sn, num_domains=args.num_domains,
                                                     image_size=args.crop_size)
            for net in list(m):
                self.optimizer[net] = FlatAdam(m[net].parameters(), lr=args.lr,
                                               betas=(args.beta1, float(args.beta2)), weight_decay=args.wd)
            if args.use_dis_content:
                m.content_discriminator = networks.ContentDiscriminator(
                    dim=m.content_encoder.output_dim, num_domains=args.num_domains)
                self.optimizer.content_discriminator = FlatAdam(
                    m.content_discriminator.parameters(), lr=args.lr / 2.5,
                    betas=(args.beta1, float(args.beta2)), weight_decay=args.wd)
            if args.gan_mode not in ('vanilla', 'lsgan', 'wgangp', 'hinge'):
                raise NotImplementedError(f'Loss {args.gan_mode} is not implemented')
            self.print_loss = ['g_adv', 'g_cls', 'l1_cc_rec']

    def _build_generator(self, args):
        m = self.model
        # --concat / --reparam are ignored exactly as in the reference (adain_model.py:17-22)
        m.style_encoder = networks.ReparameterizedStyleEncoder(
            args.input_dim, output_dim=self.latent_dim, dim=args.dim, num_domains=args.num_domains,
            norm_layer=None, activation='lrelu')
        m.decoder = networks.AdaINDecoder(args.input_dim, dim=m.content_encoder.output_dim,
                                          num_domains=args.num_domains, latent_dim=self.latent_dim,
                                          up_type=args.up_type, norm_layer=args.dec_norm,
                                          dropout=args.use_dropout)

    # ------------------------------------------------------------------ the loss layer as kernels
    # Default flags (vanilla GAN loss, single-scale discriminator, no relativistic term): every loss term is
    # ONE launch that writes the scalar into a slot of a device vector and the gradient with respect to its
    # input in the same pass (mt_loss_bce_logits / mt_loss_l1 / mt_loss_kl_normal); the phase then
    # back-propagates from explicit gradients (torch.autograd.backward(roots, grads)), so no loss graph of
    # tiny library kernels exists.  The named losses of the reference (loss sums, weights) are formed on the
    # host from the raw terms after the one read-back of the iteration.  Other --gan_mode values, --use_ragan
    # and --ms_dis keep the general torch formulation below.
    SLOTS = ('d1_fake', 'd1_real', 'd1_cls', 'd2_fake', 'd2_real', 'd2_cls', 'g_adv', 'g_cls', 'l1_self', 'l1_cc',
             'zc_sumsq', 'kl_zs', 'g_content', 'z_l1', 'gan2', 'gan2_cls', 'loss_dc')

    @property
    def fast_losses(self):
        a = self.args
        return (a.gan_mode == 'vanilla' and not a.use_ragan and not a.ms_dis
                and os.environ.get('MT_FAST_LOSS', '1') == '1')

    def _slot(self, name):
        """1-element view of the loss vector for one raw term"""
        if getattr(self, '_lbuf', None) is None:
            self._lbuf = torch.zeros(len(self.SLOTS), dtype=torch.float32, device=self.device)
        i = self.SLOTS.index(name)
        self._raw_used.add(name)
        return self._lbuf[i:i + 1]

    def _loss_vector(self):
        if not self.fast_losses:
            return super()._loss_vector()
        extra = list(self._scalars)        # terms that still came from torch (BaseModel without --reparam)
        if extra:
            vec = torch.cat([self._lbuf] + [self._scalars[k].float().reshape(1) for k in extra])
        else:
            vec = self._lbuf
        return [('raw', self.SLOTS, tuple(self._raw_used), tuple(extra))], vec

    def _named_losses(self, keys, values):
        if not keys or not isinstance(keys[0], tuple):
            return super()._named_losses(keys, values)
        _, slots, used, extra = keys[0]
        r = dict(zip(slots, values))
        r.update(zip(extra, values[len(slots):]))
        a = self.args
        out = {}
        if 'loss_dc' in used:
            out['loss_dc'] = r['loss_dc']
        if 'd2_fake' in used:
            # the reference logs the values of the LAST discriminator update (discriminator2), :216-223
            out['d_adv'] = r['d2_fake'] + r['d2_real']
            out['d_cls'] = r['d2_cls']
            out['d_total'] = out['d_adv'] + a.lambda_cls * out['d_cls']
        if 'g_adv' in used:
            if 'g_content' in used:
                out['g_content'] = r['g_content']
            out['g_adv'] = r['g_adv']
            out['g_cls'] = r['g_cls'] * a.lambda_cls_G
            out['kl_zc'] = r['zc_sumsq'] / self._n_zc * 0.01
            out['kl_zs'] = r['kl_zs'] * 0.01 if 'kl_zs' in used else r['kl_zs_t']
            out['l1_self_rec'] = r['l1_self']
            out['l1_cc_rec'] = r['l1_cc']
            out['total_g'] = (out['g_adv'] + out['g_cls'] + out['l1_self_rec'] + out['l1_cc_rec'] + out['kl_zc']
                              + out['kl_zs'] + out.get('g_content', 0.0))
        if 'gan2' in used:
            out['l1_recon_z'] = r['z_l1']
            out['gan2'] = r['gan2']
            out['gan2_cls'] = r['gan2_cls'] * a.lambda_cls_G
        return out.items()

    # ------------------------------------------------------------------ small pieces
    def gan_loss(self, pred, is_real):
        """GANLoss.forward (reference loss.py:52-64)."""
        mode = self.args.gan_mode
        if mode == 'wgangp':
            return -pred.mean() if is_real else pred.mean()
        if mode == 'lsgan':
            return F.mse_loss(pred, torch.full_like(pred, 1.0 if is_real else 0.0))
        return _bce(pred, 1.0 if is_real else 0.0)

    def get_z_random(self, bs, latent_dim):
        if self.rng is not None:
            return self.rng((bs, latent_dim)).to(self.device)
        return torch.randn(bs, latent_dim, device=self.device)

    def set_inputs(self, inputs):
        self._static_inputs(inputs['x1'], inputs['y1'], inputs['x2'], inputs['y2'])

    def Ec(self, img):
        return self.model.content_encoder.run(img, self.rng)

    def Es(self, img, c):
        """-> (z_s, mu, logvar); mu / logvar are None for the plain StyleEncoder (BaseModel w/o --reparam)"""
        out = self.model.style_encoder.run(img, c, self.rng)
        return out if self.reparam else (out, None, None)

    def G(self, z_c, z_s, c):
        return self.model.decoder.run(z_c, z_s, c)

    def _G_pair(self, first, second):
        """The two decoder calls the reference makes back to back in every phase (:159-162, :244-251,
        :345-346), each given as (content codes, style codes, class codes) tuples of B-sized pieces.
        The decoder has no cross-sample statistic (AdaIN / LayerNorm are per sample), so both calls run
        as ONE launch sequence over the stacked batch: half the launches, and the hot 3x3 convolution
        fills its persistent grid (32 images = 6.9 tile rounds on 74 CTA pairs instead of 2 x 3.5 -> 4).
        Kept as two calls under --use_dropout so that the masks are drawn call by call as in the
        reference.  Returns the per-piece outputs of each call."""
        B = first[0][0].shape[0]
        if self.args.use_dropout:
            outs = [self.G(torch.cat(zc), torch.cat(zs), torch.cat(c)) for zc, zs, c in (first, second)]
        else:
            both = self.G(torch.cat(first[0] + second[0]), torch.cat(first[1] + second[1]),
                          torch.cat(first[2] + second[2]))
            k = len(first[0]) * B
            outs = [both[:k], both[k:]]
        return [torch.split(o, B, dim=0) for o in outs]

    @contextmanager
    def _frozen(self, *nets):
        ps = [p for n in nets for p in self.model[n].parameters() if p.requires_grad]
        for p in ps:
            p.requires_grad_(False)
        try:
            yield
        finally:
            for p in ps:
                p.requires_grad_(True)

    def _const(self, value):
        """cached 0-dim fp32 device constant (kernel arguments that are read from device memory)"""
        c = self.__dict__.setdefault('_consts', {})
        if value not in c:
            c[value] = torch.tensor(float(value), dtype=torch.float32, device=self.device)
        return c[value]

    def _step(self, *nets, prepack=True):
        """all-reduce (data parallel) + Adam for the given networks (+ their bf16 weight packs, ahead of
        the next use, on the pack stream)."""
        if self.dp is not None:
            self.dp.allreduce([self.optimizer[n] for n in nets])
        blocks.join_prepack(self.ctx)              # no pack may still be reading the weights this step overwrites
        for n in nets:
            self.optimizer[n].step()
        if prepack:                        # the stepped weights have another consumer in this iteration
            blocks.prepack(*[self.model[n] for n in nets])

    # ------------------------------------------------------------------ sampling (:96-109)
    def _mem_gib(self):
        return torch.cuda.memory_reserved(0) / (1024 * 1024 * 1024)

    def forward_random(self, img, z_r, c_trg):
        start = time.time()
        img_fake = self._run_sampling('random', lambda a, z, c: self.G(self.Ec(a), z, c), img, z_r, c_trg)
        end = time.time()
        return img_fake, end - start, self._mem_gib()

    def forward_reference(self, img_src, img_ref, c_trg):
        start = time.time()
        img_fake = self._run_sampling(
            'reference', lambda a, b, c: self.G(self.Ec(a), self.Es(b, c)[0], c), img_src, img_ref, c_trg)
        end = time.time()
        return img_fake, end - start, self._mem_gib()

    def forward(self, img, c_org):
        """the three translations used for visualisation (:111-134)."""
        B = self.args.batch_size
        z_ca, z_cb = torch.split(self.Ec(img), B, dim=0)
        z_s, _, _ = self.Es(img, c_org)
        z_sa, z_sb = torch.split(z_s, B, dim=0)
        z_sr = self.get_z_random(B, self.args.latent_dim)
        cls_a, cls_b = torch.split(c_org, B, dim=0)
        f = self.G(torch.cat((z_cb, z_ca, z_cb)), torch.cat((z_sa, z_sa, z_sr)), torch.cat((cls_a,) * 3))
        img_ba, img_aa, img_br = torch.split(f, B, dim=0)
        f = self.G(torch.cat((z_ca, z_cb, z_ca)), torch.cat((z_sb, z_sb, z_sr)), torch.cat((cls_b,) * 3))
        img_ab, img_bb, img_ar = torch.split(f, B, dim=0)
        return torch.cat((img_ba, img_ab)), torch.cat((img_br, img_ar)), torch.cat((img_aa, img_bb))

    # ------------------------------------------------------------------ phase D (:146-223)
    def update_discriminator(self, img, c_org):
        B = self.args.batch_size
        with torch.no_grad():
            cls_a, cls_b = torch.split(c_org, B, dim=0)
            z_c, (z_s, _, _) = self._encode(img, c_org)
            z_ca, z_cb = torch.split(z_c, B, dim=0)
            z_sa, z_sb = torch.split(z_s, B, dim=0)
            z_sr = self.get_z_random(B, self.args.latent_dim)
            (img_ba, img_br), (img_ab, img_ar) = self._G_pair(
                ((z_cb, z_cb), (z_sa, z_sr), (cls_a, cls_a)), ((z_ca, z_ca), (z_sb, z_sr), (cls_b, cls_b)))
            img_fake = torch.cat((img_ba, img_ab))
            img_random = torch.cat((img_br, img_ar))
        # The two discriminator updates are independent (separate networks, optimizers and inputs) and
        # their tail layers leave most SMs idle: D1 runs on an auxiliary stream next to D2.  Host
        # order is the reference's (D1 then D2), so self.loss ends up with D2's values as there.
        # Their optimizer steps (gradient all-reduce over NVLink under data parallelism, fused Adam, bf16
        # weight re-packs) leave the critical path: nothing reads the discriminators' weights before the
        # adversarial branch of the generator phase (`_wait_d_steps`), so both steps run on a third stream
        # under the generator's Ec / Es / G forward.
        aux, cur = self._aux_stream(), torch.cuda.current_stream()
        if aux is not None:
            aux.wait_stream(cur)
        with (torch.cuda.stream(aux) if aux is not None else contextlib.nullcontext()):
            self.optimizer.discriminator1.zero_grad()
            self.backward_discriminator(self.model.discriminator1, img, img_fake, c_org)
            if aux is None:
                self._step('discriminator1')
        self.optimizer.discriminator2.zero_grad()
        self.backward_discriminator(self.model.discriminator2, img, img_random, c_org)
        if aux is None:
            self._step('discriminator2')
            return
        # D1's backward read tensors that belong to the main stream (img_fake ...): join before they are freed
        cur.wait_stream(aux)
        comm = self._comm_stream()
        comm.wait_stream(cur)
        with torch.cuda.stream(comm):
            self._step('discriminator1')
            self._step('discriminator2')

    def _comm_stream(self):
        if getattr(self, '_comm', None) is None:
            self._comm = torch.cuda.Stream()
        return self._comm

    def _wait_d_steps(self):
        """the current stream waits for the discriminators' optimizer steps issued by update_discriminator
        (called in front of every use of their weights; a no-op once joined)"""
        if getattr(self, '_comm', None) is not None:
            torch.cuda.current_stream().wait_stream(self._comm)

    def _fork(self, fn_main, fn_aux):
        """(fn_main(), fn_aux()) with fn_aux on the auxiliary stream.  For INDEPENDENT networks only
        (no shared parameters): autograd runs each one's backward on the stream of its forward, so
        the overlap holds in the backward pass too.  Host order is fn_main then fn_aux."""
        aux = self._aux_stream()
        if aux is None:
            return fn_main(), fn_aux()
        cur = torch.cuda.current_stream()
        aux.wait_stream(cur)          # fork point BEFORE fn_main is enqueued: aux does not wait for it
        a = fn_main()
        with torch.cuda.stream(aux):
            b = fn_aux()
        cur.wait_stream(aux)
        return a, b

    def _encode(self, img, c):
        """(Ec(img), Es(img, c)): the style encoder (no norms, small maps) leaves SMs idle that the
        content encoder's convs can use.  Ec first on the host, as in the reference (RNG draw order)."""
        return self._fork(lambda: self.Ec(img), lambda: self.Es(img, c))

    def _aux_stream(self):
        """second stream for the D1 || D2 overlap (MT_CONCURRENT_D=0 or the debug modes disable it)"""
        if os.environ.get('MT_CONCURRENT_D', '1') != '1' or ops.POISON or ops.PROFILE is not None:
            return None
        if getattr(self, '_aux', None) is None:
            self._aux = torch.cuda.Stream()
        return self._aux

    def backward_discriminator(self, netD, real, fake, c_org):
        n = real.shape[0]
        if self.fast_losses:
            tag = 'd1' if netD is self.model.discriminator1 else 'd2'
            if self.args.dis_sn:       # two calls (each advances the power iteration), reference order :207-209
                (pf, cf), (pr, cr) = netD.run(fake.detach()), netD.run(real)
                roots, halves = [pf, pr, cr], None
            else:
                pred, pred_cls = netD.run(torch.cat((fake.detach(), real)))
                pf, pr, cr = pred[:n], pred[n:], pred_cls[n:]
                roots, halves = [pred, pred_cls], (torch.empty_like(pred), torch.zeros_like(pred_cls))
            g_f = ops.loss_bce_logits(pf.detach(), 0.0, 1.0, self._slot(tag + '_fake'),
                                      dx_out=None if halves is None else halves[0][:n])
            g_r = ops.loss_bce_logits(pr.detach(), 1.0, 1.0, self._slot(tag + '_real'),
                                      dx_out=None if halves is None else halves[0][n:])
            g_c = ops.loss_bce_logits(cr.detach(), c_org, self.args.lambda_cls, self._slot(tag + '_cls'),
                                      dx_out=None if halves is None else halves[1][n:])
            torch.autograd.backward(roots, [g_f, g_r, g_c] if halves is None else list(halves))
            ops.join_side()
            return
        if self.args.dis_sn:
            # spectral norm: every reference forward runs one power iteration, so the fake and the real
            # call see different W / sigma (:207-209) -> two calls, in the reference's order
            outs_f, outs_r = netD.run(fake.detach()), netD.run(real)
            if not self.args.ms_dis:
                outs_f, outs_r = [outs_f], [outs_r]
            outs = [(torch.cat((pf, pr)), torch.cat((cf, cr))) for (pf, cf), (pr, cr) in zip(outs_f, outs_r)]
        else:
            outs = netD.run(torch.cat((fake.detach(), real)))
            if not self.args.ms_dis:
                outs = [outs]
        if self.args.ms_dis:           # backward_multi_scale_discriminator (:182-199)
            loss_d_adv, loss_d_cls = 0, 0
            for pred, pred_cls in outs:
                loss_d_adv = loss_d_adv + self.gan_loss(pred[:n], 0) + self.gan_loss(pred[n:], 1)
                loss_d_cls = loss_d_cls + _bce(pred_cls[n:], c_org)
        else:
            pred, pred_cls = outs[0]
            pred_fake, pred_real = pred[:n], pred[n:]
            if self.args.use_ragan:        # relativistic average (:206-208)
                loss_d_adv = (self.gan_loss(pred_real - torch.mean(pred_fake), 1) +
                              self.gan_loss(pred_fake - torch.mean(pred_real), 0)) / 2
            elif 'hinge' in self.args.gan_mode:
                loss_d_adv = F.relu(1.0 - pred_real).mean() + F.relu(1.0 + pred_fake).mean()
            else:
                loss_d_adv = self.gan_loss(pred_fake, 0) + self.gan_loss(pred_real, 1)
            loss_d_cls = _bce(pred_cls[n:], c_org)
        loss_d = loss_d_adv + self.args.lambda_cls * loss_d_cls
        loss_d.backward()
        ops.join_side()      # pending side-stream weight gradients land before the bucket is read
        self._scalars.update(d_adv=loss_d_adv.detach(), d_cls=loss_d_cls.detach(), d_total=loss_d.detach())

    # ------------------------------------------------------------------ content discriminator (:136-144)
    def update_content_discriminator(self, img, c_org):
        with torch.no_grad():
            z_c = self.Ec(img)
        opt = self.optimizer.content_discriminator
        opt.zero_grad()
        if self.fast_losses:
            pred = self.model.content_discriminator.run(z_c)
            torch.autograd.backward([pred], [ops.loss_bce_logits(pred.detach(), c_org, 1.0, self._slot('loss_dc'))])
        else:
            loss_d_content = _bce(self.model.content_discriminator.run(z_c), c_org)
            loss_d_content.backward()
            self._scalars.update(loss_dc=loss_d_content.detach())
        ops.join_side()      # pending side-stream weight gradients land before the bucket is read
        if self.dp is not None:
            self.dp.allreduce([opt])
        # nn.utils.clip_grad_norm_(params, 5) on the flat bucket (after the all-reduce: a global norm)
        bucket = opt.grad_bucket
        total_norm = torch.linalg.vector_norm(bucket) * opt.grad_scale
        bucket.mul_(torch.clamp(5.0 / (total_norm + 1e-6), max=1.0))
        blocks.join_prepack(self.ctx)
        opt.step()

    # ------------------------------------------------------------------ phases G1 / G2 (:225-394)
    def update_generator(self, img, c_org):
        for n in ('content_encoder', 'style_encoder', 'decoder'):
            self.optimizer[n].zero_grad()
        self.backward_generator(img, c_org)
        self._step('content_encoder', 'style_encoder', 'decoder')
        for n in ('content_encoder', 'decoder'):
            self.optimizer[n].zero_grad()
        self.backward_decoder_random(img, c_org)
        self._step('content_encoder', 'decoder', prepack=False)      # next use: the next iteration

    def _g_adv_fast(self, netD, fake, c_org, adv_slot, cls_slot):
        """generator-side adversarial + classification terms as (roots, gradients): BCE(D(fake), 1) and
        lambda_cls_G * BCE(cls(fake), c_org), losses written to their slots"""
        self._wait_d_steps()
        pred, pred_cls = netD.run(fake)
        g_a = ops.loss_bce_logits(pred.detach(), 1.0, 1.0, self._slot(adv_slot))
        g_c = ops.loss_bce_logits(pred_cls.detach(), c_org, self.args.lambda_cls_G, self._slot(cls_slot))
        return [pred, pred_cls], [g_a, g_c]

    def _g_adv(self, netD, fake, c_org, real=None, netD_real=None):
        """generator-side adversarial + classification loss.  `real` / `netD_real`: the relativistic
        branch (--use_ragan, :286-292 and :360-366) also scores the real images, in phase G2 with
        the OTHER discriminator (discriminator2 for the reals, discriminator1 for the fakes)."""
        self._wait_d_steps()
        if self.args.ms_dis:
            adv, cls = 0, 0
            for pred_fake, pred_fake_cls in netD.run(fake):
                adv = adv + self.gan_loss(pred_fake, 1)
                cls = cls + _bce(pred_fake_cls, c_org)
            return adv, cls * self.args.lambda_cls_G
        if self.args.use_ragan:
            if netD_real is netD and not self.args.dis_sn:   # one call over both halves (no batch statistics in D)
                n = fake.shape[0]
                pred, pred_cls = netD.run(torch.cat((fake, real)))
                pred_fake, pred_fake_cls, pred_real = pred[:n], pred_cls[:n], pred[n:]
            else:
                pred_real, _ = netD_real.run(real)
                pred_fake, pred_fake_cls = netD.run(fake)
            adv = (self.gan_loss(pred_real - torch.mean(pred_fake), 0) +
                   self.gan_loss(pred_fake - torch.mean(pred_real), 1)) / 2
            return adv, _bce(pred_fake_cls, c_org) * self.args.lambda_cls_G
        pred_fake, pred_fake_cls = netD.run(fake)
        adv = -pred_fake.mean() if 'hinge' in self.args.gan_mode else self.gan_loss(pred_fake, 1)
        return adv, _bce(pred_fake_cls, c_org) * self.args.lambda_cls_G

    def backward_generator(self, img, c_org):
        B = self.args.batch_size
        world = 1 if self.dp is None else self.dp.world
        cls_a, cls_b = torch.split(c_org, B, dim=0)
        z_c, (z_s, mu, logvar) = self._encode(img, c_org)
        z_ca, z_cb = torch.split(z_c, B, dim=0)
        z_sa, z_sb = torch.split(z_s, B, dim=0)
        (img_ba, img_aa), (img_ab, img_bb) = self._G_pair(
            ((z_cb, z_ca), (z_sa, z_sa), (cls_a, cls_a)), ((z_ca, z_cb), (z_sb, z_sb), (cls_b, cls_b)))
        img_fake = torch.cat((img_ba, img_ab))
        img_self = torch.cat((img_aa, img_bb))
        fast = self.fast_losses
        # the adversarial branch (D1 on the fakes) is independent of the reconstruction branch
        with self._frozen('discriminator1'):
            z_c_rec, ((z_s_rec, _, _), adv) = self._fork(
                lambda: self.Ec(img_fake),
                lambda: (self.Es(img_fake, c_org),
                         self._g_adv_fast(self.model.discriminator1, img_fake, c_org, 'g_adv', 'g_cls') if fast else
                         self._g_adv(self.model.discriminator1, img_fake, c_org, img, self.model.discriminator1)))
        z_c_rec_b, z_c_rec_a = torch.split(z_c_rec, B, dim=0)
        z_s_rec_a, z_s_rec_b = torch.split(z_s_rec, B, dim=0)
        img_recon = self.G(torch.cat((z_c_rec_a, z_c_rec_b)), torch.cat((z_s_rec_a, z_s_rec_b)),
                           torch.cat((cls_a, cls_b)))
        N, Hp, Wp, _ = z_c.shape
        n_zc = N * (Hp - 2 * CONTENT_PAD) * (Wp - 2 * CONTENT_PAD) * self.model.content_encoder.output_dim
        if fast:
            # every term: value into its slot + gradient with respect to the network output it reads
            roots, grads = adv
            if self.args.use_dis_content:      # backward_content_discriminator (:334-337)
                with self._frozen('content_discriminator'):
                    pred_c = self.model.content_discriminator.run(z_c)
                roots.append(pred_c)
                grads.append(ops.loss_bce_logits(pred_c.detach(), 1 - c_org, 1.0, self._slot('g_content')))
            roots += [img_self, img_recon]
            grads += [ops.loss_l1(img_self.detach(), img, self.args.lambda_rec, self._slot('l1_self')),
                      ops.loss_l1(img_recon.detach(), img, self.args.lambda_rec, self._slot('l1_cc'))]
            # 0.01 * mean(z_c^2) (:396-399): sum of squares into its slot, gradient 0.02 / n * z_c
            self._n_zc = n_zc
            ops.sumsq_interior(z_c.detach(), CONTENT_PAD, out=self._slot('zc_sumsq'))
            roots.append(z_c)
            grads.append(ops.scale_interior(z_c.detach(), self._const(0.02 / n_zc), CONTENT_PAD))
            extra = None
            if self.reparam:
                # a SUM over the batch (:313-314): under data parallelism it adds up over ranks while the
                # gradient all-reduce averages -> the gradient is pre-scaled by the world size
                dmu, dlv = ops.loss_kl_normal(mu.detach(), logvar.detach(), 0.01 * world, self._slot('kl_zs'))
                roots += [mu, logvar]
                grads += [dmu, dlv]
            else:                              # base_model.py:346-347: a mean, no world scaling
                extra = torch.mean(z_s.pow(2)) * 0.01
                self._scalars.update(kl_zs_t=extra.detach())
                roots.append(extra)
                grads.append(None)
            torch.autograd.backward(roots, grads)
            ops.join_side(all_streams=True)
            return
        loss_g_adv, loss_g_cls = adv
        loss_g_content = None
        if self.args.use_dis_content:          # backward_content_discriminator (:334-337)
            with self._frozen('content_discriminator'):
                loss_g_content = _bce(self.model.content_discriminator.run(z_c), 1 - c_org)
        loss_g_self = F.l1_loss(img_self, img) * self.args.lambda_rec
        loss_g_cc = F.l1_loss(img_recon, img) * self.args.lambda_rec
        loss_kl_zc = Fn.SumSqInterior.apply(z_c, CONTENT_PAD) / n_zc * 0.01
        # SUM over the batch, not a mean (:313-314): under data parallelism it must add up over
        # ranks while the gradient all-reduce averages -> pre-scale by the world size.
        if self.reparam:
            loss_kl_zs = torch.sum(1 + logvar - mu.pow(2) - logvar.exp()) * (-0.5) * 0.01
        else:                                  # base_model.py:346-347: a mean, no world scaling
            loss_kl_zs = torch.mean(z_s.pow(2)) * 0.01
        loss_g = loss_g_adv + loss_g_cls + loss_g_self + loss_g_cc + loss_kl_zc + loss_kl_zs
        if loss_g_content is not None:
            loss_g = loss_g + loss_g_content
            self._scalars.update(g_content=loss_g_content.detach())
        (loss_g + (world - 1) * loss_kl_zs if (world > 1 and self.reparam) else loss_g).backward()
        ops.join_side(all_streams=True)      # incl. the style encoder's, whose backward ran on the aux stream
        self._scalars.update(g_adv=loss_g_adv.detach(), g_cls=loss_g_cls.detach(),
                             kl_zc=loss_kl_zc.detach(), kl_zs=loss_kl_zs.detach(),
                             l1_self_rec=loss_g_self.detach(), l1_cc_rec=loss_g_cc.detach(),
                             total_g=loss_g.detach())

    def backward_decoder_random(self, img, c_org):
        B = self.args.batch_size
        cls_a, cls_b = torch.split(c_org, B, dim=0)
        z_ca, z_cb = torch.split(self.Ec(img), B, dim=0)
        z_sr = self.get_z_random(B, self.args.latent_dim)
        (img_br,), (img_ar,) = self._G_pair(((z_cb,), (z_sr,), (cls_a,)), ((z_ca,), (z_sr,), (cls_b,)))
        img_random = torch.cat((img_br, img_ar))
        # the reference scores img_random with discriminator1 under --ms_dis (:361) and with
        # discriminator2 otherwise (:378); the relativistic branch (:361-362) takes the reals from
        # discriminator2 and the fakes from discriminator1
        dname = 'discriminator1' if (self.args.ms_dis or self.args.use_ragan) else 'discriminator2'
        fast = self.fast_losses
        with self._frozen('discriminator1', 'discriminator2', 'style_encoder'):
            adv, (z_sr_rec, mu2, _) = self._fork(
                lambda: (self._g_adv_fast(self.model[dname], img_random, c_org, 'gan2', 'gan2_cls') if fast else
                         self._g_adv(self.model[dname], img_random, c_org, img, self.model.discriminator2)),
                lambda: self.Es(img_random, c_org))
        if fast:
            roots, grads = adv
            # (L1(mu_a, z) + L1(mu_b, z)) * 10 (:380-383) = 20 * mean over both halves at once
            # (base_model.py:417-420 without --reparam: the FIRST half of the re-encoded style, twice)
            rec = mu2 if self.reparam else torch.cat((z_sr_rec[:B], z_sr_rec[:B]))
            roots.append(rec)
            grads.append(ops.loss_l1(rec.detach(), torch.cat((z_sr, z_sr)), 20.0, self._slot('z_l1')))
            torch.autograd.backward(roots, grads)
            ops.join_side(all_streams=True)
            return
        loss_g_adv2, loss_g_cls2 = adv
        if self.reparam:
            mu2_a, mu2_b = torch.split(mu2, B, dim=0)
            loss_z_l1 = (F.l1_loss(mu2_a, z_sr) + F.l1_loss(mu2_b, z_sr)) * 10
        else:                                  # base_model.py:417-420 uses z_sr_a twice
            z_sr_a, _ = torch.split(z_sr_rec, B, dim=0)
            loss_z_l1 = (F.l1_loss(z_sr_a, z_sr) + F.l1_loss(z_sr_a, z_sr)) * 10
        (loss_z_l1 + loss_g_adv2 + loss_g_cls2).backward()
        ops.join_side(all_streams=True)
        self._scalars.update(l1_recon_z=loss_z_l1.detach(), gan2=loss_g_adv2.detach(),
                             gan2_cls=loss_g_cls2.detach())

    # ------------------------------------------------------------------ the iteration (:421-430)
    def optimize_parameters(self, global_iter):
        blocks.set_dropout_rng(self.rng, self.ctx)       # --use_dropout masks: replayed draws (parity) or device Philox
        if self.args.use_dis_content and global_iter % self.args.d_iter != 0:
            self._run_iteration(self._content_iteration, 'content')
        else:
            self._run_iteration(self._iteration, 'main')

    def _iteration(self):
        # weights stepped at the end of the previous iteration (all of them when an iteration is being
        # captured): packed layer by layer on the pack stream, under the first convolutions
        blocks.prepack(*self.model.values())
        self.update_discriminator(self.img, self.c_org)
        self.update_generator(self.img, self.c_org)
        self._wait_d_steps()
        blocks.join_prepack(self.ctx)

    def _content_iteration(self):
        blocks.prepack(self.model.content_encoder, self.model.content_discriminator)
        self.update_content_discriminator(self.img, self.c_org)
        blocks.join_prepack(self.ctx)

    def compute_visuals(self):
        B = self.args.batch_size
        with torch.no_grad():
            img_fake, img_random, img_self = self.forward(self.img, self.c_org)
        fa, fb = torch.split(img_fake, B, dim=0)
        ra, rb = torch.split(img_random, B, dim=0)
        sa, sb = torch.split(img_self, B, dim=0)
        row1 = torch.cat((self.img_a[0:1], fb[0:1], rb[0:1], sa[0:1]), dim=3)
        row2 = torch.cat((self.img_b[0:1], fa[0:1], ra[0:1], sb[0:1]), dim=3)
        return torch.cat((row1, row2), dim=2)
