"""TEST INFRASTRUCTURE — fp32 restatement of AdaINModel's training iteration and sampling
forwards (reference src/models/adain_model.py), on top of oracle/nets.py.  Same import rules as
oracle/nets.py.  Random draws are supplied by the caller (`draw(shape) -> tensor`) in the
reference's own order so that the CUDA path can consume identical noise.
"""
import math

import torch
import torch.nn.functional as F

from . import nets

NETS = ('content_encoder', 'style_encoder', 'decoder', 'discriminator1', 'discriminator2')


class Adam:
    """torch.optim.Adam with coupled L2 weight decay (adain_model.py:58-61), restated."""

    def __init__(self, params, lr, betas, wd, eps=1e-8):
        self.params = params
        self.lr, self.b1, self.b2, self.wd, self.eps = lr, betas[0], float(betas[1]), wd, eps
        self.t = 0
        self.m = {k: torch.zeros_like(v) for k, v in params.items()}
        self.v = {k: torch.zeros_like(v) for k, v in params.items()}

    def zero_grad(self):
        for p in self.params.values():
            p.grad = None

    @torch.no_grad()
    def step(self):
        self.t += 1
        for k, p in self.params.items():
            if p.grad is None:
                continue
            g = p.grad + self.wd * p
            self.m[k].mul_(self.b1).add_(g, alpha=1 - self.b1)
            self.v[k].mul_(self.b2).addcmul_(g, g, value=1 - self.b2)
            denom = self.v[k].sqrt() / math.sqrt(1 - self.b2 ** self.t) + self.eps
            p.addcdiv_(self.m[k], denom, value=-self.lr / (1 - self.b1 ** self.t))


def bce_logits(x, target):
    """nn.BCEWithLogitsLoss (mean) against a constant or tensor target (loss.py:52-64)."""
    t = torch.full_like(x, float(target)) if not torch.is_tensor(target) else target
    return F.binary_cross_entropy_with_logits(x, t)


class AdaINStep:
    """State + one iteration of AdaINModel / BaseModel (no VGG).
    model='adain': ReparameterizedStyleEncoder + AdaINDecoder (adain_model.py);
    model='base' : StyleEncoder or (reparam) ReparameterizedStyleEncoder, Decoder or (concat)
                   DecoderConcat (base_model.py).
    ms_dis: MultiScaleDiscriminator; use_dis_content: ContentDiscriminator with the d_iter schedule."""

    def __init__(self, params, B, lr=1e-4, betas=(0.5, 0.999), wd=1e-4, lambda_rec=10.0,
                 lambda_cls=1.0, lambda_cls_G=5.0, latent_dim=8, model='adain', reparam=True,
                 concat=False, ms_dis=False, use_dis_content=False, d_iter=3, gan_mode='vanilla',
                 use_ragan=False, up_type='transpose', use_dropout=False, enc_norm='instance',
                 dec_norm='layer', dis_norm=None):
        # spectral-norm buffers (`*_u`, `*_v`, --dis_sn) are state, not parameters
        self.P = {n: {k: v.detach().clone().requires_grad_(not k.endswith(('_u', '_v')))
                      for k, v in params[n].items()} for n in params}
        self.up_type, self.use_dropout = up_type, use_dropout
        self.enc_norm, self.dec_norm, self.dis_norm = enc_norm, dec_norm, dis_norm
        self.opt = {n: Adam(self.P[n], lr / 2.5 if n == 'content_discriminator' else lr, betas, wd)
                    for n in self.P}
        self.model, self.concat, self.ms_dis = model, concat, ms_dis
        self.reparam = True if model == 'adain' else reparam
        self.use_dis_content, self.d_iter, self.it = use_dis_content, d_iter, 0
        self.loss_dc = None
        self.gan_mode, self.use_ragan = gan_mode, use_ragan
        self.B, self.latent = B, latent_dim
        self.lrec, self.lcls, self.lclsg = lambda_rec, lambda_cls, lambda_cls_G
        self.loss = {}

    # -- adversarial terms ------------------------------------------------------------------
    def gan(self, pred, is_real):
        """GANLoss.forward (loss.py:52-64): BCE-with-logits, MSE (lsgan) or the signed mean (wgangp).
        loss.py:48 (`elif loss == 'wgangp' or 'hinge'`) leaves self.loss = None for 'hinge'; the
        hinge branches of the model never call it."""
        if 'wgangp' in self.gan_mode:
            return -pred.mean() if is_real else pred.mean()
        if self.gan_mode == 'lsgan':
            return F.mse_loss(pred, torch.full_like(pred, 1.0 if is_real else 0.0))
        return bce_logits(pred, 1 if is_real else 0)

    def d_adv(self, pred_fake, pred_real):
        """adain_model.py:206-215 (single-scale discriminator)"""
        if self.use_ragan:
            return (self.gan(pred_real - pred_fake.mean(), 1) + self.gan(pred_fake - pred_real.mean(), 0)) / 2
        if 'hinge' in self.gan_mode:
            return F.relu(1.0 - pred_real).mean() + F.relu(1.0 + pred_fake).mean()
        return self.gan(pred_fake, 0) + self.gan(pred_real, 1)

    def g_adv(self, which_fake, fake, c_org, which_real, real):
        """adain_model.py:277-302 / :359-376 -> (adversarial, classification * lambda_cls_G)"""
        if self.ms_dis:
            adv, cls = 0, 0
            for pred_fake, pred_fake_cls in self.D(which_fake, fake):
                adv = adv + self.gan(pred_fake, 1)
                cls = cls + bce_logits(pred_fake_cls, c_org)
            return adv, cls * self.lclsg
        if self.use_ragan:
            pred_real, _ = self.D(which_real, real)[0]
            pred_fake, pred_fake_cls = self.D(which_fake, fake)[0]
            adv = (self.gan(pred_real - pred_fake.mean(), 0) + self.gan(pred_fake - pred_real.mean(), 1)) / 2
            return adv, bce_logits(pred_fake_cls, c_org) * self.lclsg
        pred_fake, pred_fake_cls = self.D(which_fake, fake)[0]
        adv = -pred_fake.mean() if 'hinge' in self.gan_mode else self.gan(pred_fake, 1)
        return adv, bce_logits(pred_fake_cls, c_org) * self.lclsg

    # -- networks -------------------------------------------------------------------------
    def Ec(self, x, draw):
        C = self.P['content_encoder']['model.3.model.0.block.1.weight'].shape[0]
        n = draw((x.size(0), C, x.size(2) // 4, x.size(3) // 4)) if draw else None
        return nets.content_encoder(self.P['content_encoder'], x, n, norm=self.enc_norm)

    def Es(self, x, c, draw):
        """-> (z, mu, logvar); mu = logvar = None for the plain StyleEncoder"""
        if not self.reparam:
            return nets.style_encoder_plain(self.P['style_encoder'], x, c), None, None
        eps = draw((x.size(0), self.latent))
        return nets.style_encoder_reparam(self.P['style_encoder'], x, c, eps)

    def G(self, zc, zs, c, draw=None):
        """`draw` feeds the decoder's Dropout(0.5) masks (--use_dropout; training mode only)"""
        kw = dict(up_type=self.up_type, drop=draw if self.use_dropout else None, norm=self.dec_norm)
        if self.model == 'adain':
            return nets.adain_decoder(self.P['decoder'], zc, zs, c, **kw)
        if self.concat:
            return nets.decoder_concat(self.P['decoder'], zc, zs, c, **kw)
        return nets.decoder(self.P['decoder'], zc, zs, c, **kw)

    def D(self, which, x):
        """-> list of (real/fake map, class logits), one entry per scale"""
        if self.ms_dis:
            return nets.ms_discriminator(self.P[which], x)
        return [nets.discriminator(self.P[which], x, norm=self.dis_norm)]

    # -- phases (adain_model.py:146-180, 225-239, 241-394) ----------------------------------
    def update_discriminator(self, img, c_org, draw):
        B = self.B
        with torch.no_grad():
            cls_a, cls_b = torch.split(c_org, B)
            z_ca, z_cb = torch.split(self.Ec(img, draw), B)
            z_s, _, _ = self.Es(img, c_org, draw)
            z_sa, z_sb = torch.split(z_s, B)
            z_sr = draw((B, self.latent))
            img_ba, img_br = torch.split(self.G(torch.cat((z_cb, z_cb)), torch.cat((z_sa, z_sr)),
                                                torch.cat((cls_a, cls_a)), draw), B)
            img_ab, img_ar = torch.split(self.G(torch.cat((z_ca, z_ca)), torch.cat((z_sb, z_sr)),
                                                torch.cat((cls_b, cls_b)), draw), B)
            img_fake = torch.cat((img_ba, img_ab))
            img_random = torch.cat((img_br, img_ar))
        for which, fake in (('discriminator1', img_fake), ('discriminator2', img_random)):
            self.opt[which].zero_grad()
            d_adv, d_cls = 0, 0
            for (pred_fake, _), (pred_real, pred_real_cls) in zip(self.D(which, fake), self.D(which, img)):
                if self.ms_dis:        # backward_multi_scale_discriminator (:182-199): GANLoss only
                    d_adv = d_adv + self.gan(pred_fake, 0) + self.gan(pred_real, 1)
                else:
                    d_adv = d_adv + self.d_adv(pred_fake, pred_real)
                d_cls = d_cls + bce_logits(pred_real_cls, c_org)
            d = d_adv + self.lcls * d_cls
            d.backward()
            self.loss.update(d_adv=d_adv.item(), d_cls=d_cls.item(), d_total=d.item())
            self.opt[which].step()

    def backward_generator(self, img, c_org, draw):
        B = self.B
        cls_a, cls_b = torch.split(c_org, B)
        z_c = self.Ec(img, draw)
        z_ca, z_cb = torch.split(z_c, B)
        z_s, mu, logvar = self.Es(img, c_org, draw)
        z_sa, z_sb = torch.split(z_s, B)
        img_ba, img_aa = torch.split(self.G(torch.cat((z_cb, z_ca)), torch.cat((z_sa, z_sa)),
                                            torch.cat((cls_a, cls_a)), draw), B)
        img_ab, img_bb = torch.split(self.G(torch.cat((z_ca, z_cb)), torch.cat((z_sb, z_sb)),
                                            torch.cat((cls_b, cls_b)), draw), B)
        img_fake = torch.cat((img_ba, img_ab))
        img_self = torch.cat((img_aa, img_bb))
        z_c_rec_b, z_c_rec_a = torch.split(self.Ec(img_fake, draw), B)
        z_s_rec, _, _ = self.Es(img_fake, c_org, draw)
        z_s_rec_a, z_s_rec_b = torch.split(z_s_rec, B)
        img_recon = self.G(torch.cat((z_c_rec_a, z_c_rec_b)), torch.cat((z_s_rec_a, z_s_rec_b)),
                           torch.cat((cls_a, cls_b)), draw)
        g_content = None
        if self.use_dis_content:
            g_content = bce_logits(nets.content_discriminator(self.P['content_discriminator'], z_c), 1 - c_org)
        g_adv, g_cls = self.g_adv('discriminator1', img_fake, c_org, 'discriminator1', img)
        g_self = F.l1_loss(img_self, img) * self.lrec
        g_cc = F.l1_loss(img_recon, img) * self.lrec
        kl_zc = torch.mean(z_c ** 2) * 0.01
        if self.reparam:
            kl_zs = torch.sum(1 + logvar - mu ** 2 - logvar.exp()) * (-0.5) * 0.01
        else:
            kl_zs = torch.mean(z_s ** 2) * 0.01
        total = g_adv + g_cls + g_self + g_cc + kl_zc + kl_zs
        if g_content is not None:
            total = total + g_content
            self.loss.update(g_content=g_content.item())
        total.backward()
        self.loss.update(g_adv=g_adv.item(), g_cls=g_cls.item(), kl_zc=kl_zc.item(), kl_zs=kl_zs.item(),
                         l1_self_rec=g_self.item(), l1_cc_rec=g_cc.item(), total_g=total.item())

    def backward_decoder_random(self, img, c_org, draw):
        B = self.B
        cls_a, cls_b = torch.split(c_org, B)
        z_ca, z_cb = torch.split(self.Ec(img, draw), B)
        z_sr = draw((B, self.latent))
        img_random = torch.cat((self.G(z_cb, z_sr, cls_a, draw), self.G(z_ca, z_sr, cls_b, draw)))
        # fakes: discriminator1 under ms_dis (:350) and under use_ragan (:362, reals from
        # discriminator2), else discriminator2
        which = 'discriminator1' if (self.ms_dis or self.use_ragan) else 'discriminator2'
        g_adv2, g_cls2 = self.g_adv(which, img_random, c_org, 'discriminator2', img)
        z_rec, mu2, _ = self.Es(img_random, c_org, draw)
        if self.reparam:
            mu2_a, mu2_b = torch.split(mu2, B)
            z_l1 = (F.l1_loss(mu2_a, z_sr) + F.l1_loss(mu2_b, z_sr)) * 10
        else:                      # base_model.py:417-420: z_sr_a is used twice
            z_a, _ = torch.split(z_rec, B)
            z_l1 = (F.l1_loss(z_a, z_sr) + F.l1_loss(z_a, z_sr)) * 10
        (z_l1 + g_adv2 + g_cls2).backward()
        self.loss.update(l1_recon_z=z_l1.item(), gan2=g_adv2.item(), gan2_cls=g_cls2.item())

    def update_content_discriminator(self, img, c_org, draw):
        """adain_model.py:136-144."""
        z_c = self.Ec(img, draw)
        opt = self.opt['content_discriminator']
        opt.zero_grad()
        loss = bce_logits(nets.content_discriminator(self.P['content_discriminator'], z_c.detach()), c_org)
        loss.backward()
        self.loss_dc = loss.item()
        torch.nn.utils.clip_grad_norm_(list(self.P['content_discriminator'].values()), 5)
        opt.step()

    def optimize_parameters(self, img, c_org, draw):
        """adain_model.py:421-430."""
        it, self.it = self.it, self.it + 1
        if self.use_dis_content and it % self.d_iter != 0:
            self.update_content_discriminator(img, c_org, draw)
            return dict(self.loss)
        self.update_discriminator(img, c_org, draw)
        for n in ('content_encoder', 'style_encoder', 'decoder'):
            self.opt[n].zero_grad()
        # D1 receives (unused) gradients in this phase; the reference zeroes them at :174
        self.backward_generator(img, c_org, draw)
        for n in ('content_encoder', 'style_encoder', 'decoder'):
            self.opt[n].step()
        for n in ('content_encoder', 'decoder'):
            self.opt[n].zero_grad()
        self.backward_decoder_random(img, c_org, draw)
        for n in ('content_encoder', 'decoder'):
            self.opt[n].step()
        return dict(self.loss)

    # -- sampling (adain_model.py:96-109) ------------------------------------------------------
    @torch.no_grad()
    def forward_reference(self, img_src, img_ref, c_trg, draw):
        z_c = self.Ec(img_src, draw)
        z_s, _, _ = self.Es(img_ref, c_trg, draw)
        return self.G(z_c, z_s, c_trg)

    @torch.no_grad()
    def forward_random(self, img, z_r, c_trg, draw):
        return self.G(self.Ec(img, draw), z_r, c_trg)
