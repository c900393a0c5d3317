"""BaseModel (reference src/models/base_model.py:9-470): style injected by concatenation
(`Decoder` by default, `DecoderConcat` with --concat), style code from the plain `StyleEncoder`
(default) or the `ReparameterizedStyleEncoder` (--reparam).

The training schedule, the losses and the sampling forwards are line for line those of AdaINModel
(reference: `diff base_model.py adain_model.py` only shows the `if self.args.reparam` branches), so
this class only swaps the generator networks; the non-reparam branches (mean-square style
regulariser :346-347, the `z_sr_a`-twice latent regression :417-420) live in AdaINModel behind
`self.reparam`.
"""
from .adain_model import AdaINModel
from .core import networks


class BaseModel(AdaINModel):
    def __init__(self, args):
        self.reparam = bool(args.reparam)
        super().__init__(args)

    def _build_generator(self, args):
        m = self.model
        if args.reparam:
            m.style_encoder = networks.ReparameterizedStyleEncoder(
                args.input_dim, output_dim=self.latent_dim, dim=args.dim, num_domains=args.num_domains,
                norm_layer=None, activation='lrelu')
        else:
            m.style_encoder = networks.StyleEncoder(args.input_dim, output_dim=self.latent_dim, dim=args.dim,
                                                    num_domains=args.num_domains, activation='lrelu')
        if args.concat:
            m.decoder = networks.DecoderConcat(args.input_dim, dim=m.content_encoder.output_dim,
                                               num_domains=args.num_domains, latent_dim=self.latent_dim,
                                               up_type=args.up_type, norm_layer=args.dec_norm,
                                               dropout=args.use_dropout)
        else:
            m.decoder = networks.Decoder(args.input_dim, dim=m.content_encoder.output_dim,
                                         latent_dim=self.latent_dim, num_domains=args.num_domains,
                                         up_type=args.up_type, norm_layer=args.dec_norm,
                                         dropout=args.use_dropout)
