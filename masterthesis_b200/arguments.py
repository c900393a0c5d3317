"""The reference's command-line defaults as a plain Namespace (reference src/arguments.py:20-46,
85-118): what `train.py` / `sample.py` would hand to `models.AdaINModel(args)` when no flag is given.
Used by bench.py and the tools; the models read the same attribute names as the reference's."""
import argparse


def default_args(**over):
    d = dict(
        input_dim=3, output_dim=3, dim=64, init_type="normal", init_gain=0.02, use_dropout=False,
        num_domains=4, mode="train", concat=False, reparam=False, use_dis_content=False,
        latent_dim=8, up_type="transpose", dec_norm="layer", enc_norm="instance", batch_size=1,
        crop_size=256, resume=None, resume_opt=None, gpu_ids=[], dis_norm=None, norm_feat=False,
        lr=1e-4, wd=1e-4, beta1=0.5, beta2=0.999, lr_policy="step", n_iters=1000000, last_iter=-1,
        n_iter_decay=600000, d_iter=3, lambda_rec=10.0, lambda_cls=1.0, lambda_cls_G=5.0,
        lambda_style=5.0, gan_mode="vanilla", ms_dis=False, dis_sn=False, num_scales=3,
        use_ragan=False, lambda_perceptual=1.0, vgg_type="vgg19", vgg_loss=None,
        vgg_layers=["conv5_4"], layer_weights=[1.0], logdir="/tmp/mt_logs",
        checkpoint_dir="/tmp/mt_ckpt", display_dir="/tmp/mt_disp",
    )
    d.update(over)
    return argparse.Namespace(**d)
