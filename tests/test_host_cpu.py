"""Host-side logic of the optional variants that needs no GPU: the spectral-norm state / power
iteration of the conv parameter holders against torch.nn.utils.spectral_norm (what the reference
wraps its convs in, functions.py:113-121), and the replayable dropout masks."""
import torch
import torch.nn as nn

from masterthesis_b200.models.core import blocks
from masterthesis_b200.models.core.functions import spectral_norm
from oracle import synth


def _pair(transposed):
    torch.manual_seed(0)
    if transposed:
        ref = nn.ConvTranspose2d(5, 7, 3, 2, 1, 1)
        mine = blocks.ConvTranspose2d(5, 7, 3, 2, 1, 1)
    else:
        ref = nn.Conv2d(5, 7, 3, 1, 1)
        mine = blocks.Conv2d(5, 7, 3, 1, 1)
    ref = torch.nn.utils.spectral_norm(ref, 'weight', 1, 1e-12, 1 if transposed else 0)
    mine = spectral_norm(mine)
    assert list(mine.state_dict().keys()) == list(ref.state_dict().keys())
    mine.load_state_dict(ref.state_dict())
    return ref, mine


def test_spectral_norm_matches_torch():
    for transposed in (False, True):
        ref, mine = _pair(transposed)
        x = torch.randn(2, 5, 8, 8)
        proj = torch.randn_like(ref.weight_orig)
        for it in range(3):                      # three training-mode forwards = three power iterations
            ref(x)                               # the pre-forward hook recomputes ref.weight
            w_ref = ref.weight
            w = mine.sn_weight()
            assert torch.allclose(w, w_ref, rtol=1e-5, atol=1e-7), (transposed, it)
            assert torch.allclose(mine.weight_u, ref.weight_u, atol=1e-6)
            assert torch.allclose(mine.weight_v, ref.weight_v, atol=1e-6)
        # gradient through sigma
        g_ref, = torch.autograd.grad((w_ref * proj).sum(), ref.weight_orig)
        g, = torch.autograd.grad((w * proj).sum(), mine.weight_orig)
        assert torch.allclose(g, g_ref, rtol=1e-4, atol=1e-6)
        # eval mode: no power iteration, buffers untouched
        mine.eval()
        u0 = mine.weight_u.clone()
        mine.sn_weight()
        assert torch.equal(u0, mine.weight_u)


def test_spectral_norm_changes_the_discriminator_state_dict_like_the_reference():
    from masterthesis_b200.models.core import networks
    d = networks.Discriminator(3, 8, num_domains=4, sn=True, image_size=64)
    keys = list(d.state_dict().keys())
    assert keys[:4] == ['model.0.block.1.bias', 'model.0.block.1.weight_orig', 'model.0.block.1.weight_u',
                        'model.0.block.1.weight_v']
    assert 'conv1.weight' in keys and 'conv2.weight' in keys          # the heads are not normalised
    assert sum(p.numel() for p in d.parameters()) == sum(
        p.numel() for p in networks.Discriminator(3, 8, num_domains=4, image_size=64).parameters())


def test_dropout_mask_replays_the_reference_rule():
    draws_a, draws_b = synth.Draws(7), synth.Draws(7)
    ctx = blocks.HostContext()                                         # per-model host state: not a module global
    blocks.set_dropout_rng(draws_a, ctx)
    x = torch.zeros(2, 4, 6, 16, dtype=torch.bfloat16)                 # NHWC, 12 logical of 16 stored channels
    m = blocks.dropout_mask(x, 12, 0.5, ctx.dropout_rng)
    assert blocks.context_of(torch.nn.Dropout(0.5)).dropout_rng is None    # other models are not affected
    want = (draws_b((2, 12, 4, 6)) > 0).permute(0, 2, 3, 1)
    assert m.dtype == torch.bfloat16 and m.shape == x.shape
    assert torch.equal(m[..., :12].float(), want.float() * 2.0)
    assert torch.equal(m[..., 12:].float(), torch.zeros(2, 4, 6, 4))   # storage padding stays zero
    assert draws_a.log == [(2, 12, 4, 6)]


def test_pixelshuffle_is_rejected_with_the_reason():
    import pytest
    with pytest.raises(NotImplementedError, match="cannot run in the reference"):
        blocks.UpsampleBlock(16, 8, 3, 2, 1, 1, norm_layer='layer', activation='relu', up_type='pixelshuffle')
