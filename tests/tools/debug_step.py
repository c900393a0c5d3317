import sys, os, json
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, ROOT)
import torch
import torch.nn.functional as F
from tests import model_cases as mc
from oracle import synth, nets as onets
from masterthesis_b200.models.core.blocks import from_padded, to_padded

torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False
DEV='cuda'
what = sys.argv[1] if len(sys.argv) > 1 else 'step'
if what == 'step':
    meta, gold = mc.load_gold('step_dim8_64')
    seed, B, nd, size, dim = meta['seed'], meta['B'], meta['nd'], meta['size'], meta['dim']
    m, params = mc.make_model(dim, size, B, nd, seed, meta['shapes'])
    m.rng = synth.Draws(seed)
    for n in m.optimizer:
        opt = m.optimizer[n]; orig = opt.step
        def rec(orig=orig, n=n):
            bad = [k for k, p in m.model[n].named_parameters() if p.grad is not None and not torch.isfinite(p.grad).all()]
            print('STEP', n, 'nonfinite grads:', bad[:6], len(bad))
            return orig()
        opt.step = rec
    m.set_inputs(mc.batch(B, size, nd, seed))
    m.optimize_parameters(0)
    for k in m.loss:
        print(k, m.loss[k], float(gold[f'loss.0.{k}']))
else:
    # decoder layer-by-layer vs oracle at dim=64
    dim, size, N, nd, seed = 64, 256, 1, 4, 5
    meta = {'dim': dim, 'nd': nd, 'latent': 8, 'seed': seed}
    from masterthesis_b200.models.core import networks
    G = networks.AdaINDecoder(3, dim*4, num_domains=nd, latent_dim=8)
    shapes = {k: list(v.shape) for k, v in G.state_dict().items()}
    P = synth.synth_params(shapes, seed, 'decoder.')
    G.load_state_dict(P); G.to(DEV)
    P = {k: v.to(DEV) for k, v in P.items()}
    x = synth.synth_tensor('zc', (N, 256, 64, 64), seed).to(DEV) * 30   # unit-variance content code
    x = x.to(torch.bfloat16).float()
    z = synth.synth_tensor('z', (N, 8), seed).to(DEV) * 10
    c = synth.synth_onehot(N, nd, seed).to(DEV)
    print('x std', x.std().item())
    with torch.no_grad():
        s = G.linear(torch.cat([c, z], 1))
        out = to_padded(x, 1)
        ref = x
        for i, dec in enumerate(G.dec1):
            out = dec.run(out, s, out_pad=0 if i == 3 else 1)
            ref = onets.adain_resnet_block(P, f'dec1.{i}', ref, s)
            mine = from_padded(out, 256, 0 if i == 3 else 1)
            print('block', i, 'rel', mc.rel(mine, ref), 'ref std', ref.std().item())
        for i in range(2):
            out = G.dec2[i].run(out)
            ref = F.conv_transpose2d(ref, P[f'dec2.{i}.block.0.weight'], P[f'dec2.{i}.block.0.bias'], stride=2, padding=1, output_padding=1)
            ref = F.relu(onets.layer_norm(P, f'dec2.{i}.block.1', ref))
            print('up', i, 'rel', mc.rel(from_padded(out, ref.shape[1]), ref), 'ref std', ref.std().item())
        img = G.dec2[2].run(out)
        pre = F.conv_transpose2d(ref, P['dec2.2.block.0.weight'])
        print('pre-tanh std', pre.std().item(), 'absmax', pre.abs().max().item())
        print('img rel', mc.rel(img, torch.tanh(pre)))
