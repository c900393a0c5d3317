"""TEST INFRASTRUCTURE — drive the REFERENCE's own train.py / sample.py (oracle/_ref/src, unmodified) over this
repo's CUDA models, and exchange checkpoints with the reference's own modules.  Run as a subprocess by
tests/test_dropin_gpu.py (module aliasing is process-global); prints one `RESULT {json}` line.

    python tests/tools/dropin_driver.py train      # Trainer.run: dataset -> DataLoader -> our AdaINModel, 4 iterations
    python tests/tools/dropin_driver.py sample     # Sampler.sample_batch incl. the canonical 540x960 size
    python tests/tools/dropin_driver.py ckpt       # model_{it}.ckpt / opt_{it}.ckpt interchange, both directions

Reference call sites: src/train.py:23-59 (Trainer), src/sample.py:57-91 (Sampler.load_model / sample_batch),
src/arguments.py:56-57 (class resolution through the `models` module), src/models/model.py:70-100 (save / load).
"""
import json
import os
import sys
import tempfile
import types

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from oracle import ref_import  # noqa: E402


def alias_models():
    """what INTEGRATION.md section 1 describes: `import models` resolves to this repo's package"""
    import masterthesis_b200.models as mt_models
    sys.modules['models'] = mt_models
    for sub in ('core', 'core.networks', 'core.blocks', 'core.functions', 'core.norm', 'core.misc'):
        mod = __import__('masterthesis_b200.models.' + sub, fromlist=['x'])
        sys.modules['models.' + sub] = mod
    # modules the reference imports that are absent here and never touch arithmetic
    for name in ('matplotlib', 'matplotlib.cm', 'videoreaders'):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.modules['matplotlib'].cm = sys.modules['matplotlib.cm']
    sys.modules['videoreaders'].FrameWriter = sys.modules['videoreaders'].SVOReader = object
    if ref_import.REF_SRC not in sys.path:
        sys.path.insert(0, ref_import.REF_SRC)


def unalias_models():
    for k in list(sys.modules):
        if k == 'models' or k.startswith('models.'):
            del sys.modules[k]


def make_dataset(root, domains=4, per_domain=4, size=72):
    """dataroot/train/<domain>/*.png as PairedDataset expects (reference dataset.py:100-139); the number of
    images per domain is a multiple of the batch size: the reference (and therefore this path) splits the
    stacked batch by args.batch_size (adain_model.py:113-129) and cannot take a partial last batch"""
    from PIL import Image
    rng = np.random.RandomState(0)
    for d in range(domains):
        dd = os.path.join(root, 'train', f'domain{d}')
        os.makedirs(dd, exist_ok=True)
        for i in range(per_domain):
            Image.fromarray(rng.randint(0, 255, (size, size, 3), dtype=np.uint8)).save(os.path.join(dd, f'img{i}.png'))


def run_train():
    alias_models()
    import train as ref_train                       # the reference's own driver
    from arguments import TrainArguments
    import masterthesis_b200.models as mt_models
    tmp = tempfile.mkdtemp(prefix='mt_dropin_')
    make_dataset(tmp)
    sys.argv = ['train.py', '--dataroot', tmp, '--model', 'AdaINModel', '--num_domains', '4', '--batch_size', '2',
                '--load_size', '72', '--crop_size', '64', '--dim', '8', '--exp_dir', tmp, '--name', 'run',
                '--gpu_ids', '0', '--num_workers', '0', '--max_iter', '3', '--print_freq', '1', '--save_freq', '2',
                '--display_freq', '3']
    args = TrainArguments().parse()
    assert args.model is mt_models.AdaINModel, args.model     # resolved by NAME through the aliased module
    trainer = ref_train.Trainer()
    loader = trainer.load_dataset(args)
    model = trainer.create_model(args)
    trainer.train(args, model, loader)
    ck = sorted(os.listdir(args.checkpoint_dir))
    imgs = sorted(os.listdir(args.display_dir))
    finite = all(v == v and abs(v) < 1e6 for v in model.loss.values())
    return {'checkpoints': ck, 'images': imgs, 'loss_keys': sorted(model.loss), 'finite': finite,
            'lr': model.get_current_lr()['decoder'], 'device': str(model.device)}


def run_sample():
    alias_models()
    import sample as ref_sample                     # the reference's own driver
    import masterthesis_b200.models as mt_models
    from masterthesis_b200.arguments import default_args
    out = {}
    for (H, W), B in (((256, 256), 2), ((540, 960), 1)):       # 540x960: sample.py:50, scripts/sample.sh:2
        args = default_args(dim=64 if H == 540 else 16, crop_size=256, batch_size=B, num_domains=4, mode='test',
                            init_type='normal')
        args.model = mt_models.AdaINModel
        torch.manual_seed(0)
        model = args.model(args)
        model.initialize()
        sampler = ref_sample.Sampler.__new__(ref_sample.Sampler)     # __init__ only builds PIL transforms
        batch = (torch.rand(B, 3, H, W) * 2 - 1).to(model.device)
        z = model.get_z_random(B, args.latent_dim)
        # reference-image path of sample_batch loads a file: save one and hand over its name
        from PIL import Image
        from torchvision import transforms
        sampler.transforms = transforms.Compose([transforms.Resize((H, W)), transforms.ToTensor(),
                                                 transforms.Normalize([0.5] * 3, [0.5] * 3)])
        ref_path = os.path.join(tempfile.mkdtemp(prefix='mt_dropin_'), 'ref.png')
        Image.fromarray(np.random.RandomState(1).randint(0, 255, (80, 90, 3), dtype=np.uint8)).save(ref_path)
        for call in range(4):                       # eager warm-ups, graph capture, replay
            y1, secs, gib = sampler.sample_batch(args, model, batch, 2, z_sr=z, device=model.device)
        y2, _, _ = sampler.sample_batch(args, model, batch, 1, ref=ref_path, device=model.device)
        key = f'{H}x{W}'
        out[key + '.shape'] = list(y1.shape)
        out[key + '.finite'] = bool(torch.isfinite(y1).all() and torch.isfinite(y2).all())
        out[key + '.range_ok'] = bool(y1.abs().max() <= 1.0 and y2.abs().max() <= 1.0)
        out[key + '.types'] = [type(secs).__name__, type(gib).__name__]
        out[key + '.graph_captured'] = bool(getattr(model, '_sample_graphs', None))
    return out


def _rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-30))


def run_ckpt():
    """ours -> reference and reference -> ours, model and optimizer state, with and without `module.`"""
    from masterthesis_b200 import models as mt_models
    from masterthesis_b200.arguments import default_args
    tmp = tempfile.mkdtemp(prefix='mt_ckpt_')
    kw = dict(dim=8, crop_size=64, batch_size=1, num_domains=4, init_type='normal', checkpoint_dir=tmp)
    torch.manual_seed(0)
    ours = mt_models.AdaINModel(default_args(**kw))
    ours.initialize()
    g = torch.Generator().manual_seed(3)
    b = {'x1': torch.rand(1, 3, 64, 64, generator=g) * 2 - 1, 'x2': torch.rand(1, 3, 64, 64, generator=g) * 2 - 1,
         'y1': torch.eye(4)[[0]], 'y2': torch.eye(4)[[2]]}
    for it in range(2):
        ours.set_inputs(b)
        ours.optimize_parameters(it)
    ours.save(5)
    out = {}
    # ---- the reference's own modules load what we saved (strict key / shape match, torch.optim.Adam state)
    ref = ref_import.import_reference()
    rm = ref.AdaINModel(ref_import.default_args(**kw))
    rm.device = 'cpu'
    rm.initialize()
    rm.load(os.path.join(tmp, 'model_5.ckpt'), os.path.join(tmp, 'opt_5.ckpt'))
    out['ours_to_ref.params'] = max(_rel(dict(rm.model[n].state_dict())[k], v.cpu())
                                    for n in ours.model for k, v in ours.model[n].state_dict().items())
    st_r = rm.optimizer.decoder.state_dict()['state']
    st_o = ours.optimizer.decoder.state_dict()['state']
    out['ours_to_ref.adam_state'] = max(max(_rel(st_r[i]['exp_avg'], st_o[i]['exp_avg'].cpu()),
                                            _rel(st_r[i]['exp_avg_sq'], st_o[i]['exp_avg_sq'].cpu())) for i in st_o)
    out['ours_to_ref.adam_step'] = float(st_r[0]['step'])
    # one more optimizer step on both sides from the same gradients: the loaded state must continue identically
    torch.manual_seed(9)
    grads = [torch.randn(p.shape) * 0.01 for p in rm.model.decoder.parameters()]
    before = [p.detach().clone() for p in rm.model.decoder.parameters()]
    for p, gr in zip(rm.model.decoder.parameters(), grads):
        p.grad = gr.clone()
    rm.optimizer.decoder.step()
    ours.optimizer.decoder.zero_grad()
    for p, gr in zip(ours.model.decoder.parameters(), grads):
        p.grad.copy_(gr)
    ours.optimizer.decoder.step()
    torch.cuda.synchronize()
    # max |delta| relative to the size of the update itself
    upd = max(float((a.detach() - b0).abs().max()) for a, b0 in zip(rm.model.decoder.parameters(), before))
    dif = max(float((a.detach() - c.detach().cpu()).abs().max()) for a, c in
              zip(rm.model.decoder.parameters(), ours.model.decoder.parameters()))
    out['step_after_load.rel_to_update'] = dif / max(upd, 1e-30)
    worst = max(zip(rm.model.decoder.named_parameters(), ours.model.decoder.parameters(), before),
                key=lambda t: float((t[0][1].detach() - t[1].detach().cpu()).abs().max()))
    out['step_after_load.debug'] = [worst[0][0], float((worst[0][1].detach() - worst[2]).abs().max()),
                                    float((worst[1].detach().cpu() - worst[2]).abs().max()),
                                    str(ours.optimizer.decoder.param_groups[0]['lr']), str(rm.optimizer.decoder.param_groups[0]['lr'])]
    # ---- the reference saves (model.py:70-82); we load it, plain and with DataParallel's `module.` prefix
    rm.args.checkpoint_dir = tmp
    rm.save(7)
    mpath, opath = os.path.join(tmp, 'model_7.ckpt'), os.path.join(tmp, 'opt_7.ckpt')
    ck = torch.load(mpath)
    torch.save({n: {'module.' + k: v for k, v in sd.items()} for n, sd in ck.items()}, os.path.join(tmp, 'model_8.ckpt'))
    for tag, path in (('plain', mpath), ('module_prefix', os.path.join(tmp, 'model_8.ckpt'))):
        torch.manual_seed(1)
        m2 = mt_models.AdaINModel(default_args(**kw))
        m2.initialize()
        m2.load(path, opath)
        out[f'ref_to_ours.{tag}.params'] = max(_rel(v.cpu(), ck[n][k]) for n in m2.model
                                               for k, v in m2.model[n].state_dict().items())
        m2.set_inputs(b)
        m2.optimize_parameters(0)                  # runs from the loaded optimizer state
        out[f'ref_to_ours.{tag}.trains'] = all(v == v for v in m2.loss.values())
        out[f'ref_to_ours.{tag}.adam_step'] = float(m2.optimizer.decoder.state_dict()['state'][0]['step'])
    # ---- resume mid-run: capture the iteration graph, load a checkpoint (re-allocates the optimizer buckets), go on
    m3 = ours
    for it in range(4):
        m3.set_inputs(b)
        m3.optimize_parameters(it)
    captured = m3._graph is not None
    m3.load(mpath, opath)
    w0 = torch.cat([p.detach().flatten() for p in m3.model.decoder.parameters()]).clone()
    for it in range(4):
        m3.set_inputs(b)
        m3.optimize_parameters(it)
    torch.cuda.synchronize()
    w1 = torch.cat([p.detach().flatten() for p in m3.model.decoder.parameters()])
    out['load_after_capture.captured_before'] = captured
    out['load_after_capture.recaptured'] = m3._graph is not None
    out['load_after_capture.weights_moved'] = bool((w0 != w1).any())
    out['load_after_capture.finite'] = all(v == v for v in m3.loss.values())
    return out


if __name__ == '__main__':
    res = {'train': run_train, 'sample': run_sample, 'ckpt': run_ckpt}[sys.argv[1]]()
    print('RESULT ' + json.dumps(res))
