"""Drop-in boundary end to end (GPU only; needs oracle/_ref, the reference's unmodified sources copied by
`python -m oracle.make_ref`): the reference's own train.py / sample.py drive this repo's models through
`import models`, and checkpoints travel between the reference's modules and ours in both directions."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(what):
    from oracle import ref_import
    if not ref_import.available():
        pytest.skip("reference sources not available (oracle/_ref absent: python -m oracle.make_ref)")
    p = subprocess.run([sys.executable, os.path.join(ROOT, 'tests', 'tools', 'dropin_driver.py'), what],
                       capture_output=True, text=True, timeout=900, cwd=ROOT)
    lines = [l for l in p.stdout.splitlines() if l.startswith('RESULT ')]
    assert lines, f"driver failed (rc={p.returncode}):\n{p.stdout[-3000:]}\n{p.stderr[-3000:]}"
    return json.loads(lines[0][7:])


def test_reference_trainer_drives_the_cuda_models():
    """reference src/train.py: TrainArguments().parse() resolves --model AdaINModel to OUR class by name,
    Trainer.load_dataset / create_model / train run 4 iterations (update_lr, set_inputs, optimize_parameters,
    write_loss, print_losses, save, save_images) on a synthetic PairedDataset folder."""
    r = _run('train')
    assert r['finite'] and r['device'] == 'cuda', r
    assert {'model_0.ckpt', 'model_2.ckpt', 'model_4.ckpt', 'opt_0.ckpt', 'opt_2.ckpt', 'opt_4.ckpt'} <= set(r['checkpoints']), r
    assert 'gen_0.jpg' in r['images'] and 'gen_3.jpg' in r['images'], r
    assert set(r['loss_keys']) == {'d_adv', 'd_cls', 'd_total', 'g_adv', 'g_cls', 'kl_zc', 'kl_zs', 'l1_self_rec',
                                   'l1_cc_rec', 'total_g', 'l1_recon_z', 'gan2', 'gan2_cls'}, r


def test_reference_sampler_drives_the_cuda_models():
    """reference src/sample.py: Sampler.sample_batch (forward_random and forward_reference) at 256x256 and at
    the reference's canonical 540x960 (odd feature maps: 135 and 67 wide) without any environment switch."""
    r = _run('sample')
    for key, shape in (('256x256', [2, 3, 256, 256]), ('540x960', [1, 3, 540, 960])):
        assert r[key + '.shape'] == shape and r[key + '.finite'] and r[key + '.range_ok'], r
        assert r[key + '.types'] == ['float', 'float'] and r[key + '.graph_captured'], r


def test_checkpoint_interchange_with_the_reference():
    """model_{it}.ckpt / opt_{it}.ckpt written by us load into the reference's modules + torch.optim.Adam
    (strict keys, identical tensors, the next Adam step agrees) and the reference's files load into ours,
    with and without DataParallel's `module.` prefix; loading after the iteration graph was captured
    re-captures it (the optimizer buckets move)."""
    r = _run('ckpt')
    assert r['ours_to_ref.params'] == 0.0 and r['ours_to_ref.adam_state'] == 0.0 and r['ours_to_ref.adam_step'] == 4.0, r
    assert r['step_after_load.rel_to_update'] <= 1e-3, r
    for tag in ('plain', 'module_prefix'):
        assert r[f'ref_to_ours.{tag}.params'] == 0.0 and r[f'ref_to_ours.{tag}.trains'], r
        assert r[f'ref_to_ours.{tag}.adam_step'] == 7.0, r
    assert r['load_after_capture.captured_before'] and r['load_after_capture.recaptured'], r
    assert r['load_after_capture.weights_moved'] and r['load_after_capture.finite'], r


def test_data_parallel_gradients_equal_single_gpu():
    """SURVEY.md 8e on hardware (skipped below 2 GPUs): one iteration on 2 ranks (NCCL all-reduce of the flat
    buckets, KL pre-scale, 1/world in Adam) vs the same global batch on one GPU: gradient buckets at all seven
    optimizer steps: the five steps taken from identical weights (both discriminators, content encoder, style
    encoder, decoder: steps 0-4) cosine >= 0.9999 and relative norm difference <= 1e-5 (measured <= 1e-7: only the
    fp32 summation order differs between a 2-image and a 4-image launch); the two steps AFTER the first Adam
    updates (5, 6) cosine >= 0.99: Adam's first step is lr * g / (|g| + eps), i.e. sign-like, so last-bit
    differences in near-zero gradient elements become +-lr differences in the weights those steps start from
    (measured 0.9994 / 0.9998); losses within 1e-3 (measured <= 1e-7)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2; log committed as profiles/r2_dp_equality_2gpu.txt)")
    p = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2',
                        '--master-addr', '127.0.0.1', '--master-port', '29613',
                        os.path.join(ROOT, 'tests', 'tools', 'dp_equality.py'), '16', '64', '4'],
                       capture_output=True, text=True, timeout=900, cwd=ROOT)
    lines = [l for l in p.stdout.splitlines() if l.startswith('RESULT ')]
    assert lines, f"rc={p.returncode}\n{p.stdout[-2000:]}\n{p.stderr[-3000:]}"
    r = json.loads(lines[0][7:])
    for k, v in r.items():
        if k.startswith('cos_'):
            late = k.startswith('cos_step5') or k.startswith('cos_step6')
            assert v >= (0.99 if late else 0.9999), (k, v, r)
        elif k.startswith('relnorm_') and not (k.startswith('relnorm_step5') or k.startswith('relnorm_step6')):
            assert v <= 1e-5, (k, v, r)
        elif k.startswith('loss.'):
            assert v <= 1e-3, (k, v, r)
