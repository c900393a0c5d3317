"""Network / training-step / sampling parity of the CUDA path (GPU only).
Oracles: the reference's recorded outputs (tests/golden) and oracle/ run in fp32."""
import pytest

pytestmark = pytest.mark.gpu


def _check(res, rel_tol, cos_tol, rel_over=None):
    bad = {}
    for k, v in res.items():
        if k.startswith('emu'):
            continue
        if k.startswith('cos'):
            if not v >= cos_tol:
                bad[k] = v
        elif not v <= (rel_over or {}).get(k, rel_tol):
            bad[k] = v
    assert not bad, f"out of tolerance (rel<={rel_tol}, cos>={cos_tol}): {bad}\nall: {res}"


# Tolerances.  The CUDA path stores activations and GEMM weights in bf16 (fp32 accumulate, fp32
# statistics / losses / master weights); the reference is fp32 end to end.  Measured on B200:
#   single network, outputs  : <= 1.3e-2 of the output range (3.5e-2 for the tanh image at dim=16)
#   single network, gradients: cosine >= 0.988
# and the bf16-storage-emulated oracle (tests/emu.py) shows the same distance from fp32, i.e. the
# residual is the storage format, not the kernels.
def test_nets_small_vs_reference_golden():
    from tests.model_cases import nets_vs_golden
    _check(nets_vs_golden('nets_dim16_64'), 2e-2, 0.985, {'fake': 5e-2})


def test_nets_full_size_vs_reference_golden():
    from tests.model_cases import nets_vs_golden
    _check(nets_vs_golden('nets_dim64_256'), 2e-2, 0.985)


def test_training_step_vs_reference_golden():
    """13 losses x 2 iterations vs the reference's recorded values: <= 1e-2 (measured <= 6e-4).
    Gradients at the five optimizer steps vs the fp32 oracle: the D phases (6 layers) >= 0.99;
    the G phases back-propagate through 30+ bf16 layers (D -> G -> Ec/Es -> G -> Ec), where
    rounding noise is amplified by ReLU-mask flips, so they are required to be as close to fp32
    as the bf16-storage-emulated oracle is (within 0.05), and >= 0.8 in absolute terms."""
    from tests.model_cases import step_vs_golden
    res = step_vs_golden()
    losses = {k: v for k, v in res.items() if k.startswith('loss')}
    _check(losses, 1e-2, 0.0)
    for k, v in res.items():
        if not k.startswith('cos_step'):
            continue
        emu_v = res['emu' + k]
        if 'discriminator' in k:
            assert v >= 0.99, (k, v)
        else:
            assert v >= 0.8 and v >= emu_v - 0.05, (k, v, emu_v)


def test_training_step_at_the_benchmarked_geometry_vs_reference_golden():
    """BASELINE config 3 — dim 64, 256x256, batch 8 (16 / 32 images per call: the persistent conv kernel walks
    3.5 - 7 work items per CTA pair, split-K weight gradients over 16 / 32 images, the stacked 32-image
    decoder calls) — two iterations of optimize_parameters with the reference's replayed draws against the
    values the unmodified REFERENCE recorded on the CPU (tests/golden/step_dim64_256_B8.npz):
      * every loss of the first iteration within 1e-2 relative (measured <= 1.3e-3), of the second within
        1e-2 except kl_zs (a 0.03-sized sum over 128 values one sign-like Adam step later: <= 3e-2, measured 1.2e-2);
      * the gradients handed to each optimizer step vs the reference's recorded samples: discriminators,
        style encoder and the decoder's first step cosine >= 0.999, the decoder's second step >= 0.985
        (measured 0.994 - 0.996);
        the content encoder >= 0.9 AND no worse than the bf16-storage-emulated oracle minus 0.02 — its
        gradient is chaotic in the forward rounding: rounding ONLY the GEMM weights of the fp32 oracle to bf16
        already drops its cosine to 0.95-0.96 and keeping every gradient in fp32 does not help
        (tests/tools/precision_experiment.py);
      * the worst single parameter tensor >= 0.9."""
    from tests.model_cases import step_vs_golden
    res = step_vs_golden('step_dim64_256_B8')
    for k, v in res.items():
        if k.startswith('loss_it0.'):
            assert v <= 1e-2, (k, v)
        elif k.startswith('loss_it1.'):
            assert v <= (3e-2 if k.endswith('kl_zs') else 1e-2), (k, v)
        elif k.startswith('cosrefmin_'):
            assert v >= 0.9, (k, v, res['_worst_tensors'])
        elif k.startswith('cosref_'):
            if 'content_encoder' in k:
                emu_v = res['emucos_' + k[len('cosref_'):]]
                assert v >= 0.9 and v >= emu_v - 0.02, (k, v, emu_v)
            elif k == 'cosref_step6_decoder':
                assert v >= 0.985, (k, v)
            else:
                assert v >= 0.999, (k, v)


def test_iteration_is_reproducible_bit_for_bit_at_the_benchmarked_geometry():
    """dim 64, 256x256, batch 8: two eager runs and the CUDA-graph run of the same three iterations agree in
    every loss and every parameter BIT FOR BIT (fixed-order sums everywhere: no atomics in the statistics,
    the norm backward, the split-K weight gradients or the loss layer; the captured iteration launches the
    same kernels with the same host-computed Adam scalars as the eager one)."""
    from tests.model_cases import determinism
    res = determinism()
    assert all(v == 0.0 for v in res.values()), res


def test_sampling_at_config2_batch_vs_oracle():
    """BASELINE config 2: forward_reference / forward_random at N = 8, 256x256 vs the fp32 oracle, bounded
    like test_sampling_vs_oracle."""
    from tests.model_cases import sampling_vs_oracle
    res = sampling_vs_oracle(N=8)
    for k in ('forward_reference', 'forward_random'):
        assert res[k] <= 1e-1 and res[k] <= max(2e-2, 1.5 * res['emu_' + k]), res


def test_sampling_vs_oracle():
    """forward_reference / forward_random at 256x256 (config 2 shape) vs the fp32 oracle: the chain
    Ec -> Es -> G (25 layers) with variance-preserving synthetic weights; bounded by 1.5x the
    bf16-storage-emulated oracle's own distance from fp32, and 1e-1 absolute."""
    from tests.model_cases import sampling_vs_oracle
    res = sampling_vs_oracle()
    for k in ('forward_reference', 'forward_random'):
        assert res[k] <= 1e-1 and res[k] <= max(2e-2, 1.5 * res['emu_' + k]), res


def _exact(res):
    bad = {k: v for k, v in res.items() if v != 0.0}
    assert not bad, f"graph replay / eager re-run differ: {bad}"


def test_cuda_graph_iteration_matches_eager():
    """6 iterations (2 eager warm-up, capture, replays; learning rate changed every iteration): the
    graph-replayed losses and final weights equal the eager launch sequence exactly."""
    from tests.model_cases import graph_vs_eager
    _exact(graph_vs_eager())


def test_base_nets_vs_reference_golden():
    """StyleEncoder, Decoder, DecoderConcat (odd channel counts 44/52/26/34/17/25 at dim=16),
    ContentDiscriminator, MultiScaleDiscriminator: outputs <= 2e-2 of range (tanh images 5e-2),
    gradients cosine >= 0.985 vs the reference's recorded values."""
    from tests.model_cases import base_nets_vs_golden
    res = base_nets_vs_golden()
    # DecoderConcat at dim=16 has 17..76-channel layers: its tanh image is bounded by 1.5x the
    # bf16-storage-emulated oracle's own distance from fp32 (and 1e-1 absolute)
    assert res['decc.fake'] <= 1e-1 and res['decc.fake'] <= max(5e-2, 1.5 * res['emu_decc.fake']), res
    res.pop('decc.fake')
    _check(res, 2e-2, 0.985, {'dec.fake': 5e-2})


@pytest.mark.parametrize('name', ['step_basec_dim8_64', 'step_basecr_dim8_64', 'step_dcms_dim8_256',
                                  'step_ragan_dim8_64', 'step_hinge_dim8_64', 'step_lsgan_dim8_64',
                                  'step_wgangp_dim8_64', 'step_nearest_dim8_64', 'step_dropout_dim8_64',
                                  'step_sn_dim8_64', 'step_basec_ndsn_dim8_64', 'step_norms_dim8_128'])
def test_flagged_training_steps_vs_reference_golden(name):
    """BaseModel --concat, BaseModel --concat --reparam, AdaINModel --use_dis_content --ms_dis (4
    iterations through the d_iter schedule), --use_ragan, --gan_mode hinge / lsgan / wgangp,
    --up_type nearest, --use_dropout, --dis_sn, BaseModel --concat with the last three together, and
    --enc_norm layer --dec_norm instance --dis_norm instance:
    every loss of every iteration within 1e-2 of the reference's recorded value."""
    from tests.model_cases import flag_step_vs_golden
    res = flag_step_vs_golden(name)
    over = {}
    # the second iteration's kl_zs (0.005-sized, one sign-like Adam step after the first): 3e-2
    over.update({k: 3e-2 for k in res if k.startswith('loss_it') and not k.startswith('loss_it0') and k.endswith('kl_zs')})
    if name in ('step_hinge_dim8_64', 'step_wgangp_dim8_64'):
        # hinge / wgangp adversarial terms are SIGNED MEANS of raw logits (differences of means near zero for
        # wgangp): a relative error is ill-conditioned there; bounded by 6e-2 (absolute errors <= 1e-3)
        over.update({k: 6e-2 for k in res if k.split('.')[-1] in ('d_adv', 'd_total', 'g_adv', 'gan2')})
    if name == 'step_norms_dim8_128':
        # --dis_norm instance: un-affine InstanceNorm on the discriminator's small feature planes makes
        # the class logits (weighted x5 in g_cls / gan2_cls) the most rounding-sensitive outputs: the
        # bf16-storage-emulated ORACLE (tests/emu.py) is itself 0.74e-2 / 0.98e-2 away from the fp32
        # reference on gan2_cls here, so these keys are bounded by 2e-2
        over.update({k: 2e-2 for k in res if k.endswith('_cls')})
        # ... and 5e-2 in the SECOND iteration, which runs on weights after one Adam step (a sign-like
        # step of size lr whatever |grad|, so rounding noise in tiny gradients moves weights as far as
        # signal does): measured 1.7e-2 .. 2.2e-2 over repeated runs (the weight-gradient split-K uses
        # fp32 atomics, so the second iteration is not bit-reproducible run to run)
        over.update({k: 5e-2 for k in res if k.startswith('loss_it1.') and k.endswith('_cls')})
    _check(res, 1e-2, 0.0, over)
    # gradients of the first iteration's optimizer steps vs the reference's recorded samples (dim = 8 toys:
    # the chained generator phases are dominated by bf16 rounding noise there, see the dim-64 test)
    for k, v in res.items():
        if k.startswith('cosref_'):
            assert v >= (0.97 if 'discriminator' in k else 0.7), (k, v)     # D: 0.984 with --dis_norm instance, else >= 0.998


def test_base_default_training_step_vs_oracle():
    """BaseModel default flags (StyleEncoder + Decoder): losses of 2 iterations within 1e-2 of the oracle
    (second-iteration kl_zs = 0.01 * mean(z_s^2), a 1e-4-sized value after one Adam step: 3e-2)."""
    from tests.model_cases import base_default_step_vs_oracle
    res = base_default_step_vs_oracle()
    _check(res, 1e-2, 0.0, {k: 3e-2 for k in res if k.startswith('loss_it1') and k.endswith('kl_zs')})


def test_base_sampling_vs_oracle():
    """BASELINE config 1 (BaseModel generator forward, batch 1, 256x256) on the GPU vs the fp32 oracle,
    bounded like test_sampling_vs_oracle."""
    from tests.model_cases import base_sampling_vs_oracle
    res = base_sampling_vs_oracle()
    for k in ('forward_reference', 'forward_random'):
        assert res[k] <= 1e-1 and res[k] <= max(2e-2, 1.5 * res['emu_' + k]), res


def test_cuda_graph_with_content_discriminator_schedule():
    """--use_dis_content alternates two iteration kinds (d_iter = 3): each is captured in its own CUDA
    graph; 10 iterations (warm-up, both captures, replays) match the eager launch sequence."""
    from tests.model_cases import graph_vs_eager
    _exact(graph_vs_eager(size=256, B=1, iters=10, use_dis_content=True))


def test_cuda_graph_with_spectral_norm_and_nearest_upsampling():
    """--dis_sn (power iteration + W / sigma inside the captured iteration, u / v buffers advancing on
    every replay) and --up_type nearest: graph replays match the eager launch sequence."""
    from tests.model_cases import graph_vs_eager
    _exact(graph_vs_eager(dis_sn=True, up_type='nearest'))


def test_sampling_graph_matches_eager():
    """Test-mode sampling (forward_random) replayed as a CUDA graph vs eager launches, 5 calls with
    changing inputs (2 eager warm-ups, capture, replays): identical up to the atomics of the unfused
    plane statistics."""
    from tests.model_cases import sampling_graph_vs_eager
    _check(sampling_graph_vs_eager(), 1e-3, 0.0)
