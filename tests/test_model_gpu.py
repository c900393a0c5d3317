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


def test_sampling_vs_oracle():
    """forward_reference / forward_random at 256x256 (config 2 shape) vs the fp32 oracle: the chain
    Ec -> Es -> G (25 layers) with variance-preserving synthetic weights; bounded by 1.5x the
    bf16-storage-emulated oracle's own distance from fp32, and 1e-1 absolute."""
    from tests.model_cases import sampling_vs_oracle
    res = sampling_vs_oracle()
    for k in ('forward_reference', 'forward_random'):
        assert res[k] <= 1e-1 and res[k] <= max(2e-2, 1.5 * res['emu_' + k]), res


def test_cuda_graph_iteration_matches_eager():
    """6 iterations (2 eager warm-up, capture, replays; learning rate changed every iteration):
    the graph-replayed losses and final weights match the eager launch sequence.  Not bit-exact
    because the weight-gradient split-K uses fp32 atomics (order varies run to run): the bound is
    5e-3 beyond three times the difference between two eager runs of the same sequence."""
    from tests.model_cases import graph_vs_eager
    res = graph_vs_eager()
    _check(res, 5e-3, 0.9999)


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
    over = None
    if name == 'step_norms_dim8_128':
        # --dis_norm instance: un-affine InstanceNorm on the discriminator's small feature planes makes
        # the class logits (weighted x5 in g_cls / gan2_cls) the most rounding-sensitive outputs: the
        # bf16-storage-emulated ORACLE (tests/emu.py) is itself 0.74e-2 / 0.98e-2 away from the fp32
        # reference on gan2_cls here, so these keys are bounded by 2e-2
        over = {k: 2e-2 for k in res if k.endswith('_cls')}
        # ... and 5e-2 in the SECOND iteration, which runs on weights after one Adam step (a sign-like
        # step of size lr whatever |grad|, so rounding noise in tiny gradients moves weights as far as
        # signal does): measured 1.7e-2 .. 2.2e-2 over repeated runs (the weight-gradient split-K uses
        # fp32 atomics, so the second iteration is not bit-reproducible run to run)
        over.update({k: 5e-2 for k in res if k.startswith('loss_it1.') and k.endswith('_cls')})
    _check(res, 1e-2, 0.0, over)


def test_base_default_training_step_vs_oracle():
    """BaseModel default flags (StyleEncoder + Decoder): losses of 2 iterations within 1e-2 of the oracle."""
    from tests.model_cases import base_default_step_vs_oracle
    _check(base_default_step_vs_oracle(), 1e-2, 0.0)


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
    _check(graph_vs_eager(size=256, B=1, iters=10, use_dis_content=True), 5e-3, 0.9999)


def test_cuda_graph_with_spectral_norm_and_nearest_upsampling():
    """--dis_sn (power iteration + W / sigma inside the captured iteration, u / v buffers advancing on
    every replay) and --up_type nearest: graph replays match the eager launch sequence."""
    from tests.model_cases import graph_vs_eager
    _check(graph_vs_eager(dis_sn=True, up_type='nearest'), 5e-3, 0.9999)


def test_sampling_graph_matches_eager():
    """Test-mode sampling (forward_random) replayed as a CUDA graph vs eager launches, 5 calls with
    changing inputs (2 eager warm-ups, capture, replays): identical up to the atomics of the unfused
    plane statistics."""
    from tests.model_cases import sampling_graph_vs_eager
    _check(sampling_graph_vs_eager(), 1e-3, 0.0)
