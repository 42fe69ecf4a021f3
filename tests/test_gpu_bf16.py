"""tcgen05 / TMEM path parity, both operand formats (north_star: policy KL <= 1e-3 against the fp32 reference).
Targets: the reference's recorded outputs (tests/golden), including the trained-like "_sharp" / "_peaked" weight sets,
and the oracle on fresh states."""

import numpy as np
import pytest
import torch
from gpu_util import kl_rows, load_case, network

from alphatriangle_b200 import _lib
from alphatriangle_b200.packing import pack_states
from alphatriangle_b200.synth import edge_states, synth_states
from oracle import features_ref, model_ref

pytestmark = pytest.mark.gpu

KL_TOL = 1e-3      # north_star: policy KL(reference || ours) per state, bf16 path
TRUNK_TOL = 3e-2   # first shared-FC activation, relative to its largest magnitude

# Tolerances by operand format and by how sharp the network's outputs are. What is asserted is what was measured on the
# B200 (DESIGN.md section 2 has the table) with margin. Every operand of every MMA is rounded to the operand format once:
# 2^-9 relative for bf16's 8-bit mantissa, 2^-12 for fp16's 11 bits; through the tower, the trunk FC and the heads that
# comes to ~0.5 % rms of a logit's spread for bf16 and ~0.06 % for fp16. KL ~ (logit error)^2 / 2 and the error of E[z] both
# grow with the spread of the logits, so:
#   "plain"  small weights, logits within +-0.3, policy uniform to three digits          -- says little; kept from round 1
#   "sharp"  trained-like: policies of ~1.8 nats, best move > 0.3 on 80 % of the states, values spread over +-8
#   "peaked" near-deterministic policies (~0.4-0.6 nats), value distributions concentrated on one or two atoms
# bf16 holds north_star's KL <= 1e-3 up to about the "sharp" level (measured 0.8e-3 / 1.0e-3: at the edge) and not beyond
# (1e-2 on "peaked"); fp16 operands hold it everywhere with a wide margin. That is why precision="auto" picks fp16.
KL_TOL_BY_KIND = {
    "bf16": {"plain": KL_TOL, "sharp": 1.5e-3, "peaked": 2e-2},
    "f16": {"plain": KL_TOL, "sharp": KL_TOL, "peaked": KL_TOL},
}
# expected value, |ours - reference| on the [-10, 10] support
VALUE_TOL_BY_KIND = {
    "bf16": {"plain": 5e-3, "sharp": 0.6, "peaked": 2.0},
    "f16": {"plain": 1e-3, "sharp": 0.1, "peaked": 0.3},
}
VALUE_TOL = VALUE_TOL_BY_KIND["bf16"]
FORMATS = ["bf16", "f16"]


def kind_of(case):
    return case.rsplit("_", 1)[-1] if case.endswith(("_sharp", "_peaked")) else "plain"



def test_bf16_path_is_available_for_the_baseline_config():
    nn = network("conv_only", "bf16")
    assert nn.precision == _lib.PREC_BF16 and nn.engine.supports(_lib.PREC_BF16) and nn.engine.supports(_lib.PREC_F16)
    # "auto": fp16 operands for networks with normalisation layers (both BASELINE models), bf16 without them
    assert network("conv_only", "auto").precision == _lib.PREC_F16
    assert network("default_tf", "auto").precision == _lib.PREC_F16  # the reference's default model, transformer included
    import alphatriangle_b200 as at

    no_bn = at.NeuralNetwork(at.ModelConfig(USE_TRANSFORMER=False, USE_BATCH_NORM=False), at.EnvConfig(), None, torch.device("cuda:0"))
    assert no_bn.precision == _lib.PREC_BF16
    # configurations outside the specialisation report why and stay on the fp32 path
    odd = network("odd_gelu", "auto")
    assert odd.precision == _lib.PREC_FP32 and not odd.engine.supports(_lib.PREC_BF16)


BF16_CASES = ["conv_only", "default_tf"]  # BASELINE configs 2 and 3: conv tower alone / fused with the encoder layers
GOLDEN_CASES = BF16_CASES + ["conv_only_sharp", "conv_only_peaked", "default_tf_sharp", "default_tf_peaked"]


@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("case", BF16_CASES)
def test_trunk_activation_matches_oracle(case, fmt, golden_dir):
    mc, env, _seed, g = load_case(case, golden_dir)
    nn = network(case, fmt)
    nn.evaluate_packed(torch.from_numpy(g["packed"][:2]).cuda())  # the trunk hook runs in the format used last
    taps = []
    model_ref.forward(nn.get_weights(), mc, torch.from_numpy(g["grid"]), torch.from_numpy(g["other"]), taps)
    ref = taps[0].numpy()
    got = nn.engine.bf16_trunk(torch.from_numpy(g["packed"]).cuda()).cpu().numpy()
    scale = np.abs(ref).max()
    assert np.abs(got - ref).max() <= TRUNK_TOL * scale, (np.abs(got - ref).max(), scale)


@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("case", GOLDEN_CASES)
def test_tensor_core_path_matches_reference_golden(case, fmt, golden_dir):
    mc, env, _seed, g = load_case(case, golden_dir)
    nn = network(case, fmt)
    probs, value = nn.evaluate_packed(torch.from_numpy(g["packed"]).cuda())
    probs, value = probs.cpu().numpy(), value.cpu().numpy()
    kl = kl_rows(g["probs"], probs)
    verr = np.abs(value - g["values"])
    agree = float((probs.argmax(axis=1) == g["probs"].argmax(axis=1)).mean())
    print(f"[{fmt} parity] {case}: policy KL max {kl.max():.3e} mean {kl.mean():.3e}; value |err| max {verr.max():.3e} "
          f"mean {verr.mean():.3e}; best move agrees on {agree:.1%}")
    assert kl.max() <= KL_TOL_BY_KIND[fmt][kind_of(case)]
    assert verr.max() <= VALUE_TOL_BY_KIND[fmt][kind_of(case)]
    assert np.abs(probs.sum(axis=1) - 1.0).max() < 1e-5
    if kind_of(case) != "plain":  # the move a search would rank first is the reference's (ties aside)
        assert agree >= 0.95


@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("case", BF16_CASES)
@pytest.mark.parametrize("batch", [1, 2, 3, 127, 129, 1000])
def test_ragged_batches_match_oracle(case, batch, fmt, golden_dir):
    """Odd board counts (a half-empty pair), partial 128-board FC tiles, single states."""
    mc, env, _seed, _g = load_case(case, golden_dir)
    nn = network(case, fmt)
    states = (synth_states(batch, env, seed=9000 + batch) + edge_states(env))[:batch]
    grid, other = features_ref.extract_batch(states, mc)
    torch.set_num_threads(8)
    p_ref, v_ref = model_ref.evaluate_tensors(nn.get_weights(), mc, grid, other)
    probs, value = nn.evaluate_packed(torch.from_numpy(pack_states(states, env)).cuda())
    assert kl_rows(p_ref, probs.cpu().numpy()).max() <= KL_TOL
    assert np.abs(value.cpu().numpy() - v_ref).max() <= VALUE_TOL["plain"]


@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("case", BF16_CASES)
def test_full_batch_rows_are_independent(case, fmt, golden_dir):
    """At the BASELINE batch (4096): every board's result equals the result of the same board in a small
    batch -- no cross-board leakage through the shared-memory images, TMEM or the FC tiles."""
    mc, env, _seed, g = load_case(case, golden_dir)
    nn = network(case, fmt)
    base = g["packed"]
    packed = np.tile(base, (-(-4096 // base.shape[0]), 1))[:4096]
    probs, value = nn.evaluate_packed(torch.from_numpy(packed).cuda())
    p0, v0 = nn.evaluate_packed(torch.from_numpy(base).cuda())
    idx = np.arange(4096) % base.shape[0]
    probs, value, p0, v0 = probs.cpu().numpy(), value.cpu().numpy(), p0.cpu().numpy(), v0.cpu().numpy()
    # inside one launch the same board gives the same bits wherever it sits in the batch
    n = base.shape[0]
    assert np.array_equal(probs, probs[:n][idx])
    assert np.array_equal(value, value[:n][idx])
    # against a small batch only the split-K factor of the trunk FC differs: fp32 summation order, which can move a
    # trunk activation across a bf16 rounding boundary (one ulp = 0.4 %) before the head GEMMs -- nothing larger
    assert np.abs(probs - p0[idx]).max() <= 1e-3
    assert np.abs(value - v0[idx]).max() <= 5e-2
    assert kl_rows(p0[idx], probs).max() <= 1e-5


@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("case", BF16_CASES)
def test_repeated_boards_agree_over_many_launches(case, fmt, golden_dir):
    """300 launches of 4096 repeated boards: every copy of a board gives the same bits in every launch. (A missing
    barrier between the last convolution's residual reads and the encoder operands laid over the same bytes showed up
    as ONE board off by ~5e-5 in about one launch in seventy -- too rare for a handful of repetitions to catch.)"""
    _mc, _env, _seed, g = load_case(case, golden_dir)
    nn = network(case, fmt)
    base = g["packed"]
    n = base.shape[0]
    packed = torch.from_numpy(np.tile(base, (-(-4096 // n), 1))[:4096]).cuda()
    idx = torch.arange(4096, device="cuda") % n
    bad = torch.zeros((), dtype=torch.int64, device="cuda")
    first = None
    for _ in range(300):
        probs, value = nn.evaluate_packed(packed)
        if first is None:
            first = (probs.clone(), value.clone())
        bad += (probs != probs[:n][idx]).any(dim=1).sum() + (value != value[:n][idx]).sum()
        bad += (probs != first[0]).any(dim=1).sum() + (value != first[1]).sum()
    assert int(bad) == 0


@pytest.mark.parametrize("case", BF16_CASES)
def test_results_do_not_depend_on_which_cta_takes_which_pair(case, golden_dir):
    """Board pairs are handed out to the persistent CTAs by an atomic counter (so that a launch whose CTAs do not all get an SM
    at once -- something else is running -- does not fall into a second wave). A second engine keeps a few SMs busy from a side
    stream while the big batch is evaluated: the results must be the bits of the undisturbed evaluation."""
    import alphatriangle_b200 as at
    from golden_cases import case_configs, case_weights

    _mc, _env, _seed, g = load_case(case, golden_dir)
    nn = network(case, "f16", max_batch=8192)
    mc_kwargs, env_kwargs, _ = case_configs(case)
    hog = at.NeuralNetwork(at.ModelConfig(**mc_kwargs), at.EnvConfig(**env_kwargs), at.TrainConfig(), torch.device("cuda:0"),
                           precision="f16", max_batch=64)
    hog.set_weights(case_weights(case, hog.get_weights()))
    base = g["packed"]
    packed = torch.from_numpy(np.tile(base, (-(-8192 // base.shape[0]), 1))[:8192]).cuda()
    small = torch.from_numpy(base[:32].copy()).cuda()
    p0, v0 = nn.evaluate_packed(packed)
    h0, _ = hog.evaluate_packed(small)
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        for _ in range(400):
            h1, _ = hog.evaluate_packed(small)
    for _ in range(6):
        p1, v1 = nn.evaluate_packed(packed)
        assert torch.equal(p1, p0) and torch.equal(v1, v0)
    torch.cuda.synchronize()
    assert torch.equal(h1, h0)


def test_bf16_through_the_drop_in_api(golden_dir):
    from golden_cases import case_states

    mc, env, _seed, g = load_case("conv_only", golden_dir)
    nn = network("conv_only", "bf16")
    states = case_states("conv_only", env)[:16]
    assert nn.engine.kernel_times() is None or True  # timing is opt-in
    nn.engine.set_kernel_timing(True)
    res = nn.evaluate_batch(states)
    for i, (pol, val) in enumerate(res):
        row = np.array([pol[a] for a in range(nn.action_dim)])
        assert abs(row.sum() - 1.0) < 1e-5
        assert kl_rows(g["probs"][i : i + 1], row[None, :]).max() <= KL_TOL
        assert abs(val - float(g["values"][i])) <= VALUE_TOL["plain"]
    t = nn.engine.kernel_times()
    assert t is not None and t["tower_ms"] > 0 and t["fc_ms"] > 0
    nn.engine.set_kernel_timing(False)


# Head shapes that leave the tensor-core head kernel: (ModelConfig overrides, kernels expected after the tower and the FC)
HEAD_SHAPES = {
    # hidden widths differ from the trunk's -> heads_fused_kernel (same fusion on CUDA cores, fp32)
    "cuda_core_heads": (dict(USE_TRANSFORMER=False, FC_DIMS_SHARED=[128], POLICY_HEAD_DIMS=[64], VALUE_HEAD_DIMS=[32]), 4),
    # two shared layers, a two-layer policy head and a value head without hidden layer -> the unfused tail
    # (split-K finish, one GEMM per layer, softmax kernel, status merge)
    "unfused_tail": (dict(USE_TRANSFORMER=False, FC_DIMS_SHARED=[128, 64], POLICY_HEAD_DIMS=[64, 32], VALUE_HEAD_DIMS=[]), None),
    # the same two with the encoder layers in front
    "cuda_core_heads_tf": (dict(FC_DIMS_SHARED=[96], POLICY_HEAD_DIMS=[48], VALUE_HEAD_DIMS=[24]), 4),
}


@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("shape", sorted(HEAD_SHAPES))
def test_other_head_kernels_match_oracle(shape, fmt):
    """Every CUDA path that ships gets launched: head shapes outside heads_mma_kernel's specialisation."""
    import alphatriangle_b200 as at
    from alphatriangle_b200.synth import synth_state_dict

    overrides, want_launches = HEAD_SHAPES[shape]
    mc, env = at.ModelConfig(**overrides), at.EnvConfig()
    nn = at.NeuralNetwork(mc, env, None, torch.device("cuda:0"), precision=fmt, max_batch=256)
    nn.set_weights(synth_state_dict(nn.get_weights(), 77, sharpness=1.0))
    states = synth_states(150, env, seed=4321) + edge_states(env)
    grid, other = features_ref.extract_batch(states, mc)
    p_ref, v_ref = model_ref.evaluate_tensors(nn.get_weights(), mc, grid, other)
    probs, value = nn.evaluate_packed(torch.from_numpy(pack_states(states, env)).cuda())
    launches = nn.engine.last_launches()
    if want_launches is None:
        assert launches > 6, launches
    else:
        assert launches == want_launches, launches  # features, tower, FC, one head kernel
    kl = kl_rows(p_ref, probs.cpu().numpy())
    verr = np.abs(value.cpu().numpy() - v_ref)
    print(f"[{fmt} parity] {shape}: policy KL max {kl.max():.3e}; value |err| max {verr.max():.3e}; {launches} launches")
    assert kl.max() <= KL_TOL
    assert verr.max() <= 0.3
    assert np.abs(probs.cpu().numpy().sum(axis=1) - 1.0).max() < 1e-5


ENCODER_SHAPES = {
    "one_layer_ff128_gelu": dict(TRANSFORMER_LAYERS=1, TRANSFORMER_FC_DIM=128, ACTIVATION_FUNCTION="GELU"),  # a single hidden half
    "three_layers_ff512": dict(TRANSFORMER_LAYERS=3, TRANSFORMER_FC_DIM=512),                               # four hidden halves
    "four_layers_ff384": dict(TRANSFORMER_LAYERS=4, TRANSFORMER_FC_DIM=384),
    # tower ends on 64 channels: 1x1 projection into the residual blocks and input_proj (model.py:179-186) up to the
    # encoder's 128 -- both 1x1 layers run on the tensor cores, input_proj without activation
    "input_proj_from_64": dict(RESIDUAL_BLOCK_FILTERS=64, NUM_RESIDUAL_BLOCKS=1),
    "input_proj_from_32_no_res": dict(CONV_FILTERS=[32, 32], CONV_KERNEL_SIZES=[3, 3], CONV_STRIDES=[1, 1], CONV_PADDING=[1, 1],
                                      NUM_RESIDUAL_BLOCKS=0, TRANSFORMER_LAYERS=1),
}


@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("shape", sorted(ENCODER_SHAPES))
def test_other_encoder_shapes_match_oracle(shape, fmt):
    """The fused tower + encoder kernel outside the default 2 layers x 256: the weight ring's slab order (projection, bias
    and LayerNorm folds, hidden halves) is generated per shape on the host and must match what the issuers consume."""
    import alphatriangle_b200 as at
    from alphatriangle_b200.synth import synth_state_dict

    mc, env = at.ModelConfig(**ENCODER_SHAPES[shape]), at.EnvConfig()
    nn = at.NeuralNetwork(mc, env, None, torch.device("cuda:0"), precision=fmt, max_batch=256)
    assert nn.precision in _lib.TENSOR_CORE_PRECISIONS
    nn.set_weights(synth_state_dict(nn.get_weights(), 91, sharpness=1.0))
    states = synth_states(101, env, seed=777) + edge_states(env)
    grid, other = features_ref.extract_batch(states, mc)
    p_ref, v_ref = model_ref.evaluate_tensors(nn.get_weights(), mc, grid, other)
    probs, value = nn.evaluate_packed(torch.from_numpy(pack_states(states, env)).cuda())
    kl = kl_rows(p_ref, probs.cpu().numpy())
    verr = np.abs(value.cpu().numpy() - v_ref)
    print(f"[{fmt} parity] {shape}: policy KL max {kl.max():.3e}; value |err| max {verr.max():.3e}")
    assert kl.max() <= KL_TOL
    assert verr.max() <= 0.3


@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("case", BF16_CASES)
def test_repeated_evaluations_are_bit_identical(case, fmt, golden_dir):
    """The pipelines inside the tower kernels hand shared-memory images, accumulators and ring stages from warp to warp and
    from one board pair to the next; a missing fence or barrier shows up as run-to-run differences long before it shows up
    as a wrong answer. 40 launches at a batch that gives every CTA a different number of pairs must agree bit for bit."""
    mc, env, _seed, g = load_case(case, golden_dir)
    nn = network(case, fmt)
    base = g["packed"]
    n = 4096 + 37
    packed = torch.from_numpy(np.tile(base, (-(-n // base.shape[0]), 1))[:n]).cuda()
    big = at_network_for(nn, n)
    p0, v0 = [t.clone() for t in big.evaluate_packed(packed)]
    for _ in range(40):
        p, v = big.evaluate_packed(packed)
        assert torch.equal(p, p0) and torch.equal(v, v0)


def at_network_for(nn, batch):
    """A network with the same weights and precision whose engine holds `batch` states."""
    import alphatriangle_b200 as at

    name = _lib.PRECISION_NAMES[nn.precision]
    big = at.NeuralNetwork(nn.model_config, nn.env_config, None, nn.device, precision=name, max_batch=batch)
    big.set_weights(nn.get_weights())
    return big


def test_fp16_operands_saturate_instead_of_overflowing():
    """fp16's range ends at 65504. A network whose activations go beyond it (here: the first convolution scaled by 3e5) must
    not produce infinities or NaNs on the fp16 path -- the conversions clamp -- and the bf16 path, whose range is fp32's,
    must still track the reference. That is the trade-off precision="auto" decides by."""
    import alphatriangle_b200 as at
    from alphatriangle_b200.synth import synth_state_dict

    mc, env = at.ModelConfig(USE_TRANSFORMER=False), at.EnvConfig()
    states = synth_states(64, env, seed=31)
    packed = torch.from_numpy(pack_states(states, env)).cuda()
    sd = None
    out = {}
    for fmt in FORMATS:
        nn = at.NeuralNetwork(mc, env, None, torch.device("cuda:0"), precision=fmt, max_batch=64)
        if sd is None:
            sd = synth_state_dict(nn.get_weights(), 5, sharpness=1.0)
            sd["conv_body.0.0.weight"] = sd["conv_body.0.0.weight"] * 3e5  # BatchNorm follows, but its input is the conv output
            sd["conv_body.0.1.running_var"] = sd["conv_body.0.1.running_var"] * 0 + 1.0
            sd["conv_body.0.1.weight"] = sd["conv_body.0.1.weight"] * 0 + 1.0
        nn.set_weights(sd)
        probs, value, flags = nn.engine.evaluate_packed(packed, nn.precision)
        assert int(flags[0].item()) == 0
        assert bool(torch.isfinite(probs).all()) and bool(torch.isfinite(value).all())
        assert np.abs(probs.cpu().numpy().sum(axis=1) - 1.0).max() < 1e-5
        out[fmt] = probs.cpu().numpy()
    grid, other = features_ref.extract_batch(states, mc)
    p_ref, _v = model_ref.evaluate_tensors(sd, mc, grid, other)
    assert (out["bf16"].argmax(axis=1) == p_ref.argmax(axis=1)).mean() >= 0.9  # bf16 follows the reference through 1e5-sized activations
