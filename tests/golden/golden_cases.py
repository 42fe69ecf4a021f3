"""Configurations and seeded inputs of the golden fixtures; shared by make_golden.py and the tests."""

from __future__ import annotations

from alphatriangle_b200.synth import edge_states, synth_states

# the two BASELINE models again with trained-like weights (synth_state_dict(sharpness=...)): "_sharp" policies of about two nats
# with a clear best move on most states and expected values spread over the support, "_peaked" near-deterministic ones
SHARP_CASES = ("conv_only_sharp", "conv_only_peaked", "default_tf_sharp", "default_tf_peaked")
CASES = ("default_tf", "conv_only", "ref_test_cfg", "odd_gelu", "strided_silu", "tanh_nobn", "sigmoid_tiny") + SHARP_CASES
_SHARPNESS = {"conv_only_sharp": 2.0, "conv_only_peaked": 8.0, "default_tf_sharp": 3.0, "default_tf_peaked": 8.0}

_DIAMOND_4x5 = [(1, 4), (0, 5), (0, 5), (1, 4)]


def case_sharpness(name: str) -> float:
    return _SHARPNESS.get(name, 0.0)


def case_weights(name: str, state_dict: dict) -> dict:
    """The seeded weights of a case for a reference-keyed state_dict."""
    from alphatriangle_b200.synth import synth_state_dict

    return synth_state_dict(state_dict, case_configs(name)[2], case_sharpness(name))


def case_configs(name: str):
    """-> (ModelConfig kwargs, EnvConfig kwargs, weight seed)"""
    if name in _SHARPNESS:
        mc_kwargs, env_kwargs, seed = case_configs(name.rsplit("_", 1)[0])
        return mc_kwargs, env_kwargs, seed + 100
    if name == "default_tf":  # BASELINE config 1 / 3: the reference's defaults
        return {}, {}, 0
    if name == "conv_only":  # BASELINE config 2
        return {"USE_TRANSFORMER": False}, {}, 1
    if name == "ref_test_cfg":  # the reference's own test fixture (tests/conftest.py:30-71): 3x3, 1 slot, D truncated to 10
        return (
            dict(GRID_INPUT_CHANNELS=1, CONV_FILTERS=[4], CONV_KERNEL_SIZES=[3], CONV_STRIDES=[1], CONV_PADDING=[1],
                 NUM_RESIDUAL_BLOCKS=0, RESIDUAL_BLOCK_FILTERS=4, USE_TRANSFORMER=False, TRANSFORMER_DIM=16, TRANSFORMER_HEADS=2,
                 TRANSFORMER_LAYERS=0, TRANSFORMER_FC_DIM=32, FC_DIMS_SHARED=[8], POLICY_HEAD_DIMS=[9], VALUE_HEAD_DIMS=[1],
                 OTHER_NN_INPUT_FEATURES_DIM=10, ACTIVATION_FUNCTION="ReLU", USE_BATCH_NORM=True),
            dict(ROWS=3, COLS=3, PLAYABLE_RANGE_PER_ROW=[(0, 3), (0, 3), (0, 3)], NUM_SHAPE_SLOTS=1),
            2,
        )
    if name == "odd_gelu":  # projection before the res block, input_proj, no BN, GELU, 2 grid channels, D zero-padded
        return (
            dict(GRID_INPUT_CHANNELS=2, CONV_FILTERS=[8, 16], CONV_KERNEL_SIZES=[3, (3, 3)], CONV_STRIDES=[1, 1],
                 CONV_PADDING=[1, "same"], NUM_RESIDUAL_BLOCKS=1, RESIDUAL_BLOCK_FILTERS=32, USE_TRANSFORMER=True,
                 TRANSFORMER_DIM=16, TRANSFORMER_HEADS=2, TRANSFORMER_LAYERS=1, TRANSFORMER_FC_DIM=32, FC_DIMS_SHARED=[24, 16],
                 POLICY_HEAD_DIMS=[], VALUE_HEAD_DIMS=[12, 8], NUM_VALUE_ATOMS=11, VALUE_MIN=-3.0, VALUE_MAX=5.0,
                 OTHER_NN_INPUT_FEATURES_DIM=40, ACTIVATION_FUNCTION="GELU", USE_BATCH_NORM=False),
            dict(ROWS=4, COLS=5, PLAYABLE_RANGE_PER_ROW=_DIAMOND_4x5, NUM_SHAPE_SLOTS=2),
            3,
        )
    if name == "strided_silu":  # stride-2 convolution, 4 slots, SiLU
        return (
            dict(CONV_FILTERS=[8, 16], CONV_KERNEL_SIZES=[3, 3], CONV_STRIDES=[1, 2], CONV_PADDING=[1, 1], NUM_RESIDUAL_BLOCKS=1,
                 RESIDUAL_BLOCK_FILTERS=16, USE_TRANSFORMER=False, FC_DIMS_SHARED=[32], POLICY_HEAD_DIMS=[32], VALUE_HEAD_DIMS=[16],
                 OTHER_NN_INPUT_FEATURES_DIM=38, ACTIVATION_FUNCTION="SiLU", USE_BATCH_NORM=True),
            dict(ROWS=6, COLS=6, PLAYABLE_RANGE_PER_ROW=[(0, 6)] * 6, NUM_SHAPE_SLOTS=4),
            4,
        )
    if name == "tanh_nobn":  # unpadded 3x3 convolution (board shrinks), Tanh, no BatchNorm
        return (
            dict(CONV_FILTERS=[6], CONV_KERNEL_SIZES=[3], CONV_STRIDES=[1], CONV_PADDING=[0], NUM_RESIDUAL_BLOCKS=2,
                 RESIDUAL_BLOCK_FILTERS=6, USE_TRANSFORMER=False, FC_DIMS_SHARED=[20], POLICY_HEAD_DIMS=[20], VALUE_HEAD_DIMS=[10],
                 OTHER_NN_INPUT_FEATURES_DIM=30, ACTIVATION_FUNCTION="Tanh", USE_BATCH_NORM=False),
            dict(ROWS=5, COLS=7, PLAYABLE_RANGE_PER_ROW=[(1, 6), (0, 7), (0, 7), (0, 7), (1, 6)], NUM_SHAPE_SLOTS=3),
            5,
        )
    if name == "sigmoid_tiny":  # 1x1 board edge of the geometry, Sigmoid
        return (
            dict(CONV_FILTERS=[4], CONV_KERNEL_SIZES=[3], CONV_STRIDES=[1], CONV_PADDING=[1], NUM_RESIDUAL_BLOCKS=0,
                 RESIDUAL_BLOCK_FILTERS=4, USE_TRANSFORMER=False, FC_DIMS_SHARED=[8], POLICY_HEAD_DIMS=[8], VALUE_HEAD_DIMS=[8],
                 OTHER_NN_INPUT_FEATURES_DIM=14, ACTIVATION_FUNCTION="Sigmoid", USE_BATCH_NORM=True),
            dict(ROWS=1, COLS=2, PLAYABLE_RANGE_PER_ROW=[(0, 2)], NUM_SHAPE_SLOTS=1),
            6,
        )
    raise KeyError(name)


def case_states(name: str, env_config):
    if name in _SHARPNESS:  # same states as the base case: only the weights differ
        return case_states(name.rsplit("_", 1)[0], env_config)
    n = 48 if name in ("default_tf", "conv_only") else 20
    return synth_states(n, env_config, seed=1234 + CASES.index(name)) + edge_states(env_config)
