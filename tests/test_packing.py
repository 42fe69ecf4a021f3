import numpy as np
import pytest

import alphatriangle_b200 as at
from alphatriangle_b200.packing import PACKED_DTYPE, RECORD_BYTES, pack_states, unpack_masks
from alphatriangle_b200.synth import SynthShape, SynthState, edge_states, synth_states


def test_record_layout():
    assert PACKED_DTYPE.itemsize == RECORD_BYTES == 80
    assert PACKED_DTYPE.fields["score"][1] == 64 and PACKED_DTYPE.fields["step"][1] == 72


def test_masks_round_trip():
    env = at.EnvConfig()
    states = synth_states(32, env, seed=7) + edge_states(env)
    packed = pack_states(states, env)
    occ, death = unpack_masks(packed, env.ROWS, env.COLS)
    for i, s in enumerate(states):
        g = s.get_grid_data_np()
        assert np.array_equal(occ[i], g["occupied"]) and np.array_equal(death[i], g["death"])
    recs = packed.reshape(-1).view(PACKED_DTYPE)
    for i, s in enumerate(states):
        assert recs["score"][i] == s.game_score() or (np.isnan(recs["score"][i]) and np.isnan(s.game_score()))
        for j, shp in enumerate(s.get_shapes()):
            slot = recs["slots"][i][j]
            assert slot[0] == (1 if shp else 0)
            if shp and shp.triangles:
                assert slot[1].view(np.uint8) == len(shp.triangles)
                assert tuple(slot[4:8]) == shp.bbox()


def test_empty_and_ragged():
    env = at.EnvConfig()
    assert pack_states([], env).shape == (0, 80)
    one = synth_states(1, env)[0]
    ragged = SynthState(env, one._occupied, one._death, [None], 0.0, 0)  # fewer shapes than slots
    rec = pack_states([ragged], env).reshape(-1).view(PACKED_DTYPE)
    assert not rec["slots"][0].any()


def test_limits_are_rejected():
    big = at.EnvConfig(ROWS=12, COLS=12, PLAYABLE_RANGE_PER_ROW=[(0, 12)] * 12, NUM_SHAPE_SLOTS=3)
    with pytest.raises(ValueError, match="ROWS\\*COLS"):
        pack_states([], big)
    env = at.EnvConfig()
    one = synth_states(1, env)[0]
    far = SynthShape([(0, 0, True), (0, 300, False)])
    with pytest.raises(ValueError, match="int8"):
        pack_states([SynthState(env, one._occupied, one._death, [far, None, None], 0.0, 0)], env)
    wrong = SynthState(env, np.zeros((3, 3), bool), np.zeros((3, 3), bool), [None] * 3, 0.0, 0)
    with pytest.raises(ValueError, match="shape"):
        pack_states([wrong], env)
    with pytest.raises(ValueError, match="out must"):
        pack_states([one], env, out=np.zeros((2, 80), np.uint8))
