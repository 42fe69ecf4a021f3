"""Timeline of one board pair inside tower_tf_kernel (experiments build: make -C alphatriangle_b200/csrc experiments).

    AT_B200_LIB=$PWD/alphatriangle_b200/libat_b200_exp.so python tools/tf_timeline.py [batch]

For board 0 of the fourth pair CTA 0 processes: every hand-off between the issuing warp and the epilogue warps, in SM cycles.
Columns per step: when the issuer saw its operands, when it had issued + committed, when the epilogue warp saw the
accumulator, when it had finished its work and posted -- and the four gaps of the round trip.
"""
import ctypes as C
import os
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import alphatriangle_b200 as at  # noqa: E402
from alphatriangle_b200 import _lib  # noqa: E402
from alphatriangle_b200.packing import pack_states  # noqa: E402
from alphatriangle_b200.synth import synth_state_dict, synth_states  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
env, mc = at.EnvConfig(), at.ModelConfig()
nn = at.NeuralNetwork(mc, env, None, torch.device("cuda:0"), precision=os.environ.get("AT_PREC", "f16"), max_batch=B)
nn.set_weights(synth_state_dict(nn.get_weights(), 1))
base = pack_states(synth_states(256, env, seed=5), env)
packed = torch.from_numpy(np.tile(base, (-(-B // 256), 1))[:B]).cuda()
for _ in range(3):
    nn.engine.evaluate_packed(packed, nn.precision)
torch.cuda.synchronize()
lib = _lib.load()
EV = 256
buf = (C.c_uint64 * (32 * EV))()
lib.at_debug_read_timeline.restype = C.c_int
lib.at_debug_read_timeline.argtypes = [C.c_void_p, C.c_int]
lib.at_debug_read_timeline(buf, 32 * EV)
a = np.frombuffer(buf, dtype=np.uint64).reshape(32, EV)
tags, clk = (a >> np.uint64(48)).astype(np.int64), (a & np.uint64(0xFFFFFFFFFFFF)).astype(np.int64)
t0 = clk[tags > 0].min()


def events(w, tag):
    return [int(clk[w, i] - t0) for i in range(EV) if tags[w, i] == tag]


n_conv = 6
labels = [f"conv{l + 1}" for l in range(n_conv)]
for t in range(mc.TRANSFORMER_LAYERS):
    for h in range(mc.TRANSFORMER_HEADS):
        labels += [f"L{t} qkv{h}", f"L{t} S{h}", f"L{t} PV{h}"]
    labels += [f"L{t} attn->x"] + [f"L{t} ffn{f}" for f in range(mc.TRANSFORMER_FC_DIM // 128)] + [f"L{t} lin2"]
# issuer: operands seen (3) / committed (4). The projection of head h+1 is committed without a wait in front of it, so the
# issuer's commits are matched to the epilogue's accumulator events one to one, its waits are listed separately.
seen, commit = events(1, 3), events(1, 4)
acc = {w: events(w, 1) for w in (2, 6)}
post = {w: events(w, 2) for w in (2, 6)}
print(f"issuer A: {len(seen)} waits, {len(commit)} commits; epilogue warp 2: {len(acc[2])} accumulators, {len(post[2])} posts")
print(f"{'step':12s} {'commit':>7s} {'acc seen':>8s} {'posted(h0)':>10s} {'posted(h1)':>10s} | {'mma+commit':>10s} {'epilogue':>8s}")
for i, c in enumerate(commit):
    if i >= len(acc[2]):
        break
    a2 = acc[2][i]
    # the pair's layer-0 image was posted during the previous pair (not stamped): post i answers accumulator i
    p2 = post[2][i] if i < len(post[2]) else -1
    p6 = post[6][i] if i < len(post[6]) else -1
    name = labels[i] if i < len(labels) else "?"
    print(f"{name:12s} {c:7d} {a2:8d} {p2:10d} {p6:10d} | {a2 - c:10d} {max(p2, p6) - a2:8d}")
# issuer A in full: tag:cycle (3 operands seen, 5 weight slab seen, 4 issued + committed)
w = 1
print("issuer A all:", " ".join(f"{int(tags[w, i])}:{int(clk[w, i] - t0)}" for i in range(EV) if tags[w, i] > 0))
for w in (2, 6):
    print(f"epilogue warp {w} all:", " ".join(f"{int(tags[w, i])}:{int(clk[w, i] - t0)}" for i in range(EV) if tags[w, i] > 0)[:700])
print("issuer waits (operands seen):", " ".join(str(x) for x in seen))
print("total cycles for the pair:", max(max(v) for v in post.values() if v) - min(seen))
