"""BASELINE config 5 substitute (real trimcts / trianglengin are not installable here): the call pattern of many
concurrent self-play searches against ONE NeuralNetwork -- every game thread asks for <= 32 leaves at a time
(mcts_batch_size, alphatriangle/config/mcts_config.py:14), blocking like trimcts does -- with and without the
cross-game LeafBatcher. Prints one JSON line per variant: API-level states/s through evaluate_batch (state packing
and result-dict construction included), GPU calls made, mean batch size.

    python tools/selfplay_pattern.py [games=128] [requests_per_game=40] [leaves=32] [workload=conv_only|default_tf]
"""
import json
import sys
import threading
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import alphatriangle_b200 as at  # noqa: E402
from alphatriangle_b200.synth import synth_state_dict, synth_states  # noqa: E402

games = int(sys.argv[1]) if len(sys.argv) > 1 else 128
reqs = int(sys.argv[2]) if len(sys.argv) > 2 else 40
leaves = int(sys.argv[3]) if len(sys.argv) > 3 else 32
workload = sys.argv[4] if len(sys.argv) > 4 else "conv_only"

env = at.EnvConfig()
mc = at.ModelConfig(USE_TRANSFORMER=(workload == "default_tf"))
nn = at.NeuralNetwork(mc, env, None, torch.device("cuda:0"), precision="bf16", max_batch=4096)
nn.set_weights(synth_state_dict(nn.get_weights(), 1))
pool = synth_states(4096, env, seed=55)
nn.evaluate_batch(pool[:64])  # warm-up


def run(label, make_client, close_all):
    def game(g, client, out):
        n = 0
        for r in range(reqs):
            lo = (g * 131 + r * 17) % (len(pool) - leaves)
            res = client.evaluate_batch(pool[lo : lo + leaves])
            n += len(res)
        out.append(n)
        if hasattr(client, "close"):
            client.close()

    out = []
    clients = [make_client() for _ in range(games)]
    threads = [threading.Thread(target=game, args=(g, c, out)) for g, c in enumerate(clients)]
    t0 = time.perf_counter()
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    dt = time.perf_counter() - t0
    stats = close_all()
    total = sum(out)
    print(json.dumps({"variant": label, "workload": workload, "games": games, "leaves_per_request": leaves, "requests": games * reqs,
                      "states": total, "seconds": dt, "api_states_per_s": total / dt, **stats}))


# (1) the reference's arrangement: every search calls the network directly, 32 states per GPU call
lock = threading.Lock()


class Direct:
    def evaluate_batch(self, states):
        with lock:  # one NeuralNetwork, one stream: calls serialise exactly as they would under the GIL
            return nn.evaluate_batch(states)


run("direct, 32 leaves per GPU call", Direct, lambda: {"gpu_calls": games * reqs, "mean_batch": float(leaves)})

# (2) cross-game batching
b = at.LeafBatcher(nn, max_batch=4096, max_wait_s=0.002)


def close_b():
    b.close()
    return {"gpu_calls": b.calls, "mean_batch": b.states / max(b.calls, 1)}


run("LeafBatcher, up to 4096 states per GPU call", b.client, close_b)
