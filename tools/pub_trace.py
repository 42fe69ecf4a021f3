"""Where a publication's evaluation time goes: per-step device times around each publication and the host time of publish().

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29541 tools/pub_trace.py
"""
import os
import sys
import time
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import alphatriangle_b200 as at  # noqa: E402
from alphatriangle_b200.dist import Publisher  # noqa: E402
from alphatriangle_b200.packing import pack_states  # noqa: E402
from alphatriangle_b200.synth import synth_state_dict, synth_states  # noqa: E402

rank, local = int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
B = 8192
env, mc = at.EnvConfig(), at.ModelConfig()
nn = at.NeuralNetwork(mc, env, None, dev, precision="f16", max_batch=B)
nn.set_weights(synth_state_dict(nn.get_weights(), 1))
base = pack_states(synth_states(256, env, seed=5 + rank), env)
x = torch.from_numpy(np.tile(base, (-(-B // 256), 1))[:B]).to(dev)
pub = Publisher(nn)
pub.publish(src=0)
pub.wait()
for _ in range(5):
    nn.engine.evaluate_packed(x, nn.precision)
torch.cuda.synchronize()
dist.barrier()
n, every = 96, 32
evs = [torch.cuda.Event(enable_timing=True) for _ in range(n + 1)]
host_pub, wall = [], []
evs[0].record()
for i in range(n):
    t0 = time.perf_counter()
    if i % every == 2:
        pub.publish(src=0)
        host_pub.append((time.perf_counter() - t0) * 1e3)
    nn.engine.evaluate_packed(x, nn.precision)
    evs[i + 1].record()
    torch.cuda.current_stream(dev).synchronize()
    wall.append((time.perf_counter() - t0) * 1e3)
pub.wait()
torch.cuda.synchronize()
dev_ms = [evs[i].elapsed_time(evs[i + 1]) for i in range(n)]
med = sorted(dev_ms)[n // 2]
if rank in (0, 1):
    print(f"rank {rank}: median step {med:.3f} ms (device, event to event); host time of publish(): {[round(v, 3) for v in host_pub]}")
    for k in range(0, n, every):
        seg = dev_ms[k:k + 8]
        print(f"rank {rank}: steps {k}..{k + 7} device ms {[round(v, 2) for v in seg]} wall ms {[round(v, 2) for v in wall[k:k + 8]]}")
    print(f"rank {rank}: extra device time over the median, summed over all steps: {sum(v - med for v in dev_ms):.2f} ms for {len(host_pub)} publications", flush=True)
dist.barrier()
dist.destroy_process_group()
