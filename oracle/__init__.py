"""CPU oracle of the leaf-evaluation path -- TEST INFRASTRUCTURE, not product code.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this
package. It restates the reference algorithm (each function cites the file:line it
follows) and is pinned against the unmodified reference executed in the authoring
container: tests/golden/make_golden.py imported /root/reference, ran it on seeded states
and weights, and committed its outputs under tests/golden/. The reference's own test
suite holds no golden vectors for this path, so those fixtures are
the pin.

  features_ref.py   numpy restatement of alphatriangle/features (what the parity tests and smoke() check against)
  model_ref.py      torch-CPU fp32 restatement of AlphaTriangleNet.forward / NeuralNetwork.evaluate_batch
  features_ref.c    the feature path once more in plain C over PackedState records (make -C oracle, built by
                    __graft_entry__.build()); tests/test_oracle_c.py pins it to the same reference vectors
"""
