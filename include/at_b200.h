/*
 * at_b200.h -- C ABI of the B200 leaf-evaluation library (libat_b200.so).
 *
 * The reference (lguibr/alphatriangle) has no FFI of its own: its boundary for
 * this path is the duck-typed Python protocol `NeuralNetwork.evaluate_state /
 * evaluate_batch / get_weights / set_weights` (alphatriangle/nn/network.py:165,
 * 243,320,325) that trimcts calls. Every entry point below names the reference
 * code it stands in for. All functions return 0 on success or a negative
 * AT_E_* code; `at_last_error()` holds the message of the calling thread's last
 * failure. Nothing here throws, allocates through torch, or takes torch types:
 * pointers are raw device (d_*) or host (h_*) addresses, `stream` is a
 * cudaStream_t passed as void* (NULL = the legacy default stream).
 */
#ifndef AT_B200_H
#define AT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AT_ABI_VERSION 1

/* error codes */
#define AT_OK 0
#define AT_E_INVALID (-1)     /* bad argument / unsupported configuration */
#define AT_E_CUDA (-2)        /* a CUDA runtime call failed */
#define AT_E_STATE (-3)       /* call order violated (e.g. evaluate before weights were loaded) */
#define AT_E_UNSUPPORTED (-4) /* valid model, but the requested precision path cannot run it */

/* bits of the per-call status word written by the evaluate entry points; they
 * map one-to-one onto the reference's isfinite checks (network.py:147-154,
 * 260-274), which the Python host turns into NetworkEvaluationError. */
#define AT_FLAG_NONFINITE_GRID 1u
#define AT_FLAG_NONFINITE_OTHER 2u
#define AT_FLAG_NONFINITE_POLICY_LOGITS 4u
#define AT_FLAG_NONFINITE_VALUE_LOGITS 8u
#define AT_FLAG_NONFINITE_POLICY_PROBS 16u

/* PackedState: 80-byte record, see alphatriangle_b200/packing.py for the field map. */
#define AT_STATE_BYTES 80
#define AT_MAX_CELLS 128
#define AT_MAX_SLOTS 4
#define AT_MAX_CONVS 8
#define AT_MAX_MLP 4

enum at_activation { AT_ACT_NONE = 0, AT_ACT_RELU = 1, AT_ACT_GELU = 2, AT_ACT_SILU = 3, AT_ACT_TANH = 4, AT_ACT_SIGMOID = 5 };

/* precision of the forward pass */
enum at_precision {
  AT_PREC_FP32 = 0, /* fp32 CUDA-core path, any configuration; parity max-abs 1e-4 */
  AT_PREC_BF16 = 1, /* tcgen05/TMEM path, bf16 operands (8-bit mantissa, fp32's range), fp32 accumulate */
  AT_PREC_F16 = 2   /* the same kernels with fp16 operands (11-bit mantissa; conversions saturate at +-65504): 8x smaller
                     * rounding error on every operand, the format to use for trained networks whose activations stay
                     * in fp16's range (any network with normalisation layers); parity policy KL <= 1e-3 */
};

/* Game geometry: trianglengin.EnvConfig as read at extractor.py:25,39,50. */
typedef struct at_env_desc {
  int32_t rows, cols, slots;
} at_env_desc;

/* Network dimensions: ModelConfig (alphatriangle/config/model_config.py:17-59),
 * with CONV_* already resolved to integer (kh,kw),(sh,sw),(ph,pw). */
typedef struct at_model_desc {
  at_env_desc env;
  int32_t grid_channels;
  int32_t n_conv;
  int32_t conv_filters[AT_MAX_CONVS];
  int32_t conv_kernel[AT_MAX_CONVS][2];
  int32_t conv_stride[AT_MAX_CONVS][2];
  int32_t conv_pad[AT_MAX_CONVS][2];
  int32_t n_res_blocks;
  int32_t res_filters;
  int32_t use_transformer; /* USE_TRANSFORMER && TRANSFORMER_LAYERS > 0 */
  int32_t t_dim, t_heads, t_layers, t_ff;
  int32_t n_shared, shared_dims[AT_MAX_MLP];
  int32_t n_policy, policy_dims[AT_MAX_MLP];
  int32_t n_value, value_dims[AT_MAX_MLP];
  int32_t num_atoms;
  float v_min, v_max;
  int32_t activation; /* enum at_activation */
  int32_t use_batch_norm;
  int32_t other_dim; /* OTHER_NN_INPUT_FEATURES_DIM */
} at_model_desc;

typedef struct at_engine at_engine; /* opaque */

int at_abi_version(void);
const char* at_last_error(void);

/* ---- features ---------------------------------------------------------- */

/* extract_state_features for a batch (alphatriangle/features/extractor.py:150-171,
 * grid_features.py:7-42): PackedState[B] -> grid f32 [B,grid_channels,rows,cols]
 * and other f32 [B,other_dim] (pad / truncate per extractor.py:127-139). Bit-exact. */
int at_extract_features(const at_env_desc* env, int grid_channels, int other_dim, const void* d_states, int64_t batch,
                        float* d_grid, float* d_other, void* stream);

/* ---- model description (no GPU needed) ---------------------------------- */

/* The weight-image table for a configuration, as text: one line per float tensor of
 * the reference state_dict (module tree built at alphatriangle/nn/model.py:116-266),
 * "<key> <float offset> <numel>\n". Returns the number of bytes needed (excluding the
 * terminating NUL) or a negative AT_E_* code; writes at most `cap` bytes. */
int64_t at_model_describe(const at_model_desc* desc, char* buf, int64_t cap);

/* ---- engine ------------------------------------------------------------ */

/* NeuralNetwork.__init__ (network.py:40-67): validates the configuration, sizes
 * the weight image and allocates engine-owned device memory for batches of up to
 * `max_batch` states on CUDA device `device`. */
int at_engine_create(const at_model_desc* desc, int device, int64_t max_batch, at_engine** out);
int at_engine_destroy(at_engine* e);

/* The weight image is every floating-point tensor of the reference state_dict
 * (network.py:320-323) concatenated as fp32. The table below is the single
 * source of truth for its order: entry i is the reference key name (e.g.
 * "res_body.0.conv1.1.running_var"), its float offset and element count. Offsets are
 * multiples of 4 floats (a tensor whose size is not leaves up to three unused floats
 * behind it), so a host can keep its parameters IN the image, as 16-byte-aligned views,
 * and publish them without packing (alphatriangle_b200/flat.py). */
int64_t at_engine_weight_floats(const at_engine* e);
int at_engine_weight_entries(const at_engine* e);
int at_engine_weight_entry(const at_engine* e, int index, const char** name, int64_t* offset, int64_t* numel);

/* NeuralNetwork.set_weights (network.py:325-336) for the engine's copy: reads the
 * fp32 weight image from DEVICE memory (e.g. straight out of an NCCL receive
 * buffer), folds eval-mode BatchNorm, permutes / casts / tiles for the kernels
 * into the engine's inactive weight set and makes it the active one once the
 * work queued on `stream` completes. `version` is an opaque tag (the trainer's
 * global_step, worker_manager.py:184-187) returned by at_engine_weight_version. */
int at_engine_load_weights(at_engine* e, const float* d_image, int64_t n_floats, int64_t version, void* stream);
/* The same, for a worker that keeps evaluating while new weights arrive (the reference's workers go on playing while
 * WorkerManager.update_worker_networks pushes, alphatriangle/training/worker_manager.py:169-209): the fold / repack is
 * queued on `stream` -- typically the side stream the NCCL broadcast ran on -- into the inactive set, and evaluations on
 * other streams keep reading the OLD set, without waiting, until that work has finished; the first evaluate call after
 * that switches over. at_engine_weight_version reports the set evaluations currently read. An evaluation reads one set
 * from start to end: never a mix. */
int at_engine_publish_weights(at_engine* e, const float* d_image, int64_t n_floats, int64_t version, void* stream);
int64_t at_engine_weight_version(const at_engine* e);

/* 1 if the configuration can run on the tcgen05 bf16 path, else 0 (reason in at_last_error). */
int at_engine_supports(const at_engine* e, int precision);

/* AlphaTriangleNet.forward + softmax + expected value on prepared feature tensors
 * (model.py:268-297, network.py:258-278): grid f32 [B,grid_channels,rows,cols]
 * contiguous, other f32 [B,other_dim] -> policy probs f32 [B,action_dim],
 * expected value f32 [B]; *d_flags (uint32, device) is OR-ed with AT_FLAG_* bits. */
int at_engine_forward(at_engine* e, int precision, const float* d_grid, const float* d_other, int64_t batch, float* d_policy,
                      float* d_value, uint32_t* d_flags, void* stream);

/* NeuralNetwork.evaluate_batch at the tensor boundary (network.py:243-278 without
 * the Python dict building): PackedState[B] in device memory -> probs, values. */
int at_engine_evaluate_packed(at_engine* e, int precision, const void* d_states, int64_t batch, float* d_policy, float* d_value,
                              uint32_t* d_flags, void* stream);

/* No counterpart in the reference. at_engine_evaluate_packed for up to 128 states (four default MCTS leaf
 * batches) on the tensor-core path (in the operand format the engine served last: bf16 until an fp16 evaluation was made), for hosts that must not allocate device memory once they are running:
 *   - a self-play process that shares its GPU with the trainer, where a cudaMalloc / cudaFree in the search loop
 *     would synchronise the whole device and stall the trainer's streams, or fail once the trainer has taken the rest
 *     of the memory;
 *   - a league / old-vs-new evaluator holding many engines (one per model version): created with max_batch = 1 they
 *     keep their weights and a minimum of private scratch, and all of them evaluate through this call.
 * The buffers the kernels pass to each other (other-feature vectors, the tower's output tiles, the trunk FC's split-K
 * partial sums; about 5.1 MB at the default configuration) live in ONE 6 MiB static device array inside the library
 * instead of engine-owned allocations; the call itself allocates nothing (no cudaMalloc / cudaFree, no allocating
 * library routine). Inputs and outputs are still the caller's.
 *   - batch > 128, or a model whose scratch exceeds 6 MiB, is refused with an error (AT_E_INVALID / AT_E_UNSUPPORTED);
 *     it is not limited by the engine's max_batch.
 *   - The first call on each device looks up the array's device address (cudaGetSymbolAddress): make that first
 *     call before starting any stream capture.
 *   - Every engine and every call on a GPU share the array and it keeps the previous call's bytes. Nothing in it needs
 *     clearing (each call rewrites all rows it later reads), but two calls whose kernels overlap on the GPU -- different
 *     streams, threads or engines -- race on it: serialise them (one stream, or events).
 *   - Needs the tensor-core path and the fused-head shape (one shared FC layer, one hidden layer per head). */
int at_engine_evaluate_packed_static(at_engine* e, const void* d_states, int64_t batch, float* d_policy, float* d_value,
                                     uint32_t* d_flags, void* stream);

/* Same with HOST buffers (pinned memory recommended): copies the states in,
 * evaluates, copies probs / values / flags out and waits. This is the call a
 * host application (trimcts-style batcher) makes. Batches of up to 256 states in
 * pinned buffers skip the copies: the kernels read the records and write the
 * results through PCIe directly (the latency path of a single search's <= 32 leaves). */
int at_engine_evaluate_host(at_engine* e, int precision, const void* h_states, int64_t batch, float* h_policy, float* h_value,
                            uint32_t* h_flags, void* stream);

/* The same call split in two so that two batches can be in flight: submit queues the copies and kernels of one batch on
 * `slot` (0 or 1, each with its own device staging) and returns; wait blocks until that slot's probs / values / flags are
 * in the host buffers given at submit time. While slot 0's results drain over PCIe (5.9 MB per 4096 states) slot 1's
 * kernels run. A slot must be waited on before it is submitted to again; the host buffers must stay valid (and, for the
 * copies to be asynchronous, pinned) until the wait returns. Both slots share the engine's tower -> FC hand-off scratch:
 * submit both on the SAME `stream` (their kernels then run back to back; the overlap is between one slot's PCIe copies
 * and the other slot's kernels). The reference's evaluate_batch is synchronous
 * (alphatriangle/nn/network.py:242-318); this pair is what a batcher serving many searches calls instead. */
int at_engine_submit_host(at_engine* e, int precision, const void* h_states, int64_t batch, float* h_policy, float* h_value,
                          uint32_t* h_flags, int slot, void* stream);
int at_engine_wait_host(at_engine* e, int slot);

/* Introspection used by bench.py: kernels launched by the last evaluate call, and the
 * device time (CUDA events on the launching stream) of the two tcgen05 kernels of the
 * last bf16 evaluation; waits for that evaluation to finish. */
int64_t at_engine_last_launches(const at_engine* e);
/* Kernel timing is off by default (four event records per call that the serving path does not need): switch it on before the
 * evaluations at_engine_kernel_ms is asked about. */
int at_engine_set_kernel_timing(at_engine* e, int on);
int at_engine_kernel_ms(const at_engine* e, float* tower_ms, float* fc_ms);
#define AT_FLAG_PIPELINE_TIMEOUT 0x80000000u /* a tcgen05 pipeline wait timed out; bits 8..15 say which: 1 weight producer (ring
                                              * stage), 2 issuer (operands), 4 issuer (weight slab), 8 epilogue (accumulator),
                                              * 16/32/64 trunk FC (ring stage, slab, accumulator), 128 board-pair queue */
int at_engine_action_dim(const at_engine* e);

/* Test hook: the bf16 path up to and including the first shared FC layer
 * (tower + split-K FC + BatchNorm + activation, model.py:275-294), fp32 [B, FC_DIMS_SHARED[0]].
 * Lets the parity tests localise an error to the tensor-core kernels. */
int at_engine_bf16_trunk(at_engine* e, const void* d_states, int64_t batch, float* d_trunk, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* AT_B200_H */
