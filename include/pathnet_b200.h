/*
 * pathnet_b200.h -- C ABI of libpathnet_b200.so
 *
 * B200 (sm_100a) implementation of the PathNet-gym hot path: population-batched
 * forward + A3C loss + backward of the module super-net over each genotype's
 * active path, per-tensor-clipped freeze-masked RMSProp, and the genotype
 * tournament / mutation.  The reference (jan-hofmeier/pathnet-gym) has no FFI
 * layer: the path sits behind plain Python call signatures that end in
 * TensorFlow-1 kernels.  Each entry point below cites the reference interface
 * it replaces (file:line relative to the reference checkout).
 *
 * Conventions
 *   - return 0 on success, negative pn_status on error; the message is kept per
 *     handle (pn_last_error) and per thread for handle-less calls
 *     (pn_last_global_error).  No C++ exception crosses this boundary.
 *   - pointers named *_dev are device pointers owned by the caller, *_host are
 *     host pointers.  `stream` is a cudaStream_t passed as void* (NULL = legacy
 *     default stream).  All work is enqueued on `stream`; no call synchronises
 *     unless its comment says so.
 *   - a handle is bound to one device and is not thread-safe.
 *   - genotypes are bitmasks: masks[p*L + l] bit j  <=>  geopath[l][j] == 1.0 of
 *     path p (pathnet.py:25-30, game_ac_network.py:158-170).  M <= 32.
 *   - transitions are laid out path-major: t = p*B + b.
 */
#ifndef PATHNET_B200_H
#define PATHNET_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define PN_API __attribute__((visibility("default")))
#else
#define PN_API
#endif

#define PN_ABI_VERSION 1
#define PN_MAX_LAYERS 8
#define PN_MAX_MODULES 32

typedef enum pn_status {
    PN_OK = 0,
    PN_ERR_INVALID = -1,   /* bad argument / unsupported configuration */
    PN_ERR_CUDA = -2,      /* a CUDA runtime call failed */
    PN_ERR_STATE = -3,     /* call order (e.g. slabs not bound, genotypes not set) */
    PN_ERR_NOMEM = -4
} pn_status;

/* Super-net shape.  Mirrors the FLAGS fields and hard-coded sizes read by
 * GameACPathNetNetwork.__init__ (game_ac_network.py:124-155,180):
 * L, M, kernel_num, stride_size, feature_num=8, hidden=256, the input shape
 * and max(ACTION_SIZEZ). */
typedef struct pn_config {
    int32_t L;              /* layers incl. the linear last layer (FLAGS.L)        */
    int32_t M;              /* modules per layer (FLAGS.M), <= 32                  */
    int32_t H, W, C;        /* frame shape, NHWC (reference fork: 160,120,4)       */
    int32_t F;              /* conv output channels per module (feature_num) = 8   */
    int32_t hidden;         /* width of the linear modules = 256                   */
    int32_t A;              /* policy head width = max(ACTION_SIZEZ), <= 31        */
    int32_t kernel[PN_MAX_LAYERS];   /* FLAGS.kernel_num, conv layers 0..L-2       */
    int32_t stride[PN_MAX_LAYERS];   /* FLAGS.stride_size                          */
    int32_t P;              /* paths (genotypes) evaluated by THIS handle / GPU     */
    int32_t B_max;          /* max transitions per path per call                   */
    int32_t device;         /* CUDA device ordinal                                 */
    int32_t flags;          /* debug: bit 0 / bit 1 force the fp32 CUDA-core kernels for layer 0 / the linear modules,
                             * bit 2 the simple (untiled) kernels for the middle conv layers,
                             * bit 3 the tiled CUDA-core kernels (instead of tcgen05) for the middle conv layers;
                             * bit 4 (PN_FLAG_LSTM): the LSTM variant of the network, see pn_lstm_bind_state */
} pn_config;

/* cfg.flags bit 4: GameACPathNetLSTMNetwork (game_ac_network.py:303-521, the reference's default, constants.py:31):
 * the last layer's module sum is not divided by M and feeds a BasicLSTMCell(256) unrolled over each path's B
 * transitions (time order b = 0..B-1); the cell outputs feed the heads.  Two tensors are appended to the slab in
 * get_vars() order: kind 8 = W_lstm [512, 1024] (rows 0..255 input, 256..511 previous output; gate order i, j, f, o)
 * and kind 9 = b_lstm [1024]; neither is ever frozen.  Needs hidden = 256. */
#define PN_FLAG_LSTM 16
/* cfg.flags bit 6 (debug / comparison): keep every backward kernel on the caller's stream instead of running the middle
 * conv layers' weight-gradient kernels on the handle's side stream beside the data-gradient chain. */
#define PN_FLAG_SERIAL_BWD 64

/* One row per trainable tensor, in get_vars() order with nothing frozen
 * (game_ac_network.py:239-253): for l<L-1, j<M: W_conv[l,j], b_conv[l,j];
 * for j<M: W_lin[j], b_lin[j]; W_fc2, b_fc2, W_fc3, b_fc3.
 * kind: 0 conv W (HWIO), 1 conv b, 2 lin W [in,hidden], 3 lin b, 4 fc2 W, 5 fc2 b, (8 / 9: W_lstm / b_lstm)
 * 6 fc3 W, 7 fc3 b.  offset/length are in floats into the slabs; every offset
 * is a multiple of 32 floats (the gaps are never read or written). */
typedef struct pn_tensor_desc {
    int64_t offset;
    int64_t length;
    int32_t layer;
    int32_t module;         /* -1 for the heads */
    int32_t kind;
    int32_t reserved;
} pn_tensor_desc;

typedef struct pn_handle pn_handle;

/* ---- lifecycle -------------------------------------------------------- */

PN_API int pn_abi_version(void);

/* Replaces GameACPathNetNetwork.__init__ graph construction
 * (game_ac_network.py:114-204).  Allocates only private workspace
 * (activations, ReLU masks, gradient partials, module->path lists).
 * Synchronises the device. */
PN_API int pn_create(const pn_config* cfg, pn_handle** out);
PN_API void pn_destroy(pn_handle* h);
PN_API const char* pn_last_error(const pn_handle* h);
PN_API const char* pn_last_global_error(void);

/* Layout of the parameter slab == get_vars() order (game_ac_network.py:239-253). */
PN_API int64_t pn_slab_floats(const pn_handle* h);
PN_API int pn_num_tensors(const pn_handle* h);
PN_API int pn_tensor_table(const pn_handle* h, pn_tensor_desc* out, int capacity);
/* flattened size of the last conv output fed to the linear modules
 * (`last_lin_num`, game_ac_network.py:133-135) */
PN_API int pn_last_lin(const pn_handle* h);
/* bytes of private device workspace held by the handle */
PN_API int64_t pn_workspace_bytes(const pn_handle* h);

/* Bind caller-owned fp32 device slabs of pn_slab_floats() floats each:
 * variables, the RMSProp "rms" slot, optional "momentum" slot (NULL when the
 * momentum coefficient is 0, as in doom_pathnet.py:97-102) and the gradient
 * slab (tf.gradients output, a3c_training_thread.py:55-60).  Borrowed until
 * rebound or pn_destroy. */
PN_API int pn_bind_slabs(pn_handle* h, float* params_dev, float* ms_dev, float* mom_dev, float* grads_dev);

/* Replaces pathnet.geopath_insert: L*M scalar assigns per worker
 * (pathnet.py:20-23, doom_pathnet.py:216-221,261-266).  masks_host[P*L] are the
 * EFFECTIVE masks (genotype OR fixed path).  Rebuilds the module->path lists
 * used by the weight-gradient kernels.  Synchronises `stream`. */
PN_API int pn_set_genotypes(pn_handle* h, const uint32_t* masks_host, void* stream);

/* LSTM variant only: the cell state every path starts its sequence from, and where the state after the last
 * step goes.  Both are caller-owned device arrays [P][2][256] = (c, h) per path (initial_lstm_state0 / 1,
 * game_ac_network.py:403-406), borrowed until rebound; state_in NULL = zeros (reset_state, :441-443), state_out NULL =
 * not written.  Acting (pn_forward with B = 1) carries the state by binding last call's output as the next input
 * (run_policy_and_value, :448-456); run_value (:468-481) simply does not adopt its output; a training call binds
 * the state the rollout started from (a3c_training_thread.py:113-114,200). */
PN_API int pn_lstm_bind_state(pn_handle* h, const float* state_in_dev, float* state_out_dev);

/* Replaces set_fixed_path + get_vars_idx + the op selection in process()
 * (game_ac_network.py:236-271, a3c_training_thread.py:205-209).
 * frozen_host[num_tensors]: 1 = the tensor's update op is not run. */
PN_API int pn_set_frozen(pn_handle* h, const uint8_t* frozen_host);

/* ---- the hot path ------------------------------------------------------ */

/* Replaces sess.run([pi, v]) -- run_policy_and_value / run_policy / run_value
 * (game_ac_network.py:182-201,209-227) for all P paths at once.
 * frames_dev [P,B,H,W,C]: uint8 (value k means k/255, game_state.py:49-50) when
 * frames_are_u8 != 0, else float32.  pi_dev [P,B,A], v_dev [P,B]. */
PN_API int pn_forward(pn_handle* h, const void* frames_dev, int frames_are_u8, int B,
               float* pi_dev, float* v_dev, void* stream);

/* Replaces the training sess.run of process() up to (not including) the apply
 * ops: forward, prepare_loss (game_ac_network.py:20-59), tf.gradients
 * (a3c_training_thread.py:55-60,205-216) for every path, summed over the P
 * local paths into the bound gradient slab (overwritten, not accumulated).
 * actions_dev [P,B] int32 (the one-hot `a` feed; a NEGATIVE action marks a padded transition that contributes
 * nothing to the loss or the gradients -- see pn_nstep_returns), td_dev [P,B], r_dev [P,B].
 * loss_dev [P] (total_loss per path) may be NULL; pi_dev / v_dev may be NULL. */
PN_API int pn_train_step(pn_handle* h, const void* frames_dev, int frames_are_u8,
                  const int32_t* actions_dev, const float* td_dev, const float* r_dev, int B,
                  float entropy_beta, float* loss_dev, float* pi_dev, float* v_dev, void* stream);

/* Same as pn_train_step but every array is a HOST pointer (pinned memory is
 * fastest); host->device copies of the inputs and the device->host copy of
 * loss_host [P] are enqueued on `stream` inside the call.  This is the
 * reference-facing shape of the call: process() feeds NumPy arrays
 * (a3c_training_thread.py:210-216).  Synchronises `stream` before returning. */
PN_API int pn_train_step_host(pn_handle* h, const void* frames_host, int frames_are_u8,
                       const int32_t* actions_host, const float* td_host, const float* r_host, int B,
                       float entropy_beta, float* loss_host, void* stream);

/* Pipelined form of pn_train_step_host for callers that feed a stream of batches
 * (process() is called in a loop, doom_pathnet.py:197-201): pn_stage_inputs_host
 * enqueues the host->device copies of one batch on `copy_stream` into staging slot
 * 0 or 1 and returns; pn_train_step_staged makes `stream` wait for that slot, runs
 * the step and enqueues the device->host copy of loss_host [P] (no synchronisation:
 * read loss_host after synchronising `stream`).  Staging batch i+1 while step i
 * runs hides the PCIe transfer behind the kernels.  Host arrays must stay valid
 * (and should be pinned) until their copies have completed. */
PN_API int pn_stage_inputs_host(pn_handle* h, int slot, const void* frames_host, int frames_are_u8,
                                const int32_t* actions_host, const float* td_host, const float* r_host, int B,
                                void* copy_stream);

/* pn_stage_inputs_host for a rollout: consecutive transitions of one rollout share C-1 of the C planes of their frame
 * stacks (GameState keeps the C most recent frames: game_state.py:62-66 reset, :75-81 step), so the host ships only
 * the B + C - 1 distinct planes of every path and the stacks are rebuilt on the device:
 *     planes_host [P][B + C - 1][H][W] u8;  planes[p][0..C-1] = the channels of the rollout's first state (after a
 *     reset: C copies of the first frame), planes[p][C-1+i] = the frame observed after step i;
 *     frames[p][b][y][x][c] = planes[p][b + c][y][x].
 * 3.7x fewer bytes over PCIe at B = 32, C = 4.  Then call pn_train_step_staged as for pn_stage_inputs_host.
 * Every read of the host arrays is enqueued on `copy_stream` (once it has drained they may be reused); the stacks are
 * rebuilt on a stream of the handle's own, a group of paths at a time, while the next group's planes are in flight. */
PN_API int pn_stage_rollout_host(pn_handle* h, int slot, const uint8_t* planes_host, const int32_t* actions_host,
                                 const float* td_host, const float* r_host, int B, void* copy_stream);
PN_API int pn_train_step_staged(pn_handle* h, int slot, float entropy_beta, float* loss_host, void* stream);

/* Multi-GPU overlap (the reference has one process and no exchange).  The gradient slab is
 * ordered like the parameters: conv tensors, linear-module tensors, head tensors.  The linear modules hold
 * ~97 % of the gradient bytes and their weight-gradient kernel runs BEFORE the conv layers' backward pass, so
 * the host can start summing that region over the GPUs while the rest of the step still runs:
 *   pn_lin_grads_region    -> offset / length (in floats) of the W_lin, b_lin region inside the gradient slab
 *   pn_stream_wait_lin_grads -> `stream` waits until the most recent pn_train_step* has written that region
 * The other two regions are complete when the step's stream work is -- and all of them except the first conv layer's
 * tensors (the first *late_floats floats of the slab) already before that layer's weight-gradient kernel, the longest
 * of the step, starts:
 *   pn_stream_wait_early_grads -> `stream` waits until the most recent pn_train_step* has written every gradient
 *                                 outside [0, *late_floats)  (late_floats may be NULL)
 * No counterpart in the reference. */
PN_API int pn_lin_grads_region(const pn_handle* h, int64_t* offset_floats, int64_t* count_floats);
PN_API int pn_stream_wait_lin_grads(pn_handle* h, void* stream);
PN_API int pn_stream_wait_early_grads(pn_handle* h, void* stream, int64_t* late_floats);

/* The exchange kernels wait for the other ranks' flags for at most 2 s; a rank that gives up records that in its own
 * exchange buffer and finishes (the sums of that step are then wrong).  This synchronising call reads the record:
 * PN_OK, or PN_ERR_STATE with a message naming the wait that timed out.  Call it whenever the host synchronises
 * anyway (end of an epoch, before a checkpoint). */
PN_API int pn_peer_check(pn_handle* h, const void* my_peer_buf_dev);

/* Sum of the two small gradient regions (everything but the linear modules: conv layers + heads, ~140 KB) over the
 * GPUs of one node in ONE kernel over NVLink peer memory, instead of an all-reduce that would sit exposed at the end of
 * the step: every rank publishes its values in an exchange buffer that all ranks have mapped (symmetric memory of
 * pn_peer_buffer_bytes() bytes, zero-initialised, allocated and exchanged by the host, e.g.
 * torch.distributed._symmetric_memory), raises a flag on every peer, waits for the peers and adds the copies in rank
 * order (bit-identical on all ranks).  peer_bufs_host[r] = address of rank r's buffer in THIS process; step = 0, 1, 2, ...
 * counted PER SET OF BUFFERS (a fresh zeroed set starts again at 0), identical on all ranks.  No counterpart in the reference (single process per parameter server). */
PN_API int64_t pn_peer_buffer_bytes(const pn_handle* h);
PN_API int pn_peer_sum_small(pn_handle* h, const uint64_t* peer_bufs_host, int rank, int world, uint32_t step, void* stream);

/* pn_peer_sum_small in two parts, so that only the first conv layer's gradients are exchanged after the step's last
 * kernel: part 1 = the small regions outside [0, late_floats) (enqueue it on a stream that waits with
 * pn_stream_wait_early_grads: it then runs underneath the first layer's weight-gradient kernel), part 2 = [0, late_floats)
 * (on the compute stream after the step), part 0 = both (= pn_peer_sum_small).  Each part keeps its own flags in the
 * exchange buffers and its own step count; a host uses either part 0 or parts 1 + 2 on one set of buffers. */
PN_API int pn_peer_sum_part(pn_handle* h, const uint64_t* peer_bufs_host, int rank, int world, uint32_t step, int part,
                            void* stream);

/* All-reduce of the linear modules' gradient region (pn_lin_grads_region) in place over NVLink peer memory, for hosts
 * whose GRADIENT SLABS are themselves symmetric memory (peer_slabs_host[r] = rank r's slab base in this process,
 * peer_slabs_host[rank] must be the slab bound with pn_bind_slabs; peer_bufs_host = the exchange buffers of
 * pn_peer_sum_small, whose control words carry the flags).  Rank r adds shard r of all slabs in rank order and writes the
 * sum into shard r of every slab, between a "gradients ready" and a "sums written" barrier; bit-identical results on all
 * ranks.  Enqueue it where the gradients are final (after pn_stream_wait_lin_grads); step = 0, 1, 2, ... per set of
 * exchange buffers.  No counterpart in the reference. */
PN_API int pn_peer_allreduce_lin(pn_handle* h, const uint64_t* peer_slabs_host, const uint64_t* peer_bufs_host, int rank,
                          int world, uint32_t step, void* stream);

/* Replaces RMSPropApplier.apply_gradients for the non-frozen tensors
 * (rmsprop_applier.py:79-106): per tensor g' = clip_by_norm(g, clip_norm);
 * ms += (g'^2 - ms)(1-decay); mom = momentum*mom + lr*g'/sqrt(ms+eps);
 * var -= mom.  Reads the bound gradient slab (all-reduce it first when the
 * population is sharded over GPUs). */
PN_API int pn_rmsprop_apply(pn_handle* h, float lr, float decay, float momentum, float epsilon,
                     float clip_norm, void* stream);

/* The same update for the tensors [first_tensor, first_tensor + n_tensors) of the table only, or (complement != 0)
 * for every tensor outside that range (the update of one tensor never depends on another: rmsprop_applier.py:92-106
 * loops over (grad, var) pairs).  Lets the host update the linear modules -- whose summed gradients are ready first,
 * see pn_stream_wait_lin_grads -- on a second stream while the conv layers' backward pass still runs, and the rest
 * afterwards.  Over one step the calls must cover every tensor exactly once. */
PN_API int pn_rmsprop_apply_range(pn_handle* h, int first_tensor, int n_tensors, int complement, float lr, float decay,
                           float momentum, float epsilon, float clip_norm, void* stream);

/* Replaces the re-initialisation loop at task switch (doom_pathnet.py:289-293):
 * every non-frozen tensor <- backup_params_dev (same slab layout).  RMSProp
 * slots are not reset (as in the reference). */
PN_API int pn_reinit_unfrozen(pn_handle* h, const float* backup_params_dev, void* stream);

/* No counterpart in the reference.  The same update as pn_rmsprop_apply for hosts
 * that cannot allocate device memory when they call it (the GPU was handed over
 * fully reserved, cudaMalloc's device-wide synchronisation is not acceptable, or
 * the host forbids plugins to own device memory): no handle, no cudaMalloc.  The
 * caller passes the slabs, the tensor table (pn_tensor_desc rows, host memory) and
 * optional per-tensor frozen flags (host, may be NULL).  Scratch is a built-in
 * 64 KiB array per device: at most 128 tensors and 16384 chunks of 2048 floats
 * (~33.5 M parameters); larger inputs return PN_ERR_INVALID.  The current device
 * must own the slabs.  All calls on a device share that array: calls that may
 * overlap on the GPU race and must be serialised by the caller.  The first call on
 * each device resolves the array's address -- make it before any stream capture. */
PN_API int pn_rmsprop_apply_noalloc(float* params_dev, float* ms_dev, float* mom_dev, const float* grads_dev,
                                    const pn_tensor_desc* tensors_host, const uint8_t* frozen_host,
                                    int num_tensors, float lr, float decay, float momentum, float epsilon,
                                    float clip_norm, void* stream);

/* ---- measurement support (no counterpart in the reference) ------------- */

/* Which engine each layer of this handle runs on ("conv0=tcgen05,conv1=tcgen05,lin=tcgen05"): geometries the
 * tensor-core kernels do not take (and cfg.flags) fall back to the CUDA-core kernels, and this makes that visible.
 * Returns the length of the full string. */
PN_API int pn_engine_paths(const pn_handle* h, char* out, int capacity);

/* Kernels launched through this handle since pn_create. */
PN_API int64_t pn_launch_count(const pn_handle* h);

typedef struct pn_profile_entry {
    char name[40];
    float total_ms;         /* summed CUDA-event time of the recorded launches */
    int32_t launches;       /* launches the sum covers (<= 64 most recent)     */
} pn_profile_entry;

/* Development aid: with PN_TC_DEBUG=1 in the environment at pn_create, the
 * tensor-core layer-0 kernel records per-CTA wait cycles; copies them out. */
PN_API int pn_debug_read(pn_handle* h, long long* out_host, int capacity);

/* Test aid: the ReLU decisions the LAST pn_train_step / pn_forward call with batch B made in PathNet layer `layer`
 * (game_ac_network.py:188-192: relu(...) of every active module).  Synchronises the device and copies to host:
 *   layer < L-1 : [P*B][M][OH*OW] bytes, bit c = output channel c of that pixel was > 0
 *   layer = L-1 : [P*B][M][hidden] bytes, 1 = the unit was > 0
 * Rows of modules that are not active for a path are undefined.  Returns the number of bytes written, or a
 * negative error code (capacity too small, bad layer).  The parity tests use it to separate arithmetic
 * differences from ReLU decisions that lie below fp32 resolution. */
PN_API int64_t pn_debug_read_relu_masks(pn_handle* h, int layer, int B, uint8_t* out_host, int64_t capacity_bytes);

/* When on, every kernel launch is bracketed by cudaEventRecord on the launch
 * stream.  Resets the counters. */
PN_API int pn_profile_enable(pn_handle* h, int on);
/* Synchronises the device; returns the number of entries written (kernels
 * that ran at least once). */
PN_API int pn_profile_read(pn_handle* h, pn_profile_entry* out, int capacity);

/* ---- genotype ops (no handle needed) ----------------------------------- */
/* Counter-based RNG: Philox4x32-10, key = seed, counter =
 * (epoch, path_id, position, domain<<24 | slot); u = (x>>11) * 2^-53. */

/* Replaces pathnet.get_geopath (pathnet.py:78-87) for n_paths genotypes:
 * masks_dev[(i)*L + l], path ids path_id0 + i. */
PN_API int pn_geopath_init(uint32_t* masks_dev, int n_paths, int path_id0, int L, int M, int N,
                    uint64_t seed, uint32_t epoch, void* stream);

/* Replaces pathnet.mutation (pathnet.py:50-63), in place, one genotype per path. */
PN_API int pn_mutate(uint32_t* masks_dev, int n_paths, int path_id0, int L, int M, int N,
              uint64_t seed, uint32_t epoch, void* stream);

/* Replaces one controller pass (doom_pathnet.py:225-270) for n_groups disjoint
 * candidate groups of Bcand paths drawn from a Philox-keyed permutation of the
 * P_total paths (n_groups = 1 is exactly one reference tournament): a group
 * with any score == -1000 is skipped; else winner = first arg-max, every other
 * member <- mutation(copy(winner)), all members' scores <- -1000.
 * genotypes_dev [P_total,L] raw genotypes (in/out), scores_dev [P_total] (in/out),
 * fixed_dev [L] frozen-path masks, effective_dev [P_total,L] = genotype | fixed (out),
 * winners_dev [n_groups] winner path id or -1 (out). */
PN_API int pn_tournament(uint32_t* genotypes_dev, float* scores_dev, const uint32_t* fixed_dev,
                  uint32_t* effective_dev, int32_t* winners_dev, int P_total, int L, int M, int N,
                  int Bcand, int n_groups, uint64_t seed, uint32_t epoch, void* stream);

/* ---- rollout bookkeeping (the host code between two network calls of a rollout) ---- */

/* Replaces A3CTrainingThread.choose_action (a3c_training_thread.py:89-90), np.random.choice(range(A), p=pi), for n
 * probability rows at once: numpy's algorithm (fp64 cumsum, normalise by the last entry, searchsorted side='right')
 * on ONE uniform per row from the Philox stream, counter (epoch, id0 + row, 0, 3<<24), key = seed. */
PN_API int pn_sample_actions(const float* pi_dev, int n, int A, uint64_t seed, uint32_t epoch, uint32_t id0,
                      int32_t* actions_dev, void* stream);

/* Replaces the batch assembly of A3CTrainingThread.process (a3c_training_thread.py:136,155-180) for P paths: rewards
 * clipped to [-1,1]; R <- r_i + gamma*R backwards from bootstrap[p] (V(s_T), or 0 after a terminal step); td = R - V_i;
 * fp64 scan like the Python floats, rounded once to fp32.  rewards/values/R/td are [P,B] in rollout order;
 * lengths_dev[p] <= B (NULL: all B) is the number of transitions path p really has -- the rest are padding and get
 * R = td = 0; give them action = -1 in pn_train_step*, which then contribute nothing to loss or gradients. */
PN_API int pn_nstep_returns(const float* rewards_dev, const float* values_dev, const float* bootstrap_dev,
                     const int32_t* lengths_dev, int P, int B, double gamma, float* R_dev, float* td_dev, void* stream);

/* Replaces GameState._process_frame + the stack update (game_state.py:37-51,62-66,75-81) for n environments:
 * obs_dev [n,Hin,Win,3] u8 -> cv2.cvtColor(COLOR_BGR2GRAY) -> cv2.resize(.., (W,H)) (INTER_LINEAR), both bit-exact
 * restatements of OpenCV's 8-bit fixed-point paths -> pushed into frames_dev [n,H,W,C] u8 in place
 * (channels shift down by one, the new frame becomes channel C-1; where reset_dev[e] != 0 all C channels are set to
 * the new frame, as GameState.reset does).  The u8 value k stands for k/255, the reference's fp32 scaling. */
PN_API int pn_preprocess_frames(const uint8_t* obs_dev, int n, int Hin, int Win, uint8_t* frames_dev, int H, int W,
                         int C, const uint8_t* reset_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PATHNET_B200_H */
