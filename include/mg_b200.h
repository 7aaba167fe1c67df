/*
 * mg_b200.h -- C ABI of libmagical_b200.so, the B200-native batched MAGICAL simulator.
 *
 * This is the drop-in boundary for the per-step hot path of khanhptnk/magical-1.
 * The reference has no FFI of its own (it is pure Python over pymunk/pygame/cv2), so
 * each entry point below names the reference Python interface it stands in for
 * (paths relative to /root/reference):
 *
 *   mg_create / mg_destroy    BaseEnv.__init__ / close              magical/base_env.py:78-123, 353-356
 *                             + lores_*_entry_point wrapper stack    magical/benchmarks/__init__.py:232-266
 *   mg_set_scenes             BaseEnv.reset (space + entities built  magical/base_env.py:190-246
 *                             from host-sampled on_reset() results)  magical/benchmarks/<task>.py on_reset
 *   mg_queue_resets           the roll-out loop `while not done: env.step(a)` ... `env.reset()` of
 *                             magical/misc/benchmark_env_perf.py:12-18 and evaluation.py, batched: envs that finish on
 *                             the next step start their next episode inside that step (BaseEnv.reset, base_env.py:190-246)
 *   mg_step / mg_step_host    BaseEnv.step                           magical/base_env.py:267-307
 *                               Robot.set_action / Robot.update      magical/entities.py:435-476
 *                               pm.Space.step x10                    magical/base_env.py:248-255
 *                               score_on_end_of_traj                 magical/benchmarks/<task>.py
 *                               render('rgb_array') ego/allo         magical/base_env.py:324-343, magical/render.py:385-395
 *                               ResizeDictObservation (INTER_AREA)   magical/benchmarks/__init__.py:150-190
 *                               FlattenFrameStack/EagerDictFrameStack magical/benchmarks/__init__.py:51-147
 *   mg_render                 BaseEnv.render(mode='rgb_array')       magical/base_env.py:324-343
 *   mg_score                  score_on_end_of_traj()                 magical/benchmarks/*.py
 *   mg_debug_reward           debug_shaped_reward()                  move_to_corner.py:86-100, pick_and_place.py:114-124
 *   mg_get_bodies             body.position/.angle/.velocity reads   magical/base_env.py:358-413 (debug_print_entity_spec)
 *   mg_get_state/mg_set_state (no reference equivalent: EzPickle re-constructs, base_env.py:20-46)
 *   mg_get_status             (no reference equivalent: Chipmunk grows its arbiter hash without bound and lets
 *                             NaNs propagate; this library's fixed capacities and its NaN guard report here)
 *   mg_frames_downsample      ResizeDictObservation on stored frames magical/benchmarks/__init__.py:150-190
 *   mg_frames_stack           FlattenFrameStack / EagerDictFrameStack / ChannelsFirst replayed over a stored
 *                             trajectory (preprocess_demos_with_wrapper) magical/saved_trajectories.py:87-149
 *   mg_host_entity_overlaps   space.shape_query in the placement     magical/geom.py:226-233
 *                             sampler (host-side reset only)
 *   mg_host_palette           style.COLOURS_RGB / darken / lighten   magical/style.py:10-37
 *   mg_host_sample_scene(s)   BaseEnv.reset() up to the first render magical/base_env.py:190-246, geom.py:116-384,
 *                             (phys vars, on_reset of the 9 tasks)   benchmarks/*.py on_reset; rng = RandomState stream
 *
 * Conventions: plain pointers and sizes; no exceptions cross the boundary; every call
 * returns 0 on success or a negative MG_ERR_* code and records a message readable with
 * mg_last_error(); `stream` is a cudaStream_t passed as void* (NULL = default stream);
 * pointers named *_dev are device pointers owned by the caller, *_host are host pointers.
 * One handle per GPU; calls on one handle must be serialised by the caller.
 * There is NO CPU implementation behind this ABI: mg_create fails if no CUDA device is
 * usable.
 */
#ifndef MG_B200_H
#define MG_B200_H

#include <stddef.h>
#include <stdint.h>
#include "mg_scene.h"

#ifdef __cplusplus
extern "C" {
#endif

#define MG_OK 0
#define MG_ERR_ARG (-1)
#define MG_ERR_CUDA (-2)
#define MG_ERR_NODEVICE (-3)
#define MG_ERR_PLACEMENT (-4)   /* geom.PlacementError: an entity could not be placed (geom.py:111-113, 256-259) */

#define MG_VIEW_ALLO 0
#define MG_VIEW_EGO 1
#define MG_MAX_GROUPS 8          /* 3-channel groups in a stacked observation (8 = LoResStack allo4+ego4) */
#define MG_MAX_BODIES 16         /* robot group (6) + MG_MAX_BLOCKS */

typedef struct mg_handle mg_handle;

/* Observation layout produced by mg_step: `n_groups` RGB groups concatenated on the
 * channel axis, group g = the frame of view group_view[g] that is group_age[g] steps old
 * (0 = newest).  LoRes4E: 4 groups, view EGO, ages 3,2,1,0 (oldest first,
 * benchmarks/__init__.py:122-129). */
typedef struct mg_config {
    int32_t device;               /* CUDA ordinal */
    int32_t num_envs;
    int32_t n_groups;
    int32_t group_view[MG_MAX_GROUPS];
    int32_t group_age[MG_MAX_GROUPS];
    int32_t chw;                  /* 1: obs is [B, 3*n_groups, 96, 96] (ChannelsFirst, :193-216); 0: [B,96,96,3*n_groups] */
    int32_t reserved[5];
} mg_config;

int mg_create(const mg_config *cfg, mg_handle **out);
int mg_destroy(mg_handle *h);
const char *mg_last_error(const mg_handle *h); /* h may be NULL: error of the last failed mg_create */

/* Reset `n` envs to host-sampled scenes.  env_ids_host == NULL means envs 0..n-1.
 * The frame stacks of those envs are refilled with the first rendered frame on the next
 * mg_step / mg_reset_obs (FlattenFrameStack.reset, benchmarks/__init__.py:139-147).
 * The host arrays are copied before the call returns (they may be pageable and reused at once); the upload and the
 * initialisation kernel run on `stream` without draining it. */
int mg_set_scenes(mg_handle *h, const int32_t *env_ids_host, int n, const mg_scene *scenes_host, void *stream);

/* Queue resets for the NEXT step (auto-reset): the `n` envs are re-initialised to `scenes_host` after the physics of
 * the next mg_step / mg_step_host / mg_step_physics + mg_step_render and before its render, so that the step returns
 * reward / done / eval_score of the finished episode and the first observation of the new one (what a roll-out loop
 * that calls env.reset() on done sees).  Unlike mg_step_physics + mg_set_scenes + mg_step_render this keeps the one-call
 * step, whose parts render under the physics tail of the others.  The host arrays are copied before the call returns.
 * One call per step: a second call before a step has consumed the first, and mg_set_scenes in between, are
 * rejected with MG_ERR_ARG. */
int mg_queue_resets(mg_handle *h, const int32_t *env_ids_host, int n, const mg_scene *scenes_host, void *stream);

/* Render the current state of every env into the frame stacks and write the stacked
 * observation (what reset() returns).  obs_dev: u8 [B,96,96,3*n_groups] (or CHW). */
int mg_reset_obs(mg_handle *h, uint8_t *obs_dev, void *stream);

/* One env-step for all envs: action decode, 10 physics sub-steps, horizon/score, render,
 * 4x area downsample, frame stack.  reward/done/score may be NULL.
 * score_dev receives info['eval_score'] (0 until the last step of the episode).
 * For single-view HWC stacks the batch is cut into parts on internal streams (forked from and joined back to
 * `stream`), so that the rendering of finished parts runs under the tail of the physics of the others; the
 * results are those of mg_step_physics followed by mg_step_render.  While mg_set_timing is on, or with
 * MG_STEP_PARTS=1 in the environment, the plain sequence runs. */
int mg_step(mg_handle *h, const int32_t *actions_dev, uint8_t *obs_dev, float *reward_dev,
            uint8_t *done_dev, float *score_dev, void *stream);

/* Physics + score only (no render); used by mg_step and by callers that reset done envs
 * between the physics and the render of a step.  */
int mg_step_physics(mg_handle *h, const int32_t *actions_dev, float *reward_dev, uint8_t *done_dev,
                    float *score_dev, void *stream);
int mg_step_render(mg_handle *h, uint8_t *obs_dev, void *stream);

/* Same as mg_step with HOST buffers: H2D of actions, D2H of obs/reward/done/score and a
 * stream synchronise are inside the call (the end-to-end path). */
int mg_step_host(mg_handle *h, const int32_t *actions_host, uint8_t *obs_host, float *reward_host,
                 uint8_t *done_host, float *score_host);

/* Host assembly for mg_step_host.  What BaseEnv.step returns under the LoRes4* wrappers is the stack of the last four
 * frames (FlattenFrameStack.step, magical/benchmarks/__init__.py:122-137), and three of the four are already on the host
 * after the previous step.  With host assembly on, mg_step_host sends only the NEW 96x96x3 frame of the first `chunks`
 * of its 16 parts of the batch over PCIe (27 648 B per env instead of 110 592 B) into a pinned mirror of the frame ring
 * and `n_threads` host threads build those envs' stacked observations in obs_host while the other parts arrive as
 * finished stacks: same bytes as before, more env-steps per second when the link is the limit and host cores are free.
 * mode 0: off (default).  mode 1: the handle moves `chunks` step by step until link and host finish together (chunks
 * = starting value, 0: half).  mode 2: `chunks` fixed.  n_threads <= 0: $MG_HOST_THREADS, else all hardware threads.
 * Single-view HWC stacks only (MG_ERR_ARG otherwise).  Any render outside mg_step_host (mg_step, mg_step_render,
 * mg_reset_obs, mg_set_state) makes the next mg_step_host refetch the whole ring of the assembled parts once. */
int mg_set_host_assembly(mg_handle *h, int mode, int chunks, int n_threads);
/* current number of host-assembled parts / parts in all / host threads, and the last step's device-side copy time and
 * host-side assembly time in ms (any pointer may be NULL) */
int mg_get_host_assembly(const mg_handle *h, int *chunks_out, int *total_chunks_out, int *threads_out,
                         double *copy_ms_out, double *host_ms_out);
/* The host half alone (FlattenFrameStack over frames the caller holds, magical/benchmarks/__init__.py:122-147):
 * obs_host u8 [n_envs,96,96,3*n_groups], group g of env e = ring_host[e][(newest_slot - min(ages[g], since_reset[e])) & 3]
 * with ring_host u8 [n_envs,4,96,96,3]; since_reset[e] = frames since the env's episode began (3 = at least three). */
int mg_host_stack_frames(const uint8_t *ring_host, uint8_t *obs_host, int n_envs, int newest_slot,
                         const int32_t *ages, int n_groups, const uint8_t *since_reset, int n_threads);

/* Render one view of every env.  full_res != 0: out_dev u8 [B,384,384,3];
 * else the 4x area-averaged u8 [B,96,96,3].  Does not touch the frame stacks. */
int mg_render(mg_handle *h, int view, int full_res, uint8_t *out_dev, void *stream);

int mg_score(mg_handle *h, float *score_dev, void *stream);
int mg_debug_reward(mg_handle *h, float *reward_dev, void *stream);

/* Body poses/velocities, canonical order: the six robot bodies (MG_RB_*), then blocks in
 * scene order; unused slots are zero.  out_host: double [B][MG_MAX_BODIES][6] = x,y,angle,vx,vy,w.
 * Synchronous. */
int mg_get_bodies(mg_handle *h, double *out_host);

/* Per-env health flags raised by mg_step* since mg_create or the last call with clear != 0 (sticky across
 * resets).  status_host: int32 [B].  Synchronous (drains the device).
 *   MG_STATUS_ARB_OVERFLOW  a sub-step found more simultaneous cached contacts (live + up to 3 stamps old) than
 *                           the per-env cache holds (160 shape pairs); the extra contact was dropped, so the env
 *                           no longer follows the reference
 *   MG_STATUS_NONFINITE     a body position / angle / velocity became NaN or Inf */
#define MG_STATUS_ARB_OVERFLOW 1
#define MG_STATUS_NONFINITE 2
int mg_get_status(mg_handle *h, int32_t *status_host, int clear);

/* Opaque per-env checkpoint blobs (bodies, cached impulses, contact cache, counters). */
size_t mg_state_size(const mg_handle *h);
int mg_get_state(mg_handle *h, const int32_t *env_ids_host, int n, void *blobs_host);
int mg_set_state(mg_handle *h, const int32_t *env_ids_host, int n, const void *blobs_host);

/* mg_get_state for callers that may not allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments and byte-identical blobs as mg_get_state, but the device-side gather
 * scratch is a fixed file-scope __device__ array (about 1.25 MB, held by every GPU that loads the library until
 * exit) instead of a per-call cudaMalloc: the call issues no cudaMalloc / cudaFree / allocating library routine.
 * At most MG_NOALLOC_MAX_ENVS envs per call; a larger n is rejected with MG_ERR_ARG and a message naming the limit
 * (call it in slices).  All calls share the one scratch array: calls that overlap on the GPU (several host threads,
 * or several handles on one device, at the same time) race on it -- serialise them.  Synchronous. */
#define MG_NOALLOC_MAX_ENVS 32
int mg_get_state_noalloc(mg_handle *h, const int32_t *env_ids_host, int n, void *blobs_host);

/* Kernel launches issued through this handle so far (for bench.py's gpu_launches). */
int64_t mg_launch_count(const mg_handle *h);
/* CUDA-event time (ms) accumulated over the physics / render kernels since the last call
 * with reset != 0; which: 0 physics, 1 render.  Only counted while mg_set_timing(h,1). */
int mg_set_timing(mg_handle *h, int enable);
double mg_kernel_ms(mg_handle *h, int which, int reset, int64_t *launches_out);

/* ---- the preprocess half of the path on STORED frames (demo trajectories, saved_trajectories.py:87-149) ----
 * No handle: `device` is the CUDA ordinal; errors are readable with mg_last_error(NULL).
 * mg_frames_downsample: cv2.resize(INTER_AREA) 384 -> 96 of n_frames frames, u8 [n,384,384,3] -> [n,96,96,3]
 * (round-half-even of each 4x4 block mean, as the step kernel does).  frames_dev must be 16-byte aligned. */
int mg_frames_downsample(int device, const uint8_t *frames_dev, int64_t n_frames, uint8_t *out_dev, void *stream);
/* mg_frames_stack: frame t of the output is the channel concatenation of the 96x96 frames
 * (group_view[g], max(t - group_age[g], 0)), g = 0..n_groups-1: what a stacking wrapper returns t steps after
 * a reset that filled its deque with frame 0.  allo_lo_dev / ego_lo_dev: u8 [n_frames,96,96,3] (either may be
 * NULL if no group names it); out_dev: u8 [n_frames,96,96,3*n_groups], or [n_frames,3*n_groups,96,96] if chw. */
int mg_frames_stack(int device, const uint8_t *allo_lo_dev, const uint8_t *ego_lo_dev, int64_t n_frames, int n_groups,
                    const int32_t *group_view, const int32_t *group_age, int chw, uint8_t *out_dev, void *stream);

/* ---- host-side helpers for reset sampling (no GPU involved) ---- */

enum { MG_ENT_ROBOT = 0, MG_ENT_BLOCK = 1, MG_ENT_GOAL = 2 };

/* 1 if entity (kind, idx) of `scene`, at the pose the scene holds for it, touches an arena
 * wall or any shape of the entities selected by the masks (bit i = block i / goal i;
 * robot_active != 0 = robot).  Stands in for the shape_query loop of
 * geom.pm_randomise_pose (geom.py:226-233); touching counts (distance <= 0). */
int mg_host_entity_overlaps(const mg_scene *scene, int kind, int idx, int robot_active,
                            uint32_t block_mask, uint32_t goal_mask);

/* The 18 palette entries the renderer draws with, u8 RGB each, in the order
 * background, white, grey, dark grey, light grey (finger inner), pupil, then colour / darkened colour /
 * twice-lightened colour for red, green, blue, yellow (magical/style.py:10-37, render.py:146-149).
 * Returns the number of entries written to rgb_out[n][3]. */
int mg_host_palette(uint8_t *rgb_out);

/* ---- native host reset engine (CPU only; no device needed) ----
 * What BaseEnv.reset() does before the first render, for one env: phys-var sampling, the task's on_reset
 * and the rejection sampler of geom.pm_randomise_all_poses, consuming the env's numpy-legacy MT19937 stream
 * (np.random.RandomState) draw for draw like the reference consumes self.rng.  The caller owns the stream:
 * mt_key[624] / *mt_pos are RandomState.get_state()[1:3] and are advanced in place (set_state them back).
 * kw[] are the task's boolean keyword arguments in this order:
 *   MoveToCorner, PickAndPlace : rand_shape_colour, rand_shape_type, rand_poses, debug_reward
 *   MoveToRegion               : rand_poses_minor, rand_poses_full, rand_goal_colour
 *   MatchRegions               : rand_target_colour, rand_shape_type, rand_shape_count, rand_layout_minor, rand_layout_full
 *   MakeLine, FindDupe, FixColour : rand_colours, rand_shapes, rand_count, rand_layout_minor, rand_layout_full
 *   ClusterColour, ClusterShape   : rand_shape_colour, rand_shape_type, rand_layout_minor, rand_layout_full, rand_shape_count
 * Returns MG_OK, MG_ERR_ARG, or MG_ERR_PLACEMENT where the reference raises PlacementError (the stream has
 * then advanced exactly as far as the reference's would have). */
typedef struct mg_reset_spec {
    int32_t task;            /* MG_TASK_* */
    int32_t horizon;         /* max_episode_steps */
    int32_t rand_dynamics;
    int32_t kw[8];
    int32_t reserved;
} mg_reset_spec;

int mg_host_sample_scene(const mg_reset_spec *spec, uint32_t *mt_key, int32_t *mt_pos, mg_scene *out);
/* n envs at once on n_threads host threads: specs[n], mt_keys[n][624], mt_pos[n], out[n], status[n] */
int mg_host_sample_scenes(const mg_reset_spec *specs, int n, uint32_t *mt_keys, int32_t *mt_pos, mg_scene *out,
                          int32_t *status, int n_threads);

#ifdef __cplusplus
}
#endif
#endif /* MG_B200_H */
