/*
 * mg_kernels.h -- device-side handle and kernel launchers shared by mg_api.cu,
 * mg_physics.cu and mg_raster.cu.
 */
#pragma once
#include <cuda_runtime.h>
#include "mg_state.cuh"
#include "../../include/mg_b200.h"

#define MG_MAX_RING 4   /* frames kept per view (deepest reference stack: 4, benchmarks/__init__.py:232-306) */
#define MG_FRAME_BYTES (MG_LORES * MG_LORES * 3)

struct MgDeviceState {
    int num_envs;
    double *dstate;          /* [groups][D_NF][32] */
    int *istate;             /* [groups][I_NF][32] */
    mg_scene *scenes;        /* [B] scene each env was last reset to */
    MgPrim *prims;           /* [B][MG_MAX_PRIMS] draw list */
    int *n_prims;            /* [B] */
    double *poses;           /* [B][MG_NB][MG_POSE_NF] exported after every reset / step */
    const MgTables *tables;  /* device copy */
    /* frame stacks: ring[view] = u8 [B][MG_MAX_RING][96][96][3]; the newest slot is the same for all envs
     * (every env pushes one frame per step) and is tracked on the host */
    uint8_t *ring[2];
    /* screen-space geometry of the current poses, per view (mg_geometry_kernel -> mg_render_kernel) */
    MgScreenPrim *sprims[2]; /* [B][MG_MAX_PRIMS] */
    int2 *sverts[2];         /* [B][MG_MAX_VERTS] truncated vertices */
    double *dashv[2];        /* [B][MG_MAX_GOALS][4][2] untruncated goal-region vertices (dashed outlines) */
    /* optional order of the envs for this launch (NULL = identity): the step and render kernels take env perm[k] for
     * their k-th work slot, so that a part of the batch can be any set of envs (mg_step groups envs by contact load) */
    const int *perm;
    int *status;             /* [B] sticky MG_STATUS_* bits (NULL: not collected) */
    /* resets queued with mg_queue_resets: pending[env] = index into pending_scenes, or -1.  The geometry pre-pass of
     * the step path applies them (init_env) after the step's physics and before the env is rendered. */
    int *pending;
    const mg_scene *pending_scenes;
};
#define MG_MAX_DEVICES 64    /* CUDA ordinals the per-device launch attributes are tracked for */

struct MgObsSpec {
    int n_groups;
    int view[MG_MAX_GROUPS];
    int age[MG_MAX_GROUPS];
    int chw;
    int render_view[2];      /* which views mg_step must render */
};

void mg_launch_init(const MgDeviceState &S, const mg_scene *scenes_dev, const int32_t *env_ids_dev, int n, cudaStream_t st);
void mg_launch_mark_pending(const MgDeviceState &S, const int32_t *env_ids_dev, int n, cudaStream_t st);
/* around a graph-replayed step (whose kernel arguments are frozen): the caller's actions are copied into the handle's
 * buffer before it, and reward / done / score from the handle's buffers to the caller's after it (NULL: not wanted) */
void mg_launch_step_io(int n, const int32_t *actions_src, int32_t *actions_dst, const float *reward_src, float *reward_dst,
                       const uint8_t *done_src, uint8_t *done_dst, const float *score_src, float *score_dst, cudaStream_t st);
/* max_blocks: largest block count of any env in the batch (sizes the per-env shared-memory copy) */
void mg_launch_physics(const MgDeviceState &S, const int32_t *actions, float *reward, uint8_t *done, float *score, int max_blocks, cudaStream_t st);
void mg_launch_score(const MgDeviceState &S, float *out, int debug_reward, cudaStream_t st);
/* render `view` of every env and push the 96x96 frame into ring slot `slot`; when `fused` is given the
 * same kernel also writes the stacked HWC observation (slots[v] = newest slot of view v); envs
 * [env_begin, env_end) only, so that a host copy of one part can overlap the rendering of the next */
void mg_launch_render_push(const MgDeviceState &S, int view, int slot, const MgObsSpec *fused, const int *slots,
                           uint8_t *obs, int env_begin, int env_end, cudaStream_t st, int apply_pending = 0);
/* transform every primitive of every env into the screen space of `view` (must precede the render launches
 * of that view whenever poses or scenes changed) */
void mg_launch_geometry(const MgDeviceState &S, int view, int env_begin, int env_end, cudaStream_t st, int apply_pending = 0);
/* compose the stacked observation from the rings (any view mix, HWC or CHW) */
void mg_launch_compose(const MgDeviceState &S, const MgObsSpec &spec, const int *slots, uint8_t *obs, cudaStream_t st);
/* clear the per-env "fresh" flag after all views were pushed */
void mg_launch_clear_fresh(const MgDeviceState &S, cudaStream_t st);
/* stand-alone render of one view: full_res ? [B,384,384,3] : [B,96,96,3] */
void mg_launch_render_out(const MgDeviceState &S, int view, int full_res, uint8_t *out, cudaStream_t st);
/* perm <- the envs ordered by the number of cached contacts (ascending), one launch */
void mg_launch_order(const MgDeviceState &S, int *perm, cudaStream_t st);
void mg_launch_state_gather(const MgDeviceState &S, const int32_t *env_ids_dev, int n, double *dblobs, int *iblobs, int scatter, cudaStream_t st);
