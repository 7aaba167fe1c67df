/*
 * mg_api.cu -- the C ABI of libmagical_b200.so (include/mg_b200.h): handle, device memory,
 * stream plumbing and the order of kernel launches that makes up reset / step / render.
 * There is no CPU implementation behind these entry points.
 */
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <mutex>
#include <string>
#include <vector>
#include "mg_kernels.h"
#include "mg_tables.h"
#include "mg_hostasm.h"
#include <chrono>
#include <thread>
#include <sched.h>
#include "../../include/mg_trig.h"

#define MG_STEP_PARTS 64   /* parts of the batch mg_step runs on their own streams */
#define MG_HOST_CHUNKS 16  /* parts of the batch whose D2H copy overlaps the rendering of the next (mg_step_host) */

struct mg_handle {
    mg_config cfg;
    MgDeviceState S;
    MgObsSpec spec;
    bool fused;                 /* single-view HWC stack: composed inside the render kernel */
    int slots[2];               /* newest ring slot per view */
    MgTables *tables_dev;
    int max_blocks;          /* largest n_blocks ever installed: sizes the step kernel's shared memory */
    mg_scene *scene_stage_dev;  /* [B] staging for mg_set_scenes */
    int32_t *ids_stage_dev;     /* [B] */
    /* pinned host side of the staging, two slots: mg_set_scenes copies the caller's (possibly pageable) arrays here and
     * returns without draining the stream; a slot is reused once the init kernel that read it has run */
    mg_scene *scene_stage_host[2];
    int32_t *ids_stage_host[2];
    cudaEvent_t stage_ev[2];
    int stage_used[2], stage_slot;
    /* mg_step_host staging */
    int32_t *actions_dev;
    uint8_t *obs_dev;
    float *reward_dev, *score_dev;
    uint8_t *done_dev;
    size_t obs_bytes;
    cudaStream_t own_stream, copy_stream;   /* mg_step_host: compute stream and the stream of the chunked D2H copies */
    cudaStream_t part_stream[MG_STEP_PARTS];   /* mg_step: one stream per part of the batch (render of a part under the physics of the others) */
    cudaEvent_t part_done[MG_STEP_PARTS], fork_ev;
    int step_parts;                          /* 0/1: plain sequence */
    int step_order;                          /* group the envs of a part by contact load (mg_order_kernel) */
    int *perm_dev;                           /* [B] env order of the current mg_step */
    cudaEvent_t chunk_done[MG_HOST_CHUNKS];
    /* host-assembled part of mg_step_host (mg_set_host_assembly): chunks [0, asm_chunks) send only their NEW frame into a
     * pinned mirror of the frame ring and the host builds their stacked observations (mg_hostasm.cpp) while the other
     * chunks arrive as finished stacks */
    int asm_mode;                            /* 0 off, 1 balanced by the handle, 2 fixed */
    int asm_chunks;
    uint8_t *ring_host;                      /* pinned [B][MG_MAX_RING][MG_FRAME_BYTES], lazily allocated */
    uint8_t asm_hist_ok[MG_HOST_CHUNKS];     /* the mirror holds the chunk's three older frames */
    std::vector<uint8_t> since_reset;        /* [B] frames since the env's episode began, saturating at MG_MAX_RING - 1 */
    std::vector<int32_t> fresh_ids;          /* envs reset since the last mg_step_host */
    MgHostAsm *hostasm;
    cudaEvent_t chunk_copied[MG_HOST_CHUNKS], asm_t0, asm_t1;
    int asm_hold;                            /* steps until the balance may move again */
    double asm_last_copy_ms, asm_last_host_ms;
    /* last work this handle put on a CALLER's stream (mg_set_scenes, mg_reset_obs, mg_step*, mg_render): mg_step_host
     * runs on own_stream and must not start before it (the caller need not synchronise between the two) */
    cudaEvent_t caller_ev;
    int caller_ev_used;
    int pending_queued;                      /* mg_queue_resets ran and no step has applied its resets yet */
    /* mg_step as a CUDA graph: the fork / join of the parts is captured once per (output buffers, ring slot, scene size)
     * and replayed, so a step costs the host one graph launch instead of ~50 kernel launches */
    struct StepGraph { const void *obs; int slot, max_blocks, seen; cudaGraphExec_t exec; };
    std::vector<StepGraph> graphs;
    int step_graph;                          /* 0: launch the parts eagerly (MG_STEP_GRAPH=0) */
    /* accounting */
    int64_t launches;
    int timing;
    double ms[2];
    int64_t timed_launches[2];
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> pending[2];   /* recorded, not yet read (per kernel class) */
    std::vector<cudaEvent_t> free_events;
    std::string err;
};

static std::string g_create_err;

static int fail(mg_handle *h, int code, const std::string &msg)
{
    if (h) h->err = msg; else g_create_err = msg;
    return code;
}

#define CK(h, call)                                                                         \
    do {                                                                                    \
        cudaError_t e_ = (call);                                                            \
        if (e_ != cudaSuccess)                                                              \
            return fail(h, MG_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

/* handle-less entry points (mg_frames.cu) report through mg_last_error(NULL) */
int mg_fail_global(int code, const std::string &msg) { return fail(nullptr, code, msg); }

extern "C" const char *mg_last_error(const mg_handle *h) { return h ? h->err.c_str() : g_create_err.c_str(); }

extern "C" int mg_create(const mg_config *cfg, mg_handle **out)
{
    if (!cfg || !out) return fail(nullptr, MG_ERR_ARG, "mg_create: null argument");
    if (cfg->num_envs <= 0) return fail(nullptr, MG_ERR_ARG, "mg_create: num_envs must be positive");
    if (cfg->n_groups < 0 || cfg->n_groups > MG_MAX_GROUPS) return fail(nullptr, MG_ERR_ARG, "mg_create: bad n_groups");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0)
        return fail(nullptr, MG_ERR_NODEVICE, "mg_create: no CUDA device (this library has no CPU path)");
    if (cfg->device < 0 || cfg->device >= ndev) return fail(nullptr, MG_ERR_ARG, "mg_create: bad device ordinal");
    mg_handle *h = new mg_handle();
    h->cfg = *cfg;
    h->launches = 0; h->timing = 0; h->ms[0] = h->ms[1] = 0; h->timed_launches[0] = h->timed_launches[1] = 0;
    h->slots[0] = h->slots[1] = 0;
    h->max_blocks = 0;
    memset(&h->S, 0, sizeof(h->S));
    memset(&h->spec, 0, sizeof(h->spec));
    h->spec.n_groups = cfg->n_groups; h->spec.chw = cfg->chw;
    bool single_view = true;
    for (int g = 0; g < cfg->n_groups; g++) {
        int v = cfg->group_view[g], a = cfg->group_age[g];
        if (v < 0 || v > 1 || a < 0 || a >= MG_MAX_RING) { delete h; return fail(nullptr, MG_ERR_ARG, "mg_create: bad group spec"); }
        h->spec.view[g] = v; h->spec.age[g] = a;
        h->spec.render_view[v] = 1;
        if (v != cfg->group_view[0]) single_view = false;
    }
    h->fused = cfg->n_groups > 0 && single_view && !cfg->chw;
    const int B = cfg->num_envs;
    const size_t groups = (B + MG_GROUP - 1) / MG_GROUP;
    CK(h, cudaSetDevice(cfg->device));
    CK(h, cudaMalloc(&h->S.dstate, groups * D_NF * MG_GROUP * sizeof(double)));
    CK(h, cudaMalloc(&h->S.istate, groups * I_NF * MG_GROUP * sizeof(int)));
    CK(h, cudaMemset(h->S.dstate, 0, groups * D_NF * MG_GROUP * sizeof(double)));
    CK(h, cudaMemset(h->S.istate, 0, groups * I_NF * MG_GROUP * sizeof(int)));
    CK(h, cudaMalloc(&h->S.scenes, (size_t)B * sizeof(mg_scene)));
    CK(h, cudaMemset(h->S.scenes, 0, (size_t)B * sizeof(mg_scene)));
    CK(h, cudaMalloc(&h->S.prims, (size_t)B * MG_MAX_PRIMS * sizeof(MgPrim)));
    CK(h, cudaMalloc(&h->S.n_prims, (size_t)B * sizeof(int)));
    CK(h, cudaMemset(h->S.n_prims, 0, (size_t)B * sizeof(int)));
    CK(h, cudaMalloc(&h->S.pending, (size_t)B * sizeof(int)));
    CK(h, cudaMemset(h->S.pending, 0xff, (size_t)B * sizeof(int)));       /* -1: no reset queued */
    CK(h, cudaMalloc(&h->S.status, (size_t)B * sizeof(int)));
    CK(h, cudaMemset(h->S.status, 0, (size_t)B * sizeof(int)));
    CK(h, cudaMalloc(&h->S.poses, (size_t)B * MG_POSE_STRIDE * sizeof(double)));
    CK(h, cudaMemset(h->S.poses, 0, (size_t)B * MG_POSE_STRIDE * sizeof(double)));
    for (int v = 0; v < 2; v++) {
        CK(h, cudaMalloc(&h->S.ring[v], (size_t)B * MG_MAX_RING * MG_FRAME_BYTES));
        CK(h, cudaMemset(h->S.ring[v], 0, (size_t)B * MG_MAX_RING * MG_FRAME_BYTES));
        CK(h, cudaMalloc(&h->S.sprims[v], (size_t)B * MG_MAX_PRIMS * sizeof(MgScreenPrim)));
        CK(h, cudaMalloc(&h->S.sverts[v], (size_t)B * MG_MAX_VERTS * sizeof(int2)));
        CK(h, cudaMalloc(&h->S.dashv[v], (size_t)B * MG_MAX_GOALS * 8 * sizeof(double)));
    }
    CK(h, cudaMalloc(&h->tables_dev, sizeof(MgTables)));
    CK(h, cudaMemcpy(h->tables_dev, mg_tables_host(), sizeof(MgTables), cudaMemcpyHostToDevice));
    h->S.tables = h->tables_dev;
    h->S.num_envs = B;
    CK(h, cudaMalloc(&h->scene_stage_dev, (size_t)B * sizeof(mg_scene)));
    CK(h, cudaMalloc(&h->ids_stage_dev, (size_t)B * sizeof(int32_t)));
    h->S.pending_scenes = h->scene_stage_dev;
    h->pending_queued = 0;
    for (int k = 0; k < 2; k++) {
        CK(h, cudaMallocHost(&h->scene_stage_host[k], (size_t)B * sizeof(mg_scene)));
        CK(h, cudaMallocHost(&h->ids_stage_host[k], (size_t)B * sizeof(int32_t)));
        CK(h, cudaEventCreateWithFlags(&h->stage_ev[k], cudaEventDisableTiming));
        h->stage_used[k] = 0;
    }
    h->stage_slot = 0;
    h->obs_bytes = (size_t)B * MG_LORES * MG_LORES * 3 * (cfg->n_groups > 0 ? cfg->n_groups : 1);
    CK(h, cudaMalloc(&h->actions_dev, (size_t)B * sizeof(int32_t)));
    CK(h, cudaMalloc(&h->obs_dev, h->obs_bytes));
    CK(h, cudaMalloc(&h->reward_dev, (size_t)B * sizeof(float)));
    CK(h, cudaMalloc(&h->score_dev, (size_t)B * sizeof(float)));
    CK(h, cudaMalloc(&h->done_dev, (size_t)B));
    CK(h, cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking));
    CK(h, cudaStreamCreateWithFlags(&h->copy_stream, cudaStreamNonBlocking));
    for (int c = 0; c < MG_HOST_CHUNKS; c++) {
        CK(h, cudaEventCreateWithFlags(&h->chunk_done[c], cudaEventDisableTiming));
        CK(h, cudaEventCreateWithFlags(&h->chunk_copied[c], cudaEventDisableTiming));
        h->asm_hist_ok[c] = 0;
    }
    CK(h, cudaEventCreate(&h->asm_t0));
    CK(h, cudaEventCreate(&h->asm_t1));
    h->asm_mode = 0; h->asm_chunks = 0; h->ring_host = nullptr; h->hostasm = nullptr; h->asm_hold = 0;
    h->asm_last_copy_ms = h->asm_last_host_ms = 0;
    for (int c = 0; c < MG_STEP_PARTS; c++) {
        CK(h, cudaStreamCreateWithFlags(&h->part_stream[c], cudaStreamNonBlocking));
        CK(h, cudaEventCreateWithFlags(&h->part_done[c], cudaEventDisableTiming));
    }
    CK(h, cudaEventCreateWithFlags(&h->fork_ev, cudaEventDisableTiming));
    CK(h, cudaEventCreateWithFlags(&h->caller_ev, cudaEventDisableTiming));
    h->caller_ev_used = 0;
    {
        /* parts of >= 128 envs, 32 at most, measured best at 4096 envs (plain sequence 0.80 ms, 16 parts 0.707, 32 parts
         * 0.693, 64 parts 0.705); batches beyond 8192 envs (several waves of physics CTAs) are level from 8 to 32 parts
         * and keep 16.  MG_STEP_PARTS=1 keeps the plain sequence */
        const char *ev = getenv("MG_STEP_PARTS");
        h->step_parts = ev ? atoi(ev) : (cfg->num_envs <= 8192 ? cfg->num_envs / 128 : 16);
        if (!ev && h->step_parts > 32) h->step_parts = 32;
        if (h->step_parts > MG_STEP_PARTS) h->step_parts = MG_STEP_PARTS;
        const char *eg = getenv("MG_STEP_GRAPH");
        h->step_graph = eg ? atoi(eg) : 1;
        const char *eo = getenv("MG_STEP_ORDER");
        h->step_order = eo ? atoi(eo) : 1;
        CK(h, cudaMalloc(&h->perm_dev, (size_t)cfg->num_envs * sizeof(int)));
    }
    *out = h;
    return MG_OK;
}

extern "C" int mg_destroy(mg_handle *h)
{
    if (!h) return MG_OK;
    cudaSetDevice(h->cfg.device);
    cudaDeviceSynchronize();
    cudaFree(h->S.dstate); cudaFree(h->S.istate); cudaFree(h->S.scenes); cudaFree(h->S.prims);
    cudaFree(h->S.n_prims); cudaFree(h->S.status); cudaFree(h->S.pending); cudaFree(h->S.poses); cudaFree(h->S.ring[0]); cudaFree(h->S.ring[1]);
    for (int v = 0; v < 2; v++) { cudaFree(h->S.sprims[v]); cudaFree(h->S.sverts[v]); cudaFree(h->S.dashv[v]); }
    cudaFree(h->tables_dev); cudaFree(h->scene_stage_dev); cudaFree(h->ids_stage_dev);
    cudaFree(h->actions_dev); cudaFree(h->obs_dev); cudaFree(h->reward_dev); cudaFree(h->score_dev); cudaFree(h->done_dev);
    cudaStreamDestroy(h->own_stream); cudaStreamDestroy(h->copy_stream);
    for (int c = 0; c < MG_HOST_CHUNKS; c++) { cudaEventDestroy(h->chunk_done[c]); cudaEventDestroy(h->chunk_copied[c]); }
    cudaEventDestroy(h->asm_t0); cudaEventDestroy(h->asm_t1);
    if (h->ring_host) cudaFreeHost(h->ring_host);
    mg_hostasm_destroy(h->hostasm);
    for (int c = 0; c < MG_STEP_PARTS; c++) { cudaStreamDestroy(h->part_stream[c]); cudaEventDestroy(h->part_done[c]); }
    cudaEventDestroy(h->fork_ev); cudaEventDestroy(h->caller_ev);
    for (int k = 0; k < 2; k++) { cudaFreeHost(h->scene_stage_host[k]); cudaFreeHost(h->ids_stage_host[k]); cudaEventDestroy(h->stage_ev[k]); }
    cudaFree(h->perm_dev);
    for (auto &g : h->graphs) if (g.exec) cudaGraphExecDestroy(g.exec);
    for (int w = 0; w < 2; w++) for (auto &p : h->pending[w]) { cudaEventDestroy(p.first); cudaEventDestroy(p.second); }
    for (auto e : h->free_events) cudaEventDestroy(e);
    delete h;
    return MG_OK;
}

/* remember the tail of the work just put on a caller's stream so that mg_step_host (own_stream) can order itself behind it */
static void mark_caller(mg_handle *h, cudaStream_t st)
{
    if (st == h->own_stream) return;
    if (cudaEventRecord(h->caller_ev, st) == cudaSuccess) h->caller_ev_used = 1;
}

/* the frame rings moved without the host mirror seeing it (any render outside the host-assembled mg_step_host) */
static void hostasm_invalidate(mg_handle *h)
{
    for (int c = 0; c < MG_HOST_CHUNKS; c++) h->asm_hist_ok[c] = 0;
}
/* envs whose next frame is the first of an episode */
static void hostasm_note_resets(mg_handle *h, const int32_t *env_ids_host, int n)
{
    if (!h->asm_mode) return;
    for (int i = 0; i < n; i++) h->fresh_ids.push_back(env_ids_host ? env_ids_host[i] : i);
}

static int check_launch(mg_handle *h, const char *what)
{
    /* MG_DEBUG_SYNC=1: drain the device after every group of launches so that an asynchronous fault is reported by
     * the call that caused it (debugging only: it serialises everything) */
    static const bool debug_sync = getenv("MG_DEBUG_SYNC") && atoi(getenv("MG_DEBUG_SYNC")) != 0;
    cudaError_t e = debug_sync ? cudaDeviceSynchronize() : cudaSuccess;
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) return fail(h, MG_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
    return MG_OK;
}

/* validate and copy `n` scenes (+ ids) into the pinned staging slot and start their upload on `st`; the caller
 * launches the kernel that consumes the device staging and then calls stage_done */
static int stage_upload(mg_handle *h, const char *who, const int32_t *env_ids_host, int n, const mg_scene *scenes_host, cudaStream_t st, int *slot_out)
{
    for (int i = 0; i < n; i++) {
        if (env_ids_host && (env_ids_host[i] < 0 || env_ids_host[i] >= h->cfg.num_envs))
            return fail(h, MG_ERR_ARG, std::string(who) + ": env id out of range");
        const mg_scene &s = scenes_host[i];
        if (s.n_blocks < 0 || s.n_blocks > MG_MAX_BLOCKS || s.n_goals < 0 || s.n_goals > MG_MAX_GOALS ||
            s.task < 0 || s.task >= MG_NUM_TASKS)
            return fail(h, MG_ERR_ARG, std::string(who) + ": scene out of range");
    }
    for (int i = 0; i < n; i++) h->max_blocks = scenes_host[i].n_blocks > h->max_blocks ? scenes_host[i].n_blocks : h->max_blocks;
    const int slot = h->stage_slot ^= 1;
    if (h->stage_used[slot]) CK(h, cudaEventSynchronize(h->stage_ev[slot]));          /* long done: two calls ago */
    memcpy(h->scene_stage_host[slot], scenes_host, (size_t)n * sizeof(mg_scene));
    if (env_ids_host) memcpy(h->ids_stage_host[slot], env_ids_host, (size_t)n * sizeof(int32_t));
    /* the device side of the staging is one buffer: order this call behind the kernel that read the previous upload */
    if (h->stage_used[slot ^ 1]) CK(h, cudaStreamWaitEvent(st, h->stage_ev[slot ^ 1], 0));
    CK(h, cudaMemcpyAsync(h->scene_stage_dev, h->scene_stage_host[slot], (size_t)n * sizeof(mg_scene), cudaMemcpyHostToDevice, st));
    if (env_ids_host) CK(h, cudaMemcpyAsync(h->ids_stage_dev, h->ids_stage_host[slot], (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    *slot_out = slot;
    return MG_OK;
}

extern "C" int mg_set_scenes(mg_handle *h, const int32_t *env_ids_host, int n, const mg_scene *scenes_host, void *stream)
{
    if (!h || !scenes_host || n < 0 || n > h->cfg.num_envs) return fail(h, MG_ERR_ARG, "mg_set_scenes: bad argument");
    if (n == 0) return MG_OK;
    if (h->pending_queued) return fail(h, MG_ERR_ARG, "mg_set_scenes: resets queued with mg_queue_resets are still waiting for a step");
    cudaStream_t st = (cudaStream_t)stream;
    CK(h, cudaSetDevice(h->cfg.device));
    int slot = 0;
    int rc = stage_upload(h, "mg_set_scenes", env_ids_host, n, scenes_host, st, &slot);
    if (rc != MG_OK) return rc;
    mg_launch_init(h->S, h->scene_stage_dev, env_ids_host ? h->ids_stage_dev : nullptr, n, st);
    hostasm_note_resets(h, env_ids_host, n);
    h->launches++;
    CK(h, cudaEventRecord(h->stage_ev[slot], st));
    h->stage_used[slot] = 1;
    mark_caller(h, st);
    return check_launch(h, "mg_init_kernel");
}

/* resets for the NEXT step: uploaded and marked now, applied by that step's geometry pre-pass after its physics
 * (the staged scenes stay in place until then: one queue per step) */
extern "C" int mg_queue_resets(mg_handle *h, const int32_t *env_ids_host, int n, const mg_scene *scenes_host, void *stream)
{
    if (!h || n < 0 || n > h->cfg.num_envs || (n > 0 && !scenes_host)) return fail(h, MG_ERR_ARG, "mg_queue_resets: bad argument");
    if (n == 0) return MG_OK;
    if (h->pending_queued) return fail(h, MG_ERR_ARG, "mg_queue_resets: resets are already queued for the next step (one call per step)");
    cudaStream_t st = (cudaStream_t)stream;
    CK(h, cudaSetDevice(h->cfg.device));
    int slot = 0;
    int rc = stage_upload(h, "mg_queue_resets", env_ids_host, n, scenes_host, st, &slot);
    if (rc != MG_OK) return rc;
    mg_launch_mark_pending(h->S, env_ids_host ? h->ids_stage_dev : nullptr, n, st);
    hostasm_note_resets(h, env_ids_host, n);
    h->launches++;
    /* the device staging is read by the step that applies the resets: its end is recorded by that step (pending_done) */
    h->pending_queued = 1 + slot;
    mark_caller(h, st);
    return check_launch(h, "mg_mark_pending_kernel");
}

/* a step has launched the render work that applies the queued resets: the staging may be reused after it */
static int pending_done(mg_handle *h, cudaStream_t st)
{
    if (!h->pending_queued) return MG_OK;
    const int slot = h->pending_queued - 1;
    CK(h, cudaEventRecord(h->stage_ev[slot], st));
    h->stage_used[slot] = 1;
    h->pending_queued = 0;
    return MG_OK;
}

/* per-kernel timing: event pairs are recorded on the launching stream without synchronising and
 * read back (after a synchronise) by mg_kernel_ms, so enabling it does not serialise the step */
static void time_begin(mg_handle *h, int which, cudaStream_t st)
{
    if (!h->timing) return;
    cudaEvent_t a, b;
    if (h->free_events.size() >= 2) {
        a = h->free_events.back(); h->free_events.pop_back();
        b = h->free_events.back(); h->free_events.pop_back();
    } else {
        cudaEventCreate(&a); cudaEventCreate(&b);
    }
    cudaEventRecord(a, st);
    h->pending[which].push_back(std::make_pair(a, b));
}
static void time_end(mg_handle *h, int which, int n_launches, cudaStream_t st)
{
    if (!h->timing) return;
    cudaEventRecord(h->pending[which].back().second, st);
    h->timed_launches[which] += n_launches;
}

extern "C" int mg_step_physics(mg_handle *h, const int32_t *actions_dev, float *reward_dev, uint8_t *done_dev,
                               float *score_dev, void *stream)
{
    if (!h || !actions_dev) return fail(h, MG_ERR_ARG, "mg_step_physics: bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    CK(h, cudaSetDevice(h->cfg.device));
    time_begin(h, 0, st);
    mg_launch_physics(h->S, actions_dev, reward_dev, done_dev, score_dev, h->max_blocks, st);
    h->launches++;
    time_end(h, 0, 1, st);
    mark_caller(h, st);
    return check_launch(h, "mg_physics_kernel");
}

/* render every view the observation needs, push into the frame rings, write the stacked observation */
extern "C" int mg_step_render(mg_handle *h, uint8_t *obs_dev, void *stream)
{
    if (!h) return fail(h, MG_ERR_ARG, "mg_step_render: bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    CK(h, cudaSetDevice(h->cfg.device));
    hostasm_invalidate(h);
    int last_view = -1;
    for (int v = 0; v < 2; v++) if (h->spec.render_view[v]) { h->slots[v] = (h->slots[v] + 1) & (MG_MAX_RING - 1); last_view = v; }
    time_begin(h, 1, st);
    int nl = 0;
    for (int v = 0; v < 2; v++) {
        if (!h->spec.render_view[v]) continue;
        bool fuse = h->fused && v == last_view && obs_dev;
        mg_launch_render_push(h->S, v, h->slots[v], fuse ? &h->spec : nullptr, h->slots, obs_dev, 0, h->cfg.num_envs, st, h->pending_queued != 0);
        nl++;
    }
    time_end(h, 1, nl, st);      /* one timed unit = geometry + raster launch of one view */
    nl *= 2;
    if (!h->fused && obs_dev && h->spec.n_groups > 0) { mg_launch_compose(h->S, h->spec, h->slots, obs_dev, st); nl++; }
    mg_launch_clear_fresh(h->S, st); nl++;
    h->launches += nl;
    { int rc = pending_done(h, st); if (rc != MG_OK) return rc; }
    mark_caller(h, st);
    return check_launch(h, "mg_render_kernel");
}

extern "C" int mg_reset_obs(mg_handle *h, uint8_t *obs_dev, void *stream) { return mg_step_render(h, obs_dev, stream); }

/* mg_step with the batch cut into parts, each on its own stream: the physics launch is one wave whose slowest CTA
 * (an env wedged into a corner) takes 1.6x the mean, so most SMs idle through its tail, and the rasteriser of the parts
 * that are done fills that slack.  All physics parts are launched first (together they are the same single wave), then
 * geometry + raster of part k follow physics k on stream k.  Same kernels, same results; single-view HWC stacks only. */
/* apply_pending: the geometry pre-pass looks for queued resets (always on in a captured graph: one flag load per env) */
static int step_parts_body(mg_handle *h, const int32_t *actions_dev, uint8_t *obs_dev, float *reward_dev, uint8_t *done_dev,
                           float *score_dev, cudaStream_t st, int view, int apply_pending)
{
    const int B = h->cfg.num_envs, P = h->step_parts;
    /* envs without contacts take a third of the time of an env wedged into a corner, and a CTA (a part) is as slow as
     * its slowest env: order the envs by contact load so that the light parts finish, and render, early */
    const bool ordered = h->step_order != 0;
    if (ordered) mg_launch_order(h->S, h->perm_dev, st);
    CK(h, cudaEventRecord(h->fork_ev, st));
    int lo[MG_STEP_PARTS + 1];
    for (int k = 0; k <= P; k++) lo[k] = k == P ? B : (int)((long long)B * k / P) & ~31;   /* whole CTAs of 16 / 8 envs */
    for (int k = 0; k < P; k++) {
        cudaStream_t sk = h->part_stream[k];
        CK(h, cudaStreamWaitEvent(sk, h->fork_ev, 0));
        MgDeviceState Sk = h->S;
        Sk.num_envs = lo[k + 1] - lo[k];
        if (ordered) {
            Sk.perm = h->perm_dev + lo[k];
            mg_launch_physics(Sk, actions_dev, reward_dev, done_dev, score_dev, h->max_blocks, sk);
        } else {
            Sk.dstate += (size_t)lo[k] * D_NF; Sk.istate += (size_t)lo[k] * I_NF; Sk.poses += (size_t)lo[k] * MG_POSE_STRIDE;
            mg_launch_physics(Sk, actions_dev + lo[k], reward_dev ? reward_dev + lo[k] : nullptr, done_dev ? done_dev + lo[k] : nullptr,
                              score_dev ? score_dev + lo[k] : nullptr, h->max_blocks, sk);
        }
    }
    for (int k = 0; k < P; k++) {
        cudaStream_t sk = h->part_stream[k];
        MgDeviceState Sr = h->S;
        if (ordered) Sr.perm = h->perm_dev;
        mg_launch_render_push(Sr, view, h->slots[view], &h->spec, h->slots, obs_dev, lo[k], lo[k + 1], sk, apply_pending);
        CK(h, cudaEventRecord(h->part_done[k], sk));
        CK(h, cudaStreamWaitEvent(st, h->part_done[k], 0));
    }
    mg_launch_clear_fresh(h->S, st);
    h->launches += 3 * P + 1 + (ordered ? 1 : 0);
    return MG_OK;
}

static int step_in_parts(mg_handle *h, const int32_t *actions_dev, uint8_t *obs_dev, float *reward_dev, uint8_t *done_dev,
                         float *score_dev, cudaStream_t st)
{
    CK(h, cudaSetDevice(h->cfg.device));
    hostasm_invalidate(h);
    int view = -1;
    for (int v = 0; v < 2; v++) if (h->spec.render_view[v]) { h->slots[v] = (h->slots[v] + 1) & (MG_MAX_RING - 1); view = v; }
    static const bool debug_sync = getenv("MG_DEBUG_SYNC") && atoi(getenv("MG_DEBUG_SYNC")) != 0;
    int rc = MG_OK;
    if (!h->step_graph || debug_sync) {
        rc = step_parts_body(h, actions_dev, obs_dev, reward_dev, done_dev, score_dev, st, view, h->pending_queued != 0);
    } else {
        /* A graph's kernel arguments are frozen.  The observation buffer and the ring slot are part of the key (a caller
         * that reuses one buffer needs four graphs); everything else that changes from step to step goes through the
         * handle's own buffers: the actions are copied in before the replay and reward / done / score out after it. */
        mg_launch_step_io(h->cfg.num_envs, actions_dev, h->actions_dev, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, st);
        h->launches++;
        mg_handle::StepGraph *g = nullptr;
        for (auto &c : h->graphs)
            if (c.obs == obs_dev && c.slot == h->slots[view] && c.max_blocks == h->max_blocks) { g = &c; break; }
        if (!g) {
            if (h->graphs.size() >= 64) {          /* callers that never reuse their buffers: start over */
                for (auto &c : h->graphs) if (c.exec) cudaGraphExecDestroy(c.exec);
                h->graphs.clear();
            }
            h->graphs.push_back({ obs_dev, h->slots[view], h->max_blocks, 0, nullptr });
            g = &h->graphs.back();
        }
        if (!g->exec && g->seen++ == 0) {
            /* first use of this combination: eager (also sets the kernels' shared-memory attributes outside a capture) */
            rc = step_parts_body(h, h->actions_dev, obs_dev, h->reward_dev, h->done_dev, h->score_dev, st, view, 1);
        } else {
            if (!g->exec) {
                /* captured on the handle's own stream (the caller's may be the legacy default stream, which cannot
                 * capture) and launched on the caller's */
                cudaGraph_t graph = nullptr;
                cudaStream_t cs = h->own_stream;
                CK(h, cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal));
                const int64_t launches0 = h->launches;
                rc = step_parts_body(h, h->actions_dev, obs_dev, h->reward_dev, h->done_dev, h->score_dev, cs, view, 1);
                h->launches = launches0;
                cudaError_t e = cudaStreamEndCapture(cs, &graph);
                if (rc == MG_OK && e != cudaSuccess) rc = fail(h, MG_ERR_CUDA, std::string("cudaStreamEndCapture: ") + cudaGetErrorString(e));
                if (rc == MG_OK) {
                    e = cudaGraphInstantiate(&g->exec, graph, 0);
                    if (e != cudaSuccess) rc = fail(h, MG_ERR_CUDA, std::string("cudaGraphInstantiate: ") + cudaGetErrorString(e));
                }
                if (graph) cudaGraphDestroy(graph);
                if (rc != MG_OK) return rc;
            }
            CK(h, cudaGraphLaunch(g->exec, st));
            h->launches += 3 * h->step_parts + 1 + (h->step_order != 0 ? 1 : 0);      /* kernel nodes of the graph */
        }
        if (rc == MG_OK && (reward_dev || done_dev || score_dev)) {
            mg_launch_step_io(h->cfg.num_envs, nullptr, nullptr, h->reward_dev, reward_dev, h->done_dev, done_dev, h->score_dev, score_dev, st);
            h->launches++;
        }
    }
    if (rc != MG_OK) return rc;
    rc = pending_done(h, st);
    if (rc != MG_OK) return rc;
    mark_caller(h, st);
    return check_launch(h, "mg_step (parts)");
}

extern "C" int mg_step(mg_handle *h, const int32_t *actions_dev, uint8_t *obs_dev, float *reward_dev, uint8_t *done_dev,
                       float *score_dev, void *stream)
{
    if (h && actions_dev && obs_dev && h->fused && h->step_parts > 1 && !h->timing && h->cfg.num_envs >= 32 * h->step_parts)
        return step_in_parts(h, actions_dev, obs_dev, reward_dev, done_dev, score_dev, (cudaStream_t)stream);
    int rc = mg_step_physics(h, actions_dev, reward_dev, done_dev, score_dev, stream);
    if (rc != MG_OK) return rc;
    return mg_step_render(h, obs_dev, stream);
}

/* ------------------------------------------------------------------ mg_step_host
 * The end-to-end step.  The long pole is the observation's way to the host: 110 592 B per env over PCIe.  The batch is
 * rendered in MG_HOST_CHUNKS parts and the device-to-host copy of each part starts as soon as it is rendered.  With
 * host assembly on (mg_set_host_assembly), the first asm_chunks parts send only their NEW frame (a quarter of the bytes)
 * into a pinned mirror of the frame ring and host threads build their stacks while the remaining parts arrive as
 * finished stacks; the handle moves the boundary between the two sets until the link and the host finish together. */
extern "C" int mg_set_host_assembly(mg_handle *h, int mode, int chunks, int n_threads)
{
    if (!h || mode < 0 || mode > 2 || chunks < 0 || chunks > MG_HOST_CHUNKS) return fail(h, MG_ERR_ARG, "mg_set_host_assembly: bad argument");
    if (mode != 0 && !(h->fused && h->spec.n_groups > 0 && h->cfg.num_envs >= 4 * MG_HOST_CHUNKS))
        return fail(h, MG_ERR_ARG, "mg_set_host_assembly: needs a single-view HWC frame stack and at least 64 envs");
    CK(h, cudaSetDevice(h->cfg.device));
    if (mode != 0) {
        if (!h->ring_host) CK(h, cudaMallocHost(&h->ring_host, (size_t)h->cfg.num_envs * MG_MAX_RING * MG_FRAME_BYTES));
        if (n_threads <= 0) {
            const char *ev = getenv("MG_HOST_THREADS");
            n_threads = ev ? atoi(ev) : 0;
            if (n_threads <= 0) {
                /* the CPUs this process may run on (a container's share), not the machine's */
                cpu_set_t set;
                CPU_ZERO(&set);
                if (sched_getaffinity(0, sizeof(set), &set) == 0) n_threads = CPU_COUNT(&set);
                if (n_threads <= 0) n_threads = (int)std::thread::hardware_concurrency();
            }
            if (n_threads <= 0) n_threads = 1;
            if (n_threads > 64) n_threads = 64;
        }
        if (h->hostasm && mg_hostasm_threads(h->hostasm) != n_threads) { mg_hostasm_destroy(h->hostasm); h->hostasm = nullptr; }
        if (!h->hostasm) h->hostasm = mg_hostasm_create(n_threads);
        if (h->since_reset.empty()) h->since_reset.assign((size_t)h->cfg.num_envs, (uint8_t)(MG_MAX_RING - 1));
    }
    h->asm_mode = mode;
    h->asm_chunks = mode == 0 ? 0 : (mode == 1 && chunks == 0 ? MG_HOST_CHUNKS / 2 : chunks);
    h->asm_hold = 4;
    h->fresh_ids.clear();
    hostasm_invalidate(h);
    return MG_OK;
}

extern "C" int mg_get_host_assembly(const mg_handle *h, int *chunks_out, int *total_chunks_out, int *threads_out,
                                    double *copy_ms_out, double *host_ms_out)
{
    if (!h) return MG_ERR_ARG;
    if (chunks_out) *chunks_out = h->asm_mode ? h->asm_chunks : 0;
    if (total_chunks_out) *total_chunks_out = MG_HOST_CHUNKS;
    if (threads_out) *threads_out = h->hostasm ? mg_hostasm_threads(h->hostasm) : 0;
    if (copy_ms_out) *copy_ms_out = h->asm_last_copy_ms;
    if (host_ms_out) *host_ms_out = h->asm_last_host_ms;
    return MG_OK;
}

/* stacks from frames on the host, without a handle: what mg_step_host's workers do, for callers that keep their own
 * frame history (and for the CPU-side test of the assembler) */
extern "C" int mg_host_stack_frames(const uint8_t *ring_host, uint8_t *obs_host, int n_envs, int newest_slot,
                                    const int32_t *ages, int n_groups, const uint8_t *since_reset, int n_threads)
{
    if (!ring_host || !obs_host || n_envs < 0 || !ages || n_groups < 1 || n_groups > MG_MAX_GROUPS || !since_reset)
        return fail(nullptr, MG_ERR_ARG, "mg_host_stack_frames: bad argument");
    for (int g = 0; g < n_groups; g++)
        if (ages[g] < 0 || ages[g] >= MG_MAX_RING) return fail(nullptr, MG_ERR_ARG, "mg_host_stack_frames: bad age");
    MgHostAsm *A = mg_hostasm_create(n_threads > 0 ? n_threads : 1);
    int a[MG_MAX_GROUPS];
    for (int g = 0; g < n_groups; g++) a[g] = ages[g];
    mg_hostasm_begin(A, ring_host, obs_host, n_envs, newest_slot & (MG_MAX_RING - 1), a, n_groups, since_reset);
    mg_hostasm_publish(A, n_envs);
    mg_hostasm_finish(A);
    mg_hostasm_destroy(A);
    return MG_OK;
}

extern "C" int mg_step_host(mg_handle *h, const int32_t *actions_host, uint8_t *obs_host, float *reward_host,
                            uint8_t *done_host, float *score_host)
{
    if (!h || !actions_host) return fail(h, MG_ERR_ARG, "mg_step_host: bad argument");
    const auto t_begin = std::chrono::steady_clock::now();
    const int B = h->cfg.num_envs;
    cudaStream_t st = h->own_stream;
    CK(h, cudaSetDevice(h->cfg.device));
    /* resets / renders the caller queued on its own stream (mg_set_scenes, mg_reset_obs, mg_step*) come first */
    if (h->caller_ev_used) { CK(h, cudaStreamWaitEvent(st, h->caller_ev, 0)); h->caller_ev_used = 0; }
    const bool want_obs = obs_host && h->spec.n_groups > 0;
    const bool chunked = h->fused && want_obs && B >= 4 * MG_HOST_CHUNKS;
    int n_asm = chunked && h->asm_mode ? h->asm_chunks : 0, n_asm_envs = 0;
    if (n_asm) CK(h, cudaEventRecord(h->asm_t0, st));       /* the device-side clock of the balance starts with the step */
    CK(h, cudaMemcpyAsync(h->actions_dev, actions_host, (size_t)B * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    int rc = mg_step_physics(h, h->actions_dev, h->reward_dev, h->done_dev, h->score_dev, st);
    if (rc != MG_OK) return rc;
    if (chunked) {
        /* single-view HWC stack: render the batch in MG_HOST_CHUNKS parts and start the device-to-host copy of
         * each part as soon as it is rendered, so the PCIe transfer overlaps the rendering of the following parts */
        int view = -1;
        for (int v = 0; v < 2; v++) if (h->spec.render_view[v]) view = v;
        h->slots[view] = (h->slots[view] + 1) & (MG_MAX_RING - 1);
        const size_t env_bytes = h->obs_bytes / (size_t)B;
        if (n_asm) for (int32_t id : h->fresh_ids) if (id >= 0 && id < B) h->since_reset[id] = 0;
        h->fresh_ids.clear();
        for (int c = 0; c < MG_HOST_CHUNKS; c++) {
            const int e0 = (int)((long long)B * c / MG_HOST_CHUNKS), e1 = (int)((long long)B * (c + 1) / MG_HOST_CHUNKS);
            const bool assembled = c < n_asm;
            /* a host-assembled part only pushes its frame into the ring (no device-side stack) */
            mg_launch_render_push(h->S, view, h->slots[view], assembled ? nullptr : &h->spec, h->slots, assembled ? nullptr : h->obs_dev,
                                  e0, e1, st, h->pending_queued != 0);
            CK(h, cudaEventRecord(h->chunk_done[c], st));
            CK(h, cudaStreamWaitEvent(h->copy_stream, h->chunk_done[c], 0));
            if (assembled) {
                const size_t env_ring = (size_t)MG_MAX_RING * MG_FRAME_BYTES;
                if (h->asm_hist_ok[c]) {
                    /* the mirror has the older frames: the new ring slot of every env of the part, 27 648 B each */
                    const size_t off = (size_t)e0 * env_ring + (size_t)h->slots[view] * MG_FRAME_BYTES;
                    CK(h, cudaMemcpy2DAsync(h->ring_host + off, env_ring, h->S.ring[view] + off, env_ring, MG_FRAME_BYTES,
                                            (size_t)(e1 - e0), cudaMemcpyDeviceToHost, h->copy_stream));
                } else {
                    /* first time (or the rings moved behind the mirror's back): all four slots, as many bytes as the stacks */
                    CK(h, cudaMemcpyAsync(h->ring_host + (size_t)e0 * env_ring, h->S.ring[view] + (size_t)e0 * env_ring,
                                          (size_t)(e1 - e0) * env_ring, cudaMemcpyDeviceToHost, h->copy_stream));
                    for (int e = e0; e < e1; e++) h->since_reset[e] = MG_MAX_RING - 1;     /* the device filled a fresh env's four slots */
                    h->asm_hist_ok[c] = 1;
                }
                CK(h, cudaEventRecord(h->chunk_copied[c], h->copy_stream));
                n_asm_envs = e1;
            } else {
                h->asm_hist_ok[c] = 0;
                CK(h, cudaMemcpyAsync(obs_host + (size_t)e0 * env_bytes, h->obs_dev + (size_t)e0 * env_bytes,
                                      (size_t)(e1 - e0) * env_bytes, cudaMemcpyDeviceToHost, h->copy_stream));
            }
        }
        if (n_asm) CK(h, cudaEventRecord(h->asm_t1, h->copy_stream));
        mg_launch_clear_fresh(h->S, st);
        h->launches += 2 * MG_HOST_CHUNKS + 1;
        rc = pending_done(h, st);
        if (rc != MG_OK) return rc;
        rc = check_launch(h, "mg_render_kernel");
        if (rc != MG_OK) return rc;
    } else {
        rc = mg_step_render(h, h->obs_dev, st);          /* invalidates the host mirror */
        if (rc != MG_OK) return rc;
        h->fresh_ids.clear();
        if (want_obs) CK(h, cudaMemcpyAsync(obs_host, h->obs_dev, h->obs_bytes, cudaMemcpyDeviceToHost, st));
    }
    if (reward_host) CK(h, cudaMemcpyAsync(reward_host, h->reward_dev, (size_t)B * sizeof(float), cudaMemcpyDeviceToHost, st));
    if (done_host) CK(h, cudaMemcpyAsync(done_host, h->done_dev, (size_t)B, cudaMemcpyDeviceToHost, st));
    if (score_host) CK(h, cudaMemcpyAsync(score_host, h->score_dev, (size_t)B * sizeof(float), cudaMemcpyDeviceToHost, st));
    if (n_asm) {
        /* build the stacks of the host-assembled parts as their frames land */
        int view = -1;
        for (int v = 0; v < 2; v++) if (h->spec.render_view[v]) view = v;
        mg_hostasm_begin(h->hostasm, h->ring_host, obs_host, n_asm_envs, h->slots[view], h->spec.age, h->spec.n_groups, h->since_reset.data());
        for (int c = 0; c < n_asm; c++) {
            CK(h, cudaEventSynchronize(h->chunk_copied[c]));
            mg_hostasm_publish(h->hostasm, (int)((long long)B * (c + 1) / MG_HOST_CHUNKS));
        }
        mg_hostasm_finish(h->hostasm);
        const double host_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_begin).count();
        for (int e = 0; e < B; e++) if (h->since_reset[e] < MG_MAX_RING - 1) h->since_reset[e]++;
        CK(h, cudaStreamSynchronize(st));
        CK(h, cudaStreamSynchronize(h->copy_stream));
        float copy_ms = 0;
        CK(h, cudaEventElapsedTime(&copy_ms, h->asm_t0, h->asm_t1));
        h->asm_last_copy_ms = copy_ms; h->asm_last_host_ms = host_ms;
        /* both clocks start when the first part begins to render (the host got here within microseconds of queueing it):
         * move one part towards the side that finished first, with some hysteresis */
        if (h->asm_mode == 1 && --h->asm_hold <= 0) {
            const double slack = copy_ms - host_ms, tol = 0.04 * (copy_ms > host_ms ? copy_ms : host_ms) + 0.1;
            if (slack > tol && h->asm_chunks < MG_HOST_CHUNKS) { h->asm_chunks++; h->asm_hold = 3; }
            else if (slack < -tol && h->asm_chunks > 0) { h->asm_chunks--; h->asm_hold = 3; }
        }
        return MG_OK;
    }
    CK(h, cudaStreamSynchronize(st));
    CK(h, cudaStreamSynchronize(h->copy_stream));
    return MG_OK;
}

extern "C" int mg_render(mg_handle *h, int view, int full_res, uint8_t *out_dev, void *stream)
{
    if (!h || !out_dev || view < 0 || view > 1) return fail(h, MG_ERR_ARG, "mg_render: bad argument");
    CK(h, cudaSetDevice(h->cfg.device));
    mg_launch_render_out(h->S, view, full_res, out_dev, (cudaStream_t)stream);
    h->launches += 2;            /* geometry + raster */
    return check_launch(h, "mg_render_kernel");
}

extern "C" int mg_score(mg_handle *h, float *score_dev, void *stream)
{
    if (!h || !score_dev) return fail(h, MG_ERR_ARG, "mg_score: bad argument");
    CK(h, cudaSetDevice(h->cfg.device));
    mg_launch_score(h->S, score_dev, 0, (cudaStream_t)stream);
    h->launches++;
    return check_launch(h, "mg_score_kernel");
}

extern "C" int mg_debug_reward(mg_handle *h, float *reward_dev, void *stream)
{
    if (!h || !reward_dev) return fail(h, MG_ERR_ARG, "mg_debug_reward: bad argument");
    CK(h, cudaSetDevice(h->cfg.device));
    mg_launch_score(h->S, reward_dev, 1, (cudaStream_t)stream);
    h->launches++;
    return check_launch(h, "mg_score_kernel");
}

extern "C" int mg_get_bodies(mg_handle *h, double *out_host)
{
    if (!h || !out_host) return fail(h, MG_ERR_ARG, "mg_get_bodies: bad argument");
    const int B = h->cfg.num_envs;
    CK(h, cudaSetDevice(h->cfg.device));
    CK(h, cudaDeviceSynchronize());
    std::vector<double> poses((size_t)B * MG_POSE_STRIDE);
    CK(h, cudaMemcpy(poses.data(), h->S.poses, poses.size() * sizeof(double), cudaMemcpyDeviceToHost));
    for (int e = 0; e < B; e++)
        for (int b = 0; b < MG_MAX_BODIES; b++)
            for (int k = 0; k < 6; k++)
                out_host[((size_t)e * MG_MAX_BODIES + b) * 6 + k] = poses[(size_t)e * MG_POSE_STRIDE + (b + 1) * MG_POSE_NF + k];
    return MG_OK;
}

extern "C" int mg_get_status(mg_handle *h, int32_t *status_host, int clear)
{
    if (!h || !status_host) return fail(h, MG_ERR_ARG, "mg_get_status: bad argument");
    CK(h, cudaSetDevice(h->cfg.device));
    CK(h, cudaDeviceSynchronize());
    CK(h, cudaMemcpy(status_host, h->S.status, (size_t)h->cfg.num_envs * sizeof(int32_t), cudaMemcpyDeviceToHost));
    if (clear) CK(h, cudaMemset(h->S.status, 0, (size_t)h->cfg.num_envs * sizeof(int)));
    return MG_OK;
}

/* checkpoint blob: D_NF doubles, I_NF ints, the scene, the draw list and the frame rings of one env */
static const size_t BLOB_D = D_NF * sizeof(double);
static const size_t BLOB_I = I_NF * sizeof(int);
static const size_t BLOB_SCENE = sizeof(mg_scene);
static const size_t BLOB_PRIMS = MG_MAX_PRIMS * sizeof(MgPrim) + 8;
static const size_t BLOB_RING = 2 * MG_MAX_RING * MG_FRAME_BYTES;

extern "C" size_t mg_state_size(const mg_handle *h) { (void)h; return BLOB_D + BLOB_I + BLOB_SCENE + BLOB_PRIMS + BLOB_RING + 16; }

static int state_xfer(mg_handle *h, const int32_t *env_ids_host, int n, unsigned char *blobs, bool set)
{
    if (!h || !blobs || n < 0 || n > h->cfg.num_envs) return fail(h, MG_ERR_ARG, "mg_get/set_state: bad argument");
    if (n == 0) return MG_OK;
    CK(h, cudaSetDevice(h->cfg.device));
    CK(h, cudaDeviceSynchronize());
    const size_t sz = mg_state_size(h);
    double *dtmp; int *itmp;
    CK(h, cudaMalloc(&dtmp, (size_t)n * BLOB_D));
    CK(h, cudaMalloc(&itmp, (size_t)n * BLOB_I));
    std::vector<int32_t> ids(n);
    for (int i = 0; i < n; i++) {
        ids[i] = env_ids_host ? env_ids_host[i] : i;
        if (ids[i] < 0 || ids[i] >= h->cfg.num_envs) { cudaFree(dtmp); cudaFree(itmp); return fail(h, MG_ERR_ARG, "env id out of range"); }
    }
    CK(h, cudaMemcpy(h->ids_stage_dev, ids.data(), (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice));
    cudaMemcpyKind dir = set ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost;
    if (set)
        for (int i = 0; i < n; i++) {
            CK(h, cudaMemcpy(dtmp + (size_t)i * D_NF, blobs + i * sz, BLOB_D, dir));
            CK(h, cudaMemcpy(itmp + (size_t)i * I_NF, blobs + i * sz + BLOB_D, BLOB_I, dir));
        }
    mg_launch_state_gather(h->S, h->ids_stage_dev, n, dtmp, itmp, set ? 1 : 0, 0);
    h->launches++;
    CK(h, cudaDeviceSynchronize());
    for (int i = 0; i < n; i++) {
        unsigned char *b = blobs + i * sz;
        int e = ids[i];
        if (!set) {
            CK(h, cudaMemcpy(b, dtmp + (size_t)i * D_NF, BLOB_D, dir));
            CK(h, cudaMemcpy(b + BLOB_D, itmp + (size_t)i * I_NF, BLOB_I, dir));
        }
        unsigned char *p = b + BLOB_D + BLOB_I;
        void *dev_scene = h->S.scenes + e, *dev_prims = h->S.prims + (size_t)e * MG_MAX_PRIMS, *dev_np = h->S.n_prims + e;
        if (set) {
            CK(h, cudaMemcpy(dev_scene, p, BLOB_SCENE, dir)); p += BLOB_SCENE;
            CK(h, cudaMemcpy(dev_prims, p, MG_MAX_PRIMS * sizeof(MgPrim), dir)); p += MG_MAX_PRIMS * sizeof(MgPrim);
            CK(h, cudaMemcpy(dev_np, p, sizeof(int), dir)); p += 8;
        } else {
            CK(h, cudaMemcpy(p, dev_scene, BLOB_SCENE, dir)); p += BLOB_SCENE;
            CK(h, cudaMemcpy(p, dev_prims, MG_MAX_PRIMS * sizeof(MgPrim), dir)); p += MG_MAX_PRIMS * sizeof(MgPrim);
            CK(h, cudaMemcpy(p, dev_np, sizeof(int), dir)); p += 8;
        }
        for (int v = 0; v < 2; v++) {
            /* frames are stored oldest-first relative to the handle's current slot so a blob is portable */
            for (int a = 0; a < MG_MAX_RING; a++) {
                int slot = (h->slots[v] - (MG_MAX_RING - 1 - a)) & (MG_MAX_RING - 1);
                uint8_t *dev = h->S.ring[v] + ((size_t)e * MG_MAX_RING + slot) * MG_FRAME_BYTES;
                if (set) CK(h, cudaMemcpy(dev, p, MG_FRAME_BYTES, dir)); else CK(h, cudaMemcpy(p, dev, MG_FRAME_BYTES, dir));
                p += MG_FRAME_BYTES;
            }
        }
    }
    cudaFree(dtmp); cudaFree(itmp);
    return check_launch(h, "mg_state_copy_kernel");
}

extern "C" int mg_get_state(mg_handle *h, const int32_t *env_ids_host, int n, void *blobs_host)
{
    return state_xfer(h, env_ids_host, n, (unsigned char *)blobs_host, false);
}
extern "C" int mg_set_state(mg_handle *h, const int32_t *env_ids_host, int n, const void *blobs_host)
{
    /* restored envs may hold any number of blocks */
    if (h) { h->max_blocks = MG_MAX_BLOCKS; hostasm_invalidate(h); }
    return state_xfer(h, env_ids_host, n, (unsigned char *)blobs_host, true);
}

/* ------------------------------------------------------------------ mg_get_state_noalloc
 * The checkpoint gather for callers that may not allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc): same blobs as mg_get_state, but the device-side gather scratch that state_xfer
 * cudaMallocs per call lives in a file-scope __device__ array of fixed size instead, so this entry point calls no
 * cudaMalloc / cudaFree / allocating library routine.  Every GPU that loads the library holds the array until exit.
 * All calls share it: two calls that overlap on the GPU (two host threads, two handles on one device) race on it. */
#define MG_NOALLOC_IDS_BYTES (((size_t)MG_NOALLOC_MAX_ENVS * sizeof(int32_t) + 255) & ~(size_t)255)
#define MG_NOALLOC_D_BYTES ((size_t)MG_NOALLOC_MAX_ENVS * D_NF * sizeof(double))
#define MG_NOALLOC_I_BYTES ((size_t)MG_NOALLOC_MAX_ENVS * I_NF * sizeof(int))
#define MG_NOALLOC_BYTES (MG_NOALLOC_IDS_BYTES + MG_NOALLOC_D_BYTES + MG_NOALLOC_I_BYTES)
/* [ env ids | MAX x D_NF doubles | MAX x I_NF ints ]  (about 1.25 MB for 32 envs) */
static __device__ __align__(256) unsigned char mg_noalloc_scratch[MG_NOALLOC_BYTES];

/* variant of mg_state_copy_kernel (gather direction) whose ids and blobs live in the scratch passed in */
__global__ void mg_state_gather_noalloc_kernel(MgDeviceState S, int n, unsigned char *scratch)
{
    const int k = blockIdx.x;
    if (k >= n) return;
    const int32_t *ids = (const int32_t *)scratch;
    double *dblobs = (double *)(scratch + MG_NOALLOC_IDS_BYTES);
    int *iblobs = (int *)(scratch + MG_NOALLOC_IDS_BYTES + MG_NOALLOC_D_BYTES);
    const int env = ids[k];
    const double *d = S.dstate + mg_field_index(env, D_NF, 0);
    const int *iv = S.istate + mg_field_index(env, I_NF, 0);
    for (int f = threadIdx.x; f < D_NF; f += blockDim.x) dblobs[(size_t)k * D_NF + f] = d[(size_t)f * MG_GROUP];
    for (int f = threadIdx.x; f < I_NF; f += blockDim.x) iblobs[(size_t)k * I_NF + f] = iv[(size_t)f * MG_GROUP];
}

/* device address of the scratch, looked up once per device with that device current: in host code the array's
 * name is only a host-side stand-in for the symbol, never a device address */
static int noalloc_scratch_ptr(mg_handle *h, unsigned char **out)
{
    static unsigned char *ptr_dev[MG_MAX_DEVICES] = { nullptr };
    static std::mutex mu;
    const int dev = h->cfg.device;
    if (dev < 0 || dev >= MG_MAX_DEVICES) return fail(h, MG_ERR_ARG, "mg_get_state_noalloc: device ordinal beyond MG_MAX_DEVICES");
    std::lock_guard<std::mutex> lock(mu);
    if (!ptr_dev[dev]) {
        void *p = nullptr;
        CK(h, cudaGetSymbolAddress(&p, mg_noalloc_scratch));      /* cudaSetDevice(dev) was done by the caller */
        ptr_dev[dev] = (unsigned char *)p;
    }
    *out = ptr_dev[dev];
    return MG_OK;
}

extern "C" int mg_get_state_noalloc(mg_handle *h, const int32_t *env_ids_host, int n, void *blobs_host)
{
    if (!h || !blobs_host || n < 0) return fail(h, MG_ERR_ARG, "mg_get_state_noalloc: bad argument");
    if (n > MG_NOALLOC_MAX_ENVS) {
        char msg[160];
        snprintf(msg, sizeof msg, "mg_get_state_noalloc: n = %d exceeds MG_NOALLOC_MAX_ENVS = %d (fixed device scratch); "
                                  "call it in slices or use mg_get_state", n, MG_NOALLOC_MAX_ENVS);
        return fail(h, MG_ERR_ARG, msg);
    }
    if (n == 0) return MG_OK;
    int32_t ids[MG_NOALLOC_MAX_ENVS];
    for (int i = 0; i < n; i++) {
        ids[i] = env_ids_host ? env_ids_host[i] : i;
        if (ids[i] < 0 || ids[i] >= h->cfg.num_envs) return fail(h, MG_ERR_ARG, "mg_get_state_noalloc: env id out of range");
    }
    CK(h, cudaSetDevice(h->cfg.device));
    unsigned char *scratch = nullptr;
    int rc = noalloc_scratch_ptr(h, &scratch);
    if (rc != MG_OK) return rc;
    CK(h, cudaDeviceSynchronize());
    /* nothing in the scratch has to start at zero: the ids and every blob word read back below are written by this
     * call (ids by the copy, blobs by the kernel), whatever an earlier call left behind */
    CK(h, cudaMemcpy(scratch, ids, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice));
    mg_state_gather_noalloc_kernel<<<n, 128, 0, 0>>>(h->S, n, scratch);
    h->launches++;
    rc = check_launch(h, "mg_state_gather_noalloc_kernel");
    if (rc != MG_OK) return rc;
    CK(h, cudaDeviceSynchronize());
    const size_t sz = mg_state_size(h);
    const unsigned char *dblobs = scratch + MG_NOALLOC_IDS_BYTES, *iblobs = dblobs + MG_NOALLOC_D_BYTES;
    for (int i = 0; i < n; i++) {
        unsigned char *b = (unsigned char *)blobs_host + (size_t)i * sz;
        const int e = ids[i];
        CK(h, cudaMemcpy(b, dblobs + (size_t)i * BLOB_D, BLOB_D, cudaMemcpyDeviceToHost));
        CK(h, cudaMemcpy(b + BLOB_D, iblobs + (size_t)i * BLOB_I, BLOB_I, cudaMemcpyDeviceToHost));
        unsigned char *p = b + BLOB_D + BLOB_I;
        CK(h, cudaMemcpy(p, h->S.scenes + e, BLOB_SCENE, cudaMemcpyDeviceToHost)); p += BLOB_SCENE;
        CK(h, cudaMemcpy(p, h->S.prims + (size_t)e * MG_MAX_PRIMS, MG_MAX_PRIMS * sizeof(MgPrim), cudaMemcpyDeviceToHost)); p += MG_MAX_PRIMS * sizeof(MgPrim);
        CK(h, cudaMemcpy(p, h->S.n_prims + e, sizeof(int), cudaMemcpyDeviceToHost));
        memset(p + sizeof(int), 0, 8 - sizeof(int)); p += 8;
        for (int v = 0; v < 2; v++)
            for (int a = 0; a < MG_MAX_RING; a++) {
                const int slot = (h->slots[v] - (MG_MAX_RING - 1 - a)) & (MG_MAX_RING - 1);
                CK(h, cudaMemcpy(p, h->S.ring[v] + ((size_t)e * MG_MAX_RING + slot) * MG_FRAME_BYTES, MG_FRAME_BYTES, cudaMemcpyDeviceToHost));
                p += MG_FRAME_BYTES;
            }
        memset(p, 0, (size_t)(b + sz - p));      /* tail padding of the blob */
    }
    return MG_OK;
}

extern "C" int64_t mg_launch_count(const mg_handle *h) { return h ? h->launches : 0; }
extern "C" int mg_set_timing(mg_handle *h, int enable) { if (!h) return MG_ERR_ARG; h->timing = enable; return MG_OK; }
extern "C" double mg_kernel_ms(mg_handle *h, int which, int reset, int64_t *launches_out)
{
    if (!h || which < 0 || which > 1) return 0.0;
    for (auto &p : h->pending[which]) {
        float ms1 = 0;
        cudaEventSynchronize(p.second);
        if (cudaEventElapsedTime(&ms1, p.first, p.second) == cudaSuccess) h->ms[which] += ms1;
        h->free_events.push_back(p.first); h->free_events.push_back(p.second);
    }
    h->pending[which].clear();
    double ms = h->ms[which];
    if (launches_out) *launches_out = h->timed_launches[which];
    if (reset) { h->ms[which] = 0; h->timed_launches[which] = 0; }
    return ms;
}

/* ------------------------------------------------------------------ host-side placement predicate */

namespace {
struct HostShape { WShape s; int ent; };   /* ent: 0 walls, 1 robot, 2+i block i, 100+i goal i */

vec2 rot_of(double a) { double s, c; mg_sincos(a, &s, &c); return V(c, s); }

int entity_shapes(const MgTables *T, const mg_scene *sc, int kind, int idx, HostShape *out)
{
    int n = 0;
    if (kind == MG_ENT_ROBOT) {
        const double *r = sc->robot[MG_RB_ROBOT];
        wshape_from_tpl(T, MG_TPL_CIRCLE_ROBOT, V(r[0], r[1]), rot_of(r[2]), &out[n].s); out[n++].ent = 1;
        for (int k = 0; k < 2; k++)
            for (int part = 0; part < 2; part++) {
                const double *f = sc->robot[MG_RB_FINGER_L + k];
                wshape_from_tpl(T, MG_TPL_FINGER0 + k * 2 + part, V(f[0], f[1]), rot_of(f[2]), &out[n].s); out[n++].ent = 1;
            }
    } else if (kind == MG_ENT_BLOCK) {
        const mg_block *b = &sc->blocks[idx];
        vec2 p = V(b->x, b->y), rot = rot_of(b->angle);
        int ent = 2 + idx;
        switch (b->type) {
        case MG_SQUARE: wshape_from_tpl(T, MG_TPL_SQUARE, p, rot, &out[n].s); out[n++].ent = ent; break;
        case MG_CIRCLE: wshape_from_tpl(T, MG_TPL_CIRCLE_BLOCK, p, rot, &out[n].s); out[n++].ent = ent; break;
        case MG_STAR:
            for (int k = 0; k < MG_STAR_PIECES; k++) { wshape_from_tpl(T, MG_TPL_STAR0 + k, p, rot, &out[n].s); out[n++].ent = ent; }
            break;
        default: wshape_from_tpl(T, MG_TPL_PENT, p, rot, &out[n].s); out[n++].ent = ent; break;
        }
    } else {
        const mg_goal *g = &sc->goals[idx];
        double hw = g->w / 2, hh = g->h / 2;
        vec2 lv[4] = { V(hw, -hh), V(hw, hh), V(-hw, hh), V(-hw, -hh) };
        WShape *s = &out[n].s;
        s->kind = MG_SHAPE_POLY; s->n = 4; s->r = 0.0;
        for (int i = 0; i < 4; i++) {
            vec2 a = lv[(i + 3) % 4], b = lv[i];
            double dx = b.x - a.x, dy = b.y - a.y, nx = dy, ny = -dx;
            double inv = 1.0 / (sqrt(nx * nx + ny * ny) + MG_DBL_MIN);
            s->v[i] = V(b.x + g->cx, b.y + g->cy);
            s->nrm[i] = V(nx * inv, ny * inv);
        }
        out[n++].ent = 100 + idx;
    }
    return n;
}

bool touches(const WShape *a, const WShape *b)
{
    Manifold m;
    return (a->kind <= b->kind ? collide_shapes(a, b, &m) : collide_shapes(b, a, &m)) > 0;
}
} // namespace

extern "C" int mg_host_entity_overlaps(const mg_scene *scene, int kind, int idx, int robot_active,
                                       uint32_t block_mask, uint32_t goal_mask)
{
    const MgTables *T = mg_tables_host();
    HostShape mine[8], other[8];
    int nm = entity_shapes(T, scene, kind, idx, mine);
    for (int i = 0; i < nm; i++) {
        /* arena walls are always collidable */
        for (int w = 0; w < 4; w++) {
            WShape ws;
            wshape_from_tpl(T, MG_TPL_WALL0 + w, V(0, 0), V(1, 0), &ws);
            if (touches(&mine[i].s, &ws)) return 1;
        }
    }
    for (int pass = 0; pass < 3; pass++) {
        int count = pass == 0 ? 1 : (pass == 1 ? scene->n_blocks : scene->n_goals);
        for (int j = 0; j < count; j++) {
            int okind = pass == 0 ? MG_ENT_ROBOT : (pass == 1 ? MG_ENT_BLOCK : MG_ENT_GOAL);
            if (okind == kind && (kind == MG_ENT_ROBOT || j == idx)) continue;
            if (pass == 0 && !robot_active) continue;
            if (pass == 1 && !(block_mask >> j & 1)) continue;
            if (pass == 2 && !(goal_mask >> j & 1)) continue;
            int no = entity_shapes(T, scene, okind, j, other);
            for (int a = 0; a < nm; a++)
                for (int b = 0; b < no; b++)
                    if (touches(&mine[a].s, &other[b].s)) return 1;
        }
    }
    return 0;
}

extern "C" int mg_host_palette(uint8_t *rgb_out)
{
    const MgTables *T = mg_tables_host();
    for (int i = 0; i < MG_PAL_COUNT; i++)
        for (int k = 0; k < 3; k++) rgb_out[i * 3 + k] = T->pal[i][k];
    return MG_PAL_COUNT;
}
