/*
 * mg_envref.cuh -- access to one env's state (EnvRef and the field macros) and the reset expansion
 * (BaseEnv.reset after on_reset(): scene -> bodies, constraints, shapes, draw list), shared by the physics
 * translation unit (mg_init_kernel) and the rasteriser's geometry pre-pass, which applies resets queued with
 * mg_queue_resets between the physics and the render of a step.
 */
#pragma once
#include <cuda_runtime.h>
#include "mg_state.cuh"
#include "mg_kernels.h"
#include "../../include/mg_trig.h"

/* d / i: the env's double and int fields up to (not including) the arbiter cache -- HBM, or the step
 * kernel's shared-memory copy (bodies at d, constraints at dc); da / ia: the env's full field arrays in HBM (arbiters, goals, target) */
struct EnvRef {
    double *d; double *dc; int *i; double *da; int *ia;
    /* step kernel only: shared-memory pool of the first MG_ARB_POOL LIVE arbiters of the current sub-step (A_NF doubles
     * each, same field layout as a cache slot) and their packed a | b << 8 | count << 16 | state << 24 words */
    double *ap; int *ainfo; int apn;     /* apn: pool entries in use by this launch (0: none) */
};
#define DF(e, f) ((e).d[(size_t)(f) * MG_GROUP])
#define IF(e, f) ((e).i[(size_t)(f) * MG_GROUP])
#define DGF(e, f) ((e).da[(size_t)(f) * MG_GROUP])
#define BF(e, b, f) DF(e, D_BODY0 + (b) * B_NF + (f))
#define CF(e, c, f) ((e).dc[(size_t)((c) * C_NF + (f)) * MG_GROUP])
#define AF(e, a, f) DGF(e, D_ARB0 + (a) * A_NF + (f))
#define ACF(e, a, k, f) DGF(e, D_ARB0 + (a) * A_NF + A_CON0 + (k) * AC_NF + (f))
#define AIF(e, a, f) ((e).ia[(size_t)(I_ARB0 + (a) * 4 + (f)) * MG_GROUP])

__device__ __forceinline__ EnvRef env_ref(double *dstate, int *istate, int env)
{
    EnvRef e;
    e.d = e.da = dstate + mg_field_index(env, D_NF, 0);
    e.dc = e.d + (size_t)D_CON0 * MG_GROUP;
    e.i = e.ia = istate + mg_field_index(env, I_NF, 0);
    e.ap = nullptr; e.ainfo = nullptr; e.apn = 0;
    return e;
}

#define MG_DT ((1.0 / 8) / 10)   /* fps = 8 (benchmarks/__init__.py:435), phys_steps = 10 (base_env.py:249) */

/* ------------------------------------------------------------------ reset expansion */

static __device__ void init_body(EnvRef e, int b, int dynamic, double m, double i, double x, double y, double a)
{
    for (int f = 0; f < B_NF; f++) BF(e, b, f) = 0.0;
    BF(e, b, B_PX) = x; BF(e, b, B_PY) = y; BF(e, b, B_A) = a;
    double s, c;
    mg_sincos(a, &s, &c);
    BF(e, b, B_RX) = c; BF(e, b, B_RY) = s;
    BF(e, b, B_MINV) = dynamic ? 1.0 / m : 0.0;
    BF(e, b, B_IINV) = dynamic ? 1.0 / i : 0.0;
}

static __device__ int add_con(EnvRef e, int kind, int a, int b, double max_force, double max_bias, double bc)
{
    int c = IF(e, I_NCON);
    IF(e, I_NCON) = c + 1;
    IF(e, I_CON0 + c) = kind | (a << 8) | (b << 16);
    for (int f = 0; f < C_NF; f++) CF(e, c, f) = 0.0;
    CF(e, c, C_JMAX) = max_force * MG_DT;
    CF(e, c, C_MAXBIAS) = max_bias;
    CF(e, c, C_BC) = bc;
    return c;
}

static __device__ void add_shape(EnvRef e, int tpl, int body, int ent)
{
    int s = IF(e, I_NSHAPES);
    IF(e, I_NSHAPES) = s + 1;
    IF(e, I_SHAPE0 + s) = tpl | (body << 8) | (ent << 16);
}

static __device__ void add_prim(MgPrim *prims, int *n, int dl, int chain, int body, int fill, int outline, int mode, int aux)
{
    MgPrim p;
    p.dl = (uint8_t)dl; p.chain = (uint8_t)chain; p.body = (uint8_t)body; p.fill = (uint8_t)fill;
    p.outline = (uint8_t)outline; p.mode = (uint8_t)mode; p.aux = (uint8_t)aux; p.pad = 0;
    prims[(*n)++] = p;
}

/* Robot.setup (entities.py:238-433) */
static __device__ void setup_robot(EnvRef e, const MgTables *T, const mg_scene *sc, MgPrim *prims, int *np)
{
    const double radius = 0.2, mass = 1.0, inf = INFINITY;
    const int rb = MG_BODY_ROBOT;
    double inertia = mass * (0.5 * (0.0 * 0.0 + radius * radius) + 0.0);
    init_body(e, rb + MG_RB_ROBOT, 1, mass, inertia, sc->robot[MG_RB_ROBOT][0], sc->robot[MG_RB_ROBOT][1], sc->robot[MG_RB_ROBOT][2]);
    init_body(e, rb + MG_RB_CONTROL, 0, 0, 0, sc->robot[MG_RB_CONTROL][0], sc->robot[MG_RB_CONTROL][1], sc->robot[MG_RB_CONTROL][2]);
    add_con(e, CON_PIVOT, rb + MG_RB_CONTROL, rb + MG_RB_ROBOT, sc->phys[0], 0.0, T->bias_coef_default);
    int c = add_con(e, CON_GEAR, rb + MG_RB_CONTROL, rb + MG_RB_ROBOT, sc->phys[1], 2.5, 1.0);
    CF(e, c, C_P0) = 0.0; CF(e, c, C_P1) = 1.0;
    for (int k = 0; k < 2; k++) {
        double eye_mass = mass / 10;
        double eye_inertia = eye_mass * (0.5 * (radius * radius));
        int pb = rb + MG_RB_PUPIL_L + k;
        init_body(e, pb, 1, eye_mass, eye_inertia, sc->robot[MG_RB_PUPIL_L + k][0], sc->robot[MG_RB_PUPIL_L + k][1], sc->robot[MG_RB_PUPIL_L + k][2]);
        c = add_con(e, CON_SPRING, rb + MG_RB_ROBOT, pb, 0.001, 3.0, T->bias_coef_default);
        CF(e, c, C_P0) = T->spring_w_coef;
    }
    for (int k = 0; k < 2; k++) {
        double side = k == 0 ? -1.0 : 1.0;
        double lower = k == 0 ? -0.0 : -(M_PI / 8);
        double upper = k == 0 ? (M_PI / 8) : 0.0;
        int fb = rb + MG_RB_FINGER_L + k;
        init_body(e, fb, 1, mass / 8, T->finger_moment[k], sc->robot[MG_RB_FINGER_L + k][0], sc->robot[MG_RB_FINGER_L + k][1], sc->robot[MG_RB_FINGER_L + k][2]);
        c = add_con(e, CON_PIVOT, rb + MG_RB_ROBOT, fb, inf, inf, 1.0);
        CF(e, c, C_P0) = side * radius * 0.45; CF(e, c, C_P1) = radius * 0.1;
        c = add_con(e, CON_LIMIT, rb + MG_RB_ROBOT, fb, inf, inf, 1.0);
        CF(e, c, C_P0) = lower; CF(e, c, C_P1) = upper;
        c = add_con(e, CON_MOTOR, rb + MG_RB_ROBOT, fb, sc->phys[2], 0.0, T->bias_coef_default);
        CF(e, c, C_P0) = 0.0;
    }
    add_shape(e, MG_TPL_CIRCLE_ROBOT, rb + MG_RB_ROBOT, MG_ENTID_ROBOT);
    for (int k = 0; k < 2; k++)
        for (int part = 0; part < 2; part++)
            add_shape(e, MG_TPL_FINGER0 + k * 2 + part, rb + MG_RB_FINGER_L + k, MG_ENTID_ROBOT);
    /* graphics, entities.py:374-433 */
    for (int k = 0; k < 2; k++)
        for (int part = 0; part < 2; part++)
            add_prim(prims, np, MG_DL_FINGER_OUT0 + k * 2 + part, MG_CHAIN_BODY, rb + MG_RB_FINGER_L + k, MG_PAL_GREY, 0, 0, 0);
    for (int k = 0; k < 2; k++)
        for (int part = 0; part < 2; part++)
            add_prim(prims, np, MG_DL_FINGER_IN0 + k * 2 + part, MG_CHAIN_BODY, rb + MG_RB_FINGER_L + k, MG_PAL_GREY_LIGHT, 0, 0, 0);
    add_prim(prims, np, MG_DL_ROBOT, MG_CHAIN_BODY, rb + MG_RB_ROBOT, MG_PAL_GREY, MG_PAL_GREY_DARK, 1, 0);
    for (int k = 0; k < 2; k++) {
        add_prim(prims, np, MG_DL_EYE, MG_CHAIN_EYE, rb + MG_RB_ROBOT, MG_PAL_WHITE, 0, 0, k);
        add_prim(prims, np, MG_DL_PUPIL, MG_CHAIN_PUPIL, rb + MG_RB_PUPIL_L + k, MG_PAL_PUPIL, 0, 0, k);
    }
}

/* Shape.setup (entities.py:610-749) */
static __device__ void setup_block(EnvRef e, const MgTables *T, const mg_scene *sc, int bi, MgPrim *prims, int *np)
{
    const mg_block *bl = &sc->blocks[bi];
    const double mass = 0.5, shape_size = 0.2 * 0.6;
    int body = MG_BODY_BLOCK0 + bi, ent = MG_ENTID_BLOCK0 + bi;
    int col = MG_PAL_COLOUR0 + bl->colour, dark = MG_PAL_DARK0 + bl->colour;
    switch (bl->type) {
    case MG_SQUARE:
        init_body(e, body, 1, mass, T->square_moment, bl->x, bl->y, bl->angle);
        add_shape(e, MG_TPL_SQUARE, body, ent);
        break;
    case MG_CIRCLE:
        init_body(e, body, 1, mass, mass * (0.5 * (shape_size * shape_size)), bl->x, bl->y, bl->angle);
        add_shape(e, MG_TPL_CIRCLE_BLOCK, body, ent);
        break;
    case MG_STAR:
        init_body(e, body, 1, mass, T->star_moment, bl->x, bl->y, bl->angle);
        for (int k = 0; k < MG_STAR_PIECES; k++) add_shape(e, MG_TPL_STAR0 + k, body, ent);
        break;
    default:
        init_body(e, body, 1, mass, T->pent_moment, bl->x, bl->y, bl->angle);
        add_shape(e, MG_TPL_PENT, body, ent);
        break;
    }
    /* floor friction: entities.py:700-708 */
    add_con(e, CON_PIVOT, 0, body, sc->phys[3], 0.0, T->bias_coef_default);
    int c = add_con(e, CON_GEAR, 0, body, sc->phys[4], 0.0, T->bias_coef_default);
    CF(e, c, C_P0) = 0.0; CF(e, c, C_P1) = 1.0;
    switch (bl->type) {
    case MG_SQUARE: add_prim(prims, np, MG_DL_SQUARE, MG_CHAIN_BODY, body, col, dark, 1, 0); break;
    case MG_CIRCLE: add_prim(prims, np, MG_DL_CIRCLE, MG_CHAIN_BODY, body, col, dark, 1, 0); break;
    case MG_STAR:
        for (int k = 0; k < MG_STAR_PIECES; k++) add_prim(prims, np, MG_DL_STAR_OUT0 + k, MG_CHAIN_BODY, body, dark, 0, 0, 0);
        for (int k = 0; k < MG_STAR_PIECES; k++) add_prim(prims, np, MG_DL_STAR_IN0 + k, MG_CHAIN_BODY, body, col, 0, 0, 0);
        break;
    default: add_prim(prims, np, MG_DL_PENT, MG_CHAIN_BODY, body, col, dark, 1, 0); break;
    }
}

static __device__ void export_poses(EnvRef e, double *pose)
{
    int nb = IF(e, I_NBODIES);
    for (int b = 0; b < MG_NB; b++) {
        double *o = pose + b * MG_POSE_NF;
        if (b < nb) {
            o[0] = BF(e, b, B_PX); o[1] = BF(e, b, B_PY); o[2] = BF(e, b, B_A);
            o[3] = BF(e, b, B_VX); o[4] = BF(e, b, B_VY); o[5] = BF(e, b, B_W);
            o[6] = BF(e, b, B_RX); o[7] = BF(e, b, B_RY);
        } else {
            for (int k = 0; k < MG_POSE_NF; k++) o[k] = 0.0;
        }
    }
}

/* BaseEnv.reset after on_reset(): expand one host-sampled scene into simulation state
 * (magical/base_env.py:190-246; add order arena, goals, [robot], blocks, [robot]).  One thread. */
static __device__ void init_env(const MgDeviceState &S, const mg_scene *sc, int env)
{
    const MgTables *T = S.tables;
    EnvRef e = env_ref(S.dstate, S.istate, env);
    S.scenes[env] = *sc;
    for (int f = 0; f < I_ARB0; f++) IF(e, f) = 0;
    IF(e, I_TASK) = sc->task; IF(e, I_HORIZON) = sc->horizon;
    IF(e, I_NBLOCKS) = sc->n_blocks; IF(e, I_NGOALS) = sc->n_goals;
    IF(e, I_ROBOT_FIRST) = sc->robot_first; IF(e, I_DEBUG_REWARD) = sc->debug_reward;
    IF(e, I_TARGET_BLOCK) = sc->target_block; IF(e, I_NVALID) = sc->n_valid_targets;
    IF(e, I_NBODIES) = 1 + MG_RB_COUNT + sc->n_blocks;
    IF(e, I_FRESH) = 1;
    DGF(e, D_TARGET) = sc->target_pos[0]; DGF(e, D_TARGET + 1) = sc->target_pos[1];
    init_body(e, 0, 0, 0, 0, 0.0, 0.0, 0.0);
    MgPrim *prims = S.prims + (size_t)env * MG_MAX_PRIMS;
    int np = 0;
    for (int w = 0; w < 4; w++) add_shape(e, MG_TPL_WALL0 + w, 0, MG_ENTID_ARENA);
    add_prim(prims, &np, MG_DL_ARENA, MG_CHAIN_CAM, 0, MG_PAL_WHITE, MG_PAL_GREY, 1, 0);
    for (int g = 0; g < sc->n_goals; g++) {
        const mg_goal *go = &sc->goals[g];
        DGF(e, D_GOAL0 + g * 4 + 0) = go->cx; DGF(e, D_GOAL0 + g * 4 + 1) = go->cy;
        DGF(e, D_GOAL0 + g * 4 + 2) = go->h;  DGF(e, D_GOAL0 + g * 4 + 3) = go->w;
        IF(e, I_GOAL0 + g) = go->colour | ((go->want_block + 1) << 8);
        add_prim(prims, &np, MG_DL_GOAL, MG_CHAIN_GOAL, 0, MG_PAL_LIGHT0 + go->colour, MG_PAL_COLOUR0 + go->colour, 2, g);
    }
    for (int b = 0; b < sc->n_blocks; b++)
        IF(e, I_BLOCK0 + b) = sc->blocks[b].type | (sc->blocks[b].colour << 8) | (sc->blocks[b].role << 16);
    if (sc->robot_first) setup_robot(e, T, sc, prims, &np);
    for (int b = 0; b < sc->n_blocks; b++) setup_block(e, T, sc, b, prims, &np);
    if (!sc->robot_first) setup_robot(e, T, sc, prims, &np);
    S.n_prims[env] = np;
    export_poses(e, S.poses + (size_t)env * MG_POSE_STRIDE);
}
