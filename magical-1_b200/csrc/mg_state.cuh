/*
 * mg_state.cuh -- layout of the per-env simulation state in HBM.
 *
 * One env = one pymunk Space of the reference (magical/base_env.py:206-226): bodies,
 * constraints in add order, collision shapes in add order, Chipmunk's cached arbiters.
 * One env's fields are contiguous (array of structs): address(env, field) = env * N_FIELDS + field.
 * The step kernel gives one warp to one env, stages the env's bodies, constraints and shape list in shared
 * memory with coalesced loads, and leaves the (rarely touched) arbiter cache in HBM.
 * Doubles and 32-bit ints live in two separate arrays with the same scheme.
 */
#pragma once
#include "mg_core.cuh"

#define MG_GROUP 1   /* envs interleaved per field (1 = array of structs) */

/* body slots: 0 = static body (space.static_body / arena), 1..6 = robot group in
 * Robot.setup add order (MG_RB_*), 7.. = blocks in scene order */
#define MG_NB (1 + MG_RB_COUNT + MG_MAX_BLOCKS)
#define MG_BODY_ROBOT 1
#define MG_BODY_BLOCK0 (1 + MG_RB_COUNT)
#define MG_NCON 32                      /* 10 robot constraints + 2 per block */
#define MG_NSHAPE (4 + 5 + MG_MAX_BLOCKS * MG_STAR_PIECES)
#define MG_NARB 160                     /* cached arbiters (live + up to 3 stamps old); the oracle holds the same number
                                         * (oracle/orc.h ORC_MAX_ARBS) so that both overflow, and flag it, at the same point */

/* ---- double fields ---- */
enum { B_PX, B_PY, B_A, B_VX, B_VY, B_W, B_VBX, B_VBY, B_WB, B_RX, B_RY, B_MINV, B_IINV, B_NF };
enum {
    C_JMAX,      /* max_force * dt */
    C_MAXBIAS, C_BC /* 1 - error_bias^dt */,
    C_P0, C_P1,  /* pivot: anchor_a | gear: phase, ratio | limit: lo, hi | motor: rate | spring: w_coef */
    C_R1X, C_R1Y, C_R2X, C_R2Y,
    C_K0, C_K1, C_K2, C_K3,
    C_BIASX, C_BIASY,   /* scalar joints use C_BIASX; spring: target_wrn in C_BIASY */
    C_JX, C_JY,         /* accumulated impulse (scalar joints use C_JX) */
    C_ISUM,
    C_NF
};
enum {
    A_NX, A_NY, A_U,
    A_CON0,     /* per contact k: A_CON0 + k * AC_NF + ... */
    AC_R1X = 0, AC_R1Y, AC_R2X, AC_R2Y, AC_NMASS, AC_TMASS, AC_BIAS, AC_JN, AC_JT, AC_JBIAS, AC_NF,
    A_NF = A_CON0 + 2 * AC_NF
};
enum {
    D_BODY0 = 0,
    D_CON0 = D_BODY0 + MG_NB * B_NF,
    D_ARB0 = D_CON0 + MG_NCON * C_NF,
    D_GOAL0 = D_ARB0 + MG_NARB * A_NF,          /* per goal: cx, cy, h, w */
    D_TARGET = D_GOAL0 + MG_MAX_GOALS * 4,      /* PickAndPlace target_position (2) */
    D_NF = D_TARGET + 2
};

/* ---- int fields ---- */
enum {
    I_TASK, I_HORIZON, I_NBLOCKS, I_NGOALS, I_ROBOT_FIRST, I_DEBUG_REWARD, I_TARGET_BLOCK, I_NVALID,
    I_NBODIES,      /* body slots in use (1 + 6 + n_blocks) */
    I_NCON, I_NSHAPES, I_NARBS,
    I_STAMP, I_STEPPED /* 0 until the first space step (Chipmunk prev_dt == 0) */, I_EPISODE_STEPS,
    I_FRESH,        /* frame stacks must be refilled with the next rendered frame */
    I_STATUS,       /* MG_STATUS_* bits raised since the last reset (copied into MgDeviceState::status, which is sticky) */
    I_BLOCK0,       /* per block: type | colour << 8 | role << 16 */
    I_GOAL0 = I_BLOCK0 + MG_MAX_BLOCKS,   /* per goal: colour | (want_block + 1) << 8 */
    I_CON0 = I_GOAL0 + MG_MAX_GOALS,      /* per constraint: kind | a << 8 | b << 16 */
    I_SHAPE0 = I_CON0 + MG_NCON,          /* per shape: tpl | body << 8 | ent << 16 */
    I_ARB0 = I_SHAPE0 + MG_NSHAPE,        /* per arbiter: 4 ints, see AI_* */
    I_NF = I_ARB0 + MG_NARB * 4
};
enum {
    AI_KEY,     /* sa | sb << 8 | ba << 16 | bb << 24 (a = lower cpShapeType) */
    AI_STAMP,
    AI_FLAGS,   /* state | count << 8 | live << 16 */
    AI_HASH     /* hash0 | hash1 << 16 */
};
enum { ARB_FIRST = 0, ARB_NORMAL = 1, ARB_CACHED = 2 };
enum { CON_PIVOT = 0, CON_GEAR = 1, CON_SPRING = 2, CON_LIMIT = 3, CON_MOTOR = 4 };

/* entity ids used for the group / same-body filter */
#define MG_ENTID_ARENA 0
#define MG_ENTID_ROBOT 1
#define MG_ENTID_BLOCK0 2

/* per-env pose export consumed by the render kernel and mg_get_bodies:
 * [MG_NB][8] = x, y, angle, vx, vy, w, cos(angle), sin(angle)  (AoS, one env contiguous) */
#define MG_POSE_NF 8
#define MG_POSE_STRIDE (MG_NB * MG_POSE_NF)

MG_HD size_t mg_field_index(int env, int n_fields, int field)
{
    return ((size_t)(env / MG_GROUP) * n_fields + field) * MG_GROUP + (env % MG_GROUP);
}

/* number of draw primitives of an env is bounded by: arena 1 + goals 3 + robot 13 + blocks 10*12 */
#define MG_MAX_PRIMS 140

/* compact draw-list entry built at reset (mg_init_kernel) and interpreted by mg_raster.cu */
struct MgPrim {
    uint8_t dl;       /* MG_DL_* vertex list */
    uint8_t chain;    /* MG_CHAIN_* transform chain */
    uint8_t body;     /* body slot for the chain */
    uint8_t fill;     /* palette index */
    uint8_t outline;  /* palette index */
    uint8_t mode;     /* 0 none, 1 solid width 2, 2 dashed width 4 */
    uint8_t aux;      /* goal index (MG_CHAIN_GOAL) / eye side 0,1 (MG_CHAIN_EYE, MG_CHAIN_PUPIL) */
    uint8_t pad;
};
/* screen-space primitive written by mg_geometry_kernel for one view of one env: the rasteriser's bands read
 * these (16 B, one coalesced load per 32 primitives) instead of redoing the FP64 transform chain per band */
#define MG_MAX_VERTS 1600      /* per env and view: 15 circles of 100 + polygons; see mg_geometry_kernel */
struct MgScreenPrim {
    int off_n;        /* first vertex in the env's vertex pool << 8 | vertex count (0 = culled) */
    int miny, maxy;   /* truncated vertex rows */
    uint32_t style;   /* fill | outline << 8 | mode << 16 | aux << 24 */
};
enum { MG_CHAIN_CAM = 0, MG_CHAIN_BODY = 1, MG_CHAIN_EYE = 2, MG_CHAIN_PUPIL = 3, MG_CHAIN_GOAL = 4 };
