/*
 * orc.h -- CPU ORACLE for the MAGICAL per-step hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may build, load or call anything under oracle/.  The product
 * (magical-1_b200/) never does.
 *
 * PARITY UNPINNED.  The arithmetic of this path lives in third-party packages
 * that are neither vendored under /root/reference nor installed here:
 *   - pymunk ~=5.6.0 -> Chipmunk2D 7.0.3 (rigid bodies, joints, collision)   setup.py:33
 *   - pygame ==1.9.6 -> SDL 1.2 software surfaces (draw.polygon/lines/line)  setup.py:42
 * and the reference's own tests hold no golden vectors for it
 * (tests/test_rollout_preproc.py:17-36 asserts only the trajectory length).
 * This oracle therefore restates the *published algorithms* of those libraries
 * (cpSpaceStep.c, cpArbiter.c, cpCollision.c, the constraint files, pygame's
 * draw.c) in the order the reference's call sites drive them, and every
 * function cites the reference file:line whose behaviour it stands in for.
 * The one piece that IS pinned here is the 384->96 INTER_AREA downsample
 * (cv2 4.13 is importable; see tests/test_oracle_preproc.py).
 *
 * Plain C, IEEE double, compiled with -ffp-contract=off.
 */
#ifndef ORC_H
#define ORC_H

#include <stdint.h>
#include <stddef.h>
#include "../include/mg_scene.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct { double x, y; } vec;

#define ORC_MAXV 8          /* max vertices of one convex collision polygon */
#define ORC_MAX_BODIES 20
#define ORC_MAX_SHAPES 80
#define ORC_MAX_CONSTRAINTS 40
#define ORC_MAX_ARBS 160
#define ORC_MAX_ENTS 16
#define ORC_MAX_PRIMS 160
#define ORC_CIRCLE_RES 100  /* render.make_circle(radius, 100, ...) entities.py:379,414,420,715 */

enum { ORC_BODY_DYNAMIC = 0, ORC_BODY_KINEMATIC = 1, ORC_BODY_STATIC = 2 };
enum { ORC_SHAPE_CIRCLE = 0, ORC_SHAPE_SEGMENT = 1, ORC_SHAPE_POLY = 2 }; /* cpShapeType order */
enum { ORC_ENT_ARENA = 0, ORC_ENT_GOAL = 1, ORC_ENT_ROBOT = 2, ORC_ENT_BLOCK = 3 };
enum { ORC_C_PIVOT = 0, ORC_C_GEAR = 1, ORC_C_SPRING = 2, ORC_C_LIMIT = 3, ORC_C_MOTOR = 4 };
enum { ORC_ARB_FIRST = 0, ORC_ARB_NORMAL = 1, ORC_ARB_CACHED = 2 };

typedef struct {
    int type;
    double m, i, m_inv, i_inv;
    vec p, v;
    double a, w;
    vec v_bias;
    double w_bias;
    vec rot; /* (cos a, sin a) refreshed whenever a changes (cpBody SetTransform) */
} orc_body;

typedef struct {
    int kind;
    int body;   /* index into bodies[]; ORC_STATIC_BODY for space.static_body-like */
    int entity; /* owning entity index */
    double r;   /* circle radius / segment radius / poly bevel radius */
    double u;   /* friction */
    int sensor;
    uint32_t group, categories, mask;
    int count;             /* poly: vertex count; segment: 2 */
    vec lv[ORC_MAXV];      /* local vertices (poly: hull order; segment: a, b; circle: lv[0] = centre) */
    vec ln[ORC_MAXV];      /* local edge normals: ln[i] belongs to edge lv[i-1] -> lv[i] */
    vec tv[ORC_MAXV];      /* world-space copies, refreshed by orc_shape_update */
    vec tn[ORC_MAXV];
    double bb_l, bb_b, bb_r, bb_t;
} orc_shape;

typedef struct {
    int kind, a, b; /* body indices */
    double max_force, max_bias, error_bias;
    /* parameters */
    vec anchor_a, anchor_b;       /* pivot */
    double phase, ratio;          /* gear */
    double rest, stiffness, damping; /* rotary spring */
    double lo, hi;                /* rotary limit */
    double rate;                  /* motor */
    /* per-step scratch + warm start */
    vec r1, r2, bias_v, jacc_v;
    double k[4];
    double isum, bias, jacc, w_coef, target_wrn;
} orc_constraint;

typedef struct {
    vec r1, r2;
    double n_mass, t_mass, bounce, bias, jn_acc, jt_acc, j_bias;
    uint32_t hash;
} orc_contact;

typedef struct {
    int sa, sb;      /* shape indices (a = lower cpShapeType) */
    int ba, bb;      /* body indices */
    int count;
    orc_contact con[2];
    vec n;
    double e, u;
    int state;
    int stamp;       /* last space stamp at which the pair collided */
    int live;        /* in the solver list this step */
} orc_arbiter;

typedef struct {
    int kind;
    int first_shape, n_shapes;
    int first_body, n_bodies;
    int idx; /* block / goal index inside the scene */
} orc_entity;

/* One draw primitive = one render.Poly (render.py:202-227). */
typedef struct {
    int n;                       /* vertex count */
    const vec *lv;               /* local vertices */
    vec lv_own[4];               /* storage for per-env rectangles (goal) */
    int n_xf;                    /* transform chain, applied first-to-last to the point */
    struct { int kind; double tx, ty, rot; int body; int body_rel; } xf[4];
    uint8_t fill[3], outline[3];
    int outline_mode;            /* 0 none, 1 solid width 2, 2 dashed width 4 */
} orc_prim;

typedef struct orc_env {
    mg_scene scene;
    int n_bodies, n_shapes, n_constraints, n_arbs, n_ents, n_prims;
    orc_body bodies[ORC_MAX_BODIES];
    orc_shape shapes[ORC_MAX_SHAPES];
    orc_constraint cons[ORC_MAX_CONSTRAINTS];
    orc_arbiter arbs[ORC_MAX_ARBS];
    orc_entity ents[ORC_MAX_ENTS];
    orc_prim prims[ORC_MAX_PRIMS];
    int robot_body0;            /* index of robot body; +MG_RB_* for the others */
    int block_body[MG_MAX_BLOCKS];
    int block_ent[MG_MAX_BLOCKS];
    int goal_shape[MG_MAX_GOALS];
    int goal_ent[MG_MAX_GOALS];
    int robot_ent;
    int motor_con[2];
    /* Robot fields (entities.py:219-228, 435-453) */
    double rel_turn_angle, target_speed, target_finger_angle;
    /* Space fields */
    int stamp;
    double curr_dt;
    int iterations;
    double collision_slop, collision_bias;
    int collision_persistence;
    int episode_steps;
    int group_ctr;
} orc_env;

#define ORC_STATIC_BODY (ORC_MAX_BODIES - 1) /* slot reserved for static bodies (m_inv = i_inv = 0, p = 0) */

/* ---- geometry tables (orc_geom.c) ---- */
typedef struct {
    int n_pieces;
    int piece_n[8];
    vec piece_v[8][ORC_MAXV];
} orc_piece_set;

typedef struct {
    /* collision geometry */
    vec finger_phys[2][2][4];       /* [side][upper,fore][vertex], hull order */
    vec finger_inner[2][2][4];      /* draw-only inner quads */
    double finger_moment[2];
    vec square_phys[4];  double square_r, square_side, square_moment;
    vec square_draw[4];
    vec pent[5];         double pent_moment;      /* original (draw) order */
    vec pent_phys[5];                              /* hull order */
    vec star[10];        double star_moment;
    orc_piece_set star_outer, star_inner;          /* draw pieces (and outer = collision pieces, hull order) */
    vec circle_block[ORC_CIRCLE_RES];
    vec circle_robot[ORC_CIRCLE_RES];
    vec circle_eye[ORC_CIRCLE_RES];
    vec circle_pupil[ORC_CIRCLE_RES];
    vec arena_draw[4];
    uint8_t pal_bg[3], pal_white[3], pal_grey[3], pal_grey_dark[3], pal_grey_light[3], pal_pupil[3];
    uint8_t pal_colour[4][3], pal_colour_dark[4][3], pal_colour_light2[4][3];
} orc_geom_tables;

const orc_geom_tables *orc_geom(void);
int orc_hull_order(int n, const vec *in, vec *out); /* cpConvexHull vertex order for an already-convex CCW loop */
double orc_moment_for_poly(double m, int n, const vec *v);

/* ---- scene / space construction (orc_scene.c) ---- */
void orc_robot_default_poses(double x, double y, double angle, double out[MG_RB_COUNT][3]);
orc_env *orc_create(const mg_scene *scene);
void orc_destroy(orc_env *env);
void orc_body_set_angle(orc_body *b, double a);
void orc_shape_update(orc_env *env, int si);

/* ---- collision + dynamics (orc_collide.c, orc_phys.c) ---- */
typedef struct { int count; vec n; vec p1[2], p2[2]; uint32_t hash[2]; int sa, sb; } orc_manifold;
int orc_collide_shapes(const orc_shape *a, const orc_shape *b, orc_manifold *out);
int orc_filter_reject(const orc_shape *a, const orc_shape *b);
void orc_space_step(orc_env *env, double dt);
void orc_env_step(orc_env *env, int action);      /* BaseEnv.step minus render: set_action + 10 sub-steps */
int orc_shape_query(orc_env *env, int si, int *hits, int max_hits); /* space.shape_query(shape) */
void orc_get_bodies(const orc_env *env, double *out); /* [n_bodies][6] x,y,a,vx,vy,w */
int orc_num_bodies(const orc_env *env);

/* ---- scores (orc_score.c) ---- */
double orc_score(orc_env *env);          /* score_on_end_of_traj */
double orc_debug_reward(orc_env *env);   /* debug_shaped_reward */

/* ---- raster + preprocess (orc_raster.c) ---- */
void orc_render(orc_env *env, int view, uint8_t *out_hwc384); /* view 0 allo, 1 ego */
void orc_downsample4(const uint8_t *in_hwc, int h, int w, int c, uint8_t *out_hwc); /* INTER_AREA at exactly 4x */

#ifdef __cplusplus
}
#endif
#endif
