/*
 * orc_scene.c -- ORACLE (test infrastructure): build the pymunk Space + Viewer
 * geom list of one env from an mg_scene.
 *
 * Restates, in entity add order, the setup() methods of
 *   ArenaBoundaries (magical/entities.py:507-533), GoalRegion (:783-801),
 *   Robot (:238-433) and Shape (:610-749),
 * as called from BaseEnv.reset / add_entities (magical/base_env.py:165-176, 190-246).
 * PARITY UNPINNED (see orc.h).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "orc.h"
#include "../include/mg_trig.h"

#define CP_ALL_CATEGORIES 0xffffffffu

void orc_body_set_angle(orc_body *b, double a)
{
    /* cpBodySetAngle -> SetAngle + SetTransform: rot = cpvforangle(a).  Chipmunk calls libm here; the
     * oracle and the CUDA kernels share include/mg_trig.h (<= 1 ulp from glibc) so that both sides of a
     * parity test see bit-identical rotations. */
    b->a = a;
    mg_sincos(a, &b->rot.y, &b->rot.x);
}

static int add_body(orc_env *e, int type, double m, double i, double x, double y, double a)
{
    int bi = e->n_bodies++;
    orc_body *b = &e->bodies[bi];
    memset(b, 0, sizeof(*b));
    b->type = type;
    if (type == ORC_BODY_DYNAMIC) {
        b->m = m; b->i = i; b->m_inv = 1.0 / m; b->i_inv = 1.0 / i;
    } else {
        b->m = INFINITY; b->i = INFINITY; b->m_inv = 0.0; b->i_inv = 0.0; /* cpBodySetType */
    }
    b->p.x = x; b->p.y = y;
    orc_body_set_angle(b, a);
    return bi;
}

static orc_shape *new_shape(orc_env *e, int kind, int body, int entity, double r, double u)
{
    orc_shape *s = &e->shapes[e->n_shapes++];
    memset(s, 0, sizeof(*s));
    s->kind = kind; s->body = body; s->entity = entity; s->r = r; s->u = u;
    s->group = 0; s->categories = CP_ALL_CATEGORIES; s->mask = CP_ALL_CATEGORIES;
    return s;
}

/* cpPolyShape SetVerts: plane i keeps v0 = verts[i] and the outward normal of
 * edge verts[i-1] -> verts[i] (rperp of the edge direction, normalised). */
static void poly_set_verts(orc_shape *s, int n, const vec *v)
{
    s->count = n;
    for (int i = 0; i < n; i++) {
        vec a = v[(i - 1 + n) % n], b = v[i];
        double dx = b.x - a.x, dy = b.y - a.y;
        double nx = dy, ny = -dx; /* cpvrperp */
        double len = sqrt(nx * nx + ny * ny);
        double inv = 1.0 / (len + 2.2250738585072014e-308); /* cpvnormalize: 1/(len + CPFLOAT_MIN) */
        s->lv[i] = b;
        s->ln[i].x = nx * inv;
        s->ln[i].y = ny * inv;
    }
}

static int add_poly(orc_env *e, int body, int entity, int n, const vec *v, double r, double u)
{
    orc_shape *s = new_shape(e, ORC_SHAPE_POLY, body, entity, r, u);
    poly_set_verts(s, n, v);
    int si = e->n_shapes - 1;
    orc_shape_update(e, si);
    return si;
}

static int add_circle(orc_env *e, int body, int entity, double r, double u)
{
    orc_shape *s = new_shape(e, ORC_SHAPE_CIRCLE, body, entity, r, u);
    s->count = 1;
    s->lv[0].x = 0; s->lv[0].y = 0;
    int si = e->n_shapes - 1;
    orc_shape_update(e, si);
    return si;
}

static int add_segment(orc_env *e, int body, int entity, vec a, vec b, double r, double u)
{
    orc_shape *s = new_shape(e, ORC_SHAPE_SEGMENT, body, entity, r, u);
    s->count = 2;
    s->lv[0] = a; s->lv[1] = b;
    double dx = b.x - a.x, dy = b.y - a.y;
    double len = sqrt(dx * dx + dy * dy);
    double inv = 1.0 / (len + 2.2250738585072014e-308);
    /* cpSegmentShapeInit: n = cpvrperp(cpvnormalize(b - a)) */
    s->ln[0].x = dy * inv;
    s->ln[0].y = -(dx * inv);
    int si = e->n_shapes - 1;
    orc_shape_update(e, si);
    return si;
}

/* cpShapeUpdate / cpPolyShapeCacheData / cpCircleShapeCacheData / cpSegmentShapeCacheData:
 * world-space vertices, normals and the AABB from the body transform. */
void orc_shape_update(orc_env *e, int si)
{
    orc_shape *s = &e->shapes[si];
    const orc_body *b = &e->bodies[s->body];
    vec rot = b->rot, p = b->p;
    double l = INFINITY, r = -INFINITY, bo = INFINITY, t = -INFINITY;
    for (int i = 0; i < s->count; i++) {
        vec v = s->lv[i];
        vec w = { rot.x * v.x - rot.y * v.y + p.x, rot.y * v.x + rot.x * v.y + p.y };
        s->tv[i] = w;
        if (w.x < l) l = w.x;
        if (w.x > r) r = w.x;
        if (w.y < bo) bo = w.y;
        if (w.y > t) t = w.y;
    }
    int nn = s->kind == ORC_SHAPE_POLY ? s->count : (s->kind == ORC_SHAPE_SEGMENT ? 1 : 0);
    for (int i = 0; i < nn; i++) {
        vec n = s->ln[i];
        s->tn[i].x = rot.x * n.x - rot.y * n.y;
        s->tn[i].y = rot.y * n.x + rot.x * n.y;
    }
    s->bb_l = l - s->r; s->bb_r = r + s->r; s->bb_b = bo - s->r; s->bb_t = t + s->r;
}

static orc_constraint *new_con(orc_env *e, int kind, int a, int b)
{
    orc_constraint *c = &e->cons[e->n_constraints++];
    memset(c, 0, sizeof(*c));
    c->kind = kind; c->a = a; c->b = b;
    /* cpConstraintInit defaults */
    c->max_force = INFINITY;
    c->max_bias = INFINITY;
    c->error_bias = pow(1.0 - 0.1, 60.0);
    return c;
}

static orc_prim *new_prim(orc_env *e, int n, const vec *lv, const uint8_t *fill, const uint8_t *outline, int mode)
{
    orc_prim *p = &e->prims[e->n_prims++];
    memset(p, 0, sizeof(*p));
    p->n = n; p->lv = lv;
    memcpy(p->fill, fill, 3);
    if (outline) memcpy(p->outline, outline, 3);
    p->outline_mode = mode;
    return p;
}
enum { XF_TRANSLATE = 0, XF_BODY = 1, XF_ROT_REL = 2 };
static void prim_xf(orc_prim *p, int kind, double tx, double ty, int body, int body_rel)
{
    int k = p->n_xf++;
    p->xf[k].kind = kind; p->xf[k].tx = tx; p->xf[k].ty = ty; p->xf[k].rot = 0;
    p->xf[k].body = body; p->xf[k].body_rel = body_rel;
}

static orc_entity *new_ent(orc_env *e, int kind, int idx)
{
    orc_entity *en = &e->ents[e->n_ents++];
    en->kind = kind; en->idx = idx;
    en->first_shape = e->n_shapes; en->n_shapes = 0;
    en->first_body = e->n_bodies; en->n_bodies = 0;
    return en;
}
static void close_ent(orc_env *e, orc_entity *en)
{
    en->n_shapes = e->n_shapes - en->first_shape;
    en->n_bodies = e->n_bodies - en->first_body;
}

/* ArenaBoundaries.setup (entities.py:507-533) */
static void setup_arena(orc_env *e)
{
    const orc_geom_tables *g = orc_geom();
    orc_entity *en = new_ent(e, ORC_ENT_ARENA, 0);
    int ent = e->n_ents - 1;
    double left = -1, right = 1, bottom = -1, top = 1, rad = 1;
    vec pts[4] = { { left - rad, top + rad }, { right + rad, top + rad },
                   { right + rad, bottom - rad }, { left - rad, bottom - rad } };
    for (int i = 0; i < 4; i++) add_segment(e, ORC_STATIC_BODY, ent, pts[i], pts[(i + 1) % 4], rad, 0.8);
    orc_prim *p = new_prim(e, 4, g->arena_draw, g->pal_white, g->pal_grey, 1);
    prim_xf(p, XF_TRANSLATE, left + (right - left) / 2, bottom + (top - bottom) / 2, 0, 0);
    close_ent(e, en);
}

/* GoalRegion.setup (entities.py:783-801) */
static void setup_goal(orc_env *e, int gi)
{
    const orc_geom_tables *g = orc_geom();
    const mg_goal *go = &e->scene.goals[gi];
    orc_entity *en = new_ent(e, ORC_ENT_GOAL, gi);
    int ent = e->n_ents - 1;
    e->goal_ent[gi] = ent;
    int b = add_body(e, ORC_BODY_STATIC, 0, 0, go->cx, go->cy, 0.0);
    /* Poly.create_box(body, (w, h)) -> cpBoxShapeInit2 vertex order, radius 0 */
    double hw = go->w / 2, hh = go->h / 2;
    vec bv[4] = { { hw, -hh }, { hw, hh }, { -hw, hh }, { -hw, -hh } };
    int si = add_poly(e, b, ent, 4, bv, 0.0, 0.0);
    e->shapes[si].sensor = 1;
    e->goal_shape[gi] = si;
    orc_prim *p = new_prim(e, 4, NULL, g->pal_colour_light2[go->colour], g->pal_colour[go->colour], 2);
    double rad_h = go->h / 2, rad_w = go->w / 2; /* render.make_rect(w, h) */
    p->lv_own[0].x = -rad_w; p->lv_own[0].y = rad_h;
    p->lv_own[1].x = rad_w;  p->lv_own[1].y = rad_h;
    p->lv_own[2].x = rad_w;  p->lv_own[2].y = -rad_h;
    p->lv_own[3].x = -rad_w; p->lv_own[3].y = -rad_h;
    p->lv = NULL; /* resolved to lv_own at draw time (struct may move) */
    prim_xf(p, XF_BODY, 0, 0, b, 0);
    close_ent(e, en);
}

/* Default poses of the six robot bodies for a robot created at (x, y, angle)
 * (Robot.setup, entities.py:243-254, 267-277, 313-331). */
void orc_robot_default_poses(double x, double y, double angle, double out[MG_RB_COUNT][3])
{
    const double radius = 0.2;
    out[MG_RB_ROBOT][0] = x; out[MG_RB_ROBOT][1] = y; out[MG_RB_ROBOT][2] = angle;
    out[MG_RB_CONTROL][0] = x; out[MG_RB_CONTROL][1] = y; out[MG_RB_CONTROL][2] = angle;
    for (int k = 0; k < 2; k++) {
        out[MG_RB_PUPIL_L + k][0] = 0.0; out[MG_RB_PUPIL_L + k][1] = 0.0; out[MG_RB_PUPIL_L + k][2] = angle;
    }
    for (int k = 0; k < 2; k++) {
        double side = k == 0 ? -1.0 : 1.0;
        double lower = k == 0 ? -0.0 : -(M_PI / 8);
        double upper = k == 0 ? (M_PI / 8) : 0.0;
        double delta = k == 0 ? upper : lower;
        double rx = side * radius * 0.45, ry = radius * 0.1;
        double c = cos(angle), s = sin(angle);
        out[MG_RB_FINGER_L + k][0] = x + (rx * c - ry * s);   /* gtools.rotate_vec + add_vecs */
        out[MG_RB_FINGER_L + k][1] = y + (rx * s + ry * c);
        out[MG_RB_FINGER_L + k][2] = angle + delta;
    }
}

/* Robot.setup (entities.py:238-433) */
static void setup_robot(orc_env *e)
{
    const orc_geom_tables *g = orc_geom();
    const mg_scene *sc = &e->scene;
    const double radius = 0.2, mass = 1.0;
    orc_entity *en = new_ent(e, ORC_ENT_ROBOT, 0);
    int ent = e->n_ents - 1;
    e->robot_ent = ent;
    const double (*rp)[3] = sc->robot;

    double inertia = mass * (0.5 * (0.0 * 0.0 + radius * radius) + 0.0); /* cpMomentForCircle */
    int body = add_body(e, ORC_BODY_DYNAMIC, mass, inertia, rp[MG_RB_ROBOT][0], rp[MG_RB_ROBOT][1], rp[MG_RB_ROBOT][2]);
    e->robot_body0 = body;
    int control = add_body(e, ORC_BODY_KINEMATIC, 0, 0, rp[MG_RB_CONTROL][0], rp[MG_RB_CONTROL][1], rp[MG_RB_CONTROL][2]);

    orc_constraint *c = new_con(e, ORC_C_PIVOT, control, body);
    c->max_bias = 0; c->max_force = sc->phys[0];
    c = new_con(e, ORC_C_GEAR, control, body);
    c->phase = 0.0; c->ratio = 1.0; c->error_bias = 0.0; c->max_bias = 2.5; c->max_force = sc->phys[1];

    for (int k = 0; k < 2; k++) {
        double eye_mass = mass / 10;
        double eye_inertia = eye_mass * (0.5 * (radius * radius));
        int eb = add_body(e, ORC_BODY_DYNAMIC, eye_mass, eye_inertia,
                          rp[MG_RB_PUPIL_L + k][0], rp[MG_RB_PUPIL_L + k][1], rp[MG_RB_PUPIL_L + k][2]);
        c = new_con(e, ORC_C_SPRING, body, eb);
        c->rest = 0; c->stiffness = 0.1; c->damping = 3e-3;
        c->max_bias = 3.0; c->max_force = 0.001; /* set by the reference, unused by the spring */
    }

    int finger_body[2];
    for (int k = 0; k < 2; k++) {
        double side = k == 0 ? -1.0 : 1.0;
        double lower = k == 0 ? -0.0 : -(M_PI / 8);
        double upper = k == 0 ? (M_PI / 8) : 0.0;
        double finger_mass = mass / 8;
        int fb = add_body(e, ORC_BODY_DYNAMIC, finger_mass, g->finger_moment[k],
                          rp[MG_RB_FINGER_L + k][0], rp[MG_RB_FINGER_L + k][1], rp[MG_RB_FINGER_L + k][2]);
        finger_body[k] = fb;
        c = new_con(e, ORC_C_PIVOT, body, fb);
        /* PivotJoint(body, finger_body, finger_body.position): both anchors are the
         * pivot in body-local coordinates at setup time */
        c->anchor_a.x = side * radius * 0.45; c->anchor_a.y = radius * 0.1;
        c->anchor_b.x = 0; c->anchor_b.y = 0;
        c->error_bias = 0.0;
        c = new_con(e, ORC_C_LIMIT, body, fb);
        c->lo = lower; c->hi = upper; c->error_bias = 0.0;
        c = new_con(e, ORC_C_MOTOR, body, fb);
        c->rate = 0.0; c->max_bias = 0.0; c->max_force = sc->phys[2];
        e->motor_con[k] = e->n_constraints - 1;
    }

    int si = add_circle(e, body, ent, radius, 0.5);
    e->shapes[si].group = 1;
    for (int k = 0; k < 2; k++)
        for (int part = 0; part < 2; part++) {
            si = add_poly(e, finger_body[k], ent, 4, g->finger_phys[k][part], 0.0, 5.0);
            e->shapes[si].group = 1;
        }

    /* graphics (entities.py:374-433): outer finger quads, inner finger quads, body, eyes */
    for (int k = 0; k < 2; k++)
        for (int part = 0; part < 2; part++) {
            orc_prim *p = new_prim(e, 4, g->finger_phys[k][part], g->pal_grey, NULL, 0);
            prim_xf(p, XF_BODY, 0, 0, finger_body[k], 0);
        }
    for (int k = 0; k < 2; k++)
        for (int part = 0; part < 2; part++) {
            orc_prim *p = new_prim(e, 4, g->finger_inner[k][part], g->pal_grey_light, NULL, 0);
            prim_xf(p, XF_BODY, 0, 0, finger_body[k], 0);
        }
    orc_prim *p = new_prim(e, ORC_CIRCLE_RES, g->circle_robot, g->pal_grey, g->pal_grey_dark, 1);
    prim_xf(p, XF_BODY, 0, 0, body, 0);
    for (int k = 0; k < 2; k++) {
        double x_sign = k == 0 ? -1.0 : 1.0;
        double ex = x_sign * 0.4 * radius, ey = 0.3 * radius;
        p = new_prim(e, ORC_CIRCLE_RES, g->circle_eye, g->pal_white, NULL, 0);
        prim_xf(p, XF_TRANSLATE, ex, ey, 0, 0);
        prim_xf(p, XF_BODY, 0, 0, body, 0);
        p = new_prim(e, ORC_CIRCLE_RES, g->circle_pupil, g->pal_pupil, NULL, 0);
        prim_xf(p, XF_TRANSLATE, 0, radius * 0.07, 0, 0);
        prim_xf(p, XF_ROT_REL, 0, 0, body + MG_RB_PUPIL_L + k, body);
        prim_xf(p, XF_TRANSLATE, ex, ey, 0, 0);
        prim_xf(p, XF_BODY, 0, 0, body, 0);
    }
    e->rel_turn_angle = 0.0; e->target_speed = 0.0; e->target_finger_angle = 0.0;
    close_ent(e, en);
}

/* Shape.setup (entities.py:610-749) */
static void setup_block(orc_env *e, int bi)
{
    const orc_geom_tables *g = orc_geom();
    const mg_scene *sc = &e->scene;
    const mg_block *bl = &sc->blocks[bi];
    const double mass = 0.5, shape_size = 0.2 * 0.6;
    orc_entity *en = new_ent(e, ORC_ENT_BLOCK, bi);
    int ent = e->n_ents - 1;
    e->block_ent[bi] = ent;
    int body;
    const uint8_t *col = g->pal_colour[bl->colour], *dark = g->pal_colour_dark[bl->colour];
    switch (bl->type) {
    case MG_SQUARE:
        body = add_body(e, ORC_BODY_DYNAMIC, mass, g->square_moment, bl->x, bl->y, bl->angle);
        add_poly(e, body, ent, 4, g->square_phys, g->square_r, 0.5);
        break;
    case MG_CIRCLE: {
        double inertia = mass * (0.5 * (shape_size * shape_size));
        body = add_body(e, ORC_BODY_DYNAMIC, mass, inertia, bl->x, bl->y, bl->angle);
        add_circle(e, body, ent, shape_size, 0.5);
        break;
    }
    case MG_STAR: {
        body = add_body(e, ORC_BODY_DYNAMIC, mass, g->star_moment, bl->x, bl->y, bl->angle);
        if (e->group_ctr == 0) e->group_ctr = 999; /* Entity.generate_group_id, entities.py:60-66 */
        e->group_ctr += 1;
        for (int k = 0; k < g->star_outer.n_pieces; k++) {
            vec hv[ORC_MAXV];
            int n = orc_hull_order(g->star_outer.piece_n[k], g->star_outer.piece_v[k], hv);
            int si = add_poly(e, body, ent, n, hv, 0.0, 0.5);
            e->shapes[si].group = (uint32_t)e->group_ctr;
        }
        break;
    }
    default: /* MG_PENTAGON */
        body = add_body(e, ORC_BODY_DYNAMIC, mass, g->pent_moment, bl->x, bl->y, bl->angle);
        add_poly(e, body, ent, 5, g->pent_phys, 0.0, 0.5);
        break;
    }
    e->block_body[bi] = body;

    orc_constraint *c = new_con(e, ORC_C_PIVOT, ORC_STATIC_BODY, body);
    c->max_bias = 0; c->max_force = sc->phys[3];
    c = new_con(e, ORC_C_GEAR, ORC_STATIC_BODY, body);
    c->phase = 0.0; c->ratio = 1.0; c->max_bias = 0; c->max_force = sc->phys[4];

    orc_prim *p;
    switch (bl->type) {
    case MG_SQUARE:
        p = new_prim(e, 4, g->square_draw, col, dark, 1); prim_xf(p, XF_BODY, 0, 0, body, 0); break;
    case MG_CIRCLE:
        p = new_prim(e, ORC_CIRCLE_RES, g->circle_block, col, dark, 1); prim_xf(p, XF_BODY, 0, 0, body, 0); break;
    case MG_STAR:
        for (int k = 0; k < g->star_outer.n_pieces; k++) {
            p = new_prim(e, g->star_outer.piece_n[k], g->star_outer.piece_v[k], dark, NULL, 0);
            prim_xf(p, XF_BODY, 0, 0, body, 0);
        }
        for (int k = 0; k < g->star_inner.n_pieces; k++) {
            p = new_prim(e, g->star_inner.piece_n[k], g->star_inner.piece_v[k], col, NULL, 0);
            prim_xf(p, XF_BODY, 0, 0, body, 0);
        }
        break;
    default:
        p = new_prim(e, 5, g->pent, col, dark, 1); prim_xf(p, XF_BODY, 0, 0, body, 0); break;
    }
    close_ent(e, en);
}

orc_env *orc_create(const mg_scene *scene)
{
    orc_env *e = (orc_env *)calloc(1, sizeof(orc_env));
    if (!e) return NULL;
    e->scene = *scene;
    /* BaseEnv.reset: pm.Space(); collision_slop = 0.01; iterations = phys_iter (base_env.py:206-208);
     * remaining cpSpace defaults from cpSpaceInit. */
    e->iterations = 10;
    e->collision_slop = 0.01;
    e->collision_bias = pow(1.0 - 0.1, 60.0);
    e->collision_persistence = 3;
    e->stamp = 0;
    e->curr_dt = 0.0;
    /* space.static_body / arena body: infinite mass at the origin */
    orc_body *sb = &e->bodies[ORC_STATIC_BODY];
    memset(sb, 0, sizeof(*sb));
    sb->type = ORC_BODY_STATIC; sb->m = INFINITY; sb->i = INFINITY;
    orc_body_set_angle(sb, 0.0);

    setup_arena(e);
    for (int i = 0; i < scene->n_goals; i++) setup_goal(e, i);
    if (scene->robot_first) setup_robot(e);
    for (int i = 0; i < scene->n_blocks; i++) setup_block(e, i);
    if (!scene->robot_first) setup_robot(e);
    return e;
}

void orc_destroy(orc_env *e) { free(e); }

int orc_num_bodies(const orc_env *e) { return MG_RB_COUNT + e->scene.n_blocks; }

/* canonical order: six robot bodies, then blocks in scene order */
void orc_get_bodies(const orc_env *e, double *out)
{
    int k = 0;
    for (int i = 0; i < MG_RB_COUNT + e->scene.n_blocks; i++) {
        const orc_body *b = i < MG_RB_COUNT ? &e->bodies[e->robot_body0 + i] : &e->bodies[e->block_body[i - MG_RB_COUNT]];
        out[k++] = b->p.x; out[k++] = b->p.y; out[k++] = b->a;
        out[k++] = b->v.x; out[k++] = b->v.y; out[k++] = b->w;
    }
}
