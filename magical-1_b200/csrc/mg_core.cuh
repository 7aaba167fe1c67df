/*
 * mg_core.cuh -- shared definitions of the B200 MAGICAL simulator: 2-D vector math,
 * the constant geometry tables, and the narrow-phase collision routines.
 *
 * Everything here is __host__ __device__: the device physics kernel (mg_physics.cu) and
 * the host-side placement predicate used by reset sampling (mg_host.cu) run the very
 * same narrow phase, so a pose accepted on the host never starts inside another shape
 * on the device.
 *
 * Arithmetic is IEEE double with FMA contraction disabled (-fmad=false) and follows the
 * operation order of Chipmunk2D 7.0.3's cpCollision.c, which is what the reference
 * reaches through pymunk (magical/base_env.py:255 space.step, magical/geom.py:228
 * space.shape_query).
 */
#pragma once
#include <stdint.h>
#include <math.h>
#include <float.h>
#include "../../include/mg_scene.h"

#ifdef __CUDACC__
#define MG_NOUNROLL _Pragma("unroll 1")
#define MG_HD __host__ __device__ __forceinline__
#define MG_HDN __host__ __device__ __noinline__
#else
#define MG_NOUNROLL
#define MG_HD inline
#define MG_HDN
#endif

#define MG_DBL_MIN 2.2250738585072014e-308

struct vec2 { double x, y; };

MG_HD vec2 V(double x, double y) { vec2 r; r.x = x; r.y = y; return r; }
MG_HD vec2 vadd(vec2 a, vec2 b) { return V(a.x + b.x, a.y + b.y); }
MG_HD vec2 vsub(vec2 a, vec2 b) { return V(a.x - b.x, a.y - b.y); }
MG_HD vec2 vmul(vec2 a, double s) { return V(a.x * s, a.y * s); }
MG_HD vec2 vneg(vec2 a) { return V(-a.x, -a.y); }
MG_HD double vdot(vec2 a, vec2 b) { return a.x * b.x + a.y * b.y; }
MG_HD double vcross(vec2 a, vec2 b) { return a.x * b.y - a.y * b.x; }
MG_HD vec2 vperp(vec2 a) { return V(-a.y, a.x); }
MG_HD vec2 vrot(vec2 a, vec2 b) { return V(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
MG_HD vec2 vlerp(vec2 a, vec2 b, double t) { return vadd(vmul(a, 1.0 - t), vmul(b, t)); }
MG_HD double clamp01(double f) { return f < 0.0 ? 0.0 : (f > 1.0 ? 1.0 : f); }
/* chipmunk.h: cpfmax(a, b) = (a > b) ? a : b, cpfmin(a, b) = (a < b) ? a : b, cpfclamp = cpfmin(cpfmax(f, lo), hi) */
MG_HD double cp_max(double a, double b) { return a > b ? a : b; }
MG_HD double cp_min(double a, double b) { return a < b ? a : b; }
MG_HD double dclamp(double f, double lo, double hi) { return cp_min(cp_max(f, lo), hi); }

/* ------------------------------------------------------------------ geometry tables */

#define MG_MAXV 5           /* largest convex collision polygon (pentagon) */
#define MG_CIRCLE_RES 100   /* render.make_circle(radius, 100, ...), entities.py:379,414,420,715 */
#define MG_STAR_PIECES 6

/* collision templates */
enum {
    MG_TPL_FINGER0 = 0,     /* +side*2 + part: L upper, L fore, R upper, R fore (entities.py:362-372) */
    MG_TPL_SQUARE = 4,      /* Poly.create_box((L, L), 0.01 L), entities.py:616-631 */
    MG_TPL_PENT = 5,        /* entities.py:672-692 */
    MG_TPL_STAR0 = 6,       /* star convex pieces, entities.py:643-665 */
    MG_TPL_COUNT = MG_TPL_STAR0 + MG_STAR_PIECES,
    MG_TPL_CIRCLE_ROBOT = 32,
    MG_TPL_CIRCLE_BLOCK = 33,
    MG_TPL_WALL0 = 40       /* +0..3, ArenaBoundaries.setup, entities.py:507-522 */
};

/* draw vertex lists (render.Poly), all stored as (x, y) pairs in one pool */
enum {
    MG_DL_ARENA = 0,
    MG_DL_FINGER_OUT0 = 1,  /* +side*2+part */
    MG_DL_FINGER_IN0 = 5,
    MG_DL_ROBOT = 9,
    MG_DL_EYE = 10,
    MG_DL_PUPIL = 11,
    MG_DL_SQUARE = 12,
    MG_DL_PENT = 13,
    MG_DL_CIRCLE = 14,
    MG_DL_STAR_OUT0 = 15,   /* +piece */
    MG_DL_STAR_IN0 = 21,
    MG_DL_GOAL = 27,        /* vertices come from the goal's (w, h) */
    MG_DL_COUNT = 28
};

/* palette indices stored in the raster buffer */
enum {
    MG_PAL_BG = 0, MG_PAL_WHITE, MG_PAL_GREY, MG_PAL_GREY_DARK, MG_PAL_GREY_LIGHT, MG_PAL_PUPIL,
    MG_PAL_COLOUR0 = 6,     /* + colour */
    MG_PAL_DARK0 = 10,
    MG_PAL_LIGHT0 = 14,
    MG_PAL_COUNT = 18
};

struct MgTables {
    /* collision polygons: local vertices in Chipmunk hull order; nrm[i] is the outward
     * normal of edge v[i-1] -> v[i] */
    int tpl_n[MG_TPL_COUNT];
    double tpl_r[MG_TPL_COUNT];
    vec2 tpl_v[MG_TPL_COUNT][MG_MAXV];
    vec2 tpl_nrm[MG_TPL_COUNT][MG_MAXV];
    double tpl_bound[MG_TPL_COUNT];   /* max |v| + r: bounding-circle radius about the body origin */
    /* walls: endpoints a, b and normal; radius 1 */
    vec2 wall_a[4], wall_b[4], wall_n[4];
    /* mass properties */
    double finger_moment[2], square_moment, pent_moment, star_moment;
    /* precomputed transcendental constants so host (glibc) and device agree bit for bit */
    double err_bias_default;     /* pow(1 - 0.1, 60): cpConstraint errorBias / cpSpace collisionBias default */
    double bias_coef_default;    /* 1 - pow(err_bias_default, dt) */
    double spring_w_coef;        /* 1 - exp(-damping * dt * (robot.i_inv + pupil.i_inv)) */
    /* draw lists */
    int dl_off[MG_DL_COUNT], dl_n[MG_DL_COUNT];
    double dl_rad[MG_DL_COUNT];  /* max |vertex| of the list: pre-transform band cull in the rasteriser */
    vec2 dl_pool[4 * MG_CIRCLE_RES + 160];
    uint8_t pal[32][4];
};

/* ------------------------------------------------------------------ narrow phase */

enum { MG_SHAPE_CIRCLE = 0, MG_SHAPE_SEGMENT = 1, MG_SHAPE_POLY = 2 }; /* cpShapeType order */

/* a collision shape in world space */
struct WShape {
    int kind, n;
    double r;
    vec2 v[MG_MAXV];    /* circle: v[0] = centre; segment: v[0], v[1] */
    vec2 nrm[MG_MAXV];  /* poly: face normals; segment: nrm[0] = normal of a -> b */
};

struct Manifold {
    int count;
    vec2 n;
    vec2 p1[2], p2[2];
    uint32_t hash[2];
};

MG_HD void wshape_from_tpl(const MgTables *T, int tpl, vec2 p, vec2 rot, WShape *s)
{
    if (tpl >= MG_TPL_WALL0) {
        int w = tpl - MG_TPL_WALL0;
        s->kind = MG_SHAPE_SEGMENT; s->n = 2; s->r = 1.0;
        s->v[0] = T->wall_a[w]; s->v[1] = T->wall_b[w]; s->nrm[0] = T->wall_n[w];
    } else if (tpl >= MG_TPL_CIRCLE_ROBOT) {
        s->kind = MG_SHAPE_CIRCLE; s->n = 1;
        s->r = tpl == MG_TPL_CIRCLE_ROBOT ? 0.2 : 0.2 * 0.6;
        /* cpCircleShapeCacheData: tc = cpTransformPoint(transform, c) with c = 0 */
        s->v[0] = V(rot.x * 0.0 - rot.y * 0.0 + p.x, rot.y * 0.0 + rot.x * 0.0 + p.y);
    } else {
        int n = T->tpl_n[tpl];
        s->kind = MG_SHAPE_POLY; s->n = n; s->r = T->tpl_r[tpl];
        MG_NOUNROLL
        for (int i = 0; i < n; i++) {
            vec2 l = T->tpl_v[tpl][i], m = T->tpl_nrm[tpl][i];
            s->v[i] = V(rot.x * l.x - rot.y * l.y + p.x, rot.y * l.x + rot.x * l.y + p.y);
            s->nrm[i] = V(rot.x * m.x - rot.y * m.y, rot.y * m.x + rot.x * m.y);
        }
    }
}

MG_HD void mf_push(Manifold *m, vec2 p1, vec2 p2, uint32_t hash)
{
    m->p1[m->count] = p1; m->p2[m->count] = p2; m->hash[m->count] = hash; m->count++;
}

/* cpCollision.c CircleToCircle */
MG_HD void collide_circle_circle(const WShape *c1, const WShape *c2, Manifold *m)
{
    double mindist = c1->r + c2->r;
    vec2 delta = vsub(c2->v[0], c1->v[0]);
    double distsq = vdot(delta, delta);
    if (distsq < mindist * mindist) {
        double dist = sqrt(distsq);
        vec2 n = dist ? vmul(delta, 1.0 / dist) : V(1.0, 0.0);
        m->n = n;
        mf_push(m, vadd(c1->v[0], vmul(n, c1->r)), vadd(c2->v[0], vmul(n, -c2->r)), 0);
    }
}

/* cpCollision.c CircleToSegment; the arena segments have no neighbours, so no end-cap rejection */
MG_HD void collide_circle_segment(const WShape *c, const WShape *s, Manifold *m)
{
    vec2 seg_a = s->v[0], seg_b = s->v[1], center = c->v[0];
    vec2 seg_delta = vsub(seg_b, seg_a);
    double closest_t = clamp01(vdot(seg_delta, vsub(center, seg_a)) / vdot(seg_delta, seg_delta));
    vec2 closest = vadd(seg_a, vmul(seg_delta, closest_t));
    double mindist = c->r + s->r;
    vec2 delta = vsub(closest, center);
    double distsq = vdot(delta, delta);
    if (distsq < mindist * mindist) {
        double dist = sqrt(distsq);
        vec2 n = dist ? vmul(delta, 1.0 / dist) : s->nrm[0];
        m->n = n;
        mf_push(m, vadd(center, vmul(n, c->r)), vadd(closest, vmul(n, -s->r)), 0);
    }
}

/* face normal i of a convex shape; a segment is the 2-gon a -> b with faces -n (i=0), +n (i=1) */
MG_HD vec2 face_normal(const WShape *s, int i)
{
    if (s->kind == MG_SHAPE_SEGMENT) return i == 1 ? s->nrm[0] : vneg(s->nrm[0]);
    return s->nrm[i];
}

MG_HD vec2 seg_closest(vec2 a, vec2 b, vec2 p, double *t_out)
{
    vec2 d = vsub(b, a);
    double dd = vdot(d, d);
    double t = dd > 0.0 ? clamp01(vdot(vsub(p, a), d) / dd) : 0.0;
    *t_out = t;
    return vadd(a, vmul(d, t));
}

/* Minimum separating axis of two convex cores: what Chipmunk's GJK (separated: closest
 * features) / EPA (overlapping: shallowest face) return for convex polygons.  n points
 * from A to B, d is the signed core separation. */
MG_HDN inline void convex_axis(const WShape *A, const WShape *B, vec2 *n_out, double *d_out)
{
    const double rsum = A->r + B->r;
    double best = -INFINITY;
    vec2 axis = V(1.0, 0.0);
    MG_NOUNROLL
    for (int i = 0; i < A->n; i++) {
        vec2 fn = face_normal(A, i);
        double sep = INFINITY;
        MG_NOUNROLL
        for (int j = 0; j < B->n; j++) {
            double s = vdot(vsub(B->v[j], A->v[i]), fn);
            if (s < sep) sep = s;
        }
        if (sep > best) { best = sep; axis = fn; }
    }
    MG_NOUNROLL
    for (int j = 0; j < B->n; j++) {
        vec2 fn = face_normal(B, j);
        double sep = INFINITY;
        MG_NOUNROLL
        for (int i = 0; i < A->n; i++) {
            double s = vdot(vsub(A->v[i], B->v[j]), fn);
            if (s < sep) sep = s;
        }
        if (sep > best) { best = sep; axis = vneg(fn); }
    }
    if (best <= 0.0) { *n_out = axis; *d_out = best; return; }
    /* A face separation is a lower bound of the true distance: beyond r_a + r_b the pair cannot touch and
     * the caller discards (n, d), so the closest-feature search is skipped (same contacts as without it). */
    if (best > rsum) { *n_out = axis; *d_out = best; return; }

    double bestd2 = INFINITY;
    vec2 bn = axis;
    double bd = best;
    MG_NOUNROLL
    for (int i = 0; i < A->n; i++) {
        vec2 ea = A->v[(i - 1 + A->n) % A->n], eb = A->v[i];
        MG_NOUNROLL
        for (int j = 0; j < B->n; j++) {
            double t;
            vec2 q = seg_closest(ea, eb, B->v[j], &t);
            vec2 dl = vsub(B->v[j], q);
            double d2 = vdot(dl, dl);
            if (d2 < bestd2) {
                bestd2 = d2;
                if (t > 0.0 && t < 1.0) {
                    bn = face_normal(A, i); bd = vdot(dl, bn);
                    if (bd < 0.0) { bn = vneg(bn); bd = -bd; }
                } else {
                    double len = sqrt(d2); bd = len; bn = vmul(dl, 1.0 / (len + MG_DBL_MIN));
                }
            }
        }
    }
    MG_NOUNROLL
    for (int j = 0; j < B->n; j++) {
        vec2 ea = B->v[(j - 1 + B->n) % B->n], eb = B->v[j];
        MG_NOUNROLL
        for (int i = 0; i < A->n; i++) {
            double t;
            vec2 q = seg_closest(ea, eb, A->v[i], &t);
            vec2 dl = vsub(q, A->v[i]);
            double d2 = vdot(dl, dl);
            if (d2 < bestd2) {
                bestd2 = d2;
                if (t > 0.0 && t < 1.0) {
                    bn = vneg(face_normal(B, j)); bd = vdot(dl, bn);
                    if (bd < 0.0) { bn = vneg(bn); bd = -bd; }
                } else {
                    double len = sqrt(d2); bd = len; bn = vmul(dl, 1.0 / (len + MG_DBL_MIN));
                }
            }
        }
    }
    *n_out = bn; *d_out = bd;
}

struct SupEdge { vec2 a, b, n; uint32_t ha, hb; double r; };

/* cpCollision.c SupportEdgeForPoly / SupportEdgeForSegment */
MG_HD SupEdge support_edge(const WShape *s, vec2 n)
{
    SupEdge e;
    e.r = s->r;
    if (s->kind == MG_SHAPE_SEGMENT) {
        if (vdot(s->nrm[0], n) > 0.0) { e.a = s->v[0]; e.ha = 0; e.b = s->v[1]; e.hb = 1; e.n = s->nrm[0]; }
        else { e.a = s->v[1]; e.ha = 1; e.b = s->v[0]; e.hb = 0; e.n = vneg(s->nrm[0]); }
        return e;
    }
    int count = s->n, i1 = 0;
    double mx = -INFINITY;
    MG_NOUNROLL
    for (int i = 0; i < count; i++) {
        double d = vdot(s->v[i], n);
        if (d > mx) { mx = d; i1 = i; }
    }
    int i0 = (i1 - 1 + count) % count, i2 = (i1 + 1) % count;
    if (vdot(n, s->nrm[i1]) > vdot(n, s->nrm[i2])) {
        e.a = s->v[i0]; e.ha = (uint32_t)i0; e.b = s->v[i1]; e.hb = (uint32_t)i1; e.n = s->nrm[i1];
    } else {
        e.a = s->v[i1]; e.ha = (uint32_t)i1; e.b = s->v[i2]; e.hb = (uint32_t)i2; e.n = s->nrm[i2];
    }
    return e;
}

/* stable id of a (vertex of a, vertex of b) feature pair; plays the role of CP_HASH_PAIR */
MG_HD uint32_t feature_hash(uint32_t ha, uint32_t hb) { return 1u + ha * 16u + hb; }

/* cpCollision.c ContactPoints */
MG_HD void contact_points(SupEdge e1, SupEdge e2, vec2 n, double d, Manifold *m)
{
    double mindist = e1.r + e2.r;
    if (d <= mindist) {
        m->n = n;
        double d_e1_a = vcross(e1.a, n), d_e1_b = vcross(e1.b, n);
        double d_e2_a = vcross(e2.a, n), d_e2_b = vcross(e2.b, n);
        double e1_denom = 1.0 / (d_e1_b - d_e1_a + MG_DBL_MIN);
        double e2_denom = 1.0 / (d_e2_b - d_e2_a + MG_DBL_MIN);
        {
            vec2 p1 = vadd(vmul(n, e1.r), vlerp(e1.a, e1.b, clamp01((d_e2_b - d_e1_a) * e1_denom)));
            vec2 p2 = vadd(vmul(n, -e2.r), vlerp(e2.a, e2.b, clamp01((d_e1_a - d_e2_a) * e2_denom)));
            if (vdot(vsub(p2, p1), n) <= 0.0) mf_push(m, p1, p2, feature_hash(e1.ha, e2.hb));
        }
        {
            vec2 p1 = vadd(vmul(n, e1.r), vlerp(e1.a, e1.b, clamp01((d_e2_a - d_e1_a) * e1_denom)));
            vec2 p2 = vadd(vmul(n, -e2.r), vlerp(e2.a, e2.b, clamp01((d_e1_b - d_e2_a) * e2_denom)));
            if (vdot(vsub(p2, p1), n) <= 0.0) mf_push(m, p1, p2, feature_hash(e1.hb, e2.ha));
        }
    }
}

/* cpCollision.c PolyToPoly / SegmentToPoly */
MG_HDN inline void collide_convex_convex(const WShape *a, const WShape *b, Manifold *m)
{
    vec2 n; double d;
    convex_axis(a, b, &n, &d);
    if (d - a->r - b->r <= 0.0) contact_points(support_edge(a, n), support_edge(b, vneg(n)), n, d, m);
}

/* cpCollision.c CircleToPoly: closest point of the polygon core to the circle centre */
MG_HDN inline void collide_circle_poly(const WShape *c, const WShape *p, Manifold *m)
{
    vec2 center = c->v[0];
    int n = p->n;
    double best = -INFINITY; int bi = 0;
    MG_NOUNROLL
    for (int i = 0; i < n; i++) {
        double s = vdot(vsub(center, p->v[i]), p->nrm[i]);
        if (s > best) { best = s; bi = i; }
    }
    vec2 nrm, pb; double d;
    if (best <= 0.0) {
        nrm = vneg(p->nrm[bi]); d = best; pb = vsub(center, vmul(p->nrm[bi], best));
    } else if (best > c->r + p->r) {
        return;     /* face separation already exceeds r_c + r_p: no contact (see convex_axis) */
    } else {
        double bestd2 = INFINITY;
        nrm = vneg(p->nrm[bi]); d = best; pb = center;
        MG_NOUNROLL
        for (int i = 0; i < n; i++) {
            vec2 ea = p->v[(i - 1 + n) % n], eb = p->v[i];
            double t;
            vec2 q = seg_closest(ea, eb, center, &t);
            vec2 dl = vsub(q, center);
            double d2 = vdot(dl, dl);
            if (d2 < bestd2) {
                bestd2 = d2; pb = q;
                if (t > 0.0 && t < 1.0) { nrm = vneg(p->nrm[i]); d = vdot(dl, nrm); }
                else { double len = sqrt(d2); d = len; nrm = vmul(dl, 1.0 / (len + MG_DBL_MIN)); }
            }
        }
    }
    if (d <= c->r + p->r) {
        m->n = nrm;
        mf_push(m, vadd(center, vmul(nrm, c->r)), vadd(pb, vmul(nrm, -p->r)), 0);
    }
}

/* cpCollide: a->kind <= b->kind required (caller orders the pair). */
MG_HDN inline int collide_shapes(const WShape *a, const WShape *b, Manifold *m)
{
    m->count = 0; m->n = V(0.0, 0.0);
    if (a->kind == MG_SHAPE_CIRCLE) {
        if (b->kind == MG_SHAPE_CIRCLE) collide_circle_circle(a, b, m);
        else if (b->kind == MG_SHAPE_SEGMENT) collide_circle_segment(a, b, m);
        else collide_circle_poly(a, b, m);
    } else if (a->kind == MG_SHAPE_SEGMENT && b->kind == MG_SHAPE_SEGMENT) {
        /* all segments live on the static arena body: never collided */
    } else {
        collide_convex_convex(a, b, m);
    }
    return m->count;
}

MG_HD int tpl_kind(int tpl)
{
    return tpl >= MG_TPL_WALL0 ? MG_SHAPE_SEGMENT : (tpl >= MG_TPL_CIRCLE_ROBOT ? MG_SHAPE_CIRCLE : MG_SHAPE_POLY);
}
