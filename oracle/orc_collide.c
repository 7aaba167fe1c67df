/*
 * orc_collide.c -- ORACLE (test infrastructure): narrow-phase collision.
 *
 * Stands in for Chipmunk2D 7.0.3 cpCollision.c as reached through
 *   space.step  (magical/base_env.py:255)  and
 *   space.shape_query (magical/geom.py:228, magical/entities.py:819).
 * PARITY UNPINNED (Chipmunk is not on this box; see orc.h).
 *
 * Chipmunk finds the minimum separating axis (n, d) of two convex shapes with
 * GJK (separated) / EPA (overlapping) and then clips the two support edges
 * (ContactPoints).  For convex polygons that pair is unique except at exact
 * ties, so it is restated directly:
 *   overlapping: the face normal (of either shape) with the largest -- i.e.
 *                least negative -- separation  [= EPA's closest Minkowski edge]
 *   separated:   the closest feature pair, edge normal if the closest point is
 *                interior to an edge, else the vertex-vertex direction
 *                [= ClosestPointsNew's two branches]
 * ContactPoints / SupportEdgeForPoly / SupportEdgeForSegment and the circle
 * routines follow cpCollision.c line for line.
 */
#include <math.h>
#include <float.h>
#include "orc.h"

static inline vec vsub(vec a, vec b) { vec r = { a.x - b.x, a.y - b.y }; return r; }
static inline vec vadd(vec a, vec b) { vec r = { a.x + b.x, a.y + b.y }; return r; }
static inline vec vmul(vec a, double s) { vec r = { a.x * s, a.y * s }; return r; }
static inline double vdot(vec a, vec b) { return a.x * b.x + a.y * b.y; }
static inline double vcross(vec a, vec b) { return a.x * b.y - a.y * b.x; }
static inline vec vneg(vec a) { vec r = { -a.x, -a.y }; return r; }
static inline vec vlerp(vec a, vec b, double t) { return vadd(vmul(a, 1.0 - t), vmul(b, t)); }
static inline double clamp01(double f) { return f < 0.0 ? 0.0 : (f > 1.0 ? 1.0 : f); }

/* cpShapeFilterReject */
int orc_filter_reject(const orc_shape *a, const orc_shape *b)
{
    return (a->group != 0 && a->group == b->group) ||
           (a->categories & b->mask) == 0 || (b->categories & a->mask) == 0;
}

static void push(orc_manifold *m, vec p1, vec p2, uint32_t hash)
{
    m->p1[m->count] = p1; m->p2[m->count] = p2; m->hash[m->count] = hash; m->count++;
}

/* CircleToCircle */
static void circle_circle(const orc_shape *c1, const orc_shape *c2, orc_manifold *m)
{
    double mindist = c1->r + c2->r;
    vec delta = vsub(c2->tv[0], c1->tv[0]);
    double distsq = vdot(delta, delta);
    if (distsq < mindist * mindist) {
        double dist = sqrt(distsq);
        vec n = dist ? vmul(delta, 1.0 / dist) : (vec){ 1.0, 0.0 };
        m->n = n;
        push(m, vadd(c1->tv[0], vmul(n, c1->r)), vadd(c2->tv[0], vmul(n, -c2->r)), 0);
    }
}

/* CircleToSegment (segment neighbour tangents are zero: no end-cap rejection) */
static void circle_segment(const orc_shape *c, const orc_shape *s, orc_manifold *m)
{
    vec seg_a = s->tv[0], seg_b = s->tv[1], center = c->tv[0];
    vec seg_delta = vsub(seg_b, seg_a);
    double closest_t = clamp01(vdot(seg_delta, vsub(center, seg_a)) / vdot(seg_delta, seg_delta));
    vec closest = vadd(seg_a, vmul(seg_delta, closest_t));
    double mindist = c->r + s->r;
    vec delta = vsub(closest, center);
    double distsq = vdot(delta, delta);
    if (distsq < mindist * mindist) {
        double dist = sqrt(distsq);
        vec n = dist ? vmul(delta, 1.0 / dist) : s->tn[0];
        m->n = n;
        push(m, vadd(center, vmul(n, c->r)), vadd(closest, vmul(n, -s->r)), 0);
    }
}

/* A convex "polygon" view of a poly or segment shape.  A segment is the
 * degenerate 2-gon a -> b with the two face normals +tn, -tn. */
typedef struct { int n; const vec *v; vec nrm[ORC_MAXV]; double r; int is_seg; } cvx;

static void cvx_from(const orc_shape *s, cvx *c)
{
    c->n = s->count; c->v = s->tv; c->r = s->r;
    c->is_seg = s->kind == ORC_SHAPE_SEGMENT;
    if (c->is_seg) {
        /* face i carries the normal of edge v[i-1] -> v[i]:
         * edge b -> a (i = 0) is the far side (-tn... see below), edge a -> b (i = 1) has rperp = tn */
        c->nrm[1] = s->tn[0];
        c->nrm[0] = vneg(s->tn[0]);
    } else {
        for (int i = 0; i < s->count; i++) c->nrm[i] = s->tn[i];
    }
}

/* closest point on segment [a,b] to p, with the clamped parameter */
static vec closest_on_seg(vec a, vec b, vec p, double *t_out)
{
    vec d = vsub(b, a);
    double dd = vdot(d, d);
    double t = dd > 0.0 ? clamp01(vdot(vsub(p, a), d) / dd) : 0.0;
    *t_out = t;
    return vadd(a, vmul(d, t));
}

/* Minimum separating axis between two convex polygons: n points from A to B,
 * d is the signed separation of the polygon cores (radii not included). */
static void convex_msa(const cvx *A, const cvx *B, vec *n_out, double *d_out)
{
    double best = -INFINITY;
    vec axis = { 1.0, 0.0 };
    for (int i = 0; i < A->n; i++) {
        double sep = INFINITY;
        for (int j = 0; j < B->n; j++) {
            double s = vdot(vsub(B->v[j], A->v[i]), A->nrm[i]);
            if (s < sep) sep = s;
        }
        if (sep > best) { best = sep; axis = A->nrm[i]; }
    }
    for (int j = 0; j < B->n; j++) {
        double sep = INFINITY;
        for (int i = 0; i < A->n; i++) {
            double s = vdot(vsub(A->v[i], B->v[j]), B->nrm[j]);
            if (s < sep) sep = s;
        }
        if (sep > best) { best = sep; axis = vneg(B->nrm[j]); }
    }
    if (best <= 0.0) { *n_out = axis; *d_out = best; return; }

    /* separated: closest features (GJK result).  Scan vertex/edge pairs both ways. */
    double bestd2 = INFINITY;
    vec bn = axis;
    double bd = best;
    for (int i = 0; i < A->n; i++) {       /* edges of A (v[i-1] -> v[i]) vs vertices of B */
        vec ea = A->v[(i - 1 + A->n) % A->n], eb = A->v[i];
        for (int j = 0; j < B->n; j++) {
            double t;
            vec q = closest_on_seg(ea, eb, B->v[j], &t);
            vec dl = vsub(B->v[j], q);
            double d2 = vdot(dl, dl);
            if (d2 < bestd2) {
                bestd2 = d2;
                if (t > 0.0 && t < 1.0) { bn = A->nrm[i]; bd = vdot(dl, bn); if (bd < 0.0) { bn = vneg(bn); bd = -bd; } /* 2-gon: either side */ }
                else { double len = sqrt(d2); bd = len; bn = vmul(dl, 1.0 / (len + DBL_MIN)); }
            }
        }
    }
    for (int j = 0; j < B->n; j++) {       /* edges of B vs vertices of A */
        vec ea = B->v[(j - 1 + B->n) % B->n], eb = B->v[j];
        for (int i = 0; i < A->n; i++) {
            double t;
            vec q = closest_on_seg(ea, eb, A->v[i], &t);
            vec dl = vsub(q, A->v[i]);     /* from A to B */
            double d2 = vdot(dl, dl);
            if (d2 < bestd2) {
                bestd2 = d2;
                if (t > 0.0 && t < 1.0) { bn = vneg(B->nrm[j]); bd = vdot(dl, bn); if (bd < 0.0) { bn = vneg(bn); bd = -bd; } }
                else { double len = sqrt(d2); bd = len; bn = vmul(dl, 1.0 / (len + DBL_MIN)); }
            }
        }
    }
    *n_out = bn; *d_out = bd;
}

typedef struct { vec ap, bp; uint32_t ah, bh; double r; vec n; } edge_t;

/* PolySupportPointIndex: first vertex with the largest projection on n */
static int support_index(const cvx *c, vec n)
{
    double max = -INFINITY;
    int index = 0;
    for (int i = 0; i < c->n; i++) {
        double d = vdot(c->v[i], n);
        if (d > max) { max = d; index = i; }
    }
    return index;
}

/* SupportEdgeForPoly / SupportEdgeForSegment */
static edge_t support_edge(const cvx *c, const orc_shape *s, vec n)
{
    edge_t e;
    if (c->is_seg) {
        if (vdot(s->tn[0], n) > 0.0) {
            e.ap = s->tv[0]; e.ah = 0; e.bp = s->tv[1]; e.bh = 1; e.n = s->tn[0];
        } else {
            e.ap = s->tv[1]; e.ah = 1; e.bp = s->tv[0]; e.bh = 0; e.n = vneg(s->tn[0]);
        }
        e.r = s->r;
        return e;
    }
    int count = c->n;
    int i1 = support_index(c, n);
    int i0 = (i1 - 1 + count) % count;
    int i2 = (i1 + 1) % count;
    if (vdot(n, c->nrm[i1]) > vdot(n, c->nrm[i2])) {
        e.ap = c->v[i0]; e.ah = (uint32_t)i0; e.bp = c->v[i1]; e.bh = (uint32_t)i1; e.n = c->nrm[i1];
    } else {
        e.ap = c->v[i1]; e.ah = (uint32_t)i1; e.bp = c->v[i2]; e.bh = (uint32_t)i2; e.n = c->nrm[i2];
    }
    e.r = c->r;
    return e;
}

/* stable per-feature-pair id standing in for CP_HASH_PAIR(e1.x.hash, e2.y.hash) */
static uint32_t feature_hash(uint32_t ha, uint32_t hb) { return 1u + ha * 16u + hb; }

/* ContactPoints */
static void contact_points(edge_t e1, edge_t e2, vec n, double d, orc_manifold *m)
{
    double mindist = e1.r + e2.r;
    if (d <= mindist) {
        m->n = n;
        double d_e1_a = vcross(e1.ap, n);
        double d_e1_b = vcross(e1.bp, n);
        double d_e2_a = vcross(e2.ap, n);
        double d_e2_b = vcross(e2.bp, n);
        double e1_denom = 1.0 / (d_e1_b - d_e1_a + DBL_MIN);
        double e2_denom = 1.0 / (d_e2_b - d_e2_a + DBL_MIN);
        {
            vec p1 = vadd(vmul(n, e1.r), vlerp(e1.ap, e1.bp, clamp01((d_e2_b - d_e1_a) * e1_denom)));
            vec p2 = vadd(vmul(n, -e2.r), vlerp(e2.ap, e2.bp, clamp01((d_e1_a - d_e2_a) * e2_denom)));
            double dist = vdot(vsub(p2, p1), n);
            if (dist <= 0.0) push(m, p1, p2, feature_hash(e1.ah, e2.bh));
        }
        {
            vec p1 = vadd(vmul(n, e1.r), vlerp(e1.ap, e1.bp, clamp01((d_e2_a - d_e1_a) * e1_denom)));
            vec p2 = vadd(vmul(n, -e2.r), vlerp(e2.ap, e2.bp, clamp01((d_e1_b - d_e2_a) * e2_denom)));
            double dist = vdot(vsub(p2, p1), n);
            if (dist <= 0.0) push(m, p1, p2, feature_hash(e1.bh, e2.ah));
        }
    }
}

/* PolyToPoly / SegmentToPoly */
static void convex_convex(const orc_shape *a, const orc_shape *b, orc_manifold *m)
{
    cvx A, B;
    cvx_from(a, &A);
    cvx_from(b, &B);
    vec n; double d;
    convex_msa(&A, &B, &n, &d);
    if (d - a->r - b->r <= 0.0) {
        contact_points(support_edge(&A, a, n), support_edge(&B, b, vneg(n)), n, d, m);
    }
}

/* CircleToPoly: GJK between a point (the circle centre) and the polygon */
static void circle_poly(const orc_shape *c, const orc_shape *p, orc_manifold *m)
{
    vec center = c->tv[0];
    int n = p->count;
    double best = -INFINITY; int bi = 0;
    for (int i = 0; i < n; i++) {
        double s = vdot(vsub(center, p->tv[i]), p->tn[i]);
        if (s > best) { best = s; bi = i; }
    }
    vec nrm, pb; double d;
    if (best <= 0.0) {
        /* centre inside the polygon core: EPA picks the nearest face */
        nrm = vneg(p->tn[bi]);
        d = best;
        pb = vsub(center, vmul(p->tn[bi], best));
    } else {
        double bestd2 = INFINITY;
        nrm = vneg(p->tn[bi]); d = best; pb = center;
        for (int i = 0; i < n; i++) {
            vec ea = p->tv[(i - 1 + n) % n], eb = p->tv[i];
            double t;
            vec q = closest_on_seg(ea, eb, center, &t);
            vec dl = vsub(q, center);
            double d2 = vdot(dl, dl);
            if (d2 < bestd2) {
                bestd2 = d2; pb = q;
                if (t > 0.0 && t < 1.0) { nrm = vneg(p->tn[i]); d = vdot(dl, nrm); }
                else { double len = sqrt(d2); d = len; nrm = vmul(dl, 1.0 / (len + DBL_MIN)); }
            }
        }
    }
    if (d <= c->r + p->r) {
        m->n = nrm;
        push(m, vadd(center, vmul(nrm, c->r)), vadd(pb, vmul(nrm, -p->r)), 0);
    }
}

/* cpCollide: orders the pair by cpShapeType, dispatches, fills the manifold.
 * Returns the contact count; out->sa/out->sb say which argument became "a"
 * (0 = first argument). */
int orc_collide_shapes(const orc_shape *a, const orc_shape *b, orc_manifold *out)
{
    out->count = 0;
    out->n.x = 0; out->n.y = 0;
    out->sa = 0; out->sb = 1;
    if (a->kind > b->kind) { const orc_shape *t = a; a = b; b = t; out->sa = 1; out->sb = 0; }
    if (a->kind == ORC_SHAPE_CIRCLE) {
        if (b->kind == ORC_SHAPE_CIRCLE) circle_circle(a, b, out);
        else if (b->kind == ORC_SHAPE_SEGMENT) circle_segment(a, b, out);
        else circle_poly(a, b, out);
    } else if (a->kind == ORC_SHAPE_SEGMENT && b->kind == ORC_SHAPE_SEGMENT) {
        /* SegmentToSegment never arises: all segments sit on the static arena body */
    } else {
        convex_convex(a, b, out);
    }
    return out->count;
}
