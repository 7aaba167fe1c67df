/*
 * orc_geom.c -- ORACLE (test infrastructure): constant geometry + palette tables.
 *
 * Restates the vertex generators of magical/geom.py and the geometry halves of
 * Robot.setup / Shape.setup / ArenaBoundaries.setup in magical/entities.py.
 * PARITY UNPINNED for the parts that go through Chipmunk (hull order, moments,
 * star decomposition); see orc.h.
 */
#include <math.h>
#include <string.h>
#include "orc.h"

/* pymunk Vec2d.rotated(angle): (x cos - y sin, x sin + y cos); used by
 * geom.py:41-62 and entities.py:201-208. */
static vec v_rotated(vec v, double ang)
{
    double c = cos(ang), s = sin(ang);
    vec r = { v.x * c - v.y * s, v.x * s + v.y * c };
    return r;
}

/* geom.rect_verts(w, h): counter-clockwise from top right (geom.py:101-108) */
static void rect_verts(double w, double h, vec out[4])
{
    out[0].x = w / 2;  out[0].y = h / 2;
    out[1].x = -w / 2; out[1].y = h / 2;
    out[2].x = -w / 2; out[2].y = -h / 2;
    out[3].x = w / 2;  out[3].y = -h / 2;
}

/* entities.make_finger_vertices (entities.py:193-214) */
static void make_finger_vertices(double upper_arm_len, double forearm_len, double thickness,
                                 double side_sign, vec upper[4], vec fore[4])
{
    double up_shift = upper_arm_len / 2;
    vec fa[4];
    rect_verts(thickness, upper_arm_len, upper);
    rect_verts(thickness, forearm_len, fa);
    vec upper_start = { side_sign * thickness / 2, upper_arm_len / 2 };
    vec forearm_offset_unrot = { -side_sign * thickness / 2, forearm_len / 2 };
    double rot_angle = side_sign * M_PI / 8;
    vec off = v_rotated(forearm_offset_unrot, rot_angle);
    vec forearm_trans = { upper_start.x + off.x, upper_start.y + off.y };
    forearm_trans.y += up_shift;
    for (int i = 0; i < 4; i++) {
        vec r = v_rotated(fa[i], rot_angle);
        fore[i].x = r.x + forearm_trans.x;
        fore[i].y = r.y + forearm_trans.y;
    }
    for (int i = 0; i < 4; i++) upper[i].y += up_shift;
}

/* cpMomentForPoly(m, count, verts, offset=(0,0), r) -- Chipmunk 7.0.3 chipmunk.c.
 * The bevel radius is ignored by that version ("TODO account for radius").
 * Call sites: entities.py:314-315 (finger, 8-vertex loop), :653 (star hull), :687 (pentagon). */
double orc_moment_for_poly(double m, int n, const vec *v)
{
    double sum1 = 0.0, sum2 = 0.0;
    for (int i = 0; i < n; i++) {
        vec v1 = v[i], v2 = v[(i + 1) % n];
        double a = v2.x * v1.y - v2.y * v1.x; /* cpvcross(v2, v1) */
        double b = (v1.x * v1.x + v1.y * v1.y) + (v1.x * v2.x + v1.y * v2.y) + (v2.x * v2.x + v2.y * v2.y);
        sum1 += a * b;
        sum2 += a;
    }
    return (m * sum1) / (6.0 * sum2);
}

/* pm.Poly(body, verts) runs cpConvexHull over the vertices (cpPolyShapeNew),
 * which returns them counter-clockwise starting from the min-x (ties: min-y)
 * vertex found by cpLoopIndexes.  Every polygon the reference builds is already
 * convex, so the hull is a rotation (and, for clockwise input, a reversal) of
 * the input loop.  Call sites: entities.py:366, 661, 692. */
int orc_hull_order(int n, const vec *in, vec *out)
{
    vec tmp[ORC_MAXV * 2];
    /* drop a duplicated closing vertex (convex_decomposition polylines are closed) */
    if (n > 1 && in[0].x == in[n - 1].x && in[0].y == in[n - 1].y) n--;
    double area2 = 0.0;
    for (int i = 0; i < n; i++) {
        vec a = in[i], b = in[(i + 1) % n];
        area2 += a.x * b.y - a.y * b.x;
    }
    for (int i = 0; i < n; i++) tmp[i] = (area2 >= 0.0) ? in[i] : in[n - 1 - i];
    int start = 0;
    for (int i = 1; i < n; i++) {
        if (tmp[i].x < tmp[start].x || (tmp[i].x == tmp[start].x && tmp[i].y < tmp[start].y)) start = i;
    }
    for (int i = 0; i < n; i++) out[i] = tmp[(start + i) % n];
    return n;
}

/* geom.compute_star_verts (geom.py:49-63) */
static void star_verts(int n_points, double out_r, double in_r, vec *out)
{
    vec ov = { 0, out_r }, iv = { 0, in_r };
    for (int k = 0; k < n_points; k++) {
        double out_angle = k * 2 * M_PI / n_points;
        out[2 * k] = v_rotated(ov, out_angle);
        double in_angle = (2 * k + 1) * M_PI / n_points;
        out[2 * k + 1] = v_rotated(iv, in_angle);
    }
}

/* Stand-in for autogeom.convex_decomposition(star + star[:1], 0) (entities.py:650-651, 720-721).
 * Chipmunk's deepest-notch/Steiner-point decomposition cannot be reproduced here
 * (source absent); the union of pieces is the same star.  We use the 5 tip
 * triangles followed by the inner pentagon, each wound counter-clockwise. */
static void star_pieces(const vec s[10], orc_piece_set *ps)
{
    ps->n_pieces = 6;
    for (int k = 0; k < 5; k++) {
        ps->piece_n[k] = 3;
        ps->piece_v[k][0] = s[(2 * k + 9) % 10];
        ps->piece_v[k][1] = s[2 * k];
        ps->piece_v[k][2] = s[(2 * k + 1) % 10];
    }
    ps->piece_n[5] = 5;
    for (int k = 0; k < 5; k++) ps->piece_v[5][k] = s[2 * k + 1];
}

static void make_circle(double radius, vec *out)
{
    /* render.make_circle (render.py:27-32) */
    for (int i = 0; i < ORC_CIRCLE_RES; i++) {
        double ang = 2 * M_PI * i / ORC_CIRCLE_RES;
        out[i].x = cos(ang) * radius;
        out[i].y = sin(ang) * radius;
    }
}

static void make_rect(double width, double height, vec out[4])
{
    /* render.make_rect (render.py:13-24) */
    double rad_h = height / 2, rad_w = width / 2;
    out[0].x = -rad_w; out[0].y = rad_h;
    out[1].x = rad_w;  out[1].y = rad_h;
    out[2].x = rad_w;  out[2].y = -rad_h;
    out[3].x = -rad_w; out[3].y = -rad_h;
}

static void set3(uint8_t *d, int r, int g, int b) { d[0] = (uint8_t)r; d[1] = (uint8_t)g; d[2] = (uint8_t)b; }

static orc_geom_tables G;
static int G_ready = 0;

const orc_geom_tables *orc_geom(void)
{
    if (G_ready) return &G;
    const double ROBOT_RAD = 0.2;              /* base_env.py:62 */
    const double SHAPE_RAD = ROBOT_RAD * 0.6;  /* base_env.py:64 */
    const double ROBOT_LINE_THICKNESS = 0.01;  /* style.py:27 */
    const double SHAPE_LINE_THICKNESS = 0.015; /* style.py:26 */

    /* fingers: entities.py:280-304, 362-372 */
    double thick = 0.25 * ROBOT_RAD, upper = 1.1 * ROBOT_RAD, lower = 0.7 * ROBOT_RAD;
    for (int s = 0; s < 2; s++) {
        double side = s == 0 ? -1.0 : 1.0;
        vec up[4], fo[4], loop8[8];
        make_finger_vertices(upper, lower, thick, side, up, fo);
        for (int i = 0; i < 4; i++) { loop8[i] = up[i]; loop8[4 + i] = fo[i]; }
        G.finger_moment[s] = orc_moment_for_poly(1.0 / 8, 8, loop8); /* finger_mass = mass / 8, entities.py:313-315 */
        orc_hull_order(4, up, G.finger_phys[s][0]);
        orc_hull_order(4, fo, G.finger_phys[s][1]);
        vec iu[4], ifo[4];
        make_finger_vertices(upper - ROBOT_LINE_THICKNESS * 2, lower - ROBOT_LINE_THICKNESS * 2,
                             thick - ROBOT_LINE_THICKNESS * 2, side, iu, ifo);
        for (int i = 0; i < 4; i++) {
            G.finger_inner[s][0][i].x = iu[i].x;  G.finger_inner[s][0][i].y = iu[i].y + ROBOT_LINE_THICKNESS;
            G.finger_inner[s][1][i].x = ifo[i].x; G.finger_inner[s][1][i].y = ifo[i].y + ROBOT_LINE_THICKNESS;
        }
    }

    /* square: entities.py:616-631 (cpBoxShapeInit2 vertex order r,b / r,t / l,t / l,b) and :713 */
    double side_len = sqrt(M_PI) * SHAPE_RAD;
    double hw = side_len / 2;
    G.square_side = side_len;
    G.square_r = 0.01 * side_len;
    G.square_phys[0].x = hw;  G.square_phys[0].y = -hw;
    G.square_phys[1].x = hw;  G.square_phys[1].y = hw;
    G.square_phys[2].x = -hw; G.square_phys[2].y = hw;
    G.square_phys[3].x = -hw; G.square_phys[3].y = -hw;
    /* shape.mass = 0.5 -> body takes m and m * cpMomentForPoly(1, box verts) (cpPolyShapeMassInfo) */
    G.square_moment = 0.5 * orc_moment_for_poly(1.0, 4, G.square_phys);
    make_rect(side_len, side_len, G.square_draw);

    /* pentagon: entities.py:672-692, geom.py:13-22, 35-46 */
    {
        int n = 5;
        double p_n = M_PI / n;
        double sl = 1.0 * (2 * SHAPE_RAD * sqrt(p_n * tan(p_n)));
        double radius = sl / (2 * sin(M_PI / n));
        double step = 2 * M_PI / n;
        vec first = { 0, radius };
        for (int k = 0; k < n; k++) G.pent[k] = v_rotated(first, k * step);
        G.pent_moment = orc_moment_for_poly(0.5, n, G.pent);
        orc_hull_order(n, G.pent, G.pent_phys);
    }

    /* star: entities.py:643-665, 716-727 */
    {
        double out_r = 1.3 * SHAPE_RAD, in_r = 0.5 * out_r;
        star_verts(5, out_r, in_r, G.star);
        vec hull_in[5], hull[5];
        for (int k = 0; k < 5; k++) hull_in[k] = G.star[2 * k];
        orc_hull_order(5, hull_in, hull);
        G.star_moment = orc_moment_for_poly(0.5, 5, hull);
        star_pieces(G.star, &G.star_outer);
        vec shortv[10];
        star_verts(5, out_r - SHAPE_LINE_THICKNESS, in_r - SHAPE_LINE_THICKNESS, shortv);
        star_pieces(shortv, &G.star_inner);
    }

    make_circle(SHAPE_RAD, G.circle_block);          /* entities.py:715 */
    make_circle(ROBOT_RAD, G.circle_robot);          /* entities.py:379 */
    make_circle(0.2 * ROBOT_RAD, G.circle_eye);      /* entities.py:414 */
    make_circle(0.12 * ROBOT_RAD, G.circle_pupil);   /* entities.py:420 */
    make_rect(2.0, 2.0, G.arena_draw);               /* entities.py:524-526 */

    /* palette: round(255 * c) of style.COLOURS_RGB and its darken/lighten variants
     * (render.py:146-149, style.py:28-37); values pinned by tests/golden/constants.json,
     * which is generated from the reference's own style.py. */
    set3(G.pal_bg, 231, 231, 234);          /* lighten_rgb(grey, 4), base_env.py:199 */
    set3(G.pal_white, 255, 255, 255);
    set3(G.pal_grey, 162, 163, 175);
    set3(G.pal_grey_dark, 144, 145, 159);
    set3(G.pal_grey_light, 231, 231, 234);
    set3(G.pal_pupil, 26, 26, 26);
    set3(G.pal_colour[MG_RED], 245, 129, 165);   set3(G.pal_colour_dark[MG_RED], 243, 94, 141);    set3(G.pal_colour_light2[MG_RED], 250, 190, 209);
    set3(G.pal_colour[MG_GREEN], 195, 208, 130); set3(G.pal_colour_dark[MG_GREEN], 183, 198, 105); set3(G.pal_colour_light2[MG_GREEN], 224, 231, 191);
    set3(G.pal_colour[MG_BLUE], 135, 185, 211);  set3(G.pal_colour_dark[MG_BLUE], 110, 170, 202);  set3(G.pal_colour_light2[MG_BLUE], 194, 219, 233);
    set3(G.pal_colour[MG_YELLOW], 254, 213, 123);set3(G.pal_colour_dark[MG_YELLOW], 254, 201, 86); set3(G.pal_colour_light2[MG_YELLOW], 254, 234, 188);
    G_ready = 1;
    return &G;
}
