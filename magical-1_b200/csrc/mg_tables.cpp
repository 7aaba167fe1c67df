/*
 * mg_tables.cpp -- host-side construction of the constant geometry / palette tables.
 *
 * The values are what the reference's Python computes once per entity in
 *   magical/geom.py:13-63, 101-108      (regular polygons, stars, rect_verts)
 *   magical/entities.py:193-214         (make_finger_vertices)
 *   magical/entities.py:280-331, 362-372 (finger sizes / moments / shapes)
 *   magical/entities.py:610-749         (block shapes and their draw geometry)
 *   magical/render.py:13-36             (make_rect / make_circle / make_square)
 *   magical/style.py:1-40               (palette, darken / lighten)
 * evaluated with the host libm so they equal the numbers CPython produces.
 * pymunk's Poly() re-orders vertices through cpConvexHull; hull_order() reproduces that
 * for the already-convex loops used here.
 */
#include <math.h>
#include <string.h>
#include <algorithm>
#include "mg_core.cuh"
#include "mg_tables.h"

namespace {

const double ROBOT_RAD = 0.2;               /* base_env.py:62 */
const double SHAPE_RAD = ROBOT_RAD * 0.6;   /* base_env.py:64 */
const double ROBOT_LINE_THICKNESS = 0.01;   /* style.py:27 */
const double SHAPE_LINE_THICKNESS = 0.015;  /* style.py:26 */

vec2 rotated(vec2 v, double ang)            /* pymunk Vec2d.rotated */
{
    double c = cos(ang), s = sin(ang);
    return V(v.x * c - v.y * s, v.x * s + v.y * c);
}

void rect_verts(double w, double h, vec2 *o) /* geom.py:101-108 */
{
    o[0] = V(w / 2, h / 2); o[1] = V(-w / 2, h / 2); o[2] = V(-w / 2, -h / 2); o[3] = V(w / 2, -h / 2);
}

void finger_vertices(double upper_len, double fore_len, double thick, double side, vec2 *upper, vec2 *fore)
{
    double up_shift = upper_len / 2;
    vec2 fa[4];
    rect_verts(thick, upper_len, upper);
    rect_verts(thick, fore_len, fa);
    vec2 upper_start = V(side * thick / 2, upper_len / 2);
    vec2 fore_off = V(-side * thick / 2, fore_len / 2);
    double ang = side * M_PI / 8;
    vec2 off = rotated(fore_off, ang);
    vec2 trans = V(upper_start.x + off.x, upper_start.y + off.y);
    trans.y += up_shift;
    for (int i = 0; i < 4; i++) { vec2 r = rotated(fa[i], ang); fore[i] = V(r.x + trans.x, r.y + trans.y); }
    for (int i = 0; i < 4; i++) upper[i].y += up_shift;
}

/* cpMomentForPoly (offset 0; Chipmunk 7.0.3 ignores the bevel radius) */
double moment_for_poly(double m, int n, const vec2 *v)
{
    double s1 = 0.0, s2 = 0.0;
    for (int i = 0; i < n; i++) {
        vec2 v1 = v[i], v2 = v[(i + 1) % n];
        double a = v2.x * v1.y - v2.y * v1.x;
        double b = (v1.x * v1.x + v1.y * v1.y) + (v1.x * v2.x + v1.y * v2.y) + (v2.x * v2.x + v2.y * v2.y);
        s1 += a * b; s2 += a;
    }
    return (m * s1) / (6.0 * s2);
}

/* counter-clockwise, starting from the min-x (then min-y) vertex */
int hull_order(int n, const vec2 *in, vec2 *out)
{
    vec2 tmp[16];
    if (n > 1 && in[0].x == in[n - 1].x && in[0].y == in[n - 1].y) n--;
    double area2 = 0.0;
    for (int i = 0; i < n; i++) { vec2 a = in[i], b = in[(i + 1) % n]; area2 += a.x * b.y - a.y * b.x; }
    for (int i = 0; i < n; i++) tmp[i] = area2 >= 0.0 ? in[i] : in[n - 1 - i];
    int start = 0;
    for (int i = 1; i < n; i++)
        if (tmp[i].x < tmp[start].x || (tmp[i].x == tmp[start].x && tmp[i].y < tmp[start].y)) start = i;
    for (int i = 0; i < n; i++) out[i] = tmp[(start + i) % n];
    return n;
}

void set_tpl(MgTables &T, int tpl, int n, const vec2 *hull, double r)
{
    T.tpl_n[tpl] = n; T.tpl_r[tpl] = r;
    double bound = 0.0;
    for (int i = 0; i < n; i++) bound = std::max(bound, sqrt(hull[i].x * hull[i].x + hull[i].y * hull[i].y));
    T.tpl_bound[tpl] = bound + r;
    for (int i = 0; i < n; i++) {
        vec2 a = hull[(i - 1 + n) % n], b = hull[i];
        double dx = b.x - a.x, dy = b.y - a.y;
        double nx = dy, ny = -dx;                      /* cpvrperp */
        double inv = 1.0 / (sqrt(nx * nx + ny * ny) + MG_DBL_MIN);
        T.tpl_v[tpl][i] = b;
        T.tpl_nrm[tpl][i] = V(nx * inv, ny * inv);
    }
}

void star_verts(int n_points, double out_r, double in_r, vec2 *o) /* geom.py:49-63 */
{
    for (int k = 0; k < n_points; k++) {
        o[2 * k] = rotated(V(0, out_r), k * 2 * M_PI / n_points);
        o[2 * k + 1] = rotated(V(0, in_r), (2 * k + 1) * M_PI / n_points);
    }
}

/* Convex pieces of the star standing in for autogeom.convex_decomposition (entities.py:650-651):
 * five tip triangles, then the inner pentagon. */
void star_pieces(const vec2 *s, int *pn, vec2 pv[MG_STAR_PIECES][MG_MAXV])
{
    for (int k = 0; k < 5; k++) {
        pn[k] = 3;
        pv[k][0] = s[(2 * k + 9) % 10]; pv[k][1] = s[2 * k]; pv[k][2] = s[(2 * k + 1) % 10];
    }
    pn[5] = 5;
    for (int k = 0; k < 5; k++) pv[5][k] = s[2 * k + 1];
}

int dl_cursor = 0;
void set_dl(MgTables &T, int id, int n, const vec2 *v)
{
    T.dl_off[id] = dl_cursor; T.dl_n[id] = n;
    double rad = 0.0;
    for (int i = 0; i < n; i++) {
        T.dl_pool[dl_cursor++] = v[i];
        rad = std::max(rad, sqrt(v[i].x * v[i].x + v[i].y * v[i].y));
    }
    T.dl_rad[id] = rad;
}

void circle_verts(double radius, vec2 *o) /* render.make_circle, render.py:27-32 */
{
    for (int i = 0; i < MG_CIRCLE_RES; i++) {
        double ang = 2 * M_PI * i / MG_CIRCLE_RES;
        o[i] = V(cos(ang) * radius, sin(ang) * radius);
    }
}

void make_rect(double w, double h, vec2 *o) /* render.make_rect, render.py:13-24 */
{
    double rh = h / 2, rw = w / 2;
    o[0] = V(-rw, rh); o[1] = V(rw, rh); o[2] = V(rw, -rh); o[3] = V(-rw, -rh);
}

/* colorsys.rgb_to_hls / hls_to_rgb (CPython Lib/colorsys.py) and style.py:10-23 */
void rgb_to_hls(const double *c, double *hls)
{
    double r = c[0], g = c[1], b = c[2];
    double maxc = std::max(r, std::max(g, b)), minc = std::min(r, std::min(g, b));
    double sumc = maxc + minc, rangec = maxc - minc;
    double l = sumc / 2.0;
    if (minc == maxc) { hls[0] = 0.0; hls[1] = l; hls[2] = 0.0; return; }
    double s = l <= 0.5 ? rangec / sumc : rangec / (2.0 - sumc);
    double rc = (maxc - r) / rangec, gc = (maxc - g) / rangec, bc = (maxc - b) / rangec, h;
    if (r == maxc) h = bc - gc;
    else if (g == maxc) h = 2.0 + rc - bc;
    else h = 4.0 + gc - rc;
    h = fmod(h / 6.0, 1.0);
    if (h < 0) h += 1.0;                    /* Python % */
    hls[0] = h; hls[1] = l; hls[2] = s;
}
double hls_v(double m1, double m2, double hue)
{
    hue = fmod(hue, 1.0);
    if (hue < 0) hue += 1.0;
    if (hue < 1.0 / 6.0) return m1 + (m2 - m1) * hue * 6.0;
    if (hue < 0.5) return m2;
    if (hue < 2.0 / 3.0) return m1 + (m2 - m1) * (2.0 / 3.0 - hue) * 6.0;
    return m1;
}
void hls_to_rgb(const double *hls, double *c)
{
    double h = hls[0], l = hls[1], s = hls[2];
    if (s == 0.0) { c[0] = c[1] = c[2] = l; return; }
    double m2 = l <= 0.5 ? l * (1.0 + s) : l + s - (l * s);
    double m1 = 2.0 * l - m2;
    c[0] = hls_v(m1, m2, h + 1.0 / 3.0); c[1] = hls_v(m1, m2, h); c[2] = hls_v(m1, m2, h - 1.0 / 3.0);
}
void darken(const double *c, double *o)
{
    double hls[3]; rgb_to_hls(c, hls);
    hls[1] = std::max(0.0, hls[1] * 0.9);
    hls_to_rgb(hls, o);
}
void lighten(const double *c, double times, double *o)
{
    double hls[3]; rgb_to_hls(c, hls);
    double mult = pow(1.4, times);
    hls[1] = 1 - (1 - hls[1]) / mult;
    hls_to_rgb(hls, o);
}
void set_pal(MgTables &T, int idx, const double *c)
{
    for (int k = 0; k < 3; k++) T.pal[idx][k] = (uint8_t)nearbyint(c[k] * 255); /* np.round: half to even */
    T.pal[idx][3] = 0;
}

MgTables build()
{
    MgTables T;
    memset(&T, 0, sizeof(T));
    dl_cursor = 0;

    /* fingers */
    double thick = 0.25 * ROBOT_RAD, upper = 1.1 * ROBOT_RAD, lower = 0.7 * ROBOT_RAD;
    vec2 fin_out[2][2][4], fin_in[2][2][4];
    for (int s = 0; s < 2; s++) {
        double side = s == 0 ? -1.0 : 1.0;
        vec2 up[4], fo[4], loop8[8];
        finger_vertices(upper, lower, thick, side, up, fo);
        for (int i = 0; i < 4; i++) { loop8[i] = up[i]; loop8[4 + i] = fo[i]; }
        T.finger_moment[s] = moment_for_poly(1.0 / 8, 8, loop8);   /* entities.py:313-315 */
        hull_order(4, up, fin_out[s][0]);
        hull_order(4, fo, fin_out[s][1]);
        set_tpl(T, MG_TPL_FINGER0 + s * 2 + 0, 4, fin_out[s][0], 0.0);
        set_tpl(T, MG_TPL_FINGER0 + s * 2 + 1, 4, fin_out[s][1], 0.0);
        vec2 iu[4], ifo[4];
        finger_vertices(upper - ROBOT_LINE_THICKNESS * 2, lower - ROBOT_LINE_THICKNESS * 2,
                        thick - ROBOT_LINE_THICKNESS * 2, side, iu, ifo);
        for (int i = 0; i < 4; i++) {
            fin_in[s][0][i] = V(iu[i].x, iu[i].y + ROBOT_LINE_THICKNESS);
            fin_in[s][1][i] = V(ifo[i].x, ifo[i].y + ROBOT_LINE_THICKNESS);
        }
    }

    /* square (cpBoxShapeInit2 order: (r,b) (r,t) (l,t) (l,b)) */
    double side_len = sqrt(M_PI) * SHAPE_RAD, hw = side_len / 2;
    vec2 sq[4] = { V(hw, -hw), V(hw, hw), V(-hw, hw), V(-hw, -hw) };
    set_tpl(T, MG_TPL_SQUARE, 4, sq, 0.01 * side_len);
    T.square_moment = 0.5 * moment_for_poly(1.0, 4, sq);           /* shape.mass = 0.5, entities.py:629 */

    /* pentagon */
    vec2 pent[5], pent_hull[5];
    {
        int n = 5;
        double p_n = M_PI / n;
        double sl = 1.0 * (2 * SHAPE_RAD * sqrt(p_n * tan(p_n)));  /* geom.py:13-22 */
        double radius = sl / (2 * sin(M_PI / n));                  /* geom.py:35-46 */
        double step = 2 * M_PI / n;
        for (int k = 0; k < n; k++) pent[k] = rotated(V(0, radius), k * step);
        T.pent_moment = moment_for_poly(0.5, n, pent);
        hull_order(n, pent, pent_hull);
        set_tpl(T, MG_TPL_PENT, 5, pent_hull, 0.0);
    }

    /* star */
    vec2 star[10], star_in[10];
    int sp_n[MG_STAR_PIECES], spi_n[MG_STAR_PIECES];
    vec2 sp_v[MG_STAR_PIECES][MG_MAXV], spi_v[MG_STAR_PIECES][MG_MAXV];
    {
        double out_r = 1.3 * SHAPE_RAD, in_r = 0.5 * out_r;
        star_verts(5, out_r, in_r, star);
        vec2 tips[5], hull[5];
        for (int k = 0; k < 5; k++) tips[k] = star[2 * k];
        hull_order(5, tips, hull);
        T.star_moment = moment_for_poly(0.5, 5, hull);             /* entities.py:652-654 */
        star_pieces(star, sp_n, sp_v);
        for (int k = 0; k < MG_STAR_PIECES; k++) {
            vec2 hv[MG_MAXV];
            int n = hull_order(sp_n[k], sp_v[k], hv);
            set_tpl(T, MG_TPL_STAR0 + k, n, hv, 0.0);
        }
        star_verts(5, out_r - SHAPE_LINE_THICKNESS, in_r - SHAPE_LINE_THICKNESS, star_in);
        star_pieces(star_in, spi_n, spi_v);
    }

    /* arena walls: Segment radius 1 through the corners (+-2, +-2), entities.py:507-522 */
    {
        double l = -1, r = 1, b = -1, t = 1, rad = 1;
        vec2 pts[4] = { V(l - rad, t + rad), V(r + rad, t + rad), V(r + rad, b - rad), V(l - rad, b - rad) };
        for (int i = 0; i < 4; i++) {
            vec2 a = pts[i], c = pts[(i + 1) % 4];
            double dx = c.x - a.x, dy = c.y - a.y;
            double inv = 1.0 / (sqrt(dx * dx + dy * dy) + MG_DBL_MIN);
            T.wall_a[i] = a; T.wall_b[i] = c;
            T.wall_n[i] = V(dy * inv, -(dx * inv));    /* cpvrperp(cpvnormalize(b - a)) */
        }
    }

    /* constants with transcendental functions */
    const double dt = (1.0 / 8) / 10;                  /* base_env.py:248-251 */
    T.err_bias_default = pow(1.0 - 0.1, 60.0);
    T.bias_coef_default = 1.0 - pow(T.err_bias_default, dt);
    {
        double i_robot = 1.0 * (0.5 * (0.0 * 0.0 + ROBOT_RAD * ROBOT_RAD) + 0.0);
        double i_pupil = (1.0 / 10) * (0.5 * (ROBOT_RAD * ROBOT_RAD));
        double moment = 1.0 / i_robot + 1.0 / i_pupil;
        T.spring_w_coef = 1.0 - exp(-3e-3 * dt * moment);
    }

    /* draw lists */
    vec2 tmp[MG_CIRCLE_RES];
    make_rect(2.0, 2.0, tmp); set_dl(T, MG_DL_ARENA, 4, tmp);
    for (int s = 0; s < 2; s++) for (int p = 0; p < 2; p++) set_dl(T, MG_DL_FINGER_OUT0 + s * 2 + p, 4, fin_out[s][p]);
    for (int s = 0; s < 2; s++) for (int p = 0; p < 2; p++) set_dl(T, MG_DL_FINGER_IN0 + s * 2 + p, 4, fin_in[s][p]);
    circle_verts(ROBOT_RAD, tmp); set_dl(T, MG_DL_ROBOT, MG_CIRCLE_RES, tmp);
    circle_verts(0.2 * ROBOT_RAD, tmp); set_dl(T, MG_DL_EYE, MG_CIRCLE_RES, tmp);
    circle_verts(0.12 * ROBOT_RAD, tmp); set_dl(T, MG_DL_PUPIL, MG_CIRCLE_RES, tmp);
    make_rect(side_len, side_len, tmp); set_dl(T, MG_DL_SQUARE, 4, tmp);
    set_dl(T, MG_DL_PENT, 5, pent);
    circle_verts(SHAPE_RAD, tmp); set_dl(T, MG_DL_CIRCLE, MG_CIRCLE_RES, tmp);
    for (int k = 0; k < MG_STAR_PIECES; k++) set_dl(T, MG_DL_STAR_OUT0 + k, sp_n[k], sp_v[k]);
    for (int k = 0; k < MG_STAR_PIECES; k++) set_dl(T, MG_DL_STAR_IN0 + k, spi_n[k], spi_v[k]);
    T.dl_off[MG_DL_GOAL] = dl_cursor; T.dl_n[MG_DL_GOAL] = 4; T.dl_rad[MG_DL_GOAL] = 0.0;

    /* palette (style.py:28-37; entities.py:380-382, 415, 421, 527-528, 717-718, 793-795; base_env.py:199) */
    double grey[3] = { 162 / 255.0, 163 / 255.0, 175 / 255.0 }, c[3];
    double white[3] = { 1.0, 1.0, 1.0 }, pupil[3] = { 0.1, 0.1, 0.1 };
    lighten(grey, 4, c); set_pal(T, MG_PAL_BG, c);
    set_pal(T, MG_PAL_WHITE, white);
    set_pal(T, MG_PAL_GREY, grey);
    darken(grey, c); set_pal(T, MG_PAL_GREY_DARK, c);
    lighten(grey, 4, c); set_pal(T, MG_PAL_GREY_LIGHT, c);
    set_pal(T, MG_PAL_PUPIL, pupil);
    const double base[4][3] = { /* SHAPE_COLOURS order: red, green, blue, yellow */
        { 0xEE / 255.0, 0x1F / 255.0, 0x60 / 255.0 }, { 0x85 / 255.0, 0x94 / 255.0, 0x38 / 255.0 },
        { 0x3B / 255.0, 0x7E / 255.0, 0xA1 / 255.0 }, { 0xFD / 255.0, 0xB5 / 255.0, 0x15 / 255.0 } };
    for (int k = 0; k < 4; k++) {
        double col[3];
        lighten(base[k], 1.7, col);
        set_pal(T, MG_PAL_COLOUR0 + k, col);
        darken(col, c); set_pal(T, MG_PAL_DARK0 + k, c);
        lighten(col, 2, c); set_pal(T, MG_PAL_LIGHT0 + k, c);
    }
    return T;
}

} // namespace

const MgTables *mg_tables_host()
{
    static MgTables T = build();
    return &T;
}
