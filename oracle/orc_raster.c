/*
 * orc_raster.c -- ORACLE (test infrastructure): the software renderer and the
 * 4x area downsample.
 *
 * Python half restated from magical/render.py (Transform :39-119, Stack :122-134,
 * Geom.render :151-158, Poly._render/draw_outline :212-287, ego_cam_matrix
 * :290-304, Viewer.set_bounds/set_cam_follow/render :339-395) and
 * magical/base_env.py:309-343.  The C half stands in for pygame 1.9.6 draw.c
 * (draw_fillpoly, clip_and_draw_line_width, clipline, drawline, drawhorzline,
 * drawvertline) on a 32-bit software surface.  PARITY UNPINNED for the pygame
 * half (pygame is not installed here; see orc.h).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "orc.h"
#include "../include/mg_trig.h"

#define RES MG_RES

/* ------------------------------------------------------------ 3x3 affine (last row 0 0 1) */
typedef struct { double a, b, tx, c, d, ty; } aff;

/* numpy `A @ B` for two such matrices; each entry is (p*q + r*s) + t*u */
static aff aff_mul(aff A, aff B)
{
    aff C;
    C.a = (A.a * B.a + A.b * B.c) + A.tx * 0.0;
    C.b = (A.a * B.b + A.b * B.d) + A.tx * 0.0;
    C.tx = (A.a * B.tx + A.b * B.ty) + A.tx * 1.0;
    C.c = (A.c * B.a + A.d * B.c) + A.ty * 0.0;
    C.d = (A.c * B.b + A.d * B.d) + A.ty * 0.0;
    C.ty = (A.c * B.tx + A.d * B.ty) + A.ty * 1.0;
    return C;
}

/* Transform.__init__: translation @ rotation @ scaling (render.py:87-97) */
static aff transform_make(double tx, double ty, double rot, double sx, double sy)
{
    aff T = { 1.0, 0.0, tx, 0.0, 1.0, ty };
    double cs, sn;   /* np.cos / np.sin in the reference; shared mg_sincos here (see orc_scene.c) */
    mg_sincos(rot, &sn, &cs);
    aff R = { cs, -sn, 0.0, sn, cs, 0.0 };
    aff S = { sx, 0.0, 0.0, 0.0, sy, 0.0 };
    return aff_mul(aff_mul(T, R), S);
}

/* Viewer.pygame_transform = T(0, height) @ S(1, -1) (render.py:329-330) */
static aff pygame_transform(void)
{
    return aff_mul(transform_make(0, RES, 0, 1, 1), transform_make(0, 0, 0, 1.0, -1.0));
}

/* BaseEnv._use_allo_cam + Viewer.set_bounds (base_env.py:318-322, render.py:339-347) */
static aff allo_camera(void)
{
    const double zoom = 1.02; /* style.ARENA_ZOOM_OUT */
    double left = -1 * zoom, right = 1 * zoom, bottom = -1 * zoom, top = 1 * zoom;
    double scale_x = RES / (right - left), scale_y = RES / (top - bottom);
    aff cam = transform_make(-left * scale_x, -bottom * scale_y, 0.0, scale_x, scale_y);
    return aff_mul(pygame_transform(), cam);
}

/* BaseEnv._use_ego_cam + Viewer.set_cam_follow + ego_cam_matrix
 * (base_env.py:309-316, render.py:349-371, 290-304) */
static aff ego_camera(double cx, double cy, double rotation)
{
    const double zoom = 1.02;
    double world_h = 2 * zoom, world_w = 2 * zoom;
    double scale_x = RES / world_w, scale_y = RES / world_h;
    aff scale = transform_make(0, 0, 0, scale_x, scale_y);
    aff tr1 = transform_make(world_w * 0.5, world_h * 0.15, 0, 1, 1);
    aff rot = transform_make(0, 0, -rotation, 1, 1);
    aff tr2 = transform_make(-cx, -cy, 0, 1, 1);
    aff m = aff_mul(rot, tr2);
    m = aff_mul(tr1, m);
    m = aff_mul(scale, m);
    return aff_mul(pygame_transform(), m);
}

/* ------------------------------------------------------------ surface */
typedef struct { uint8_t *px; } surf; /* RES x RES x 3 */

static inline void set_at(surf *s, int x, int y, const uint8_t *c)
{
    uint8_t *p = s->px + ((size_t)y * RES + x) * 3;
    p[0] = c[0]; p[1] = c[1]; p[2] = c[2];
}

/* drawhorzline / drawvertline: inclusive on both ends, either order */
static void drawhorzline(surf *s, const uint8_t *c, int x1, int y, int x2)
{
    if (x1 > x2) { int t = x1; x1 = x2; x2 = t; }
    for (int x = x1; x <= x2; x++) set_at(s, x, y, c);
}
static void drawvertline(surf *s, const uint8_t *c, int x, int y1, int y2)
{
    if (y1 > y2) { int t = y1; y1 = y2; y2 = t; }
    for (int y = y1; y <= y2; y++) set_at(s, x, y, c);
}

/* drawhorzlineclip with clip rect = whole surface */
static void drawhorzlineclip(surf *s, const uint8_t *c, int x1, int y, int x2)
{
    if (y < 0 || y >= RES) return;
    if (x2 < x1) { int t = x1; x1 = x2; x2 = t; }
    if (x1 < 0) x1 = 0;
    if (x2 > RES - 1) x2 = RES - 1;
    if (x2 < 0 || x1 >= RES) return;
    if (x1 == x2) set_at(s, x1, y, c);
    else drawhorzline(s, c, x1, y, x2);
}

static int cmp_int(const void *a, const void *b) { return *(const int *)a - *(const int *)b; }

/* draw_fillpoly (pygame 1.9.6 draw.c) */
static void draw_fillpoly(surf *s, const int *vx, const int *vy, int n, const uint8_t *color)
{
    int polyints[ORC_CIRCLE_RES + 2];
    int miny = vy[0], maxy = vy[0];
    for (int i = 1; i < n; i++) {
        if (vy[i] < miny) miny = vy[i];
        if (vy[i] > maxy) maxy = vy[i];
    }
    /* rows outside the surface are clipped away by drawhorzlineclip; skip them up front */
    int ylo = miny < 0 ? 0 : miny, yhi = maxy > RES - 1 ? RES - 1 : maxy;
    for (int y = ylo; y <= yhi; y++) {
        int ints = 0;
        for (int i = 0; i < n; i++) {
            int ind1, ind2, x1, y1, x2, y2;
            if (!i) { ind1 = n - 1; ind2 = 0; }
            else { ind1 = i - 1; ind2 = i; }
            y1 = vy[ind1]; y2 = vy[ind2];
            if (y1 < y2) { x1 = vx[ind1]; x2 = vx[ind2]; }
            else if (y1 > y2) { y2 = vy[ind1]; y1 = vy[ind2]; x2 = vx[ind1]; x1 = vx[ind2]; }
            else if (miny == maxy) {
                /* special case: polygon only 1 pixel high */
                int minx = vx[0], maxx = vx[0];
                for (int k = 1; k < n; k++) {
                    if (vx[k] < minx) minx = vx[k];
                    if (vx[k] > maxx) maxx = vx[k];
                }
                polyints[ints++] = minx;
                polyints[ints++] = maxx;
                break;
            }
            else continue;
            if (((y >= y1) && (y < y2)) || ((y == maxy) && (y > y1) && (y <= y2))) {
                polyints[ints++] = (y - y1) * (x2 - x1) / (y2 - y1) + x1;
            }
        }
        qsort(polyints, ints, sizeof(int), cmp_int);
        for (int i = 0; i + 1 < ints; i += 2) drawhorzlineclip(s, color, polyints[i], y, polyints[i + 1]);
    }
}

/* clipline: Cohen-Sutherland with a float slope (pygame 1.9.6 draw.c) */
#define LEFT_EDGE 0x1
#define RIGHT_EDGE 0x2
#define BOTTOM_EDGE 0x4
#define TOP_EDGE 0x8
static int encode(int x, int y, int left, int top, int right, int bottom)
{
    int code = 0;
    if (x < left) code |= LEFT_EDGE;
    if (x > right) code |= RIGHT_EDGE;
    if (y < top) code |= TOP_EDGE;
    if (y > bottom) code |= BOTTOM_EDGE;
    return code;
}
static int clipline(int *pts, int left, int top, int right, int bottom)
{
    int x1 = pts[0], y1 = pts[1], x2 = pts[2], y2 = pts[3];
    int code1, code2, draw = 0, swaptmp;
    float m;
    while (1) {
        code1 = encode(x1, y1, left, top, right, bottom);
        code2 = encode(x2, y2, left, top, right, bottom);
        if (!(code1 | code2)) { draw = 1; break; }
        else if (code1 & code2) break;
        else {
            if (!code1) {
                swaptmp = x2; x2 = x1; x1 = swaptmp;
                swaptmp = y2; y2 = y1; y1 = swaptmp;
                swaptmp = code2; code2 = code1; code1 = swaptmp;
            }
            if (x2 != x1) m = (y2 - y1) / (float)(x2 - x1);
            else m = 1.0f;
            if (code1 & LEFT_EDGE) { y1 += (int)((left - x1) * m); x1 = left; }
            else if (code1 & RIGHT_EDGE) { y1 += (int)((right - x1) * m); x1 = right; }
            else if (code1 & BOTTOM_EDGE) { if (x2 != x1) x1 += (int)((bottom - y1) / m); y1 = bottom; }
            else if (code1 & TOP_EDGE) { if (x2 != x1) x1 += (int)((top - y1) / m); y1 = top; }
        }
    }
    if (draw) { pts[0] = x1; pts[1] = y1; pts[2] = x2; pts[3] = y2; }
    return draw;
}

/* drawline: pygame 1.9.x integer DDA (deltas are |d|+1) */
static void drawline(surf *s, const uint8_t *c, int x1, int y1, int x2, int y2)
{
    int deltax = x2 - x1, deltay = y2 - y1;
    int signx = deltax < 0 ? -1 : 1, signy = deltay < 0 ? -1 : 1;
    deltax = signx * deltax + 1;
    deltay = signy * deltay + 1;
    int px = x1, py = y1;
    int stepx_major = signx, stepy_major = 0, stepx_minor = 0, stepy_minor = signy;
    if (deltax < deltay) {
        int t = deltax; deltax = deltay; deltay = t;
        stepx_major = 0; stepy_major = signy; stepx_minor = signx; stepy_minor = 0;
    }
    int y = 0;
    for (int x = 0; x < deltax; x++) {
        set_at(s, px, py, c);
        px += stepx_major; py += stepy_major;
        y += deltay;
        if (y >= deltax) { y -= deltax; px += stepx_minor; py += stepy_minor; }
    }
}

static int clip_and_draw_line(surf *s, const uint8_t *c, int *pts)
{
    if (!clipline(pts, 0, 0, RES - 1, RES - 1)) return 0;
    if (pts[1] == pts[3]) drawhorzline(s, c, pts[0], pts[1], pts[2]);
    else if (pts[0] == pts[2]) drawvertline(s, c, pts[0], pts[1], pts[3]);
    else drawline(s, c, pts[0], pts[1], pts[2], pts[3]);
    return 1;
}

/* clip_and_draw_line_width */
static void clip_and_draw_line_width(surf *s, const uint8_t *c, int width, const int *pts)
{
    int xinc = 0, yinc = 0, newpts[4];
    if (abs(pts[0] - pts[2]) > abs(pts[1] - pts[3])) yinc = 1;
    else xinc = 1;
    memcpy(newpts, pts, sizeof(int) * 4);
    clip_and_draw_line(s, c, newpts);
    for (int loop = 1; loop < width; loop += 2) {
        newpts[0] = pts[0] + xinc * (loop / 2 + 1);
        newpts[1] = pts[1] + yinc * (loop / 2 + 1);
        newpts[2] = pts[2] + xinc * (loop / 2 + 1);
        newpts[3] = pts[3] + yinc * (loop / 2 + 1);
        clip_and_draw_line(s, c, newpts);
        if (loop + 1 < width) {
            newpts[0] = pts[0] - xinc * (loop / 2 + 1);
            newpts[1] = pts[1] - yinc * (loop / 2 + 1);
            newpts[2] = pts[2] - xinc * (loop / 2 + 1);
            newpts[3] = pts[3] - yinc * (loop / 2 + 1);
            clip_and_draw_line(s, c, newpts);
        }
    }
}

/* np.arange(start, stop, step) for doubles: length ceil((stop-start)/step),
 * values start + i * ((start + step) - start)  (numpy DOUBLE_fill) */
static int np_arange(double start, double stop, double step, double *out, int cap)
{
    double len_f = ceil((stop - start) / step);
    int len = len_f <= 0 ? 0 : (int)len_f;
    if (len > cap) len = cap;
    if (len > 0) out[0] = start;
    if (len > 1) out[1] = start + step;
    double delta = (start + step) - start;
    for (int i = 2; i < len; i++) out[i] = start + i * delta;
    return len;
}

/* Poly.draw_outline (render.py:229-287) with radius = 1 */
static void draw_outline(surf *s, double ax, double ay, double bx, double by, const uint8_t *c, int dashed)
{
    if (dashed) {
        enum { CAP = 256 };
        double xs[CAP], ys[CAP];
        int nx, ny;
        double x1 = ax, y1 = ay, x2 = bx, y2 = by;
        double dl = 10;
        if (x1 == x2) {
            ny = np_arange(y1, y2, y1 < y2 ? dl : -dl, ys, CAP);
            nx = ny;
            for (int i = 0; i < nx; i++) xs[i] = x1;
        } else if (y1 == y2) {
            nx = np_arange(x1, x2, x1 < x2 ? dl : -dl, xs, CAP);
            ny = nx;
            for (int i = 0; i < ny; i++) ys[i] = y1;
        } else {
            double a = fabs(x2 - x1), b = fabs(y2 - y1);
            double cc = rint(sqrt(a * a + b * b)); /* Python round(): half to even */
            double dx = dl * a / cc, dy = dl * b / cc;
            nx = np_arange(x1, x2, x1 < x2 ? dx : -dx, xs, CAP);
            ny = np_arange(y1, y2, y1 < y2 ? dy : -dy, ys, CAP);
        }
        int n = nx < ny ? nx : ny;
        for (int k = 0; 2 * k + 1 < n; k++) {
            int pts[4] = { (int)rint(xs[2 * k + 1]), (int)rint(ys[2 * k + 1]),
                           (int)rint(xs[2 * k]), (int)rint(ys[2 * k]) };
            clip_and_draw_line_width(s, c, 4, pts);
        }
    } else {
        /* r = round(max(1, 2)) = 2 -> pygame.draw.lines(..., [p1, p2], 2); the r > 2 branch is dead */
        int pts[4] = { (int)ax, (int)ay, (int)bx, (int)by }; /* TwoIntsFromObj truncates */
        clip_and_draw_line_width(s, c, 2, pts);
    }
}

/* ------------------------------------------------------------ Viewer.render */
void orc_render(orc_env *e, int view, uint8_t *out)
{
    const orc_geom_tables *g = orc_geom();
    surf s = { out };
    for (int i = 0; i < RES * RES; i++) memcpy(out + (size_t)i * 3, g->pal_bg, 3); /* Viewer._clear */
    const orc_body *robot = &e->bodies[e->robot_body0];
    aff cam = view == 0 ? allo_camera() : ego_camera(robot->p.x, robot->p.y, robot->a);
    static double px[ORC_CIRCLE_RES], py[ORC_CIRCLE_RES];
    int ix[ORC_CIRCLE_RES + 1], iy[ORC_CIRCLE_RES + 1];
    double fx[ORC_CIRCLE_RES], fy[ORC_CIRCLE_RES];
    (void)px; (void)py;
    for (int pi = 0; pi < e->n_prims; pi++) {
        const orc_prim *p = &e->prims[pi];
        const vec *lv = p->lv ? p->lv : p->lv_own;
        aff M = cam;
        for (int k = p->n_xf - 1; k >= 0; k--) { /* Geom.render: push transforms in reverse */
            aff X;
            if (p->xf[k].kind == 0) X = transform_make(p->xf[k].tx, p->xf[k].ty, 0.0, 1, 1);
            else if (p->xf[k].kind == 1) {
                const orc_body *b = &e->bodies[p->xf[k].body];
                X = transform_make(b->p.x, b->p.y, b->a, 1, 1);          /* pre_draw: entities.py:478-487, 751-754, 865-868 */
            } else {
                double rel = e->bodies[p->xf[k].body].a - e->bodies[p->xf[k].body_rel].a; /* entities.py:489-491 */
                X = transform_make(0, 0, rel, 1, 1);
            }
            M = aff_mul(M, X);
        }
        int n = p->n;
        for (int i = 0; i < n; i++) {
            fx[i] = (M.a * lv[i].x + M.b * lv[i].y) + M.tx;
            fy[i] = (M.c * lv[i].x + M.d * lv[i].y) + M.ty;
            ix[i] = (int)fx[i];
            iy[i] = (int)fy[i];
        }
        ix[n] = ix[0]; iy[n] = iy[0]; /* ps += [ps[0]] (render.py:213-214) */
        draw_fillpoly(&s, ix, iy, n + 1, p->fill);
        if (p->outline_mode)
            for (int i = 0; i < n; i++) {
                int j = (i + 1) % n;
                draw_outline(&s, fx[i], fy[i], fx[j], fy[j], p->outline, p->outline_mode == 2);
            }
    }
}

/* cv2.resize(x, (w/4, h/4), interpolation=cv2.INTER_AREA) at an exact integer
 * factor of 4: per channel, round-half-to-even of the 16-pixel mean
 * (benchmarks/__init__.py:175-190; rule pinned against cv2 in tests). */
void orc_downsample4(const uint8_t *in, int h, int w, int c, uint8_t *out)
{
    int oh = h / 4, ow = w / 4;
    for (int y = 0; y < oh; y++)
        for (int x = 0; x < ow; x++)
            for (int k = 0; k < c; k++) {
                int sum = 0;
                for (int dy = 0; dy < 4; dy++)
                    for (int dx = 0; dx < 4; dx++)
                        sum += in[((size_t)(4 * y + dy) * w + (4 * x + dx)) * c + k];
                int q = sum >> 4, r = sum & 15;
                if (r > 8 || (r == 8 && (q & 1))) q++;
                out[((size_t)y * ow + x) * c + k] = (uint8_t)q;
            }
}
