/*
 * mg_raster.cu -- fused raster -> 4x area downsample -> frame-stack kernel.
 *
 * Stands in for BaseEnv.render('rgb_array') (magical/base_env.py:324-343), Viewer.render
 * and Poly._render / draw_outline (magical/render.py:212-287, 385-395) on a pygame 1.9.6
 * software surface, ResizeDictObservation's cv2.resize(INTER_AREA) at exactly 4x
 * (magical/benchmarks/__init__.py:150-190) and FlattenFrameStack / EagerDictFrameStack
 * (magical/benchmarks/__init__.py:51-147).
 *
 * Work decomposition: one WARP renders one 32-row band of one (env, view) into a
 * palette-index buffer in shared memory (painter's algorithm: later primitives overwrite
 * earlier ones, exactly like drawing on the surface), then reduces the band to 8 rows of
 * the 96x96 frame with round-half-even of the 16-sample sum, stores them into the env's
 * frame ring with 16-byte stores and, when the observation is a stack of frames of this
 * view, also writes the band of the stacked uint8 HWC observation.  Bands never share
 * state, so there is no CTA-level synchronisation after the palette load.
 *
 * pygame rules restated (draw.c of pygame 1.9.6): polygon vertices truncated to int,
 * scanline fill with integer-division crossings, width-2 / width-4 lines as offset copies
 * of an integer DDA line after Cohen-Sutherland clipping with a float slope.
 */
#include <cuda_runtime.h>
#include <stddef.h>
#include <mutex>
#include "mg_state.cuh"
#include "mg_kernels.h"
#include "../../include/mg_trig.h"
#include "mg_envref.cuh"

#define RES MG_RES
#ifndef BAND_ROWS
#define BAND_ROWS 32
#endif
#define N_BANDS (RES / BAND_ROWS)
#define ROW_STRIDE 400                       /* 25 x 16 B: rows are 16-byte aligned and the 16-byte stores of a quarter
                                              * warp (8 lanes = 8 rows, 100 words apart) fall on distinct banks */
#ifndef WARPS_PER_BAND
#define WARPS_PER_BAND 2      /* warps (of one CTA) that paint one band together: the kernel is bound by the latency of a
                               * band's instruction stream at the 15 bands per SM that shared memory allows, so a second
                               * warp on the same band buffer doubles the resident warps and shortens every band */
#endif
#define NT (32 * WARPS_PER_BAND)              /* threads per band = per CTA */
#define LPR (NT / BAND_ROWS)                  /* neighbouring lanes that share one row in the polygon fills */
#define LO_ROWS (BAND_ROWS / 4)
#define LO_BAND_BYTES (LO_ROWS * MG_LORES * 3)   /* 2304 */
#define FULL 0xffffffffu
#ifndef RENDER_ROLLED
#define RENDER_ROLLED 1
#endif
#if RENDER_ROLLED
#define RNOUNROLL _Pragma("unroll 1")     /* rolled loops: the kernel is bound by instruction fetch, not by loop overhead */
#else
#define RNOUNROLL
#endif
/* per-loop unroll factors of the few loops whose bodies are short and latency-bound (everything else stays rolled) */
#define R_STR2(x) #x
#define R_STR(x) R_STR2(x)
#define R_UNROLL(n) _Pragma(R_STR(unroll n))
#ifndef R_UNROLL_DS
#define R_UNROLL_DS 3      /* downsample pass 1: quads per thread in flight.  One warp per band (six quads per lane, 96
                            * registers, 15 warps per SM): 1 was best, every unrolling cost more in code size than it gave.
                            * Two warps per band: a thread has three quads and all three in flight gain 4 % */
#endif
#ifndef R_UNROLL_BG
#define R_UNROLL_BG 1      /* background clear */
#endif
#ifndef R_UNROLL_SPAN
#define R_UNROLL_SPAN 1    /* 16-byte stores of a span fill */
#endif
#ifndef RENDER_MIN_CTAS
#define RENDER_MIN_CTAS (WARPS_PER_BAND == 1 ? 20 : 15)   /* register cap: 96 registers for one warp per band, 64 for two */
#endif
#ifndef RENDER_REREAD
#define RENDER_REREAD 0     /* INPLACE: 1 = the stack assembly reads the packed band back instead of keeping the thread's quads in
                             * registers (needed at six quads per lane; with three, the registers are there: 2 % faster) */
#endif
#ifndef RENDER_INPLACE
#define RENDER_INPLACE 1    /* fused LoRes4 kernel: downsampled band kept inside the band buffer (see the downsample) */
#endif
#ifndef RENDER_PREFETCH
#define RENDER_PREFETCH 0   /* 1 = fetch the next hit's vertices while the current one is painted: +4 % with one warp per band; with
                             * two, the partner warp covers that latency and the registers the prefetch holds cost 2 % */
#endif

struct aff { double a, b, tx, c, d, ty; };   /* 3x3 homogeneous matrix, last row 0 0 1 */

/* numpy `A @ B` on such matrices: every entry is (p*q + r*s) + t*u */
__device__ __forceinline__ aff aff_mul(aff A, aff B)
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
/* Transform(translation, rotation) with unit scale (render.py:87-97): T @ R @ S is exactly this */
__device__ __forceinline__ aff aff_tr(double tx, double ty, double cs, double sn)
{
    aff X = { cs, -sn, tx, sn, cs, ty };
    return X;
}
__device__ __forceinline__ aff aff_scale(double sx, double sy)
{
    aff X = { sx, 0.0, 0.0, 0.0, sy, 0.0 };
    return X;
}
/* Viewer.pygame_transform = T(0, H) @ S(1, -1) (render.py:329-330) */
__device__ __forceinline__ aff pygame_transform()
{
    return aff_mul(aff_tr(0.0, (double)RES, 1.0, 0.0), aff_scale(1.0, -1.0));
}
/* BaseEnv._use_allo_cam + Viewer.set_bounds (base_env.py:318-322, render.py:339-347) */
__device__ aff allo_camera()
{
    const double zoom = 1.02;   /* style.ARENA_ZOOM_OUT */
    double left = -1 * zoom, right = 1 * zoom, bottom = -1 * zoom, top = 1 * zoom;
    double sx = RES / (right - left), sy = RES / (top - bottom);
    aff cam = aff_mul(aff_mul(aff_tr(-left * sx, -bottom * sy, 1.0, 0.0), aff_tr(0.0, 0.0, 1.0, 0.0)), aff_scale(sx, sy));
    return aff_mul(pygame_transform(), cam);
}
/* BaseEnv._use_ego_cam + Viewer.set_cam_follow + ego_cam_matrix (base_env.py:309-316, render.py:290-304, 349-371) */
__device__ aff ego_camera(double cx, double cy, double cs, double sn)
{
    const double zoom = 1.02;
    double world_h = 2 * zoom, world_w = 2 * zoom;
    double sx = RES / world_w, sy = RES / world_h;
    aff scale = aff_scale(sx, sy);
    aff tr1 = aff_tr(world_w * 0.5, world_h * 0.15, 1.0, 0.0);
    aff rot = aff_tr(0.0, 0.0, cs, -sn);         /* rotation by -theta: cos(-t) = cos t, sin(-t) = -sin t */
    aff tr2 = aff_tr(-cx, -cy, 1.0, 0.0);
    aff m = aff_mul(rot, tr2);
    m = aff_mul(tr1, m);
    m = aff_mul(scale, m);
    return aff_mul(pygame_transform(), m);
}

struct WarpSmem {
    alignas(16) uint8_t buf[BAND_ROWS * ROW_STRIDE];     /* palette indices of the band */
    int vx[MG_CIRCLE_RES + 4], vy[MG_CIRCLE_RES + 4];
    uint32_t mq_more[128 * WARPS_PER_BAND > 2 * (MG_CIRCLE_RES + 4) ? 128 * WARPS_PER_BAND - 2 * (MG_CIRCLE_RES + 4) : 1];   /* vx + vy + this = the downsample's mixed-block queues */
    double fx[4], fy[4];                     /* float vertices (dashed outlines only) */
    int span_lo[BAND_ROWS], span_hi[BAND_ROWS], span_cnt[BAND_ROWS];   /* per-row crossings (edge-parallel fill) */
    alignas(16) uint8_t lo[LO_BAND_BYTES];   /* downsampled band, HWC.  Last: the fused LoRes4 kernel (VARIANT 1) keeps the
                                              * downsampled band inside `buf` and is launched without this array */
};

/* the threads of a band meet: a warp-level fence for one warp per band, the CTA barrier otherwise */
__device__ __forceinline__ void band_sync()
{
    if (WARPS_PER_BAND == 1) __syncwarp(); else __syncthreads();
}

/* ---------------------------------------------------------------- lines (pygame draw.c) */

/* clipline: Cohen-Sutherland against [0, RES-1]^2 with a float slope */
__device__ __forceinline__ int clip_encode(int x, int y)
{
    int code = 0;
    if (x < 0) code |= 1;
    if (x > RES - 1) code |= 2;
    if (y < 0) code |= 8;
    if (y > RES - 1) code |= 4;
    return code;
}
__device__ bool clipline(int &x1, int &y1, int &x2, int &y2)
{
    const int left = 0, right = RES - 1, top = 0, bottom = RES - 1;
    while (true) {
        int code1 = clip_encode(x1, y1), code2 = clip_encode(x2, y2);
        if (!(code1 | code2)) return true;
        if (code1 & code2) return false;
        if (!code1) {
            int t = x2; x2 = x1; x1 = t;
            t = y2; y2 = y1; y1 = t;
            t = code2; code2 = code1; code1 = t;
        }
        float m = 1.0f;
        if (x2 != x1) m = __fdiv_rn((float)(y2 - y1), (float)(x2 - x1));
        if (code1 & 1) { y1 += (int)__fmul_rn((float)(left - x1), m); x1 = left; }
        else if (code1 & 2) { y1 += (int)__fmul_rn((float)(right - x1), m); x1 = right; }
        else if (code1 & 4) { if (x2 != x1) x1 += (int)__fdiv_rn((float)(bottom - y1), m); y1 = bottom; }
        else if (code1 & 8) { if (x2 != x1) x1 += (int)__fdiv_rn((float)(top - y1), m); y1 = top; }
    }
}

/* One clipped line = pixels k = 0..len-1 of pygame's integer DDA (drawline / drawhorzline /
 * drawvertline produce the same set): the major axis advances by one per pixel, the minor
 * axis by floor(k * minor_len / major_len). */
struct Line { int x1, y1, sx, sy, dxp, dyp, len; bool xmajor; float rmaj; };

/* 1 / x to within 1 ulp in one MUFU instruction; every use below corrects its quotient (div_small) or keeps a margin */
__device__ __forceinline__ float rcp_fast(float x)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
/* floor(a / b) for 0 <= a < 2^24, 0 < b, a / b < 2^20: one float multiply and a +-1 correction instead of an integer
 * divide (rb = 1 / b to within a few ulp: the estimate is off by less than one) */
__device__ __forceinline__ int div_small(int a, int b, float rb)
{
    int q = (int)((float)a * rb);
    int r = a - q * b;
    if (r < 0) q--; else if (r >= b) q++;
    return q;
}
/* C integer division (truncation toward zero) of num by den > 0 */
__device__ __forceinline__ int div_trunc(int num, int den)
{
    float rb = rcp_fast((float)den);
    return num >= 0 ? div_small(num, den, rb) : -div_small(-num, den, rb);
}

__device__ __forceinline__ bool line_setup(int x1, int y1, int x2, int y2, Line &L)
{
    if (!clipline(x1, y1, x2, y2)) return false;
    if (y1 == y2 && x1 > x2) { int t = x1; x1 = x2; x2 = t; }   /* drawhorzline / drawvertline order: same pixels */
    int dx = x2 - x1, dy = y2 - y1;
    L.sx = dx < 0 ? -1 : 1; L.sy = dy < 0 ? -1 : 1;
    L.dxp = L.sx * dx + 1; L.dyp = L.sy * dy + 1;
    L.x1 = x1; L.y1 = y1;
    L.xmajor = !(L.dxp < L.dyp);
    L.len = L.xmajor ? L.dxp : L.dyp;
    L.rmaj = rcp_fast((float)L.len);
    return true;
}
__device__ __forceinline__ void line_pixel(const Line &L, int k, int &x, int &y)
{
    if (L.xmajor) { x = L.x1 + L.sx * k; y = L.dyp == 1 ? L.y1 : L.y1 + L.sy * div_small(k * L.dyp, L.dxp, L.rmaj); }
    else { y = L.y1 + L.sy * k; x = L.dxp == 1 ? L.x1 : L.x1 + L.sx * div_small(k * L.dxp, L.dyp, L.rmaj); }
}
__device__ __forceinline__ void put_px(uint8_t *buf, int y0, int x, int y, uint8_t c)
{
    unsigned r = (unsigned)(y - y0);
    if (r < BAND_ROWS) buf[r * ROW_STRIDE + x] = c;
}

/* offsets of clip_and_draw_line_width: width 2 -> 0,+1; width 4 -> 0,+1,-1,+2 */
__device__ __forceinline__ int width_offset(int copy) { return copy == 0 ? 0 : (copy == 1 ? 1 : (copy == 2 ? -1 : 2)); }

/* One batch of up to 32 line jobs, lane = job (has == false: no job).  Every lane clips and sets up its own line;
 * lines longer than LONG_LINE pixels are then handed to the whole warp one after the other (pixels spread over the
 * lanes), the short ones are drawn by their own lanes side by side.  All copies of one outline carry the same
 * colour, so the order in which their pixels land does not matter.  The only copy of the line code in the kernel
 * (the kernel is bound by instruction fetch).
 * dup: the job also draws the line's second copy of a width-2 outline, the same pixels one to the right (dup == 1) or
 * one down (dup == 2).  Only for lines that clipline leaves alone in both positions (the caller checks the end points),
 * where the two copies are the same DDA run: one set-up for both. */
#ifndef LONG_LINE
#define LONG_LINE 12
#endif
__device__ __noinline__ void draw_line_jobs(uint8_t *buf, int y0, int lane, bool has, int x1, int y1, int x2, int y2, uint8_t c, int dup)
{
    Line L = { 0, 0, 1, 1, 1, 1, 0, true, 1.0f };      /* len 0: nothing to draw */
    const int ey = dup == 2 ? 1 : 0;                   /* the copy below reaches one row further */
    if (has && !(max(y1, y2) + ey < y0 || min(y1, y2) >= y0 + BAND_ROWS)) line_setup(x1, y1, x2, y2, L);   /* sets len only if the clip keeps the line */
    /* pixels k0 .. k1 of the line can fall into this band: a superset (put_px still tests the row), so the picture
     * is the same, but a long steep line costs a band only its own rows instead of its whole length */
    int k0 = 0, k1 = L.len - 1;
    if (L.len > 0) {
        /* minor-axis (x-major) or pixel (y-major) index range whose row lies in [y0 - ey, y0 + BAND_ROWS) */
        int ta = (y0 - ey - L.y1) * L.sy, tb = (y0 + BAND_ROWS - 1 - L.y1) * L.sy;
        if (ta > tb) { int t = ta; ta = tb; tb = t; }
        if (!L.xmajor) { k0 = max(k0, ta); k1 = min(k1, tb); }
        else if (L.dyp > 1) {
            const float r = (float)L.dxp * rcp_fast((float)L.dyp);
            k0 = max(k0, (int)((float)max(ta, 0) * r) - 2);
            k1 = min(k1, (int)((float)(min(tb, L.dyp) + 1) * r) + 2);
        }
    }
    const int ox = dup == 1 ? 1 : 0;
    const int cnt = k1 - k0 + 1;
    unsigned longs = __ballot_sync(FULL, cnt > LONG_LINE);
    const int packed = (L.sx & 3) | (L.sy & 3) << 2 | (int)L.xmajor << 4 | dup << 5;
    while (longs) {
        const int src = __ffs(longs) - 1;
        longs &= longs - 1;
        Line M;
        M.x1 = __shfl_sync(FULL, L.x1, src); M.y1 = __shfl_sync(FULL, L.y1, src);
        M.dxp = __shfl_sync(FULL, L.dxp, src); M.dyp = __shfl_sync(FULL, L.dyp, src);
        M.rmaj = __shfl_sync(FULL, L.rmaj, src);
        const int m0 = __shfl_sync(FULL, k0, src), m1 = __shfl_sync(FULL, k1, src);
        const int pk = __shfl_sync(FULL, packed, src);
        M.sx = (pk & 3) == 1 ? 1 : -1; M.sy = ((pk >> 2) & 3) == 1 ? 1 : -1; M.xmajor = (pk >> 4) & 1;
        const int mdup = pk >> 5;
        RNOUNROLL
        for (int k = m0 + lane; k <= m1; k += 32) {
            int x, y;
            line_pixel(M, k, x, y);
            put_px(buf, y0, x, y, c);
            if (mdup) put_px(buf, y0, x + (mdup == 1), y + (mdup == 2), c);
        }
    }
    if (cnt <= LONG_LINE && cnt > 0) {
        /* the lane walks its own short line: floor(k * minor / major) carried from pixel to pixel (minor <= major, so
         * at most one carry per step) instead of one division per pixel; same pixels as line_pixel */
        const int minor = L.xmajor ? L.dyp : L.dxp;
        const int q0 = k0 == 0 ? 0 : div_small(k0 * minor, L.len, L.rmaj);
        int rem = k0 * minor - q0 * L.len;
        int x = L.x1 + L.sx * (L.xmajor ? k0 : q0), y = L.y1 + L.sy * (L.xmajor ? q0 : k0);
        const int xk = L.xmajor ? L.sx : 0, yk = L.xmajor ? 0 : L.sy, xq = L.xmajor ? 0 : L.sx, yq = L.xmajor ? L.sy : 0;
        RNOUNROLL
        for (int k = k0; k <= k1; k++) {
            put_px(buf, y0, x, y, c);
            if (dup) put_px(buf, y0, x + ox, y + ey, c);
            x += xk; y += yk; rem += minor;
            if (rem >= L.len) { rem -= L.len; x += xq; y += yq; }
        }
    }
}

/* np.arange(start, stop, step)[i] for doubles (numpy fills start, start+step, then start + i*delta) */
__device__ __forceinline__ int arange_len(double start, double stop, double step)
{
    double len_f = ceil((stop - start) / step);
    return len_f <= 0 ? 0 : (int)len_f;
}
__device__ __forceinline__ double arange_at(double start, double step, int i)
{
    if (i == 0) return start;
    if (i == 1) return start + step;
    double delta = (start + step) - start;
    return start + i * delta;
}

/* ---------------------------------------------------------------- one primitive (render.Poly) */

/* Poly.draw_outline(dashed=True) for one edge (render.py:232-255): dash d joins arange point 2d+1 to 2d */
__device__ void draw_dashed_edge(uint8_t *buf, int y0, int tid, double x1, double y1, double x2, double y2, uint8_t c)
{
    const double dl = 10;
    int n;
    double sx_ = 0, sy_ = 0;   /* steps; 0 = constant coordinate */
    if (x1 == x2) {
        sy_ = y1 < y2 ? dl : -dl;
        n = arange_len(y1, y2, sy_);
    } else if (y1 == y2) {
        sx_ = x1 < x2 ? dl : -dl;
        n = arange_len(x1, x2, sx_);
    } else {
        double a = fabs(x2 - x1), b = fabs(y2 - y1);
        double cc = rint(sqrt(a * a + b * b));
        double dx = dl * a / cc, dy = dl * b / cc;
        sx_ = x1 < x2 ? dx : -dx;
        sy_ = y1 < y2 ? dy : -dy;
        int nx = arange_len(x1, x2, sx_), ny = arange_len(y1, y2, sy_);
        n = nx < ny ? nx : ny;
    }
    int n_dash = n / 2;
    /* thread = (dash, copy) */
    const int lane = tid & 31;
    RNOUNROLL
    for (int job0 = 0; job0 < n_dash * 4; job0 += NT) {
        const int job = job0 + tid;
        const bool has = job < n_dash * 4;
        int d = job >> 2, copy = job & 3;
        double xa = sx_ != 0 ? arange_at(x1, sx_, 2 * d + 1) : x1, ya = sy_ != 0 ? arange_at(y1, sy_, 2 * d + 1) : y1;
        double xb = sx_ != 0 ? arange_at(x1, sx_, 2 * d) : x1, yb = sy_ != 0 ? arange_at(y1, sy_, 2 * d) : y1;
        int px1 = (int)rint(xa), py1 = (int)rint(ya), px2 = (int)rint(xb), py2 = (int)rint(yb);
        int off = width_offset(copy);
        if (abs(px1 - px2) > abs(py1 - py2)) { py1 += off; py2 += off; } else { px1 += off; px2 += off; }
        draw_line_jobs(buf, y0, lane, has, px1, py1, px2, py2, c, 0);
    }
}

/* drawhorzlineclip: inclusive span [lo, hi] of one row, clipped to the surface; word stores inside.
 * The LPR lanes that share a row (part = 0 .. LPR-1) split the span at word boundaries. */
__device__ __forceinline__ void fill_span(uint8_t *row, int lo, int hi, uint8_t c, int part)
{
    if (hi < 0 || lo > RES - 1) return;
    lo = max(lo, 0); hi = min(hi, RES - 1);
    if (LPR > 1) {
        int chunk = (((hi - lo + 1) + LPR - 1) / LPR + 3) & ~3;
        int a = lo + part * chunk;
        if (part > 0) a &= ~3;
        int b = (lo + (part + 1) * chunk) & ~3;
        lo = a; hi = part == LPR - 1 ? hi : min(hi, b - 1);
        if (lo > hi) return;
    }
    /* the lane owns its row (or its part of it), so the words at the two ends are read, merged under a byte mask and
     * written back; whole words in between, 16 bytes at a time where aligned */
    const uint32_t c4 = c * 0x01010101u;
    uint32_t *wrow = (uint32_t *)row;
    const int a = lo >> 2, b = hi >> 2;
    const uint32_t ma = 0xffffffffu << (8 * (lo & 3)), mb = 0xffffffffu >> (8 * (3 - (hi & 3)));
    if (a == b) {
        const uint32_t m = ma & mb;
        wrow[a] = (wrow[a] & ~m) | (c4 & m);
        return;
    }
    wrow[a] = (wrow[a] & ~ma) | (c4 & ma);
    wrow[b] = (wrow[b] & ~mb) | (c4 & mb);
    int w = a + 1;
    RNOUNROLL
    while (w < b && (w & 3)) wrow[w++] = c4;
    const uint4 c16 = make_uint4(c4, c4, c4, c4);
    R_UNROLL(R_UNROLL_SPAN)
    for (; w + 3 < b; w += 4) *(uint4 *)(wrow + w) = c16;
    RNOUNROLL
    for (; w < b; w++) wrow[w] = c4;
}

/* same fill for polygons with many edges (the 100-gon circles): threads take edges, every edge adds its
 * crossings to the rows of the band it spans; then the row's lanes fill it.  Same spans as the row-parallel form. */
__device__ __forceinline__ void fill_poly_band_edges(WarpSmem *W, int y0, int tid, int row, int n, int miny, int maxy, int &lo, int &hi, int &cnt)
{
    if (tid < BAND_ROWS) { W->span_lo[tid] = 0x7fffffff; W->span_hi[tid] = -0x7fffffff - 1; W->span_cnt[tid] = 0; }
    band_sync();
    if (miny == maxy) {
        /* polygon one pixel high: span = x extent */
        const unsigned r = (unsigned)(miny - y0);
        if (r < BAND_ROWS) {
            RNOUNROLL
            for (int i = tid; i < n; i += NT) { atomicMin(&W->span_lo[r], W->vx[i]); atomicMax(&W->span_hi[r], W->vx[i]); }
            if (tid == 0) W->span_cnt[r] = 2;
        }
    } else {
        RNOUNROLL
        for (int i = tid; i < n; i += NT) {
            int xa = W->vx[i], ya = W->vy[i], xb = W->vx[i + 1], yb = W->vy[i + 1];
            int x1, y1, x2, y2;
            if (ya < yb) { x1 = xa; y1 = ya; x2 = xb; y2 = yb; }
            else if (ya > yb) { x1 = xb; y1 = yb; x2 = xa; y2 = ya; }
            else continue;
            /* rows with (y >= y1 && y < y2) || (y == maxy && y > y1 && y <= y2) */
            int ylast = y2 == maxy ? y2 : y2 - 1;
            int ya_ = max(y1, y0), yb_ = min(ylast, y0 + BAND_ROWS - 1);
            float rb = rcp_fast((float)(y2 - y1));
            RNOUNROLL
            for (int y = ya_; y <= yb_; y++) {
                int num = (y - y1) * (x2 - x1);
                int x = (num >= 0 ? div_small(num, y2 - y1, rb) : -div_small(-num, y2 - y1, rb)) + x1;
                atomicMin(&W->span_lo[y - y0], x);
                atomicMax(&W->span_hi[y - y0], x);
                atomicAdd(&W->span_cnt[y - y0], 1);
            }
        }
    }
    band_sync();
    lo = W->span_lo[row]; hi = W->span_hi[row]; cnt = W->span_cnt[row];
}

/* pygame.draw.polygon fill of rows [y0, y0 + BAND_ROWS): LPR neighbouring lanes own row y0 + tid / LPR; each takes every
 * LPR-th edge, a butterfly joins their crossings, then they split the span.  n is uniform over the band's threads. */
__device__ void fill_poly_band(WarpSmem *W, int y0, int tid, int n, int miny, int maxy, uint8_t c)
{
    const int row = tid / LPR, part = tid % LPR;
    int y = y0 + row;
    int lo = 0x7fffffff, hi = -0x7fffffff - 1, cnt = 0;
    if (n > 8) fill_poly_band_edges(W, y0, tid, row, n, miny, maxy, lo, hi, cnt);
    else {
        if (y < miny || y > maxy) cnt = 0;
        else if (miny == maxy) {
            /* polygon one pixel high: span = x extent */
            RNOUNROLL
            for (int i = 0; i < n; i++) { lo = min(lo, W->vx[i]); hi = max(hi, W->vx[i]); }
            cnt = 2;
        } else {
            RNOUNROLL
            for (int i = part; i < n; i += LPR) {
                const int xa = W->vx[i], ya = W->vy[i], xb = W->vx[i + 1], yb = W->vy[i + 1];
                int x1, y1, x2, y2;
                if (ya < yb) { x1 = xa; y1 = ya; x2 = xb; y2 = yb; }
                else if (ya > yb) { x1 = xb; y1 = yb; x2 = xa; y2 = ya; }
                else continue;
                if ((y >= y1 && y < y2) || (y == maxy && y > y1 && y <= y2)) {
                    int x = div_trunc((y - y1) * (x2 - x1), y2 - y1) + x1;
                    lo = min(lo, x); hi = max(hi, x); cnt++;
                }
            }
        }
#pragma unroll
        for (int d = 1; d < LPR; d <<= 1) {
            lo = min(lo, __shfl_xor_sync(FULL, lo, d)); hi = max(hi, __shfl_xor_sync(FULL, hi, d));
            cnt += __shfl_xor_sync(FULL, cnt, d);
        }
    }
    if (cnt < 2) return;
    fill_span(W->buf + row * ROW_STRIDE, lo, hi, c, part);
}

__device__ __forceinline__ uint32_t pal_sum16(const uint32_t *pal, uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3)
{
    /* 16 palette ids -> per-channel sums (r and b share one register) -> round-half-even(sum / 16),
     * packed r | g<<8 | b<<16 */
    uint32_t rb = 0, gs = 0;
    uint32_t ws[4] = { w0, w1, w2, w3 };
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) {
            uint32_t p = pal[(ws[i] >> (8 * j)) & 31];
            rb += p & 0x00ff00ffu; gs += (p >> 8) & 255u;
        }
    uint32_t out = 0;
    uint32_t s[3] = { rb & 0xffffu, gs, rb >> 16 };
#pragma unroll
    for (int k = 0; k < 3; k++) {
        uint32_t q = s[k] >> 4, r = s[k] & 15;
        if (r > 8 || (r == 8 && (q & 1))) q++;
        out |= q << (8 * k);
    }
    return out;
}

/* ---- bulk asynchronous copies (sm_90+ / sm_100a copy engine): a finished 96 x 8 band (2304 contiguous bytes in shared
 * memory, 2304 contiguous bytes in HBM) goes out as ONE cp.async.bulk instead of a per-lane loop of 16-byte stores.
 * The issuing lane fences the warp's generic-proxy writes for the async proxy, commits the group, and waits for it
 * before the warp exits. */
__device__ __forceinline__ void bulk_store_fence()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void bulk_store(void *dst_gmem, const void *src_smem, unsigned bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 :: "l"(dst_gmem), "r"((unsigned)__cvta_generic_to_shared(src_smem)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_store_wait() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

struct RenderArgs {
    int view;
    int mode;            /* 0: push into the frame ring (+ optional fused stack); 1: lo-res out; 2: full-res out */
    uint8_t *out;        /* mode 1: [B,96,96,3]; mode 2: [B,384,384,3] */
    int slot;            /* ring slot that receives the new frame (mode 0) */
    /* fused stack: obs[B,96,96,3*n_groups]; group g = frame of view gview[g] in ring slot gslot[g] */
    uint8_t *obs;
    int n_groups;
    int gview[MG_MAX_GROUPS], gslot[MG_MAX_GROUPS], gage[MG_MAX_GROUPS];
    int env_begin, env_end;   /* envs [env_begin, env_end) of the batch */
};

/* ---------------------------------------------------------------- geometry pass
 * One warp per env: builds each primitive's transform chain (Viewer.render -> Geom.render ->
 * stack.apply_current_matrix, render.py:116-160, 373-392), transforms the vertices in FP64 and truncates them the way
 * pygame converts its float arguments to int.  The band rasterisers then only see integers.
 * The kernel sits between an env's physics and its raster on the step's critical path, so it is laid out for latency:
 * lane = primitive for the transform chains (32 chains side by side instead of one after the other), then lane = vertex
 * over the chunk's vertices in one flat loop (the owning primitive found by a search of the chunk's prefix sums), row and
 * column extents by shared-memory atomics.  Every vertex is the same arithmetic as in a serial walk of the draw list. */
#define GEO_WARPS 4
struct GeoChunk {                       /* one chunk of 32 primitives of one env */
    aff M[32];
    double gw[32], gh[32];              /* goal regions: width and height (their vertices are made here) */
    int pre[33];                        /* exclusive prefix sums of the vertex counts */
    int ext[32][4];                     /* miny, maxy, minx, maxx */
    MgPrim P[32];
};
__global__ void __launch_bounds__(GEO_WARPS * 32) mg_geometry_kernel(MgDeviceState S, int view, int env_begin, int env_end, int apply_pending)
{
    __shared__ GeoChunk chunks[GEO_WARPS];
    const MgTables *T = S.tables;
    const int lane = threadIdx.x & 31;
    GeoChunk *C = &chunks[threadIdx.x >> 5];
    const int slot = env_begin + blockIdx.x * GEO_WARPS + (threadIdx.x >> 5);
    if (slot >= env_end) return;
    const int env = S.perm ? S.perm[slot] : slot;
    /* a reset queued for this env (mg_queue_resets): this step's physics is done with the old episode (stream order),
     * so the env is re-initialised here, right before the first frame of the new episode is drawn */
    if (apply_pending && S.pending) {
        const int p = S.pending[env];
        if (p >= 0) {
            if (lane == 0) { init_env(S, S.pending_scenes + p, env); S.pending[env] = -1; }
            __syncwarp();
        }
    }
    const double *pose = S.poses + (size_t)env * MG_POSE_STRIDE;
    const double *rp = pose + MG_BODY_ROBOT * MG_POSE_NF;
    const aff cam = view == MG_VIEW_ALLO ? allo_camera() : ego_camera(rp[0], rp[1], rp[6], rp[7]);
    const aff cam_robot = aff_mul(cam, aff_tr(rp[0], rp[1], rp[6], rp[7]));
    const MgPrim *prims = S.prims + (size_t)env * MG_MAX_PRIMS;
    MgScreenPrim *sprims = S.sprims[view] + (size_t)env * MG_MAX_PRIMS;
    int2 *sverts = S.sverts[view] + (size_t)env * MG_MAX_VERTS;
    double *dashv = S.dashv[view] + (size_t)env * MG_MAX_GOALS * 8;
    const int np = S.n_prims[env];
    int off = 0;                        /* first vertex of the chunk in the env's pool: every primitive keeps its place,
                                         * visible or not (the pool holds the largest scene's whole draw list) */
    for (int base = 0; base < np; base += 32) {
        const int pi = base + lane;
        /* lane = primitive: its transform chain */
        int n = 0;
        MgPrim P = { 0, 0, 0, 0, 0, 0, 0, 0 };
        if (pi < np) {
            P = prims[pi];
            n = T->dl_n[P.dl];
            aff M;
            double gw = 0, gh = 0;
            switch (P.chain) {
            case MG_CHAIN_CAM: M = cam; break;
            case MG_CHAIN_BODY: {
                const double *b = pose + P.body * MG_POSE_NF;
                M = P.body == MG_BODY_ROBOT ? cam_robot : aff_mul(cam, aff_tr(b[0], b[1], b[6], b[7]));
                break;
            }
            case MG_CHAIN_EYE:
            case MG_CHAIN_PUPIL: {
                const double radius = 0.2;
                double x_sign = P.aux == 0 ? -1.0 : 1.0;
                M = aff_mul(cam_robot, aff_tr(x_sign * 0.4 * radius, 0.3 * radius, 1.0, 0.0));
                if (P.chain == MG_CHAIN_PUPIL) {
                    double rel = pose[P.body * MG_POSE_NF + 2] - rp[2];   /* pupil.angle - robot.angle, entities.py:489-491 */
                    double sn, cs;
                    mg_sincos(rel, &sn, &cs);
                    M = aff_mul(M, aff_tr(0.0, 0.0, cs, sn));
                    M = aff_mul(M, aff_tr(0.0, radius * 0.07, 1.0, 0.0));
                }
                break;
            }
            default: { /* MG_CHAIN_GOAL */
                const mg_goal *g = &S.scenes[env].goals[P.aux];
                M = aff_mul(cam, aff_tr(g->cx, g->cy, 1.0, 0.0));
                gw = g->w; gh = g->h;
                break;
            }
            }
            C->M[lane] = M; C->gw[lane] = gw; C->gh[lane] = gh; C->P[lane] = P;
        }
        /* the chunk's vertex ranges */
        int incl = n;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(FULL, incl, d); if (lane >= d) incl += o; }
        int total = __shfl_sync(FULL, incl, 31);
        if (off + total > MG_MAX_VERTS) { n = 0; incl = 0; total = 0; }      /* cannot happen for scenes mg_set_scenes accepts */
        C->pre[lane] = incl - n;
        if (lane == 31) C->pre[32] = total;
        C->ext[lane][0] = 0x7fffffff; C->ext[lane][1] = -0x7fffffff - 1; C->ext[lane][2] = 0x7fffffff; C->ext[lane][3] = -0x7fffffff - 1;
        __syncwarp();
        /* lane = vertex over the whole chunk: transform + truncate (stack.apply_current_matrix, then pygame's int
         * conversion) */
        {
            /* a lane's vertices v, v + 32, ... stay with one primitive for a while (circles have 100): its matrix and
             * style stay in registers and its extents are merged into shared memory only when the lane moves on */
            int q = -1, q_end = 0, q_pre = 0;
            aff M = cam;
            MgPrim Pq = { 0, 0, 0, 0, 0, 0, 0, 0 };
            const vec2 *lv = nullptr;
            double rw = 0, rh = 0;
            int y_lo = 0x7fffffff, y_hi = -0x7fffffff - 1, x_lo = 0x7fffffff, x_hi = -0x7fffffff - 1;
            for (int v = lane; v < total; v += 32) {
                if (v >= q_end) {
                    if (q >= 0) {
                        atomicMin(&C->ext[q][0], y_lo); atomicMax(&C->ext[q][1], y_hi);
                        atomicMin(&C->ext[q][2], x_lo); atomicMax(&C->ext[q][3], x_hi);
                    }
                    q = 0;              /* last primitive whose range starts at or before v */
#pragma unroll
                    for (int d = 16; d > 0; d >>= 1) if (C->pre[q + d] <= v) q += d;
                    q_pre = C->pre[q]; q_end = C->pre[q + 1];
                    Pq = C->P[q]; M = C->M[q];
                    lv = T->dl_pool + T->dl_off[Pq.dl];
                    rw = C->gw[q] / 2; rh = C->gh[q] / 2;
                    y_lo = x_lo = 0x7fffffff; y_hi = x_hi = -0x7fffffff - 1;
                }
                const int i = v - q_pre;
                double x, y;
                if (Pq.chain == MG_CHAIN_GOAL) {
                    x = (i == 0 || i == 3) ? -rw : rw;
                    y = (i < 2) ? rh : -rh;
                } else { x = lv[i].x; y = lv[i].y; }
                const double fx = (M.a * x + M.b * y) + M.tx;
                const double fy = (M.c * x + M.d * y) + M.ty;
                const int ix = (int)fx, iy = (int)fy;
                sverts[off + v] = make_int2(ix, iy);
                if (Pq.mode == 2 && i < 4) { dashv[Pq.aux * 8 + 2 * i] = fx; dashv[Pq.aux * 8 + 2 * i + 1] = fy; }
                y_lo = min(y_lo, iy); y_hi = max(y_hi, iy); x_lo = min(x_lo, ix); x_hi = max(x_hi, ix);
            }
            if (q >= 0) {
                atomicMin(&C->ext[q][0], y_lo); atomicMax(&C->ext[q][1], y_hi);
                atomicMin(&C->ext[q][2], x_lo); atomicMax(&C->ext[q][3], x_hi);
            }
        }
        __syncwarp();
        /* lane = primitive again: the header the bands read */
        if (pi < np) {
            const int miny = C->ext[lane][0], maxy = C->ext[lane][1], minx = C->ext[lane][2], maxx = C->ext[lane][3];
            /* nothing of a primitive whose vertices all lie more than 4 pixels outside the surface can reach it
             * (fills are clipped per row, outline copies are shifted by at most 2 and clipped) */
            const bool visible = n > 0 && maxy >= -4 && miny <= RES + 3 && maxx >= -4 && minx <= RES + 3;
            MgScreenPrim sp;
            sp.off_n = visible ? ((off + C->pre[lane]) << 8 | n) : 0;
            sp.miny = miny; sp.maxy = maxy;
            sp.style = (uint32_t)P.fill | (uint32_t)P.outline << 8 | (uint32_t)P.mode << 16 | (uint32_t)P.aux << 24;
            sprims[pi] = sp;
        }
        off += total;
        __syncwarp();                   /* the chunk's scratch is reused */
    }
}

/* VARIANT selects what happens to the finished band, so that each instantiation carries only its own tail
 * (the kernel is instruction-cache sensitive): 0 push into the ring; 1 push + fused 4-group HWC stack;
 * 2 push + fused generic HWC stack; 3 lo-res frame out; 4 full-res frame out */
template <int VARIANT>
__global__ void __launch_bounds__(NT, RENDER_MIN_CTAS) mg_render_kernel(MgDeviceState S, RenderArgs A)
{
    constexpr int MODE = VARIANT <= 2 ? 0 : VARIANT - 2;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    __shared__ uint32_t pal[32];
    const MgTables *T = S.tables;
    /* r | g << 8 | b << 16 (MgTables::pal rows are r g b 0); first read after the barrier that follows the background fill */
    static_assert(offsetof(MgTables, pal) % 4 == 0, "palette rows load as words");
    if (threadIdx.x < 32) pal[threadIdx.x] = ((const uint32_t *)T->pal)[threadIdx.x];
    /* one band per CTA; its NT threads: tid spreads data-parallel work, lane / wid matter where a warp acts as one */
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    int env = A.env_begin + (int)(blockIdx.x / N_BANDS);
    const int band = (int)(blockIdx.x % N_BANDS);
    if (env >= A.env_end) return;
    if (S.perm) env = S.perm[env];
    WarpSmem *W = (WarpSmem *)smem_raw;
    const int y0 = band * BAND_ROWS;

    /* the fused stack reads three older frames of this band from the ring at the very end: start pulling
     * their lines towards L2 now (18 lines of 128 B per frame band) */
    if (VARIANT == 2 && tid < LO_BAND_BYTES / 128) {
#pragma unroll
        for (int g = 0; g < MG_MAX_GROUPS; g++) {
            if (g >= A.n_groups || (A.gview[g] == A.view && A.gage[g] == 0)) continue;
            const uint8_t *p = S.ring[A.gview[g]] + ((size_t)env * MG_MAX_RING + A.gslot[g]) * MG_FRAME_BYTES
                               + (size_t)band * LO_BAND_BYTES + tid * 128;
            asm volatile("prefetch.global.L2 [%0];" :: "l"(p));
        }
    }
    /* primitives in draw order, 32 headers per coalesced load; each lane tests its own primitive against
     * the band and the band's threads then walk the hits in order (painter's algorithm).  Every warp of the band holds
     * the same headers, so the walk is uniform over the CTA without an exchange. */
    const MgScreenPrim *sprims = S.sprims[A.view] + (size_t)env * MG_MAX_PRIMS;
    const int2 *sverts = S.sverts[A.view] + (size_t)env * MG_MAX_VERTS;
    const int np = S.n_prims[env];
    /* needed after the painting (does the env start an episode with this frame?): asked for now, so that the answer has
     * long arrived by then */
    int fresh_word = 0;
    if (MODE == 0) fresh_word = S.istate[mg_field_index(env, I_NF, I_FRESH)];
    /* the first 32 headers are on their way while the background is filled */
    const MgScreenPrim no_prim = { 0, 0x7fffffff, -0x7fffffff - 1, 0 };
    MgScreenPrim first = sprims[lane];            /* MG_MAX_PRIMS >= 32 entries exist; the load does not wait for np */
    if (lane >= np) first = no_prim;
    /* Viewer._clear: background */
    {
        const uint32_t bg4 = MG_PAL_BG * 0x01010101u;
        const uint4 bg16 = make_uint4(bg4, bg4, bg4, bg4);
        R_UNROLL(R_UNROLL_BG)
        for (int i = tid; i < BAND_ROWS * ROW_STRIDE / 16; i += NT) ((uint4 *)W->buf)[i] = bg16;
    }
    band_sync();

    constexpr int VERT_IT = (MG_CIRCLE_RES + 1 + NT - 1) / NT;      /* vertex slots per thread: n + 1 entries, n <= MG_CIRCLE_RES */
    for (int base = 0; base < np; base += 32) {
        MgScreenPrim mine = base == 0 ? first : no_prim;
        if (base > 0 && base + lane < np) mine = sprims[base + lane];
        /* outlines reach at most 2 pixels beyond the vertex rows */
        unsigned hits = __ballot_sync(FULL, (mine.off_n & 255) != 0 && mine.maxy + 2 >= y0 && mine.miny - 2 < y0 + BAND_ROWS);
        /* vertices of the next hit are fetched while the current one is painted (HBM/L2 latency off the serial path) */
        int2 pre_v[VERT_IT];
#define LOAD_VERTS(SRC)                                                                        \
        {                                                                                      \
            const int on_ = __shfl_sync(FULL, mine.off_n, (SRC));                              \
            const int nn_ = on_ & 255;                                                         \
            const int2 *pv_ = sverts + (on_ >> 8);                                             \
            _Pragma("unroll") for (int k = 0; k < VERT_IT; k++) {                              \
                const int i = tid + NT * k;                                                    \
                if (i <= nn_) pre_v[k] = pv_[i == nn_ ? 0 : i];                                \
            }                                                                                  \
        }
#if RENDER_PREFETCH
        if (hits) LOAD_VERTS(__ffs(hits) - 1)
#endif
      while (hits) {
        const int src = __ffs(hits) - 1;
        hits &= hits - 1;
        const int off_n = __shfl_sync(FULL, mine.off_n, src);
        const int miny = __shfl_sync(FULL, mine.miny, src), maxy = __shfl_sync(FULL, mine.maxy, src);
        const uint32_t style = __shfl_sync(FULL, mine.style, src);
        const int n = off_n & 255;
        struct { uint8_t fill, outline, mode, aux; } P = { (uint8_t)style, (uint8_t)(style >> 8), (uint8_t)(style >> 16), (uint8_t)(style >> 24) };
#if !RENDER_PREFETCH
        LOAD_VERTS(src)
#endif
#pragma unroll
        for (int k = 0; k < VERT_IT; k++) {
            const int i = tid + NT * k;
            if (i <= n) { W->vx[i] = pre_v[k].x; W->vy[i] = pre_v[k].y; }     /* entry n closes the loop */
        }
        if (P.mode == 2 && tid < 8) {
            double d = S.dashv[A.view][((size_t)env * MG_MAX_GOALS + P.aux) * 8 + tid];
            if (tid & 1) W->fy[tid >> 1] = d; else W->fx[tid >> 1] = d;
        }
        band_sync();
#if RENDER_PREFETCH
        if (hits) LOAD_VERTS(__ffs(hits) - 1)
#endif
        {
            fill_poly_band(W, y0, tid, n, miny, maxy, P.fill);
            band_sync();
            if (P.mode == 1) {
                /* solid: pygame.draw.lines(width=2) per edge = base line + copy shifted by +1 on the minor axis.
                 * Edges whose rows (+1 for the shifted copy) miss the band are dropped first, so that a 100-gon that
                 * spans three or four bands costs each of them only the batches of its own edges.  An edge whose two
                 * copies lie inside the surface (clipline changes neither) is ONE job that draws both; the others are
                 * two jobs, listed one after the other.  Every warp of the band builds the whole list (the same bytes,
                 * so no exchange is needed) and then takes its share of the jobs. */
                uint8_t *jobs = (uint8_t *)W->span_lo;             /* <= 2 * MG_CIRCLE_RES entries of the 384 bytes of span_*:
                                                                    * edge | both << 7 */
                int n_jobs = 0;
                {
                    const unsigned lt = (1u << lane) - 1;
                    RNOUNROLL
                    for (int i0 = 0; i0 < n; i0 += 32) {
                        const int i = i0 + lane;
                        bool ok = false, two = false;
                        if (i < n) {
                            const int xa = W->vx[i], ya = W->vy[i], xb = W->vx[i + 1], yb = W->vy[i + 1];
                            ok = max(ya, yb) + 1 >= y0 && min(ya, yb) < y0 + BAND_ROWS;
                            two = ok && ((unsigned)min(xa, xb) > RES - 2 || (unsigned)max(xa, xb) > RES - 2
                                         || (unsigned)min(ya, yb) > RES - 2 || (unsigned)max(ya, yb) > RES - 2);
                        }
                        const unsigned m = __ballot_sync(FULL, ok), m2 = __ballot_sync(FULL, two);
                        if (ok) {
                            const int at = n_jobs + __popc(m & lt) + __popc(m2 & lt);
                            jobs[at] = (uint8_t)(two ? i : i | 128);
                            if (two) jobs[at + 1] = (uint8_t)i;
                        }
                        n_jobs += __popc(m) + __popc(m2);
                    }
                    __syncwarp();
                }
                /* a warp's share: whole batches of 32 when there are many jobs, else an even split (long lines are drawn
                 * by the warp that holds them, pixels over its lanes) */
                const int per_warp = n_jobs >= NT ? 32 : (n_jobs + WARPS_PER_BAND - 1) / WARPS_PER_BAND;
                RNOUNROLL
                for (int job0 = wid * per_warp; job0 < n_jobs; job0 += WARPS_PER_BAND * per_warp) {
                    const int job = job0 + lane;
                    const bool has = lane < per_warp && job < n_jobs;
                    const int e = has ? jobs[job] : 0, i = e & 127;
                    const int copy = has && job > 0 && jobs[job - 1] == e && !(e & 128);   /* second entry of a pair */
                    int x1 = W->vx[i], y1 = W->vy[i], x2 = W->vx[i + 1], y2 = W->vy[i + 1];
                    const bool shift_y = abs(x1 - x2) > abs(y1 - y2);
                    if (shift_y) { y1 += copy; y2 += copy; } else { x1 += copy; x2 += copy; }
                    draw_line_jobs(W->buf, y0, lane, has, x1, y1, x2, y2, P.outline, (e & 128) ? (shift_y ? 2 : 1) : 0);
                }
            } else if (P.mode == 2) {
                RNOUNROLL
                for (int i = 0; i < n; i++) {
                    int j = (i + 1) % n;
                    draw_dashed_edge(W->buf, y0, tid, W->fx[i], W->fy[i], W->fx[j], W->fy[j], P.outline);
                }
            }
        }
        band_sync();
      }
    }

    if (MODE == 2) {
        /* full resolution: surfarray.array3d(...).swapaxes(0, 1) -> u8 [384, 384, 3] (render.py:394-395) */
        uint32_t *dst = (uint32_t *)(A.out + ((size_t)env * RES + y0) * RES * 3);
        for (int g = tid; g < BAND_ROWS * (RES / 4); g += NT) {
            int r = g / (RES / 4), q = g % (RES / 4);
            uint32_t w = *(const uint32_t *)(W->buf + r * ROW_STRIDE + q * 4);
            uint32_t c0 = pal[w & 31], c1 = pal[(w >> 8) & 31], c2 = pal[(w >> 16) & 31], c3 = pal[(w >> 24) & 31];
            uint32_t *o = dst + (size_t)g * 3;
            o[0] = c0 | (c1 << 24);
            o[1] = (c1 >> 8) | (c2 << 16);
            o[2] = (c2 >> 16) | (c3 << 8);
        }
        return;
    }

    /* 4-group stack: the three older frames of this band are needed at the very end; load them into
     * registers now so that their HBM/L2 latency hides behind the downsample (9 words per quad, 192 / NT quads per thread) */
    constexpr int PRE_IT = LO_ROWS * MG_LORES / 4 / NT;
    uint32_t pre[VARIANT == 1 ? PRE_IT : 1][3][3];
    bool fresh = false, use_pre = false;
    if (MODE == 0) fresh = fresh_word != 0;
    if (VARIANT == 1) {
        use_pre = !fresh && A.gview[3] == A.view && A.gage[3] == 0;
#pragma unroll
        for (int g = 0; g < 3; g++) use_pre = use_pre && !(A.gview[g] == A.view && A.gage[g] == 0);
        if (use_pre) {
#pragma unroll
            for (int g = 0; g < 3; g++) {
                const uint32_t *p = (const uint32_t *)(S.ring[A.gview[g]] + ((size_t)env * MG_MAX_RING + A.gslot[g]) * MG_FRAME_BYTES
                                                       + (size_t)band * LO_BAND_BYTES);
#pragma unroll
                for (int it = 0; it < PRE_IT; it++) {
                    const int q = tid + NT * it;
                    pre[it][g][0] = __ldg(p + q * 3); pre[it][g][1] = __ldg(p + q * 3 + 1); pre[it][g][2] = __ldg(p + q * 3 + 2);
                }
            }
        }
    }

    /* cv2.resize(INTER_AREA) at exactly 4x: round-half-even of the mean of each 4x4 block.  A lane takes four
     * neighbouring blocks (a quad: 16 x 4 pixels -> 4 x RGB = three words).  Most quads are one colour (floor,
     * background, the inside of a shape): those are one table look-up in pass 1; the others are queued and done 32 at
     * a time in pass 2, whose blocks of more than one colour are in turn queued and summed densely, so that neither a
     * quad with an edge in it nor a mixed block drags its 31 neighbours through the slow path.
     * INPLACE (the fused LoRes4 kernel): the three words of quad (oy, ox4) overwrite the first 12 bytes of the quad's
     * own pixels (row 4 oy, byte 16 ox4 of `buf`) as one 16-byte store, after the last read of those pixels, and the
     * stack assembly below picks them up from there with one 16-byte load: no separate low-res array in shared
     * memory (15 resident CTAs per SM instead of 13). */
    constexpr bool INPLACE = VARIANT == 1 && RENDER_INPLACE;
    {
        static_assert(offsetof(WarpSmem, vy) == offsetof(WarpSmem, vx) + sizeof(W->vx) && offsetof(WarpSmem, mq_more) == offsetof(WarpSmem, vy) + sizeof(W->vy)
                      && sizeof(uint32_t) * 128 * WARPS_PER_BAND <= sizeof(W->vx) + sizeof(W->vy) + sizeof(W->mq_more), "mixed-block queues = vx + vy + mq_more");
        static_assert(LO_ROWS * MG_LORES / 4 <= 3 * BAND_ROWS * (int)sizeof(int) && 2 * MG_CIRCLE_RES <= 3 * BAND_ROWS * (int)sizeof(int) && MG_CIRCLE_RES < 128, "quad / outline job lists fit span_*");
        const unsigned lt_mask = (1u << lane) - 1;
        uint32_t *lo_w = (uint32_t *)W->lo;
        /* every warp of the band keeps its own quad list and its own queue: passes 1 and 2 need no CTA barrier */
        uint8_t *quads = (uint8_t *)W->span_lo + wid * (LO_ROWS * MG_LORES / 4 / WARPS_PER_BAND);   /* span_* (384 bytes) are dead by now; <= 192 entries */
        uint32_t *mq = (uint32_t *)W->vx + wid * 128;          /* vx, vy are dead by now: the <= 128 mixed blocks of one batch of
                                                                * quads, block id in, colour out */
        int n_quads = 0;
        R_UNROLL(R_UNROLL_DS)
        for (int q = tid; q < LO_ROWS * MG_LORES / 4; q += NT) {   /* 192 = 6 full warps of quads, 24 quads per block row */
            const int oy = q / (MG_LORES / 4), ox4 = q - oy * (MG_LORES / 4);
            uint8_t *base = W->buf + (4 * oy) * ROW_STRIDE + 16 * ox4;
            const uint4 r0 = *(const uint4 *)base, r1 = *(const uint4 *)(base + ROW_STRIDE);
            const uint4 r2 = *(const uint4 *)(base + 2 * ROW_STRIDE), r3 = *(const uint4 *)(base + 3 * ROW_STRIDE);
            const uint32_t a = r0.x;
            const uint32_t diff = ((a ^ r0.y) | (a ^ r0.z) | (a ^ r0.w)) | ((a ^ r1.x) | (a ^ r1.y) | (a ^ r1.z) | (a ^ r1.w))
                                  | ((a ^ r2.x) | (a ^ r2.y) | (a ^ r2.z) | (a ^ r2.w)) | ((a ^ r3.x) | (a ^ r3.y) | (a ^ r3.z) | (a ^ r3.w))
                                  | (a ^ ((a & 255) * 0x01010101u));
            const unsigned m = __ballot_sync(FULL, diff != 0);
            if (diff) quads[n_quads + __popc(m & lt_mask)] = (uint8_t)q;
            else {
                const uint32_t c = pal[a & 31];
                if (INPLACE) *(uint4 *)base = make_uint4(c | c << 24, c >> 8 | c << 16, c >> 16 | c << 8, 0);
                else {
                    lo_w[3 * q] = c | c << 24;
                    lo_w[3 * q + 1] = c >> 8 | c << 16;
                    lo_w[3 * q + 2] = c >> 16 | c << 8;
                }
            }
            n_quads += __popc(m);
        }
        __syncwarp();
        /* Pass 2, the quads with more than one colour, 32 at a time: one-colour blocks by look-up, the mixed ones of the
         * batch go through the queue (lane = block) and come back as colours */
        RNOUNROLL
        for (int k0 = 0; k0 < n_quads; k0 += 32) {
            const bool has = k0 + lane < n_quads;
            const int q = has ? quads[k0 + lane] : 0;
            const int oy = q / (MG_LORES / 4), ox4 = q - oy * (MG_LORES / 4);
            uint8_t *base = W->buf + (4 * oy) * ROW_STRIDE + 16 * ox4;
            uint32_t w0[4];
            int slot[4];                                        /* queue entry of block b, -1: one colour */
            int n_mixed = 0;
            {
                const uint4 r0 = *(const uint4 *)base, r1 = *(const uint4 *)(base + ROW_STRIDE);
                const uint4 r2 = *(const uint4 *)(base + 2 * ROW_STRIDE), r3 = *(const uint4 *)(base + 3 * ROW_STRIDE);
                const uint32_t w1[4] = { r1.x, r1.y, r1.z, r1.w }, w2[4] = { r2.x, r2.y, r2.z, r2.w }, w3[4] = { r3.x, r3.y, r3.z, r3.w };
                w0[0] = r0.x; w0[1] = r0.y; w0[2] = r0.z; w0[3] = r0.w;
#pragma unroll
                for (int b = 0; b < 4; b++) {
                    const bool mix = has && !(w0[b] == w1[b] && w0[b] == w2[b] && w0[b] == w3[b] && w0[b] == (w0[b] & 255) * 0x01010101u);
                    const unsigned m = __ballot_sync(FULL, mix);
                    slot[b] = mix ? n_mixed + __popc(m & lt_mask) : -1;
                    if (mix) mq[slot[b]] = (uint32_t)(oy * MG_LORES + 4 * ox4 + b);
                    n_mixed += __popc(m);
                }
            }
            __syncwarp();
            RNOUNROLL
            for (int k = lane; k < n_mixed; k += 32) {
                const int o = (int)mq[k];
                const int my = o / MG_LORES, mx = o - my * MG_LORES;
                const uint8_t *mb = W->buf + (4 * my) * ROW_STRIDE + 4 * mx;
                const uint32_t v0 = *(const uint32_t *)mb, v1 = *(const uint32_t *)(mb + ROW_STRIDE);
                const uint32_t v2 = *(const uint32_t *)(mb + 2 * ROW_STRIDE), v3 = *(const uint32_t *)(mb + 3 * ROW_STRIDE);
                mq[k] = pal_sum16(pal, v0, v1, v2, v3);
            }
            __syncwarp();
            uint32_t rgb[4];
#pragma unroll
            for (int b = 0; b < 4; b++) rgb[b] = slot[b] >= 0 ? mq[slot[b]] : pal[w0[b] & 31];
            if (has) {
                if (INPLACE) *(uint4 *)base = make_uint4(rgb[0] | rgb[1] << 24, rgb[1] >> 8 | rgb[2] << 16, rgb[2] >> 16 | rgb[3] << 8, 0);
                else {
                    lo_w[3 * q] = rgb[0] | rgb[1] << 24;
                    lo_w[3 * q + 1] = rgb[1] >> 8 | rgb[2] << 16;
                    lo_w[3 * q + 2] = rgb[2] >> 16 | rgb[3] << 8;
                }
            }
            __syncwarp();                                       /* the queue is reused by the next batch */
        }
    }
    band_sync();

    const size_t band_off = (size_t)band * LO_BAND_BYTES;
    if (MODE == 1) {
        if (tid == 0) {
            bulk_store_fence();
            bulk_store(A.out + (size_t)env * MG_FRAME_BYTES + band_off, W->lo, LO_BAND_BYTES);
            bulk_store_commit();
            bulk_store_wait();
        }
        return;
    }

    /* mode 0: push into the frame ring.  A freshly reset env fills every slot with its first
     * frame (FlattenFrameStack.reset / EagerDictFrameStack.reset, benchmarks/__init__.py:66-74, 139-147) */
    uint8_t *ring = S.ring[A.view] + (size_t)env * MG_MAX_RING * MG_FRAME_BYTES;
    /* INPLACE: every lane takes its six quads of the new frame into registers (the stack assembly needs them there
     * anyway), then the warp packs them to the front of the now dead band buffer, from where they leave like W->lo */
    uint32_t mine[INPLACE ? PRE_IT : 1][3];
    const uint8_t *lo_src = W->lo;
    if (INPLACE) {
#pragma unroll
        for (int it = 0; it < PRE_IT; it++) {
            const int q = tid + NT * it;
            const int oy = q / (MG_LORES / 4), ox4 = q - oy * (MG_LORES / 4);
            const uint4 m = *(const uint4 *)(W->buf + (4 * oy) * ROW_STRIDE + 16 * ox4);
            mine[it][0] = m.x; mine[it][1] = m.y; mine[it][2] = m.z;
        }
        band_sync();
#pragma unroll
        for (int it = 0; it < PRE_IT; it++) {
            uint32_t *p = (uint32_t *)W->buf + (tid + NT * it) * 3;
            p[0] = mine[it][0]; p[1] = mine[it][1]; p[2] = mine[it][2];
        }
        band_sync();
        lo_src = W->buf;
    }
    if (tid == 0) {
        bulk_store_fence();
        RNOUNROLL
        for (int s = 0; s < MG_MAX_RING; s++)
            if (fresh || s == A.slot) bulk_store(ring + (size_t)s * MG_FRAME_BYTES + band_off, lo_src, LO_BAND_BYTES);
        bulk_store_commit();          /* in flight under the stack assembly below; waited for before the warp exits */
    }
    if (VARIANT == 0) { if (tid == 0) bulk_store_wait(); return; }

    /* fused frame stack: obs[env, 8 rows of this band, 96, 3 * n_groups], channel group g taken from
     * the ring slot the host computed for (view, age).  Frames of this view that are being written
     * by this very launch (age 0, or any age of a fresh env) come from shared memory. */
    const int G = A.n_groups, C = 3 * G;
    uint32_t *obs_w = (uint32_t *)(A.obs + ((size_t)env * MG_LORES * MG_LORES + (size_t)band * LO_ROWS * MG_LORES) * C);
    if (VARIANT == 1) {
        /* LoRes4E / LoRes4A: 4 frames x RGB = 12 bytes per pixel; a lane turns 4 pixels of each frame
         * (3 words each) into 48 contiguous output bytes = three 16-byte stores.  INPLACE: the lane's quads of the new
         * frame are already in registers (`mine`). */
        const uint32_t *srcp[4];
        bool cur[4];
#pragma unroll
        for (int g = 0; g < 4; g++) {
            cur[g] = A.gview[g] == A.view && (A.gage[g] == 0 || fresh);
            if (cur[g]) srcp[g] = (const uint32_t *)lo_src;
            else srcp[g] = (const uint32_t *)(S.ring[A.gview[g]] + ((size_t)env * MG_MAX_RING + A.gslot[g]) * MG_FRAME_BYTES + band_off);
        }
#pragma unroll
        for (int it = 0; it < PRE_IT; it++) {
            const int q = tid + NT * it;
            uint32_t src[4][3];
            if (use_pre) {
#pragma unroll
                for (int g = 0; g < 3; g++) { src[g][0] = pre[it][g][0]; src[g][1] = pre[it][g][1]; src[g][2] = pre[it][g][2]; }
                if (INPLACE && !RENDER_REREAD) { src[3][0] = mine[it][0]; src[3][1] = mine[it][1]; src[3][2] = mine[it][2]; }
                else { src[3][0] = srcp[3][q * 3]; src[3][1] = srcp[3][q * 3 + 1]; src[3][2] = srcp[3][q * 3 + 2]; }
            } else {
#pragma unroll
                for (int g = 0; g < 4; g++) {
                    if (INPLACE && !RENDER_REREAD && cur[g]) { src[g][0] = mine[it][0]; src[g][1] = mine[it][1]; src[g][2] = mine[it][2]; }
                    else { src[g][0] = srcp[g][q * 3]; src[g][1] = srcp[g][q * 3 + 1]; src[g][2] = srcp[g][q * 3 + 2]; }
                }
            }
            uint4 *dst = (uint4 *)(obs_w + (size_t)q * 12);
#pragma unroll
            for (int half = 0; half < 3; half++) {
                uint32_t o[4];
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    /* word w of the 12 output words = bytes of frames `part` and `part + 1` at pixel px; win(g) is the
                     * 32-bit window of frame g that starts at the pixel's first byte (one funnel shift), and one byte
                     * permute picks r g b | r, g b | r g, b | r g b out of two windows */
                    const int w = half * 4 + k, px = w / 3, part = w % 3;
                    const int lo_w = (3 * px) >> 2, sh = 8 * ((3 * px) & 3);
                    const uint32_t wa = sh ? __funnelshift_r(src[part][lo_w], src[part][lo_w < 2 ? lo_w + 1 : 2], sh) : src[part][lo_w];
                    const uint32_t wb = sh ? __funnelshift_r(src[part + 1][lo_w], src[part + 1][lo_w < 2 ? lo_w + 1 : 2], sh) : src[part + 1][lo_w];
                    o[k] = __byte_perm(wa, wb, part == 0 ? 0x4210 : (part == 1 ? 0x5421 : 0x6542));
                }
                dst[half] = make_uint4(o[0], o[1], o[2], o[3]);
            }
        }
    } else
    for (int q = tid; q < LO_ROWS * MG_LORES / 4; q += NT) {      /* q = group of 4 consecutive pixels */
        uint32_t src[MG_MAX_GROUPS][3];
        for (int g = 0; g < G; g++) {
            const uint32_t *p;
            if (A.gview[g] == A.view && (A.gage[g] == 0 || fresh)) p = (const uint32_t *)W->lo + q * 3;
            else p = (const uint32_t *)(S.ring[A.gview[g]] + ((size_t)env * MG_MAX_RING + A.gslot[g]) * MG_FRAME_BYTES + band_off) + q * 3;
            src[g][0] = p[0]; src[g][1] = p[1]; src[g][2] = p[2];
        }
        /* output byte (pixel px, channel ch) = byte (3 * px + ch % 3) of group ch / 3 */
        for (int w = 0; w < 3 * G; w++) {
            uint32_t word = 0;
#pragma unroll
            for (int k = 0; k < 4; k++) {
                int byte = w * 4 + k;
                int px = byte / C, ch = byte % C;
                int g = ch / 3, sb = 3 * px + ch % 3;
                word |= ((src[g][sb >> 2] >> (8 * (sb & 3))) & 255u) << (8 * k);
            }
            obs_w[(size_t)q * 3 * G + w] = word;
        }
    }
    if (tid == 0) bulk_store_wait();
}

/* generic stack composer: any mix of views, HWC or CHW (LoResStack, LoRes3EA, LoResCHW*) */
__global__ void mg_compose_kernel(MgDeviceState S, MgObsSpec spec, int slot0, int slot1, uint8_t *obs)
{
    int env = blockIdx.x;
    const int G = spec.n_groups, C = 3 * G;
    uint8_t *o = obs + (size_t)env * MG_LORES * MG_LORES * C;
    for (int idx = threadIdx.x; idx < MG_LORES * MG_LORES * C; idx += blockDim.x) {
        int pix, ch;
        if (spec.chw) { ch = idx / (MG_LORES * MG_LORES); pix = idx % (MG_LORES * MG_LORES); }
        else { pix = idx / C; ch = idx % C; }
        int g = ch / 3, v = spec.view[g];
        int slot = ((v == 0 ? slot0 : slot1) - spec.age[g]) & (MG_MAX_RING - 1);
        o[idx] = S.ring[v][((size_t)env * MG_MAX_RING + slot) * MG_FRAME_BYTES + pix * 3 + ch % 3];
    }
}

__global__ void mg_clear_fresh_kernel(MgDeviceState S)
{
    int env = blockIdx.x * blockDim.x + threadIdx.x;
    if (env < S.num_envs) S.istate[mg_field_index(env, I_NF, I_FRESH)] = 0;
}

void mg_launch_geometry(const MgDeviceState &S, int view, int env_begin, int env_end, cudaStream_t st, int apply_pending)
{
    if (env_end <= env_begin) return;
    mg_geometry_kernel<<<(env_end - env_begin + GEO_WARPS - 1) / GEO_WARPS, GEO_WARPS * 32, 0, st>>>(S, view, env_begin, env_end, apply_pending);
}

static void launch_render(const MgDeviceState &S, const RenderArgs &A, cudaStream_t st, int apply_pending = 0)
{
    if (A.env_end <= A.env_begin) return;
    mg_launch_geometry(S, A.view, A.env_begin, A.env_end, st, apply_pending);
    /* per-device function attribute (see mg_launch_physics) */
    static bool attr_set_dev[MG_MAX_DEVICES] = { false };
    static std::mutex attr_mu;
    int dev_ = 0;
    cudaGetDevice(&dev_);
    std::lock_guard<std::mutex> attr_lock(attr_mu);
    bool &attr_set = attr_set_dev[dev_ < 0 || dev_ >= MG_MAX_DEVICES ? 0 : dev_];
#ifndef RENDER_EXTRA_SMEM
#define RENDER_EXTRA_SMEM 0       /* dev builds: bytes of unused shared memory per CTA, to time the kernel at lower occupancy */
#endif
    size_t smem = sizeof(WarpSmem) + RENDER_EXTRA_SMEM;
    if (!attr_set) {
        cudaFuncSetAttribute(mg_render_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaFuncSetAttribute(mg_render_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaFuncSetAttribute(mg_render_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaFuncSetAttribute(mg_render_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        cudaFuncSetAttribute(mg_render_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        attr_set = true;
    }
    /* one CTA of NT threads per band */
    const int grid = (A.env_end - A.env_begin) * N_BANDS;
    int variant = A.mode == 0 ? (!A.obs ? 0 : (A.n_groups == 4 ? 1 : 2)) : A.mode + 2;
    /* the fused LoRes4 kernel keeps the downsampled band inside the band buffer: launched without WarpSmem::lo */
    const size_t smem1 = RENDER_INPLACE ? offsetof(WarpSmem, lo) + RENDER_EXTRA_SMEM : smem;
    switch (variant) {
    case 0: mg_render_kernel<0><<<grid, NT, smem, st>>>(S, A); break;
    case 1: mg_render_kernel<1><<<grid, NT, smem1, st>>>(S, A); break;
    case 2: mg_render_kernel<2><<<grid, NT, smem, st>>>(S, A); break;
    case 3: mg_render_kernel<3><<<grid, NT, smem, st>>>(S, A); break;
    default: mg_render_kernel<4><<<grid, NT, smem, st>>>(S, A); break;
    }
}

void mg_launch_render_push(const MgDeviceState &S, int view, int slot, const MgObsSpec *fused, const int *slots,
                           uint8_t *obs, int env_begin, int env_end, cudaStream_t st, int apply_pending)
{
    RenderArgs A;
    memset(&A, 0, sizeof(A));
    A.view = view; A.mode = 0; A.slot = slot; A.env_begin = env_begin; A.env_end = env_end;
    if (fused && obs) {
        A.obs = obs; A.n_groups = fused->n_groups;
        for (int g = 0; g < fused->n_groups; g++) {
            A.gview[g] = fused->view[g]; A.gage[g] = fused->age[g];
            A.gslot[g] = (slots[fused->view[g]] - fused->age[g]) & (MG_MAX_RING - 1);
        }
    }
    launch_render(S, A, st, apply_pending);
}

void mg_launch_render_out(const MgDeviceState &S, int view, int full_res, uint8_t *out, cudaStream_t st)
{
    RenderArgs A;
    memset(&A, 0, sizeof(A));
    A.view = view; A.mode = full_res ? 2 : 1; A.out = out; A.env_begin = 0; A.env_end = S.num_envs;
    launch_render(S, A, st);
}

void mg_launch_compose(const MgDeviceState &S, const MgObsSpec &spec, const int *slots, uint8_t *obs, cudaStream_t st)
{
    mg_compose_kernel<<<S.num_envs, 256, 0, st>>>(S, spec, slots[0], slots[1], obs);
}

void mg_launch_clear_fresh(const MgDeviceState &S, cudaStream_t st)
{
    mg_clear_fresh_kernel<<<(S.num_envs + 127) / 128, 128, 0, st>>>(S);
}
