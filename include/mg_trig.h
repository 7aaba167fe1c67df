/*
 * mg_trig.h -- one sin/cos for host and device.
 *
 * The reference reaches sin/cos through libm (Chipmunk's cpvforangle in cpBodySetAngle, numpy in
 * render.Transform, magical/render.py:87-97).  glibc and the CUDA math library differ in the last
 * bit for some arguments, and contact-rich rigid-body trajectories amplify a last-bit difference.
 * This header gives the device kernels, the host-side reset code and the CPU oracle the same
 * function, built from IEEE add/mul only (Cody-Waite reduction with a 3-part pi/2 and the
 * fdlibm-style minimax kernels), so all of them produce bit-identical rotations when compiled
 * without FMA contraction (-fmad=false / -ffp-contract=off).  Error is below 1 ulp for the
 * arguments that occur here (|x| < 2^20).
 */
#ifndef MG_TRIG_H
#define MG_TRIG_H

#include <math.h>

#ifdef __CUDACC__
#define MG_TRIG_FN __host__ __device__ __forceinline__
#else
#define MG_TRIG_FN static inline
#endif

MG_TRIG_FN double mg_ksin(double x, double y)
{
    const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03,
                 S3 = -1.98412698298579493134e-04, S4 = 2.75573137070700676789e-06,
                 S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
    double z = x * x;
    double v = z * x;
    double r = S2 + z * (S3 + z * (S4 + z * (S5 + z * S6)));
    return x - ((z * (0.5 * y - v * r) - y) - v * S1);
}

MG_TRIG_FN double mg_kcos(double x, double y)
{
    const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03,
                 C3 = 2.48015872894767294178e-05, C4 = -2.75573143513906633035e-07,
                 C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
    double z = x * x;
    double r = z * (C1 + z * (C2 + z * (C3 + z * (C4 + z * (C5 + z * C6)))));
    double ax = fabs(x);
    if (ax < 0.3) return 1.0 - (0.5 * z - (z * r - x * y));
    double qx = ax > 0.78125 ? 0.28125 : (double)(float)(0.25 * ax);
    double hz = 0.5 * z - qx;
    double a = 1.0 - qx;
    return a - (hz - (z * r - x * y));
}

MG_TRIG_FN void mg_sincos(double x, double *s_out, double *c_out)
{
    const double invpio2 = 6.36619772367581382433e-01;   /* 0x1.45f306dc9c883p-1 */
    const double pio2_1 = 1.5707963267341256;            /* first 33 bits of pi/2   0x1.921fb54400000p+0 */
    const double pio2_2 = 6.077100506303966e-11;         /* next 33 bits            0x1.0b4611a600000p-34 */
    const double pio2_2t = 2.0222662487959506e-21;       /* pi/2 - (pio2_1+pio2_2)  0x1.3198a2e037073p-69 */
    double ax = fabs(x);
    int n = (int)(ax * invpio2 + 0.5);
    double fn = (double)n;
    double t = ax - fn * pio2_1;
    double w = fn * pio2_2;
    double r = t - w;
    w = fn * pio2_2t - ((t - r) - w);
    double y0 = r - w;
    double y1 = (r - y0) - w;
    double ks = mg_ksin(y0, y1), kc = mg_kcos(y0, y1);
    double s, c;
    switch (n & 3) {
    case 0: s = ks; c = kc; break;
    case 1: s = kc; c = -ks; break;
    case 2: s = -ks; c = -kc; break;
    default: s = -kc; c = ks; break;
    }
    *s_out = x < 0.0 ? -s : s;
    *c_out = c;
}

#endif /* MG_TRIG_H */
