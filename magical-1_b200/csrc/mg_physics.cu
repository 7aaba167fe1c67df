/*
 * mg_physics.cu -- reset expansion, the per-step rigid-body kernel and the score kernels.
 *
 * One group of PHYS_LANES lanes owns one env (one pymunk Space of the reference) and runs, in Chipmunk's
 * order, everything BaseEnv.step does before rendering (magical/base_env.py:267-292):
 *   Robot.set_action                      magical/entities.py:435-453
 *   10 x ( Robot.update ; Space.step )    magical/base_env.py:248-255, magical/entities.py:455-476
 *   episode counter / done / score        magical/base_env.py:284-292
 * Space.step is Chipmunk2D 7.0.3's cpSpaceStep as configured by base_env.py:206-208
 * (iterations 10, collision_slop 0.01, no gravity, damping 1): position integration,
 * narrow phase over all shape pairs in add order, arbiter cache with warm starting,
 * pre-step, cached impulses, 10 sequential-impulse iterations over arbiters then
 * constraints in add order.  All arithmetic is IEEE double without FMA contraction.
 *
 * The env's bodies, constraints and shape list are staged in shared memory for the whole step; items that
 * do not depend on each other (bodies, shape pairs, pre-steps) are spread over the group's lanes, and the
 * order-dependent half of every sub-step runs on one lane per env, the lanes of all envs of a CTA side by
 * side in warp 0 (space_step_parallel; solve_lead / solve_helper).  Every scalar goes through the same operations
 * in the same order as in a serial cpSpaceStep, so the result does not depend on the mapping.
 */
#include <cuda_runtime.h>
#include <stdlib.h>
#include <stddef.h>
#include <stdio.h>
#include <mutex>
#include "mg_state.cuh"
#include "mg_kernels.h"
#include "../../include/mg_trig.h"

#include "mg_envref.cuh"

/* BaseEnv.reset after on_reset(): expand host-sampled scenes into simulation state (init_env, mg_envref.cuh) */
__global__ void mg_init_kernel(MgDeviceState S, const mg_scene *scenes, const int32_t *env_ids, int n)
{
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    init_env(S, &scenes[k], env_ids ? env_ids[k] : k);
}

/* mg_queue_resets: pending[env] <- index of the env's new scene in the staging array; the geometry pre-pass of the
 * next step applies it after that step's physics */
__global__ void mg_mark_pending_kernel(MgDeviceState S, const int32_t *env_ids, int n)
{
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) S.pending[env_ids ? env_ids[k] : k] = k;
}

/* ------------------------------------------------------------------ body helpers (chipmunk_private.h) */

__device__ __forceinline__ vec2 body_p(EnvRef e, int b) { return V(BF(e, b, B_PX), BF(e, b, B_PY)); }
__device__ __forceinline__ vec2 body_v(EnvRef e, int b) { return V(BF(e, b, B_VX), BF(e, b, B_VY)); }
__device__ __forceinline__ vec2 body_rot(EnvRef e, int b) { return V(BF(e, b, B_RX), BF(e, b, B_RY)); }

__device__ __forceinline__ vec2 relative_velocity(EnvRef e, int a, int b, vec2 r1, vec2 r2)
{
    vec2 v1 = vadd(body_v(e, a), vmul(vperp(r1), BF(e, a, B_W)));
    vec2 v2 = vadd(body_v(e, b), vmul(vperp(r2), BF(e, b, B_W)));
    return vsub(v2, v1);
}
__device__ __forceinline__ void apply_impulse(EnvRef e, int b, vec2 j, vec2 r)
{
    double minv = BF(e, b, B_MINV), iinv = BF(e, b, B_IINV);
    BF(e, b, B_VX) = BF(e, b, B_VX) + j.x * minv;
    BF(e, b, B_VY) = BF(e, b, B_VY) + j.y * minv;
    BF(e, b, B_W) = BF(e, b, B_W) + iinv * vcross(r, j);
}
__device__ __forceinline__ void apply_impulses(EnvRef e, int a, int b, vec2 r1, vec2 r2, vec2 j)
{
    apply_impulse(e, a, vneg(j), r1);
    apply_impulse(e, b, j, r2);
}
__device__ __forceinline__ void apply_bias_impulse(EnvRef e, int b, vec2 j, vec2 r)
{
    double minv = BF(e, b, B_MINV), iinv = BF(e, b, B_IINV);
    BF(e, b, B_VBX) = BF(e, b, B_VBX) + j.x * minv;
    BF(e, b, B_VBY) = BF(e, b, B_VBY) + j.y * minv;
    BF(e, b, B_WB) = BF(e, b, B_WB) + iinv * vcross(r, j);
}
__device__ __forceinline__ double k_scalar_body(EnvRef e, int b, vec2 r, vec2 n)
{
    double rcn = vcross(r, n);
    return BF(e, b, B_MINV) + BF(e, b, B_IINV) * rcn * rcn;
}
__device__ __forceinline__ double k_scalar(EnvRef e, int a, int b, vec2 r1, vec2 r2, vec2 n)
{
    return k_scalar_body(e, a, r1, n) + k_scalar_body(e, b, r2, n);
}
__device__ __forceinline__ vec2 vclamp_len(vec2 v, double len)
{
    if (vdot(v, v) > len * len) {
        double l = sqrt(vdot(v, v));
        return vmul(vmul(v, 1.0 / (l + MG_DBL_MIN)), len);
    }
    return v;
}
__device__ __forceinline__ void set_angle(EnvRef e, int b, double a)
{
    double s, c;
    mg_sincos(a, &s, &c);
    BF(e, b, B_A) = a; BF(e, b, B_RX) = c; BF(e, b, B_RY) = s;
}

/* ------------------------------------------------------------------ joints (cpPivotJoint.c, cpGearJoint.c,
 * cpDampedRotarySpring.c, cpRotaryLimitJoint.c, cpSimpleMotor.c) */

__device__ void con_prestep(EnvRef e, int c)
{
    const double dt = MG_DT;
    int info = IF(e, I_CON0 + c);
    int kind = info & 255, a = (info >> 8) & 255, b = (info >> 16) & 255;
    switch (kind) {
    case CON_PIVOT: {
        vec2 anchor_a = V(CF(e, c, C_P0), CF(e, c, C_P1));
        vec2 r1 = vrot(anchor_a, body_rot(e, a));
        vec2 r2 = vrot(V(0.0, 0.0), body_rot(e, b));
        double m_sum = BF(e, a, B_MINV) + BF(e, b, B_MINV);
        double k11 = m_sum, k12 = 0.0, k21 = 0.0, k22 = m_sum;
        double ai = BF(e, a, B_IINV);
        double r1xsq = r1.x * r1.x * ai, r1ysq = r1.y * r1.y * ai, r1nxy = -r1.x * r1.y * ai;
        k11 += r1ysq; k12 += r1nxy; k21 += r1nxy; k22 += r1xsq;
        double bi = BF(e, b, B_IINV);
        double r2xsq = r2.x * r2.x * bi, r2ysq = r2.y * r2.y * bi, r2nxy = -r2.x * r2.y * bi;
        k11 += r2ysq; k12 += r2nxy; k21 += r2nxy; k22 += r2xsq;
        double det_inv = 1.0 / (k11 * k22 - k12 * k21);
        CF(e, c, C_K0) = k22 * det_inv; CF(e, c, C_K1) = -k12 * det_inv;
        CF(e, c, C_K2) = -k21 * det_inv; CF(e, c, C_K3) = k11 * det_inv;
        CF(e, c, C_R1X) = r1.x; CF(e, c, C_R1Y) = r1.y; CF(e, c, C_R2X) = r2.x; CF(e, c, C_R2Y) = r2.y;
        vec2 delta = vsub(vadd(body_p(e, b), r2), vadd(body_p(e, a), r1));
        vec2 bias = vclamp_len(vmul(delta, -CF(e, c, C_BC) / dt), CF(e, c, C_MAXBIAS));
        CF(e, c, C_BIASX) = bias.x; CF(e, c, C_BIASY) = bias.y;
        break;
    }
    case CON_GEAR: {
        double phase = CF(e, c, C_P0), ratio = CF(e, c, C_P1), mb = CF(e, c, C_MAXBIAS);
        double ratio_inv = 1.0 / ratio;
        CF(e, c, C_ISUM) = 1.0 / (BF(e, a, B_IINV) * ratio_inv + ratio * BF(e, b, B_IINV));
        CF(e, c, C_BIASX) = dclamp(-CF(e, c, C_BC) * (BF(e, b, B_A) * ratio - BF(e, a, B_A) - phase) / dt, -mb, mb);
        break;
    }
    case CON_SPRING: {
        double moment = BF(e, a, B_IINV) + BF(e, b, B_IINV);
        CF(e, c, C_ISUM) = 1.0 / moment;
        CF(e, c, C_BIASY) = 0.0; /* target_wrn */
        double j_spring = (((BF(e, a, B_A) - BF(e, b, B_A)) - 0.0) * 0.1) * dt; /* rest 0, stiffness 0.1 (entities.py:273) */
        CF(e, c, C_JX) = j_spring;
        BF(e, a, B_W) = BF(e, a, B_W) - j_spring * BF(e, a, B_IINV);
        BF(e, b, B_W) = BF(e, b, B_W) + j_spring * BF(e, b, B_IINV);
        break;
    }
    case CON_LIMIT: {
        double lo = CF(e, c, C_P0), hi = CF(e, c, C_P1), mb = CF(e, c, C_MAXBIAS);
        double dist = BF(e, b, B_A) - BF(e, a, B_A);
        double pdist = 0.0;
        if (dist > hi) pdist = hi - dist;
        else if (dist < lo) pdist = lo - dist;
        CF(e, c, C_ISUM) = 1.0 / (BF(e, a, B_IINV) + BF(e, b, B_IINV));
        double bias = dclamp(-CF(e, c, C_BC) * pdist / dt, -mb, mb);
        CF(e, c, C_BIASX) = bias;
        if (!bias) CF(e, c, C_JX) = 0.0;
        break;
    }
    case CON_MOTOR:
        CF(e, c, C_ISUM) = 1.0 / (BF(e, a, B_IINV) + BF(e, b, B_IINV));
        break;
    }
}

__device__ void con_apply_cached(EnvRef e, int c, double dt_coef)
{
    int info = IF(e, I_CON0 + c);
    int kind = info & 255, a = (info >> 8) & 255, b = (info >> 16) & 255;
    switch (kind) {
    case CON_PIVOT:
        apply_impulses(e, a, b, V(CF(e, c, C_R1X), CF(e, c, C_R1Y)), V(CF(e, c, C_R2X), CF(e, c, C_R2Y)),
                       vmul(V(CF(e, c, C_JX), CF(e, c, C_JY)), dt_coef));
        break;
    case CON_GEAR: {
        double j = CF(e, c, C_JX) * dt_coef;
        BF(e, a, B_W) = BF(e, a, B_W) - j * BF(e, a, B_IINV) * (1.0 / CF(e, c, C_P1));
        BF(e, b, B_W) = BF(e, b, B_W) + j * BF(e, b, B_IINV);
        break;
    }
    case CON_SPRING: break;
    case CON_LIMIT:
    case CON_MOTOR: {
        double j = CF(e, c, C_JX) * dt_coef;
        BF(e, a, B_W) = BF(e, a, B_W) - j * BF(e, a, B_IINV);
        BF(e, b, B_W) = BF(e, b, B_W) + j * BF(e, b, B_IINV);
        break;
    }
    }
}

__device__ void con_apply_impulse(EnvRef e, int c)
{
    int info = IF(e, I_CON0 + c);
    int kind = info & 255, a = (info >> 8) & 255, b = (info >> 16) & 255;
    double j_max = CF(e, c, C_JMAX);
    switch (kind) {
    case CON_PIVOT: {
        vec2 r1 = V(CF(e, c, C_R1X), CF(e, c, C_R1Y)), r2 = V(CF(e, c, C_R2X), CF(e, c, C_R2Y));
        vec2 vr = relative_velocity(e, a, b, r1, r2);
        vec2 d = vsub(V(CF(e, c, C_BIASX), CF(e, c, C_BIASY)), vr);
        vec2 j = V(d.x * CF(e, c, C_K0) + d.y * CF(e, c, C_K1), d.x * CF(e, c, C_K2) + d.y * CF(e, c, C_K3));
        vec2 j_old = V(CF(e, c, C_JX), CF(e, c, C_JY));
        vec2 j_acc = vclamp_len(vadd(j_old, j), j_max);
        CF(e, c, C_JX) = j_acc.x; CF(e, c, C_JY) = j_acc.y;
        apply_impulses(e, a, b, r1, r2, vsub(j_acc, j_old));
        break;
    }
    case CON_GEAR: {
        double ratio = CF(e, c, C_P1);
        double wr = BF(e, b, B_W) * ratio - BF(e, a, B_W);
        double j = (CF(e, c, C_BIASX) - wr) * CF(e, c, C_ISUM);
        double j_old = CF(e, c, C_JX);
        double j_acc = dclamp(j_old + j, -j_max, j_max);
        CF(e, c, C_JX) = j_acc;
        j = j_acc - j_old;
        BF(e, a, B_W) = BF(e, a, B_W) - j * BF(e, a, B_IINV) * (1.0 / ratio);
        BF(e, b, B_W) = BF(e, b, B_W) + j * BF(e, b, B_IINV);
        break;
    }
    case CON_SPRING: {
        double wrn = BF(e, a, B_W) - BF(e, b, B_W);
        double w_damp = (CF(e, c, C_BIASY) - wrn) * CF(e, c, C_P0);
        CF(e, c, C_BIASY) = wrn + w_damp;
        double j_damp = w_damp * CF(e, c, C_ISUM);
        CF(e, c, C_JX) = CF(e, c, C_JX) + j_damp;
        BF(e, a, B_W) = BF(e, a, B_W) + j_damp * BF(e, a, B_IINV);
        BF(e, b, B_W) = BF(e, b, B_W) - j_damp * BF(e, b, B_IINV);
        break;
    }
    case CON_LIMIT: {
        double bias = CF(e, c, C_BIASX);
        if (!bias) return;
        double wr = BF(e, b, B_W) - BF(e, a, B_W);
        double j = -(bias + wr) * CF(e, c, C_ISUM);
        double j_old = CF(e, c, C_JX), j_acc;
        if (bias < 0.0) j_acc = dclamp(j_old + j, 0.0, j_max);
        else j_acc = dclamp(j_old + j, -j_max, 0.0);
        CF(e, c, C_JX) = j_acc;
        j = j_acc - j_old;
        BF(e, a, B_W) = BF(e, a, B_W) - j * BF(e, a, B_IINV);
        BF(e, b, B_W) = BF(e, b, B_W) + j * BF(e, b, B_IINV);
        break;
    }
    case CON_MOTOR: {
        double wr = BF(e, b, B_W) - BF(e, a, B_W) + CF(e, c, C_P0);
        double j = -wr * CF(e, c, C_ISUM);
        double j_old = CF(e, c, C_JX);
        double j_acc = dclamp(j_old + j, -j_max, j_max);
        CF(e, c, C_JX) = j_acc;
        j = j_acc - j_old;
        BF(e, a, B_W) = BF(e, a, B_W) - j * BF(e, a, B_IINV);
        BF(e, b, B_W) = BF(e, b, B_W) + j * BF(e, b, B_IINV);
        break;
    }
    }
}

/* ------------------------------------------------------------------ collision + arbiters (cpSpaceStep.c, cpArbiter.c) */

__device__ __forceinline__ double shape_friction(int tpl)
{
    if (tpl >= MG_TPL_WALL0) return 0.8;                       /* entities.py:521 */
    if (tpl < MG_TPL_SQUARE) return 5.0;                       /* fingers, entities.py:369 */
    return 0.5;                                                /* robot body :358, blocks :696-698 */
}

__device__ __forceinline__ double shape_bound(const MgTables *T, int tpl)
{
    if (tpl >= MG_TPL_CIRCLE_ROBOT) return tpl == MG_TPL_CIRCLE_ROBOT ? 0.2 : 0.2 * 0.6;
    return T->tpl_bound[tpl];
}

/* l, b, r, t of shape template tpl on a body at p with rotation rot (cpShapeCacheBB), computed in registers:
 * the same vertex transform as wshape_from_tpl without materialising the shape */
__device__ __noinline__ void tpl_world_bb(const MgTables *T, int tpl, vec2 p, vec2 rot, double *bb)
{
    if (tpl >= MG_TPL_WALL0) {
        int w = tpl - MG_TPL_WALL0;
        vec2 a = T->wall_a[w], c = T->wall_b[w];
        bb[0] = fmin(a.x, c.x) - 1.0; bb[1] = fmin(a.y, c.y) - 1.0; bb[2] = fmax(a.x, c.x) + 1.0; bb[3] = fmax(a.y, c.y) + 1.0;
    } else if (tpl >= MG_TPL_CIRCLE_ROBOT) {
        double r = tpl == MG_TPL_CIRCLE_ROBOT ? 0.2 : 0.2 * 0.6;
        bb[0] = p.x - r; bb[1] = p.y - r; bb[2] = p.x + r; bb[3] = p.y + r;
    } else {
        const int n = T->tpl_n[tpl];
        const double r = T->tpl_r[tpl];
        double l = INFINITY, rr = -INFINITY, bo = INFINITY, t = -INFINITY;
        MG_NOUNROLL
        for (int i = 0; i < n; i++) {
            vec2 lv = T->tpl_v[tpl][i];
            double x = rot.x * lv.x - rot.y * lv.y + p.x, y = rot.y * lv.x + rot.x * lv.y + p.y;
            l = fmin(l, x); rr = fmax(rr, x); bo = fmin(bo, y); t = fmax(t, y);
        }
        bb[0] = l - r; bb[1] = bo - r; bb[2] = rr + r; bb[3] = t + r;
    }
}

/* ---- live arbiters of the current sub-step.
 * The arbiter cache (MG_NARB slots per env) lives in HBM: it is large, and only envs in contact touch it.  But the
 * sequential-impulse iterations read and write every live contact ten times per sub-step, and a global store drops the
 * line from L1, so with the cache slot as working storage every iteration paid L2 latency on jn / jt / jBias.  The first
 * MG_ARB_POOL live arbiters of a sub-step therefore work in a shared-memory pool (same A_NF-double layout as a slot):
 * cpArbiterUpdate writes r1 / r2 / warm-started jn, jt / n / u there, the pre-step and the iterations use it, and after
 * the last iteration jn / jt -- the only per-contact fields a later step reads back (warm start by contact hash) -- are
 * copied to the slot.  Live arbiters beyond the pool work in their slot as before.  Same arithmetic either way. */
#ifndef MG_ARB_POOL
#define MG_ARB_POOL 8
#endif
static_assert(MG_GROUP == 1, "arbiter working storage is addressed as a contiguous A_NF-double block");
struct ArbRef { double *p; int a, b, count, state; };

__device__ __forceinline__ ArbRef arb_ref(EnvRef e, const unsigned char *order, int q)
{
    ArbRef r;
    if (q < e.apn) {
        const int w = e.ainfo[q];
        r.p = e.ap + (size_t)q * A_NF;
        r.a = w & 255; r.b = (w >> 8) & 255; r.count = (w >> 16) & 255; r.state = (w >> 24) & 255;
    } else {
        const int s = order[q];
        const int key = AIF(e, s, AI_KEY), fl = AIF(e, s, AI_FLAGS);
        r.p = &AF(e, s, 0);
        r.a = (key >> 16) & 255; r.b = (key >> 24) & 255; r.count = (fl >> 8) & 255; r.state = fl & 255;
    }
    return r;
}
#define APF(r, f) ((r).p[(f)])
#define APC(r, k, f) ((r).p[A_CON0 + (k) * AC_NF + (f)])

/* cpArbiterUpdate for the colliding shapes sa, sb (already in cpShapeType order, infos ia, ib) with
 * manifold m; returns the arbiter slot or -1.  Serial per env: slots are handed out in discovery order.
 * q: index of this arbiter among the live ones of the sub-step (selects the working storage). */
__device__ int arbiter_update(EnvRef e, int sa, int sb, int ia, int ib, const Manifold &m, int stamp, int q)
{
    int ba = (ia >> 8) & 255, bb = (ib >> 8) & 255;
    int key = sa | (sb << 8);
    int n_arbs = IF(e, I_NARBS), slot = -1;
    MG_NOUNROLL
    for (int k = 0; k < n_arbs; k++)
        if ((AIF(e, k, AI_KEY) & 0xffff) == key) { slot = k; break; }
    int state = ARB_FIRST, old_count = 0;
    uint32_t old_hash[2] = { 0, 0 };
    double old_jn[2] = { 0, 0 }, old_jt[2] = { 0, 0 };
    if (slot < 0) {
        if (n_arbs >= MG_NARB) { IF(e, I_STATUS) = IF(e, I_STATUS) | MG_STATUS_ARB_OVERFLOW; return -1; }   /* flagged, never silent */
        slot = n_arbs;
        IF(e, I_NARBS) = n_arbs + 1;
    } else {
        int fl = AIF(e, slot, AI_FLAGS);
        state = fl & 255; old_count = (fl >> 8) & 255;
        uint32_t hh = (uint32_t)AIF(e, slot, AI_HASH);
        old_hash[0] = hh & 0xffff; old_hash[1] = hh >> 16;
        MG_NOUNROLL
        for (int k = 0; k < old_count; k++) { old_jn[k] = ACF(e, slot, k, AC_JN); old_jt[k] = ACF(e, slot, k, AC_JT); }
    }
    const bool pooled = q < e.apn;
    double *wp = pooled ? e.ap + (size_t)q * A_NF : &AF(e, slot, 0);
    vec2 pa = body_p(e, ba), pb = body_p(e, bb);
    MG_NOUNROLL
    for (int k = 0; k < m.count; k++) {
        vec2 r1 = vsub(m.p1[k], pa), r2 = vsub(m.p2[k], pb);
        double jn = 0.0, jt = 0.0;
        MG_NOUNROLL
        for (int o = 0; o < old_count; o++)
            if (m.hash[k] == old_hash[o]) { jn = old_jn[o]; jt = old_jt[o]; }
        double *c = wp + A_CON0 + k * AC_NF;
        c[AC_R1X] = r1.x; c[AC_R1Y] = r1.y; c[AC_R2X] = r2.x; c[AC_R2Y] = r2.y;
        c[AC_JN] = jn; c[AC_JT] = jt;
    }
    wp[A_NX] = m.n.x; wp[A_NY] = m.n.y;
    wp[A_U] = shape_friction(ia & 255) * shape_friction(ib & 255);
    if (state == ARB_CACHED) state = ARB_FIRST;
    AIF(e, slot, AI_KEY) = key | (ba << 16) | (bb << 24);
    AIF(e, slot, AI_STAMP) = stamp;
    AIF(e, slot, AI_FLAGS) = state | (m.count << 8) | (1 << 16);
    AIF(e, slot, AI_HASH) = (int)(m.hash[0] | (m.count > 1 ? m.hash[1] << 16 : 0u));
    if (pooled) e.ainfo[q] = ba | (bb << 8) | (m.count << 16) | (state << 24);
    return slot;
}

/* after the last iteration of a sub-step: jn / jt of the pooled arbiters go back to their cache slots (warm start) */
__device__ __forceinline__ void arb_pool_writeback(EnvRef e, const unsigned char *order, int n_live)
{
    const int n = n_live < e.apn ? n_live : e.apn;
    MG_NOUNROLL
    for (int q = 0; q < n; q++) {
        const int s = order[q], count = (e.ainfo[q] >> 16) & 255;
        const double *c = e.ap + (size_t)q * A_NF + A_CON0;
        MG_NOUNROLL
        for (int k = 0; k < count; k++) { ACF(e, s, k, AC_JN) = c[k * AC_NF + AC_JN]; ACF(e, s, k, AC_JT) = c[k * AC_NF + AC_JT]; }
    }
}

__device__ void arb_prestep(EnvRef e, const unsigned char *order, int q, double slop, double bias_c)
{
    const double dt = MG_DT;
    const ArbRef r = arb_ref(e, order, q);
    const int a = r.a, b = r.b;
    vec2 n = V(APF(r, A_NX), APF(r, A_NY));
    vec2 body_delta = vsub(body_p(e, b), body_p(e, a));
    MG_NOUNROLL
    for (int k = 0; k < r.count; k++) {
        vec2 r1 = V(APC(r, k, AC_R1X), APC(r, k, AC_R1Y)), r2 = V(APC(r, k, AC_R2X), APC(r, k, AC_R2Y));
        APC(r, k, AC_NMASS) = 1.0 / k_scalar(e, a, b, r1, r2, n);
        APC(r, k, AC_TMASS) = 1.0 / k_scalar(e, a, b, r1, r2, vperp(n));
        double dist = vdot(vadd(vsub(r2, r1), body_delta), n);
        APC(r, k, AC_BIAS) = -bias_c * cp_min(0.0, dist + slop) / dt;
        APC(r, k, AC_JBIAS) = 0.0;
        /* bounce = normal_relative_velocity * e with e = 0 for every shape */
    }
}

__device__ void arb_apply_cached(EnvRef e, const unsigned char *order, int q, double dt_coef)
{
    const ArbRef r = arb_ref(e, order, q);
    if (r.state == ARB_FIRST) return;
    const int a = r.a, b = r.b;
    vec2 n = V(APF(r, A_NX), APF(r, A_NY));
    MG_NOUNROLL
    for (int k = 0; k < r.count; k++) {
        vec2 r1 = V(APC(r, k, AC_R1X), APC(r, k, AC_R1Y)), r2 = V(APC(r, k, AC_R2X), APC(r, k, AC_R2Y));
        vec2 j = vrot(n, V(APC(r, k, AC_JN), APC(r, k, AC_JT)));
        apply_impulses(e, a, b, r1, r2, vmul(j, dt_coef));
    }
}

__device__ void arb_apply_impulse(EnvRef e, const unsigned char *order, int q)
{
    const ArbRef r = arb_ref(e, order, q);
    const int a = r.a, b = r.b;
    vec2 n = V(APF(r, A_NX), APF(r, A_NY));
    double friction = APF(r, A_U);
    MG_NOUNROLL
    for (int k = 0; k < r.count; k++) {
        double n_mass = APC(r, k, AC_NMASS);
        vec2 r1 = V(APC(r, k, AC_R1X), APC(r, k, AC_R1Y)), r2 = V(APC(r, k, AC_R2X), APC(r, k, AC_R2Y));
        vec2 vb1 = vadd(V(BF(e, a, B_VBX), BF(e, a, B_VBY)), vmul(vperp(r1), BF(e, a, B_WB)));
        vec2 vb2 = vadd(V(BF(e, b, B_VBX), BF(e, b, B_VBY)), vmul(vperp(r2), BF(e, b, B_WB)));
        vec2 vr = relative_velocity(e, a, b, r1, r2);
        double vbn = vdot(vsub(vb2, vb1), n);
        double vrn = vdot(vr, n);
        double vrt = vdot(vr, vperp(n));
        double jbn = (APC(r, k, AC_BIAS) - vbn) * n_mass;
        double jbn_old = APC(r, k, AC_JBIAS);
        double j_bias = cp_max(jbn_old + jbn, 0.0);
        APC(r, k, AC_JBIAS) = j_bias;
        double jn = -(0.0 + vrn) * n_mass;
        double jn_old = APC(r, k, AC_JN);
        double jn_acc = cp_max(jn_old + jn, 0.0);
        APC(r, k, AC_JN) = jn_acc;
        double jt_max = friction * jn_acc;
        double jt = -vrt * APC(r, k, AC_TMASS);
        double jt_old = APC(r, k, AC_JT);
        double jt_acc = dclamp(jt_old + jt, -jt_max, jt_max);
        APC(r, k, AC_JT) = jt_acc;
        vec2 jb = vmul(n, j_bias - jbn_old);
        apply_bias_impulse(e, a, vneg(jb), r1);
        apply_bias_impulse(e, b, jb, r2);
        apply_impulses(e, a, b, r1, r2, vrot(n, V(jn_acc - jn_old, jt_acc - jt_old)));
    }
}

__device__ void arb_copy(EnvRef e, int dst, int src)
{
    MG_NOUNROLL
    for (int f = 0; f < A_NF; f++) AF(e, dst, f) = AF(e, src, f);
    MG_NOUNROLL
    for (int f = 0; f < 4; f++) AIF(e, dst, f) = AIF(e, src, f);
}


#ifndef PHYS_DBG_SKIP
#define PHYS_DBG_SKIP 0    /* timing experiments only: 1 skips the serial half, 2 the parallel half, 3 the pair loop, 4 boxes + pair loop, 5 constraint pre-steps, 6 position integration */
#endif
#ifndef PHYS_LANES
#define PHYS_LANES 16
#endif
/* PHYS_LANES: lanes that share one env in the parallel phases */
#ifndef PHYS_THREADS
#define PHYS_THREADS 256
#endif
#define PHYS_ENVS (PHYS_THREADS / PHYS_LANES)      /* envs per CTA; lanes 0..PHYS_ENVS-1 of warp 0 run their serial halves */

/* cpSpaceStep(space, dt), first half, one group of PHYS_LANES lanes per env: the phases with independent items (bodies, shape
 * pairs, pre-steps) spread over the lanes.  Whatever depends on Chipmunk's order runs on one lane in that
 * order (arbiter cache updates here, the rest in solve_lead / solve_helper), so the result is the same bit pattern a
 * serial cpSpaceStep produces.  `order` = shared scratch [MG_NARB + 8]; returns the number of live arbiters. */
/* ------------------------------------------------------------------ group-cooperative narrow phase
 * The lanes of an env's group search the faces / feature pairs of ONE shape pair together.  Every candidate
 * (a face separation, a closest-feature distance) is computed by the same expression as in the serial
 * collide_* of mg_core.cuh, and the winner is the one the serial loops keep: the first strict maximum /
 * minimum in loop order (ties go to the lower loop index).  The results are therefore bit-identical. */
#define GROUP_FOR_XOR for (int off_ = PHYS_LANES / 2; off_ > 0; off_ >>= 1)

/* first maximum: (v, k) over the group, lowest k among equal v */
__device__ __forceinline__ void group_argmax(unsigned gmask, double &v, int &k)
{
    GROUP_FOR_XOR {
        const double v2 = __shfl_xor_sync(gmask, v, off_);
        const int k2 = __shfl_xor_sync(gmask, k, off_);
        if (v2 > v || (v2 == v && k2 < k)) { v = v2; k = k2; }
    }
}
__device__ __forceinline__ void group_argmin(unsigned gmask, double &v, int &k)
{
    GROUP_FOR_XOR {
        const double v2 = __shfl_xor_sync(gmask, v, off_);
        const int k2 = __shfl_xor_sync(gmask, k, off_);
        if (v2 < v || (v2 == v && k2 < k)) { v = v2; k = k2; }
    }
}

/* wshape_from_tpl with one vertex per lane */
__device__ void coop_wshape(const MgTables *T, int tpl, vec2 p, vec2 rot, WShape *s, int gl)
{
    if (tpl >= MG_TPL_CIRCLE_ROBOT) {
        if (gl == 0) wshape_from_tpl(T, tpl, p, rot, s);
        return;
    }
    const int n = T->tpl_n[tpl];
    if (gl == 0) { s->kind = MG_SHAPE_POLY; s->n = n; s->r = T->tpl_r[tpl]; }
    MG_NOUNROLL
    for (int i = gl; i < n; i += PHYS_LANES) {
        vec2 l = T->tpl_v[tpl][i], m = T->tpl_nrm[tpl][i];
        s->v[i] = V(rot.x * l.x - rot.y * l.y + p.x, rot.y * l.x + rot.x * l.y + p.y);
        s->nrm[i] = V(rot.x * m.x - rot.y * m.y, rot.y * m.x + rot.x * m.y);
    }
}

/* collide_convex_convex (PolyToPoly / SegmentToPoly): lane = face, then lane = (edge, vertex) pair; the lane
 * that owns the decisive feature pair writes the manifold */
__device__ void coop_convex_convex(const WShape *A, const WShape *B, Manifold *m, int gl, unsigned gmask)
{
    const int An = A->n, Bn = B->n;
    const double rsum = A->r + B->r;
    /* faces of A, then faces of B, in the serial loop order */
    double best = -INFINITY;
    int kbest = 0x7fffffff;
    MG_NOUNROLL
    for (int f = gl; f < An + Bn; f += PHYS_LANES) {
        const bool from_a = f < An;
        const WShape *P = from_a ? A : B, *Q = from_a ? B : A;
        const int fi = from_a ? f : f - An;
        const vec2 fn = face_normal(P, fi), pv = P->v[fi];
        double sep = INFINITY;
        MG_NOUNROLL
        for (int j = 0; j < Q->n; j++) {
            double sj = vdot(vsub(Q->v[j], pv), fn);
            if (sj < sep) sep = sj;
        }
        if (sep > best) { best = sep; kbest = f; }
    }
    group_argmax(gmask, best, kbest);
    const vec2 axis = kbest < An ? face_normal(A, kbest) : vneg(face_normal(B, kbest - An));
    if (best <= 0.0 || best > rsum) {
        /* overlapping cores (shallowest face), or too far apart to touch */
        if (gl == 0 && best - A->r - B->r <= 0.0) contact_points(support_edge(A, axis), support_edge(B, vneg(axis)), axis, best, m);
        return;
    }
    /* closest features: items k = (edge i of A, vertex j of B), then (edge j of B, vertex i of A) */
    double bestd2 = INFINITY, my_t = 0.0;
    int kmin = 0x7fffffff;
    vec2 my_dl = V(0.0, 0.0);
    const int half = An * Bn;
    MG_NOUNROLL
    for (int k = gl; k < 2 * half; k += PHYS_LANES) {
        vec2 ea, eb, pt;
        const bool first = k < half;
        if (first) {
            const int i = k / Bn, j = k - i * Bn;
            ea = A->v[(i - 1 + An) % An]; eb = A->v[i]; pt = B->v[j];
        } else {
            const int kk = k - half;
            const int j = kk / An, i = kk - j * An;
            ea = B->v[(j - 1 + Bn) % Bn]; eb = B->v[j]; pt = A->v[i];
        }
        double t;
        const vec2 q = seg_closest(ea, eb, pt, &t);
        const vec2 dl = first ? vsub(pt, q) : vsub(q, pt);
        const double d2 = vdot(dl, dl);
        if (d2 < bestd2) { bestd2 = d2; kmin = k; my_t = t; my_dl = dl; }
    }
    const int my_k = kmin;
    group_argmin(gmask, bestd2, kmin);
    if (my_k == kmin && kmin != 0x7fffffff) {
        vec2 bn; double bd;
        if (my_t > 0.0 && my_t < 1.0) {
            bn = kmin < half ? face_normal(A, kmin / Bn) : vneg(face_normal(B, (kmin - half) / An));
            bd = vdot(my_dl, bn);
            if (bd < 0.0) { bn = vneg(bn); bd = -bd; }
        } else {
            double len = sqrt(bestd2); bd = len; bn = vmul(my_dl, 1.0 / (len + MG_DBL_MIN));
        }
        if (bd - A->r - B->r <= 0.0) contact_points(support_edge(A, bn), support_edge(B, vneg(bn)), bn, bd, m);
    }
}

/* collide_circle_poly: lane = face, then lane = edge */
__device__ void coop_circle_poly(const WShape *c, const WShape *p, Manifold *m, int gl, unsigned gmask)
{
    const vec2 center = c->v[0];
    const int n = p->n;
    double best = -INFINITY;
    int bi = 0x7fffffff;
    MG_NOUNROLL
    for (int i = gl; i < n; i += PHYS_LANES) {
        double sv = vdot(vsub(center, p->v[i]), p->nrm[i]);
        if (sv > best) { best = sv; bi = i; }
    }
    group_argmax(gmask, best, bi);
    const double rsum = c->r + p->r;
    if (best <= 0.0) {
        if (gl == 0) {
            vec2 nrm = vneg(p->nrm[bi]);
            vec2 pb = vsub(center, vmul(p->nrm[bi], best));
            if (best <= rsum) { m->n = nrm; mf_push(m, vadd(center, vmul(nrm, c->r)), vadd(pb, vmul(nrm, -p->r)), 0); }
        }
        return;
    }
    if (best > rsum) return;
    double bestd2 = INFINITY, my_t = 0.0;
    int imin = 0x7fffffff;
    vec2 my_dl = V(0.0, 0.0), my_q = center;
    MG_NOUNROLL
    for (int i = gl; i < n; i += PHYS_LANES) {
        double t;
        const vec2 q = seg_closest(p->v[(i - 1 + n) % n], p->v[i], center, &t);
        const vec2 dl = vsub(q, center);
        const double d2 = vdot(dl, dl);
        if (d2 < bestd2) { bestd2 = d2; imin = i; my_t = t; my_dl = dl; my_q = q; }
    }
    const int my_i = imin;
    group_argmin(gmask, bestd2, imin);
    if (my_i == imin && imin != 0x7fffffff) {
        vec2 nrm; double d;
        if (my_t > 0.0 && my_t < 1.0) { nrm = vneg(p->nrm[imin]); d = vdot(my_dl, nrm); }
        else { double len = sqrt(bestd2); d = len; nrm = vmul(my_dl, 1.0 / (len + MG_DBL_MIN)); }
        if (d <= rsum) { m->n = nrm; mf_push(m, vadd(center, vmul(nrm, c->r)), vadd(my_q, vmul(nrm, -p->r)), 0); }
    }
}

/* collide_shapes (cpCollide) for the whole group: a->kind <= b->kind; returns the contact count (uniform) */
__device__ int coop_collide(const WShape *a, const WShape *b, Manifold *m, int gl, unsigned gmask)
{
    if (gl == 0) { m->count = 0; m->n = V(0.0, 0.0); }
    __syncwarp(gmask);
    if (a->kind == MG_SHAPE_CIRCLE) {
        if (b->kind == MG_SHAPE_POLY) coop_circle_poly(a, b, m, gl, gmask);
        else if (gl == 0) { if (b->kind == MG_SHAPE_CIRCLE) collide_circle_circle(a, b, m); else collide_circle_segment(a, b, m); }
    } else if (!(a->kind == MG_SHAPE_SEGMENT && b->kind == MG_SHAPE_SEGMENT)) {
        coop_convex_convex(a, b, m, gl, gmask);
    }
    __syncwarp(gmask);
    return m->count;
}


/* ---- section timers (-DMG_SECTION_TIMERS, tools/section_timers.py): clock64 deltas of the phases of a CTA's step, taken
 * on thread 0 (lane 0 of the warp that runs the serial halves), summed over the ten sub-steps.  Not in product builds. */
#ifdef MG_SECTION_TIMERS
#define MG_NTIMERS 16
#define MG_TIMER_CTAS 8192
__device__ long long mg_phys_timers[MG_TIMER_CTAS * MG_NTIMERS];
__device__ __forceinline__ void tm_add(long long *tm, int k, long long &t)
{
    const long long now = clock64();
    if (tm) tm[k] += now - t;
    t = now;
}
#define TM_DECL(tm_) long long tm_t_ = clock64(); long long *tm_p_ = (tm_)
#define TM_ADD(k_) tm_add(tm_p_, (k_), tm_t_)
enum { TM_PARALLEL, TM_BARRIER, TM_SERIAL, TM_SPRING, TM_CACHED, TM_ARB, TM_BLOCKCON, TM_ROBOTCON, TM_TOTAL, TM_STAGE, TM_REGS,
       TM_P_INTEG, TM_P_BB, TM_P_PAIRS, TM_P_PRESTEP };
extern "C" int mg_debug_read_timers(long long *out_host, int n_ctas)
{
    if (n_ctas > MG_TIMER_CTAS) n_ctas = MG_TIMER_CTAS;
    cudaDeviceSynchronize();
    cudaError_t e = cudaMemcpyFromSymbol(out_host, mg_phys_timers, (size_t)n_ctas * MG_NTIMERS * sizeof(long long));
    static long long zeros[MG_TIMER_CTAS * MG_NTIMERS];
    cudaMemcpyToSymbol(mg_phys_timers, zeros, sizeof(zeros));
    return e == cudaSuccess ? MG_NTIMERS : -1;
}
#else
#define TM_DECL(tm_)
#define TM_ADD(k_)
#endif

#ifndef PHYS_LEAD_ROTATE
#define PHYS_LEAD_ROTATE 1
#endif
#define NARROW_SLOTS 1
/* shape i of the current row, and shape j + manifold of up to NARROW_SLOTS pairs in flight */
struct NarrowScratch { WShape si; struct { WShape sj; Manifold mf; } slot[NARROW_SLOTS]; };

/* returns the number of live arbiters; bit 8: one of them is on a block's body */
__device__ int space_step_parallel(EnvRef e, const MgTables *T, int lane, unsigned gmask, unsigned char *order, NarrowScratch *scratch, double *bbs, long long *tm = nullptr)
{
    int on_block = 0;
    TM_DECL(tm);
    const int gl = lane & (PHYS_LANES - 1);     /* lane within the env's group */
    const double dt = MG_DT;
    const int prev_stamp = IF(e, I_STAMP);
    const int stamp = prev_stamp + 1;
    __syncwarp(gmask);

    /* cpSpaceArbiterSetFilter of the previous step (same effect as running it at the end of
     * that step) merged with this step's arbiter-list reset */
    if (gl == 0) {
        IF(e, I_STAMP) = stamp;
        int n_arbs = IF(e, I_NARBS), w = 0;
        MG_NOUNROLL
        for (int q = 0; q < n_arbs; q++) {
            int fl = AIF(e, q, AI_FLAGS);
            int ticks = prev_stamp - AIF(e, q, AI_STAMP);
            int state = fl & 255;
            if ((fl >> 16) & 1) state = ARB_NORMAL;
            if (ticks >= 1 && state != ARB_CACHED) state = ARB_CACHED;
            if (ticks >= 3) continue;                     /* collision_persistence = 3 */
            AIF(e, q, AI_FLAGS) = state | (fl & 0xff00);  /* live cleared */
            if (w != q) arb_copy(e, w, q);
            w++;
        }
        IF(e, I_NARBS) = w;
    }

    /* cpBodyUpdatePosition for dynamic and kinematic bodies: lane = body */
    const int nb = IF(e, I_NBODIES);
    MG_NOUNROLL
    for (int b = 1 + gl; b < (PHYS_DBG_SKIP == 6 ? 0 : nb); b += PHYS_LANES) {
        BF(e, b, B_PX) = BF(e, b, B_PX) + (BF(e, b, B_VX) + BF(e, b, B_VBX)) * dt;
        BF(e, b, B_PY) = BF(e, b, B_PY) + (BF(e, b, B_VY) + BF(e, b, B_VBY)) * dt;
        set_angle(e, b, BF(e, b, B_A) + (BF(e, b, B_W) + BF(e, b, B_WB)) * dt);
        BF(e, b, B_VBX) = 0.0; BF(e, b, B_VBY) = 0.0; BF(e, b, B_WB) = 0.0;
    }
    __syncwarp(gmask);

    TM_ADD(TM_P_INTEG);
    /* narrow phase over shape pairs in add order (i < j): for shape i the env's lanes take consecutive j,
     * run the bounding test and the narrow phase in parallel, then the hits update the arbiter cache one
     * after the other in j order -- the order a serial double loop discovers them in */
    int n_live = 0;
    const int ns = IF(e, I_NSHAPES);
    /* cpShapeCacheBB once per shape and sub-step (lane = shape): the pair loop below only compares boxes */
    MG_NOUNROLL
    for (int s = gl; s < (PHYS_DBG_SKIP == 4 ? 0 : ns); s += PHYS_LANES) {
        const int info = IF(e, I_SHAPE0 + s);
        const int body = (info >> 8) & 255;
        tpl_world_bb(T, info & 255, body_p(e, body), body_rot(e, body), bbs + 4 * s);
    }
    __syncwarp(gmask);
    TM_ADD(TM_P_BB);
    MG_NOUNROLL
    for (int i = 0; i + 1 < (PHYS_DBG_SKIP == 3 || PHYS_DBG_SKIP == 4 ? 0 : ns); i++) {
        const int info_i = IF(e, I_SHAPE0 + i);
        const int tpl_i = info_i & 255, body_i = (info_i >> 8) & 255, ent_i = info_i >> 16;
        const double bi0 = bbs[4 * i], bi1 = bbs[4 * i + 1], bi2 = bbs[4 * i + 2], bi3 = bbs[4 * i + 3];
        bool si_built = false;
        MG_NOUNROLL
        for (int j0 = i + 1; j0 < ns; j0 += PHYS_LANES) {
            const int j = j0 + gl;
            int info_j = 0;
            bool cand = false;
            if (j < ns) {
                info_j = IF(e, I_SHAPE0 + j);
                /* same body, ShapeFilter group (robot: 1, star pieces), static-static */
                cand = ent_i != (info_j >> 16);
                if (cand) {
                    /* bounding boxes (Chipmunk's own broadphase criterion): shapes whose boxes are apart by
                     * more than rounding noise cannot produce a contact */
                    const double *bj = bbs + 4 * j;
                    const double eps = 1e-9;
                    if (bi0 > bj[2] + eps || bj[0] > bi2 + eps || bi1 > bj[3] + eps || bj[1] > bi3 + eps) cand = false;
                }
            }
            /* narrow phase for the surviving pairs of this row, up to NARROW_SLOTS at a time on their own
             * lanes, then cpArbiterUpdate for the hits one after the other in j order (the order a serial
             * double loop discovers them in).  The shapes live in the env's shared scratch: per-lane copies
             * would sit in local memory, which has next to no L1 left beside the env state. */
            unsigned todo = __ballot_sync(gmask, cand);
            if (PHYS_DBG_SKIP == 8) todo = 0;
            if (todo && !si_built) {
                coop_wshape(T, tpl_i, body_p(e, body_i), body_rot(e, body_i), &scratch->si, gl);
                si_built = true;
            }
            /* the surviving pairs of this row one after the other in j order (the order a serial double loop
             * discovers them in), each searched by all lanes of the group; a hit goes through cpArbiterUpdate
             * on the group's first lane */
            while (todo) {
                const int src = __ffs(todo) - 1;
                todo &= todo - 1;
                const int info_s = __shfl_sync(gmask, info_j, src);
                const int js = j0 + (src & (PHYS_LANES - 1));
                const int tpl_s = info_s & 255, body_s = (info_s >> 8) & 255;
                const WShape *Si = &scratch->si;
                WShape *Sj = &scratch->slot[0].sj;
                Manifold *mf = &scratch->slot[0].mf;
                coop_wshape(T, tpl_s, body_p(e, body_s), body_rot(e, body_s), Sj, gl);
                __syncwarp(gmask);
                const bool swapped = Si->kind > Sj->kind;            /* cpSpaceCollideShapes orders by cpShapeType */
                int cnt = swapped ? coop_collide(Sj, Si, mf, gl, gmask) : coop_collide(Si, Sj, mf, gl, gmask);
                if (PHYS_DBG_SKIP == 9) cnt = 0;
                if (cnt) {
                    if (gl == 0) {
                        int slot = swapped ? arbiter_update(e, js, i, info_s, info_i, *mf, stamp, n_live)
                                           : arbiter_update(e, i, js, info_i, info_s, *mf, stamp, n_live);
                        order[MG_NARB] = (unsigned char)(slot >= 0);
                        if (slot >= 0) order[n_live] = (unsigned char)slot;
                    }
                    __syncwarp(gmask);
                    n_live += order[MG_NARB];
                    if (body_i >= MG_BODY_BLOCK0 || body_s >= MG_BODY_BLOCK0) on_block = 256;
                }
                __syncwarp(gmask);
            }
        }
    }

    TM_ADD(TM_P_PAIRS);
    /* pre-step: arbiters and constraints only write their own fields ... */
    const double slop = 0.01;                       /* base_env.py:207 */
    const double bias_c = T->bias_coef_default;     /* 1 - collision_bias^dt */
    const int nc = IF(e, I_NCON);
    MG_NOUNROLL
    for (int q = gl; q < n_live; q += PHYS_LANES) arb_prestep(e, order, q, slop, bias_c);
    MG_NOUNROLL
    for (int c = gl; c < (PHYS_DBG_SKIP == 5 ? 0 : nc); c += PHYS_LANES)
        if ((IF(e, I_CON0 + c) & 255) != CON_SPRING) con_prestep(e, c);
    __syncwarp(gmask);
    TM_ADD(TM_P_PRESTEP);
    return PHYS_DBG_SKIP == 7 ? 0 : (n_live ? n_live | on_block : 0);
}

/* ---- the robot's ten constraints with the six robot bodies' velocities in registers.
 * Robot.setup adds them in a fixed order on fixed bodies (entities.py:238-372), so inside the iteration loop
 * their body operands can be registers instead of shared-memory round trips, and with no stores to body
 * memory in between the compiler is free to load the next constraint's coefficients early.  Every formula is
 * the one in con_apply_impulse, operation for operation. */
struct BodyV { double vx, vy, w, minv, iinv; };

template <int B, int N>
__device__ __forceinline__ void reg_apply_impulse(BodyV (&r)[N], vec2 j, vec2 rr)
{
    r[B].vx = r[B].vx + j.x * r[B].minv;
    r[B].vy = r[B].vy + j.y * r[B].minv;
    r[B].w = r[B].w + r[B].iinv * vcross(rr, j);
}

template <int A, int B, int N>
__device__ __forceinline__ void reg_pivot(BodyV (&r)[N], double *cf)
{
    const double j_max = cf[C_JMAX];
    vec2 r1 = V(cf[C_R1X], cf[C_R1Y]), r2 = V(cf[C_R2X], cf[C_R2Y]);
    vec2 v1 = vadd(V(r[A].vx, r[A].vy), vmul(vperp(r1), r[A].w));
    vec2 v2 = vadd(V(r[B].vx, r[B].vy), vmul(vperp(r2), r[B].w));
    vec2 vr = vsub(v2, v1);
    vec2 d = vsub(V(cf[C_BIASX], cf[C_BIASY]), vr);
    vec2 j = V(d.x * cf[C_K0] + d.y * cf[C_K1], d.x * cf[C_K2] + d.y * cf[C_K3]);
    vec2 j_old = V(cf[C_JX], cf[C_JY]);
    vec2 j_acc = vclamp_len(vadd(j_old, j), j_max);
    cf[C_JX] = j_acc.x; cf[C_JY] = j_acc.y;
    vec2 dj = vsub(j_acc, j_old);
    reg_apply_impulse<A>(r, vneg(dj), r1);
    reg_apply_impulse<B>(r, dj, r2);
}

template <int A, int B, int N>
__device__ __forceinline__ void reg_gear(BodyV (&r)[N], double *cf)
{
    const double j_max = cf[C_JMAX];
    double ratio = cf[C_P1];
    double wr = r[B].w * ratio - r[A].w;
    double j = (cf[C_BIASX] - wr) * cf[C_ISUM];
    double j_old = cf[C_JX];
    double j_acc = dclamp(j_old + j, -j_max, j_max);
    cf[C_JX] = j_acc;
    j = j_acc - j_old;
    r[A].w = r[A].w - j * r[A].iinv * (1.0 / ratio);
    r[B].w = r[B].w + j * r[B].iinv;
}

template <int A, int B>
__device__ __forceinline__ void reg_spring(BodyV (&r)[MG_RB_COUNT], double *cf)
{
    double wrn = r[A].w - r[B].w;
    double w_damp = (cf[C_BIASY] - wrn) * cf[C_P0];
    cf[C_BIASY] = wrn + w_damp;
    double j_damp = w_damp * cf[C_ISUM];
    cf[C_JX] = cf[C_JX] + j_damp;
    r[A].w = r[A].w + j_damp * r[A].iinv;
    r[B].w = r[B].w - j_damp * r[B].iinv;
}

template <int A, int B>
__device__ __forceinline__ void reg_limit(BodyV (&r)[MG_RB_COUNT], double *cf)
{
    const double j_max = cf[C_JMAX];
    double bias = cf[C_BIASX];
    if (!bias) return;
    double wr = r[B].w - r[A].w;
    double j = -(bias + wr) * cf[C_ISUM];
    double j_old = cf[C_JX], j_acc;
    if (bias < 0.0) j_acc = dclamp(j_old + j, 0.0, j_max);
    else j_acc = dclamp(j_old + j, -j_max, 0.0);
    cf[C_JX] = j_acc;
    j = j_acc - j_old;
    r[A].w = r[A].w - j * r[A].iinv;
    r[B].w = r[B].w + j * r[B].iinv;
}

template <int A, int B>
__device__ __forceinline__ void reg_motor(BodyV (&r)[MG_RB_COUNT], double *cf)
{
    const double j_max = cf[C_JMAX];
    double wr = r[B].w - r[A].w + cf[C_P0];
    double j = -wr * cf[C_ISUM];
    double j_old = cf[C_JX];
    double j_acc = dclamp(j_old + j, -j_max, j_max);
    cf[C_JX] = j_acc;
    j = j_acc - j_old;
    r[A].w = r[A].w - j * r[A].iinv;
    r[B].w = r[B].w + j * r[B].iinv;
}

/* ---- the same formulas with the coefficients of a constraint loaded ahead of their use.
 * The chain through the robot body is latency-bound, and between two constraints sit divergent branches (clamps, the
 * limit joint's early return) that the compiler does not move loads across: each constraint then starts with a
 * shared-memory round trip on the critical path.  robot_constraints_pass therefore loads constraint k + 1's coefficients
 * (PivotC / ScalarC) before it applies constraint k.  Values and operation order are those of reg_pivot ... reg_motor. */
struct PivotC { double jmax, r1x, r1y, r2x, r2y, bx, by, k0, k1, k2, k3, jx, jy, ma, ia, mb, ib; };
struct ScalarC { double jmax, p, bias, isum, jx, ia, ib; };   /* p: gear ratio | spring w_coef | motor rate;  bias: spring target_wrn */
/* (ma, ia, mb, ib: inverse mass / inertia of the constraint's two bodies.  They are constants of the step, so they ride
 * with the prefetched coefficients instead of occupying registers through the whole iteration loop.) */
struct BodyVel { double vx, vy, w; };

/* bd: the env's body block in shared memory; A, B: robot-group body indices (MG_RB_*) */
template <int A, int B>
__device__ __forceinline__ PivotC ld_pivot(const double *cf, const double *bd)
{
    const double *ba = bd + (MG_BODY_ROBOT + A) * B_NF, *bb = bd + (MG_BODY_ROBOT + B) * B_NF;
    PivotC c = { cf[C_JMAX], cf[C_R1X], cf[C_R1Y], cf[C_R2X], cf[C_R2Y], cf[C_BIASX], cf[C_BIASY],
                 cf[C_K0], cf[C_K1], cf[C_K2], cf[C_K3], cf[C_JX], cf[C_JY], ba[B_MINV], ba[B_IINV], bb[B_MINV], bb[B_IINV] };
    return c;
}
template <int A, int B>
__device__ __forceinline__ ScalarC ld_scalar(const double *cf, const double *bd, int kind)
{
    const double *ba = bd + (MG_BODY_ROBOT + A) * B_NF, *bb = bd + (MG_BODY_ROBOT + B) * B_NF;
    ScalarC c;
    c.jmax = cf[C_JMAX]; c.isum = cf[C_ISUM]; c.jx = cf[C_JX]; c.ia = ba[B_IINV]; c.ib = bb[B_IINV];
    c.p = kind == CON_GEAR ? cf[C_P1] : cf[C_P0];
    c.bias = kind == CON_SPRING ? cf[C_BIASY] : cf[C_BIASX];
    return c;
}

template <int A, int B>
__device__ __forceinline__ void ap_pivot(BodyVel (&r)[MG_RB_COUNT], const PivotC &c, double *cf)
{
    vec2 r1 = V(c.r1x, c.r1y), r2 = V(c.r2x, c.r2y);
    vec2 v1 = vadd(V(r[A].vx, r[A].vy), vmul(vperp(r1), r[A].w));
    vec2 v2 = vadd(V(r[B].vx, r[B].vy), vmul(vperp(r2), r[B].w));
    vec2 vr = vsub(v2, v1);
    vec2 d = vsub(V(c.bx, c.by), vr);
    vec2 j = V(d.x * c.k0 + d.y * c.k1, d.x * c.k2 + d.y * c.k3);
    vec2 j_old = V(c.jx, c.jy);
    vec2 j_acc = vclamp_len(vadd(j_old, j), c.jmax);
    cf[C_JX] = j_acc.x; cf[C_JY] = j_acc.y;
    vec2 dj = vsub(j_acc, j_old), ndj = vneg(dj);
    r[A].vx = r[A].vx + ndj.x * c.ma;
    r[A].vy = r[A].vy + ndj.y * c.ma;
    r[A].w = r[A].w + c.ia * vcross(r1, ndj);
    r[B].vx = r[B].vx + dj.x * c.mb;
    r[B].vy = r[B].vy + dj.y * c.mb;
    r[B].w = r[B].w + c.ib * vcross(r2, dj);
}
/* inv_ratio = 1.0 / ratio, the division hoisted out of the iteration loop by the caller (same value every time) */
template <int A, int B>
__device__ __forceinline__ void ap_gear(BodyVel (&r)[MG_RB_COUNT], const ScalarC &c, double *cf, double inv_ratio)
{
    double ratio = c.p;
    double wr = r[B].w * ratio - r[A].w;
    double j = (c.bias - wr) * c.isum;
    double j_old = c.jx;
    double j_acc = dclamp(j_old + j, -c.jmax, c.jmax);
    cf[C_JX] = j_acc;
    j = j_acc - j_old;
    r[A].w = r[A].w - j * c.ia * inv_ratio;
    r[B].w = r[B].w + j * c.ib;
}
template <int A, int B>
__device__ __forceinline__ void ap_spring(BodyVel (&r)[MG_RB_COUNT], const ScalarC &c, double *cf)
{
    double wrn = r[A].w - r[B].w;
    double w_damp = (c.bias - wrn) * c.p;
    cf[C_BIASY] = wrn + w_damp;
    double j_damp = w_damp * c.isum;
    cf[C_JX] = c.jx + j_damp;
    r[A].w = r[A].w + j_damp * c.ia;
    r[B].w = r[B].w - j_damp * c.ib;
}
template <int A, int B>
__device__ __forceinline__ void ap_limit(BodyVel (&r)[MG_RB_COUNT], const ScalarC &c, double *cf)
{
    double bias = c.bias;
    if (!bias) return;
    double wr = r[B].w - r[A].w;
    double j = -(bias + wr) * c.isum;
    double j_old = c.jx, j_acc;
    if (bias < 0.0) j_acc = dclamp(j_old + j, 0.0, c.jmax);
    else j_acc = dclamp(j_old + j, -c.jmax, 0.0);
    cf[C_JX] = j_acc;
    j = j_acc - j_old;
    r[A].w = r[A].w - j * c.ia;
    r[B].w = r[B].w + j * c.ib;
}
template <int A, int B>
__device__ __forceinline__ void ap_motor(BodyVel (&r)[MG_RB_COUNT], const ScalarC &c, double *cf)
{
    double wr = r[B].w - r[A].w + c.p;
    double j = -wr * c.isum;
    double j_old = c.jx;
    double j_acc = dclamp(j_old + j, -c.jmax, c.jmax);
    cf[C_JX] = j_acc;
    j = j_acc - j_old;
    r[A].w = r[A].w - j * c.ia;
    r[B].w = r[B].w + j * c.ib;
}

/* one sequential-impulse pass over the robot's constraints c0 .. c0+9 (add order of setup_robot) with the next
 * constraint's coefficients loaded before the current one is applied */
__device__ __forceinline__ void robot_constraints_pass(BodyVel (&r)[MG_RB_COUNT], double *cf0, const double *bd, double gear_inv_ratio)
{
    const PivotC c0 = ld_pivot<MG_RB_CONTROL, MG_RB_ROBOT>(cf0 + 0 * C_NF, bd);
    const ScalarC c1 = ld_scalar<MG_RB_CONTROL, MG_RB_ROBOT>(cf0 + 1 * C_NF, bd, CON_GEAR);
    ap_pivot<MG_RB_CONTROL, MG_RB_ROBOT>(r, c0, cf0 + 0 * C_NF);
    const ScalarC c2 = ld_scalar<MG_RB_ROBOT, MG_RB_PUPIL_L>(cf0 + 2 * C_NF, bd, CON_SPRING);
    ap_gear<MG_RB_CONTROL, MG_RB_ROBOT>(r, c1, cf0 + 1 * C_NF, gear_inv_ratio);
    const ScalarC c3 = ld_scalar<MG_RB_ROBOT, MG_RB_PUPIL_R>(cf0 + 3 * C_NF, bd, CON_SPRING);
    ap_spring<MG_RB_ROBOT, MG_RB_PUPIL_L>(r, c2, cf0 + 2 * C_NF);
    const PivotC c4 = ld_pivot<MG_RB_ROBOT, MG_RB_FINGER_L>(cf0 + 4 * C_NF, bd);
    ap_spring<MG_RB_ROBOT, MG_RB_PUPIL_R>(r, c3, cf0 + 3 * C_NF);
    const ScalarC c5 = ld_scalar<MG_RB_ROBOT, MG_RB_FINGER_L>(cf0 + 5 * C_NF, bd, CON_LIMIT);
    ap_pivot<MG_RB_ROBOT, MG_RB_FINGER_L>(r, c4, cf0 + 4 * C_NF);
    const ScalarC c6 = ld_scalar<MG_RB_ROBOT, MG_RB_FINGER_L>(cf0 + 6 * C_NF, bd, CON_MOTOR);
    ap_limit<MG_RB_ROBOT, MG_RB_FINGER_L>(r, c5, cf0 + 5 * C_NF);
    const PivotC c7 = ld_pivot<MG_RB_ROBOT, MG_RB_FINGER_R>(cf0 + 7 * C_NF, bd);
    ap_motor<MG_RB_ROBOT, MG_RB_FINGER_L>(r, c6, cf0 + 6 * C_NF);
    const ScalarC c8 = ld_scalar<MG_RB_ROBOT, MG_RB_FINGER_R>(cf0 + 8 * C_NF, bd, CON_LIMIT);
    ap_pivot<MG_RB_ROBOT, MG_RB_FINGER_R>(r, c7, cf0 + 7 * C_NF);
    const ScalarC c9 = ld_scalar<MG_RB_ROBOT, MG_RB_FINGER_R>(cf0 + 9 * C_NF, bd, CON_MOTOR);
    ap_limit<MG_RB_ROBOT, MG_RB_FINGER_R>(r, c8, cf0 + 8 * C_NF);
    ap_motor<MG_RB_ROBOT, MG_RB_FINGER_R>(r, c9, cf0 + 9 * C_NF);
}

/* con_apply_cached for the robot's ten constraints on the register copies (same formulas, operation for operation) */
template <int A, int B>
__device__ __forceinline__ void reg_cached_pivot(BodyVel (&r)[MG_RB_COUNT], const double *cf, const double *bd, double dt_coef)
{
    const double *ba = bd + (MG_BODY_ROBOT + A) * B_NF, *bb = bd + (MG_BODY_ROBOT + B) * B_NF;
    const vec2 j = vmul(V(cf[C_JX], cf[C_JY]), dt_coef), nj = vneg(j);
    const vec2 r1 = V(cf[C_R1X], cf[C_R1Y]), r2 = V(cf[C_R2X], cf[C_R2Y]);
    r[A].vx = r[A].vx + nj.x * ba[B_MINV];
    r[A].vy = r[A].vy + nj.y * ba[B_MINV];
    r[A].w = r[A].w + ba[B_IINV] * vcross(r1, nj);
    r[B].vx = r[B].vx + j.x * bb[B_MINV];
    r[B].vy = r[B].vy + j.y * bb[B_MINV];
    r[B].w = r[B].w + bb[B_IINV] * vcross(r2, j);
}
template <int A, int B>
__device__ __forceinline__ void reg_cached_gear(BodyVel (&r)[MG_RB_COUNT], const double *cf, const double *bd, double dt_coef)
{
    const double *ba = bd + (MG_BODY_ROBOT + A) * B_NF, *bb = bd + (MG_BODY_ROBOT + B) * B_NF;
    const double j = cf[C_JX] * dt_coef;
    r[A].w = r[A].w - j * ba[B_IINV] * (1.0 / cf[C_P1]);
    r[B].w = r[B].w + j * bb[B_IINV];
}
template <int A, int B>
__device__ __forceinline__ void reg_cached_scalar(BodyVel (&r)[MG_RB_COUNT], const double *cf, const double *bd, double dt_coef)   /* limit, motor */
{
    const double *ba = bd + (MG_BODY_ROBOT + A) * B_NF, *bb = bd + (MG_BODY_ROBOT + B) * B_NF;
    const double j = cf[C_JX] * dt_coef;
    r[A].w = r[A].w - j * ba[B_IINV];
    r[B].w = r[B].w + j * bb[B_IINV];
}
__device__ __forceinline__ void robot_cached_pass(BodyVel (&r)[MG_RB_COUNT], const double *cf0, const double *bd, double dt_coef)
{
    reg_cached_pivot<MG_RB_CONTROL, MG_RB_ROBOT>(r, cf0 + 0 * C_NF, bd, dt_coef);
    reg_cached_gear<MG_RB_CONTROL, MG_RB_ROBOT>(r, cf0 + 1 * C_NF, bd, dt_coef);
    /* 2, 3: damped rotary springs have no cached impulse */
    reg_cached_pivot<MG_RB_ROBOT, MG_RB_FINGER_L>(r, cf0 + 4 * C_NF, bd, dt_coef);
    reg_cached_scalar<MG_RB_ROBOT, MG_RB_FINGER_L>(r, cf0 + 5 * C_NF, bd, dt_coef);
    reg_cached_scalar<MG_RB_ROBOT, MG_RB_FINGER_L>(r, cf0 + 6 * C_NF, bd, dt_coef);
    reg_cached_pivot<MG_RB_ROBOT, MG_RB_FINGER_R>(r, cf0 + 7 * C_NF, bd, dt_coef);
    reg_cached_scalar<MG_RB_ROBOT, MG_RB_FINGER_R>(r, cf0 + 8 * C_NF, bd, dt_coef);
    reg_cached_scalar<MG_RB_ROBOT, MG_RB_FINGER_R>(r, cf0 + 9 * C_NF, bd, dt_coef);
}

__device__ __forceinline__ void robot_vel_load(EnvRef e, BodyVel (&r)[MG_RB_COUNT])
{
#pragma unroll
    for (int k = 0; k < MG_RB_COUNT; k++) {
        const int b = MG_BODY_ROBOT + k;
        r[k].vx = BF(e, b, B_VX); r[k].vy = BF(e, b, B_VY); r[k].w = BF(e, b, B_W);
    }
}
__device__ __forceinline__ void robot_vel_store(EnvRef e, const BodyVel (&r)[MG_RB_COUNT])
{
#pragma unroll
    for (int k = 0; k < MG_RB_COUNT; k++) {
        const int b = MG_BODY_ROBOT + k;
        BF(e, b, B_VX) = r[k].vx; BF(e, b, B_VY) = r[k].vy; BF(e, b, B_W) = r[k].w;
    }
}


/* con_apply_cached for the floor-friction pair of one block (PivotJoint c, GearJoint c + 1 between the static body and
 * block body b, Shape.setup, entities.py:700-708), block side only: the static body has zero inverse mass and inertia,
 * so every impulse leaves its velocity at +0.0 */
__device__ __forceinline__ void block_pair_cached(EnvRef e, int c, int b, double dt_coef)
{
    const double *cf = e.dc + (size_t)c * C_NF;
    const double minv = BF(e, b, B_MINV), iinv = BF(e, b, B_IINV);
    const vec2 j = vmul(V(cf[C_JX], cf[C_JY]), dt_coef);
    const vec2 r2 = V(cf[C_R2X], cf[C_R2Y]);
    BF(e, b, B_VX) = BF(e, b, B_VX) + j.x * minv;
    BF(e, b, B_VY) = BF(e, b, B_VY) + j.y * minv;
    double w = BF(e, b, B_W) + iinv * vcross(r2, j);
    const double jg = cf[C_NF + C_JX] * dt_coef;
    BF(e, b, B_W) = w + jg * iinv;
}

/* one sequential-impulse application of that pair: block_constraints_pass for a single block */
__device__ __forceinline__ void block_pair_pass(EnvRef e, int c, int b)
{
    BodyV r[2];
    r[0].vx = BF(e, 0, B_VX); r[0].vy = BF(e, 0, B_VY); r[0].w = BF(e, 0, B_W);
    r[0].minv = BF(e, 0, B_MINV); r[0].iinv = BF(e, 0, B_IINV);
    r[1].vx = BF(e, b, B_VX); r[1].vy = BF(e, b, B_VY); r[1].w = BF(e, b, B_W);
    r[1].minv = BF(e, b, B_MINV); r[1].iinv = BF(e, b, B_IINV);
    reg_pivot<0, 1>(r, e.dc + (size_t)c * C_NF);
    reg_gear<0, 1>(r, e.dc + (size_t)(c + 1) * C_NF);
    BF(e, b, B_VX) = r[1].vx; BF(e, b, B_VY) = r[1].vy; BF(e, b, B_W) = r[1].w;
}

/* The order-dependent rest of cpSpaceStep for all envs of a CTA, run by the whole CTA.
 * Constraints that share no dynamic body commute exactly (each reads and writes only its own bodies' velocities and its
 * own accumulators), and in these scenes the constraint graph is a star: the robot's ten constraints touch only the six
 * robot bodies, and every block has one PivotJoint + GearJoint pair to the static body.  So the robot chains of the
 * CTA's envs run side by side in warp 0 (lane k = env k, straight-line register code as before) while each block's pair
 * runs on its own lane of the other warps (lane = (env, block)); every body still sees its own updates in Chipmunk's
 * order.  Contacts can couple any two bodies, so they stay one ordered chain per env on the lead warp, fenced from the
 * constraint passes by CTA barriers -- only in CTAs where some env has a live contact ON A BLOCK (`any_live`,
 * CTA-uniform): contacts among the robot's bodies, the walls and nothing else (a robot against a wall, the common case)
 * touch no body a helper lane touches, so they commute with the blocks' pairs like the robot's constraints do. */
/* The two roles are separate out-of-line functions so that each gets its own register allocation (the robot chain
 * keeps six bodies' velocities and masses in registers through all ten iterations: inlined into the kernel it spilled
 * inside the constraint pass).  Both pass the same CTA barriers the same number of times (warp-uniform roles). */
/* (The env's shared-memory copy is addressed from the kernel's shared array inside the functions, not through pointer
 * arguments: behind a call boundary pointer arguments are generic addresses, and every body / constraint access of the
 * chain would become a generic load instead of an LDS.) */
struct PhysSmemFixed;
__device__ __forceinline__ EnvRef smem_env_ref(int k, int envs, int max_blocks, int pool_n, double *da, int *ia);
__device__ __forceinline__ const unsigned char *smem_env_order(int k, int envs, int max_blocks, int pool_n);

template <int ENVS>
__device__ __noinline__ void solve_lead(const bool solver, const int k, const int max_blocks, const int pool_n, double *da, int *ia,
                                        const int any_live, long long *tm)
{
    TM_DECL(tm);
    const EnvRef es = smem_env_ref(k, ENVS, max_blocks, pool_n, da, ia);
    const unsigned char *order_s = smem_env_order(k, ENVS, max_blocks, pool_n);
    int n_live = 0, c_robot = 0;
    double dt_coef = 0.0;
    BodyVel r[MG_RB_COUNT];
    if (solver) {
        n_live = order_s[MG_NARB + 1];
        c_robot = IF(es, I_ROBOT_FIRST) ? 0 : IF(es, I_NCON) - 10;
        /* cpDampedRotarySpring's pre-step applies the spring torque to both bodies: the robot's two springs, in add
         * order.  No other pre-step reads angular velocity, so running these after the parallel ones changes nothing. */
        con_prestep(es, c_robot + 2);
        con_prestep(es, c_robot + 3);
        TM_ADD(TM_SPRING);
        /* cpBodyUpdateVelocity: gravity 0, damping 1, no forces -> identity */
        dt_coef = IF(es, I_STEPPED) ? 1.0 : 0.0;     /* prev_dt == 0 ? 0 : dt / prev_dt */
        MG_NOUNROLL
        for (int q = 0; q < n_live; q++) arb_apply_cached(es, order_s, q, dt_coef);
    }
    if (any_live) __syncthreads();
    double *cf_robot = es.dc + (size_t)c_robot * C_NF;
    if (solver) {
        robot_vel_load(es, r);
        robot_cached_pass(r, cf_robot, es.d, dt_coef);
    }
    TM_ADD(TM_CACHED);
    if (any_live) __syncthreads();
    const double gear_inv_ratio = solver ? 1.0 / cf_robot[1 * C_NF + C_P1] : 1.0;     /* GearJoint control -> robot */
    MG_NOUNROLL
    for (int it = 0; it < 10; it++) {               /* space.iterations = phys_iter = 10 */
        if (n_live) {
            /* contacts may involve any body: give them the current velocities and take theirs back */
            robot_vel_store(es, r);
            MG_NOUNROLL
            for (int q = 0; q < n_live; q++) arb_apply_impulse(es, order_s, q);
            robot_vel_load(es, r);
        }
        if (any_live) __syncthreads();
        TM_ADD(TM_ARB);
        if (solver) robot_constraints_pass(r, cf_robot, es.d, gear_inv_ratio);
        TM_ADD(TM_ROBOTCON);
        if (any_live) { __syncthreads(); TM_ADD(TM_BLOCKCON); }
    }
    if (solver) { robot_vel_store(es, r); arb_pool_writeback(es, order_s, n_live); }
    TM_ADD(TM_REGS);
}

template <int ENVS>
__device__ __noinline__ void solve_helper(const bool helper, const int k, const int hb, const int max_blocks, const int pool_n, const int any_live)
{
    const EnvRef eh = smem_env_ref(k, ENVS, max_blocks, pool_n, nullptr, nullptr);
    if (any_live) __syncthreads();
    const int hc = helper ? (IF(eh, I_ROBOT_FIRST) ? 10 : 0) + 2 * hb : 0;
    const int hbody = MG_BODY_BLOCK0 + hb;
    if (helper) block_pair_cached(eh, hc, hbody, IF(eh, I_STEPPED) ? 1.0 : 0.0);
    if (any_live) __syncthreads();
    MG_NOUNROLL
    for (int it = 0; it < 10; it++) {
        if (any_live) __syncthreads();
        if (helper) block_pair_pass(eh, hc, hbody);
        if (any_live) __syncthreads();
    }
}

/* ------------------------------------------------------------------ scores (magical/benchmarks/*.py score_on_end_of_traj) */

__device__ __forceinline__ vec2 block_pos(EnvRef e, int bi) { return body_p(e, MG_BODY_BLOCK0 + bi); }
__device__ __forceinline__ vec2 robot_pos(EnvRef e) { return body_p(e, MG_BODY_ROBOT); }

/* goal sensor shape: Poly.create_box(body, (w, h)) on a static body at (cx, cy) (entities.py:783-791) */
__device__ void goal_shape(EnvRef e, int g, WShape *s, double *bb)
{
    double cx = DGF(e, D_GOAL0 + g * 4 + 0), cy = DGF(e, D_GOAL0 + g * 4 + 1);
    double h = DGF(e, D_GOAL0 + g * 4 + 2), w = DGF(e, D_GOAL0 + g * 4 + 3);
    double hw = w / 2, hh = h / 2;
    vec2 lv[4] = { V(hw, -hh), V(hw, hh), V(-hw, hh), V(-hw, -hh) };
    s->kind = MG_SHAPE_POLY; s->n = 4; s->r = 0.0;
    double l = INFINITY, r = -INFINITY, bo = INFINITY, t = -INFINITY;
    for (int i = 0; i < 4; i++) {
        vec2 a = lv[(i + 3) % 4], b = lv[i];
        double dx = b.x - a.x, dy = b.y - a.y;
        double nx = dy, ny = -dx;
        double inv = 1.0 / (sqrt(nx * nx + ny * ny) + MG_DBL_MIN);
        vec2 ln = V(nx * inv, ny * inv);
        vec2 wv = V(1.0 * b.x - 0.0 * b.y + cx, 0.0 * b.x + 1.0 * b.y + cy);
        s->v[i] = wv;
        s->nrm[i] = V(1.0 * ln.x - 0.0 * ln.y, 0.0 * ln.x + 1.0 * ln.y);
        l = fmin(l, wv.x); r = fmax(r, wv.x); bo = fmin(bo, wv.y); t = fmax(t, wv.y);
    }
    bb[0] = l; bb[1] = bo; bb[2] = r; bb[3] = t;
}

/* GoalRegion.get_overlapping_ents(com_overlap=True) over the block index (entities.py:803-863):
 * bit bi set iff every shape of block bi touches the goal box and the block's body position
 * lies in the goal's AABB. */
__device__ unsigned goal_overlaps(EnvRef e, const MgTables *T, int g)
{
    WShape G;
    double bb[4];
    goal_shape(e, g, &G, bb);
    unsigned mask = 0;
    int ns = IF(e, I_NSHAPES), nbk = IF(e, I_NBLOCKS);
    for (int bi = 0; bi < nbk; bi++) {
        vec2 p = block_pos(e, bi);
        bool com_in = bb[0] <= p.x && bb[2] >= p.x && bb[1] <= p.y && bb[3] >= p.y;
        bool all = com_in;
        if (com_in) {
            for (int s = 0; s < ns && all; s++) {
                int info = IF(e, I_SHAPE0 + s);
                if ((info >> 16) != MG_ENTID_BLOCK0 + bi) continue;
                WShape S;
                int body = (info >> 8) & 255;
                wshape_from_tpl(T, info & 255, body_p(e, body), body_rot(e, body), &S);
                Manifold m;
                int cnt = S.kind <= G.kind ? collide_shapes(&S, &G, &m) : collide_shapes(&G, &S, &m);
                if (cnt == 0) all = false;
            }
        }
        if (all) mask |= 1u << bi;
    }
    return mask;
}

__device__ double score_move_to_corner(EnvRef e)                 /* move_to_corner.py:67-77 */
{
    vec2 p = robot_pos(e);
    double dx = -1.0 - p.x, dy = 1.0 - p.y;
    double dist = sqrt(dx * dx + dy * dy);
    double succeed_dist = sqrt(2.0) / 2, furthest_dist = sqrt(2.0);
    double drange = furthest_dist - succeed_dist;
    return fmin(1.0, fmax(0.0, furthest_dist - dist) / drange);
}

__device__ double score_move_to_region(EnvRef e)                 /* move_to_region.py:85-94 */
{
    WShape G; double bb[4];
    goal_shape(e, 0, &G, bb);
    vec2 p = robot_pos(e);
    double maxd = -INFINITY;
    for (int i = 0; i < 4; i++) maxd = fmax(maxd, (p.x - G.v[i].x) * G.nrm[i].x + (p.y - G.v[i].y) * G.nrm[i].y);
    return maxd <= 0.0 ? 1.0 : 0.0;
}

__device__ __forceinline__ int block_role(EnvRef e, int bi) { return (IF(e, I_BLOCK0 + bi) >> 16) & 255; }

__device__ double score_match_regions(EnvRef e, const MgTables *T) /* match_regions.py:193-213 */
{
    unsigned ov = goal_overlaps(e, T, 0);
    int n_t = 0, n_ot = 0, n_od = 0, n_o = 0, nbk = IF(e, I_NBLOCKS);
    for (int i = 0; i < nbk; i++) {
        int role = block_role(e, i);
        if (role == MG_ROLE_TARGET) n_t++;
        if (ov >> i & 1) { n_o++; if (role == MG_ROLE_TARGET) n_ot++; if (role == MG_ROLE_DISTRACTOR) n_od++; }
    }
    double target_frac_done = (double)n_ot / (double)n_t;
    double contamination = n_o == 0 ? 0.0 : (double)n_od / (double)n_o;
    return target_frac_done * (1 - contamination);
}

__device__ double score_find_dupe(EnvRef e, const MgTables *T)    /* find_dupe.py:203-216 */
{
    unsigned ov = goal_overlaps(e, T, 0);
    int n_ot = 0, n_od = 0, n_o = 0, nbk = IF(e, I_NBLOCKS);
    for (int i = 0; i < nbk; i++) {
        if (!(ov >> i & 1)) continue;
        n_o++;
        int role = block_role(e, i);
        if (role == MG_ROLE_TARGET) n_ot++;
        if (role == MG_ROLE_DISTRACTOR) n_od++;
    }
    double have_two = n_ot >= 2 ? 1.0 : 0.0;
    double contamination = n_o == 0 ? 0.0 : (double)n_od / (double)n_o;
    return have_two * (1 - contamination);
}

__device__ double score_fix_colour(EnvRef e, const MgTables *T)   /* fix_colour.py:193-202 */
{
    int ng = IF(e, I_NGOALS);
    for (int g = 0; g < ng; g++) {
        unsigned ov = goal_overlaps(e, T, g);
        int want = ((IF(e, I_GOAL0 + g) >> 8) & 255) - 1;
        bool ok = want < 0 ? ov == 0 : ov == (1u << want);
        if (!ok) return 0.0;
    }
    return 1.0;
}

/* make_line.longest_line (make_line.py:31-71) */
__device__ int longest_line(const vec2 *pts, int npts, double inlier_dist, double max_sep)
{
    int best = npts < 1 ? npts : 1;
    for (int i = 0; i < npts - 1; i++)
        for (int j = i + 1; j < npts; j++) {
            vec2 pi = pts[i];
            vec2 off_j = vsub(pts[j], pi);
            double nrm = sqrt(off_j.x * off_j.x + off_j.y * off_j.y);
            vec2 u = V(off_j.x / nrm, off_j.y / nrm);
            double inl[MG_MAX_BLOCKS];
            int n_in = 0;
            for (int k = 0; k < npts; k++) {
                vec2 off = vsub(pts[k], pi);
                double proj = off.x * u.x + off.y * u.y;
                double rx = off.x - proj * u.x, ry = off.y - proj * u.y;
                if (sqrt(rx * rx + ry * ry) <= inlier_dist) inl[n_in++] = proj;
            }
            if (n_in <= best) continue;
            for (int a = 1; a < n_in; a++) {          /* np.sort */
                double x = inl[a]; int b = a - 1;
                while (b >= 0 && inl[b] > x) { inl[b + 1] = inl[b]; b--; }
                inl[b + 1] = x;
            }
            int max_run = 0, run = 0;
            for (int k = 0; k + 1 < n_in; k++) {
                if (fabs(inl[k + 1] - inl[k]) <= max_sep) { run++; if (run > max_run) max_run = run; }
                else run = 0;
            }
            max_run += 1;
            if (max_run > best) best = max_run;
        }
    return best;
}

__device__ double score_make_line(EnvRef e)                      /* make_line.py:142-152 */
{
    const double SHAPE_RAD = 0.2 * 0.6;
    vec2 pts[MG_MAX_BLOCKS];
    int n = IF(e, I_NBLOCKS);
    for (int i = 0; i < n; i++) pts[i] = block_pos(e, i);
    int line_len = longest_line(pts, n, SHAPE_RAD * 1.5, SHAPE_RAD * 3.5);
    int min_len = n - 2 > 2 ? n - 2 : 2;
    int num = line_len - min_len > 0 ? line_len - min_len : 0;
    return (double)num / (double)(n - min_len);
}

/* cluster.py:166-216; characteristic values in np.unique (string-sorted) order */
__device__ double score_cluster(EnvRef e, int by_type)
{
    const int colour_rank[4] = { 2, 1, 0, 3 };   /* red, green, blue, yellow -> blue<green<red<yellow */
    const int type_rank[4] = { 2, 1, 3, 0 };     /* square, pentagon, star, circle -> circle<pentagon<square<star */
    int n = IF(e, I_NBLOCKS);
    int label[MG_MAX_BLOCKS], present[4] = { 0, 0, 0, 0 };
    for (int i = 0; i < n; i++) {
        int info = IF(e, I_BLOCK0 + i);
        label[i] = by_type ? type_rank[info & 255] : colour_rank[(info >> 8) & 255];
        present[label[i]] = 1;
    }
    int idx_of[4], nvals = 0;
    for (int r = 0; r < 4; r++) idx_of[r] = present[r] ? nvals++ : -1;
    vec2 cent[4];
    for (int r = 0; r < 4; r++) {
        if (!present[r]) continue;
        double sx = 0.0, sy = 0.0; int cnt = 0;
        for (int i = 0; i < n; i++) if (label[i] == r) { vec2 p = block_pos(e, i); sx += p.x; sy += p.y; cnt++; }
        cent[idx_of[r]] = V(sx / cnt, sy / cnt);
    }
    int n_blocks = 0, n_correct = 0;
    for (int r = 0; r < 4; r++) {
        if (!present[r]) continue;
        for (int i = 0; i < n; i++) {
            if (label[i] != r) continue;
            n_blocks++;
            vec2 p = block_pos(e, i);
            double true_sse = 0.0, nearest_bad = INFINITY;
            for (int c = 0; c < nvals; c++) {
                double dx = p.x - cent[c].x, dy = p.y - cent[c].y;
                double sse = dx * dx + dy * dy;
                if (c == idx_of[r]) true_sse = sse;
                else if (sse < nearest_bad) nearest_bad = sse;
            }
            n_correct += sqrt(true_sse) < sqrt(nearest_bad) - 2.0 * true_sse ? 1 : 0;
        }
    }
    double frac = (double)n_correct / (double)(n_blocks > 1 ? n_blocks : 1);
    return fmax(frac - 0.75, 0.0) / (1 - 0.75);
}

__device__ double score_pick_and_place(EnvRef e)                 /* pick_and_place.py:88-102 */
{
    const double SHAPE_RAD = 0.2 * 0.6;
    double best = -1e9;
    int nv = IF(e, I_NVALID);
    for (int k = 0; k < nv; k++) {
        vec2 sp = block_pos(e, IF(e, I_TARGET_BLOCK));
        double dx = DGF(e, D_TARGET) - sp.x, dy = DGF(e, D_TARGET + 1) - sp.y;
        double dist = sqrt(dx * dx + dy * dy);
        double furthest = sqrt(2.0);
        double drange = furthest - SHAPE_RAD;
        best = fmax(best, fmin(1.0, fmax(0.0, furthest - dist) / drange));
    }
    return best;
}

__device__ __noinline__ double env_score(EnvRef e, const MgTables *T)
{
    switch (IF(e, I_TASK)) {
    case MG_TASK_MOVE_TO_CORNER: return score_move_to_corner(e);
    case MG_TASK_MOVE_TO_REGION: return score_move_to_region(e);
    case MG_TASK_MATCH_REGIONS: return score_match_regions(e, T);
    case MG_TASK_MAKE_LINE: return score_make_line(e);
    case MG_TASK_FIND_DUPE: return score_find_dupe(e, T);
    case MG_TASK_FIX_COLOUR: return score_fix_colour(e, T);
    case MG_TASK_CLUSTER_COLOUR: return score_cluster(e, 0);
    case MG_TASK_CLUSTER_SHAPE: return score_cluster(e, 1);
    case MG_TASK_PICK_AND_PLACE: return score_pick_and_place(e);
    }
    return 0.0;
}

/* move_to_corner.py:86-100, pick_and_place.py:114-124 */
__device__ __noinline__ double env_debug_reward(EnvRef e)
{
    vec2 rp = robot_pos(e);
    int task = IF(e, I_TASK);
    if (task == MG_TASK_MOVE_TO_CORNER) {
        vec2 sp = block_pos(e, 0);
        double dx = sp.x - 0.0, dy = sp.y - 1.0;
        double shape_to_corner = sqrt(dx * dx + dy * dy);
        double rx = rp.x - sp.x, ry = rp.y - sp.y;
        double robot_to_shape = sqrt(rx * rx + ry * ry);
        double shaping = -shape_to_corner / 5 - fmax(robot_to_shape, 0.2) / 20;
        return shaping + score_move_to_corner(e);
    }
    if (task == MG_TASK_PICK_AND_PLACE) {
        const double SHAPE_RAD = 0.2 * 0.6;
        vec2 sp = block_pos(e, IF(e, I_TARGET_BLOCK));
        double dx = sp.x - DGF(e, D_TARGET), dy = sp.y - DGF(e, D_TARGET + 1);
        double shape_to_target = sqrt(dx * dx + dy * dy);
        double rx = rp.x - sp.x, ry = rp.y - sp.y;
        double robot_to_shape = sqrt(rx * rx + ry * ry);
        double shaping = -shape_to_target / 5 - fmax(robot_to_shape, SHAPE_RAD) / 10;
        return shaping + score_pick_and_place(e);
    }
    return 0.0;
}

/* ------------------------------------------------------------------ kernels */

/* BaseEnv.step up to (not including) render */

/* the part of one env's state the step touches all the time, in shared memory.  The double block is sized
 * for the largest env of the batch (max_blocks) so that small scenes leave most of the SM's L1 to the tables
 * and the arbiter cache: [ (7 + max_blocks) bodies | (10 + 2 max_blocks) constraints | world boxes of 9 + 6 max_blocks shapes ] */
struct PhysSmemFixed {
    int i[(I_ARB0 + 3) & ~3];             /* counters, block / goal / constraint / shape infos */
    unsigned char order[MG_NARB + 8];     /* live arbiter slots in discovery order (+ flag bytes) */
    int ainfo[MG_ARB_POOL];               /* bodies / count / state of the pooled live arbiters */
    NarrowScratch narrow;                 /* the pair of shapes in the narrow phase and their manifold */
};
#define PHYS_TABLE_BYTES ((int)((offsetof(MgTables, dl_off) + 15) & ~15))

/* shapes of the largest env: robot circle, 4 finger quads, 4 walls, and up to MG_STAR_PIECES per block */
__host__ __device__ inline int phys_max_shapes(int max_blocks) { return 4 + 5 + max_blocks * MG_STAR_PIECES; }
__host__ __device__ inline int phys_smem_doubles(int max_blocks, int pool_n)
{
    /* stride = 1 (mod 16) doubles: the solver lanes of neighbouring envs read the same field of their own
     * env at the same time, and this keeps those reads on different shared-memory banks */
    int n = (1 + MG_RB_COUNT + max_blocks) * B_NF + (10 + 2 * max_blocks) * C_NF + 4 * phys_max_shapes(max_blocks) + pool_n * A_NF;
    return ((n + 15) & ~15) + 1;
}

/* env k of the CTA in the kernel's dynamic shared memory: [ tables | ENVS x nd doubles | ENVS x PhysSmemFixed ] */
extern __shared__ __align__(16) unsigned char phys_smem_raw[];
__device__ __forceinline__ EnvRef smem_env_ref(int k, int envs, int max_blocks, int pool_n, double *da, int *ia)
{
    const int nd = phys_smem_doubles(max_blocks, pool_n), con_off = (1 + MG_RB_COUNT + max_blocks) * B_NF;
    const int pool_off = con_off + (10 + 2 * max_blocks) * C_NF + 4 * phys_max_shapes(max_blocks);
    double *smd = (double *)(phys_smem_raw + PHYS_TABLE_BYTES);
    PhysSmemFixed *sm = (PhysSmemFixed *)(smd + (size_t)envs * nd);
    EnvRef e;
    e.d = smd + (size_t)k * nd; e.dc = e.d + con_off; e.i = sm[k].i; e.da = da; e.ia = ia;
    e.ap = e.d + pool_off; e.ainfo = sm[k].ainfo; e.apn = pool_n;
    return e;
}
__device__ __forceinline__ const unsigned char *smem_env_order(int k, int envs, int max_blocks, int pool_n)
{
    const int nd = phys_smem_doubles(max_blocks, pool_n);
    double *smd = (double *)(phys_smem_raw + PHYS_TABLE_BYTES);
    return ((PhysSmemFixed *)(smd + (size_t)envs * nd))[k].order;
}

template <int PHYS_THREADS_T, bool PERM>
__global__ void __launch_bounds__(PHYS_THREADS_T, 512 / PHYS_THREADS_T) mg_physics_kernel(MgDeviceState S, const int32_t *actions, float *reward,
                                                                uint8_t *done_out, float *score_out, int max_blocks, int pool_n)
{
    constexpr int ENVS = PHYS_THREADS_T / PHYS_LANES;     /* envs per CTA; lanes 0..ENVS-1 of warp 0 run their serial halves */
    const int nd = phys_smem_doubles(max_blocks, pool_n), con_off = (1 + MG_RB_COUNT + max_blocks) * B_NF;
    const int bb_off = con_off + (10 + 2 * max_blocks) * C_NF;     /* world boxes of the env's shapes, 4 doubles each */
    const int pool_off = bb_off + 4 * phys_max_shapes(max_blocks); /* working storage of the first MG_ARB_POOL live arbiters */
    /* the collision half of the tables (templates, walls, constants) in shared memory: with the env state
     * resident there is little L1 left, and every table read from HBM/L2 is a ~700-cycle stall on a serial path */
    {
        const double *src = (const double *)S.tables;
        double *dst = (double *)phys_smem_raw;
        for (int k = threadIdx.x; k < PHYS_TABLE_BYTES / 8; k += PHYS_THREADS_T) dst[k] = src[k];
    }
    double *smd = (double *)(phys_smem_raw + PHYS_TABLE_BYTES);                /* [ENVS][nd] */
    PhysSmemFixed *sm = (PhysSmemFixed *)(smd + (size_t)ENVS * nd);            /* [ENVS] */
    const int lane = threadIdx.x & 31;
    const int grp = threadIdx.x / PHYS_LANES, gl = threadIdx.x % PHYS_LANES;
    const unsigned gmask = (unsigned)((1ull << PHYS_LANES) - 1) << (lane / PHYS_LANES * PHYS_LANES);
    const int env0 = blockIdx.x * ENVS;         /* work slots env0 .. env0 + ENVS - 1 of this launch */
    const bool active = env0 + grp < S.num_envs;   /* idle groups of the last CTA still meet the barriers */
    const int env = PERM ? S.perm[active ? env0 + grp : env0] : (active ? env0 + grp : env0);
    const MgTables *T = (const MgTables *)phys_smem_raw;   /* only the fields before dl_off are valid */
#ifdef MG_SECTION_TIMERS
    long long *tm_cta = threadIdx.x == 0 && blockIdx.x < MG_TIMER_CTAS ? mg_phys_timers + (size_t)blockIdx.x * MG_NTIMERS : nullptr;
    const long long tm_start = clock64();
#else
    long long *tm_cta = nullptr;
#endif
    TM_DECL(tm_cta);
    const EnvRef g = env_ref(S.dstate, S.istate, env);
    PhysSmemFixed *W = &sm[grp];
    /* Which warp leads (runs the CTA's serial chains).  The CTAs resident on one SM occupy hardware warp slots a
     * multiple of the CTA's warp count apart, and a slot's scheduler is slot % 4: with warp 0 leading everywhere, the
     * lead warps of an SM's CTAs would all sit on one scheduler and share its FP64 pipe (one DADD per 2.3 cycles and
     * warp) while the other three idle through the serial half.  The CTA in slot group k leads with the warp whose
     * slot is k mod 4 instead. */
    __shared__ int lead_warp_sh;
#if PHYS_LEAD_ROTATE && !defined(MG_SECTION_TIMERS)
    if (threadIdx.x == 0) {
        unsigned hw;
        asm volatile("mov.u32 %0, %%warpid;" : "=r"(hw));
        lead_warp_sh = (int)(((hw / (PHYS_THREADS_T / 32)) - hw) & 3);
    }
#else
    if (threadIdx.x == 0) lead_warp_sh = 0;       /* the section timers live on thread 0 */
#endif
    __syncthreads();
    const int warp = threadIdx.x >> 5, lead_warp = lead_warp_sh;
    const bool is_lead = warp == lead_warp;
    double *Wd = smd + (size_t)grp * nd;

    /* stage the env: ints first (they hold the counts), then the bodies and constraints in use */
    int nb = 0, nc = 0;
    if (active) {
        for (int f = gl; f < I_ARB0; f += PHYS_LANES) W->i[f] = g.i[f];
        __syncwarp(gmask);
        nb = W->i[I_NBODIES]; nc = W->i[I_NCON];
        for (int f = gl; f < nb * B_NF; f += PHYS_LANES) Wd[f] = g.d[D_BODY0 + f];
        for (int f = gl; f < nc * C_NF; f += PHYS_LANES) Wd[con_off + f] = g.dc[f];
        __syncwarp(gmask);
    }
    EnvRef e;
    e.d = Wd; e.dc = Wd + con_off; e.i = W->i; e.da = g.da; e.ia = g.ia;
    e.ap = Wd + pool_off; e.ainfo = W->ainfo; e.apn = pool_n;
    /* lane k of warp 0 runs the robot chain and the contacts of env k of this CTA (solve_lead): it needs that env's
     * arbiter cache in HBM */
    const bool solver = is_lead && lane < ENVS && env0 + lane < S.num_envs;
    EnvRef gs = g;
    if (solver) gs = env_ref(S.dstate, S.istate, PERM ? S.perm[env0 + lane] : env0 + lane);
    /* the other warps: lane = (env k of the CTA, block hb of that env) runs that block's floor-friction pair */
    constexpr int SLOTS_PER_WARP = 32 / ENVS;
    static_assert((PHYS_THREADS_T / 32 - 1) * SLOTS_PER_WARP >= MG_MAX_BLOCKS, "a helper lane per block");
    const int hk = lane % ENVS, hb = (warp - (warp > lead_warp ? 1 : 0)) * SLOTS_PER_WARP + lane / ENVS;   /* helper warps count past the lead */
    EnvRef eh = e;
    eh.d = smd + (size_t)hk * nd; eh.dc = eh.d + con_off; eh.i = sm[hk].i;
    bool helper = false;          /* known once the env is staged (block count) */
    /* ACTION_ID_TO_FLAGS: id = ud + 3 * lr + 9 * grip (entities.py:148-190); Robot.set_action (:435-453) */
    int action = active ? actions[env] : 0;
    action = action < 0 ? 0 : (action >= MG_NUM_ACTIONS ? MG_NUM_ACTIONS - 1 : action);
    int ud = action % 3, lr = (action / 3) % 3, grip = action / 9;
    const double radius = 0.2;
    double rel_turn_angle = 0.0, target_speed = 0.0;
    if (ud == 1) target_speed += 4.0 * radius;
    if (ud == 2) target_speed -= 3.0 * radius;
    if (lr == 1) rel_turn_angle += 1.5;
    if (lr == 2) rel_turn_angle -= 1.5;
    double target_finger_angle = grip == 0 ? M_PI / 8 : -0.0;

    __syncthreads();              /* every env of the CTA is staged: the helper lanes read another group's counters */
    helper = !is_lead && env0 + hk < S.num_envs && hb < IF(eh, I_NBLOCKS);
    TM_ADD(TM_STAGE);
    const int rb = MG_BODY_ROBOT;
    const int motor_l = active ? (IF(e, I_ROBOT_FIRST) ? 6 : 2 * IF(e, I_NBLOCKS) + 6) : 0;  /* robot constraints 0..9: motors are 6 and 9 */
    for (int sub = 0; sub < 10; sub++) {
        int n_live_sub = 0;
        if (active) {
            if (gl == 0) {
                /* Robot.update (entities.py:455-476) */
                set_angle(e, rb + MG_RB_CONTROL, BF(e, rb, B_A) + rel_turn_angle);
                vec2 vel = vrot(body_rot(e, rb), V(0.0, target_speed));
                BF(e, rb + MG_RB_CONTROL, B_VX) = vel.x; BF(e, rb + MG_RB_CONTROL, B_VY) = vel.y;
                for (int k = 0; k < 2; k++) {
                    double side = k == 0 ? -1.0 : 1.0;
                    double rel_angle = BF(e, rb + MG_RB_FINGER_L + k, B_A) - BF(e, rb, B_A);
                    double angle_error = rel_angle + side * target_finger_angle;
                    double target_rate = fmax(-1.0, fmin(1.0, angle_error * 10));
                    if (fabs(target_rate) < 1e-4) target_rate = 0.0;
                    CF(e, motor_l + 3 * k, C_P0) = target_rate;
                }
            }
            __syncwarp(gmask);
#if PHYS_DBG_SKIP != 2
            int n_live = space_step_parallel(e, T, lane, gmask, W->order, &W->narrow, Wd + bb_off, tm_cta);
#else
            int n_live = 0;
#endif
            n_live_sub = n_live >> 8;         /* live contacts on a block: the only ones the helper lanes must be fenced from */
            n_live &= 255;
            if (gl == 0) W->order[MG_NARB + 1] = (unsigned char)n_live;
        }
        /* the order-dependent half of the sub-step for all envs of the CTA: robot chains and contacts side by side in
         * warp 0, the blocks' floor-friction pairs on the other warps (solve_lead / solve_helper) */
        TM_ADD(TM_PARALLEL);
        const int any_live = __syncthreads_or(n_live_sub != 0);
        TM_ADD(TM_BARRIER);
        if (is_lead) solve_lead<ENVS>(solver, lane & (ENVS - 1), max_blocks, pool_n, gs.da, gs.ia, any_live, tm_cta);
        else solve_helper<ENVS>(helper, hk, hb, max_blocks, pool_n, any_live);
        TM_ADD(TM_SERIAL);
        __syncthreads();
        if (active && gl == 0) IF(e, I_STEPPED) = 1;      /* after the barrier: the helper lanes read it for dt_coef */
    }
#ifdef MG_SECTION_TIMERS
    if (tm_cta) tm_cta[TM_TOTAL] += clock64() - tm_start;
#endif
    if (!active) return;

    const int steps = IF(e, I_EPISODE_STEPS) + 1;
    const bool done = steps >= IF(e, I_HORIZON);
    __syncwarp(gmask);
    if (gl == 0) IF(e, I_EPISODE_STEPS) = steps;
    __syncwarp(gmask);

    /* write the env back */
    for (int f = gl; f < I_ARB0; f += PHYS_LANES) g.i[f] = W->i[f];
    for (int f = gl; f < nb * B_NF; f += PHYS_LANES) g.d[D_BODY0 + f] = Wd[f];
    for (int f = gl; f < nc * C_NF; f += PHYS_LANES) g.dc[f] = Wd[con_off + f];
    /* pose export for the rasteriser and mg_get_bodies: lane = body */
    double *pose = S.poses + (size_t)env * MG_POSE_STRIDE;
    bool finite = true;      /* a NaN / Inf body state is flagged per env, never propagated silently */
    for (int b = gl; b < MG_NB; b += PHYS_LANES) {
        double *o = pose + b * MG_POSE_NF;
        if (b < nb) {
            o[0] = BF(e, b, B_PX); o[1] = BF(e, b, B_PY); o[2] = BF(e, b, B_A);
            o[3] = BF(e, b, B_VX); o[4] = BF(e, b, B_VY); o[5] = BF(e, b, B_W);
            o[6] = BF(e, b, B_RX); o[7] = BF(e, b, B_RY);
            finite = finite && isfinite(o[0]) && isfinite(o[1]) && isfinite(o[2]) && isfinite(o[3]) && isfinite(o[4]) && isfinite(o[5]);
        } else {
            for (int k = 0; k < MG_POSE_NF; k++) o[k] = 0.0;
        }
    }
    const bool any_bad = __ballot_sync(gmask, !finite) != 0;
    __syncwarp(gmask);
    if (gl == 0) {
        const int status = IF(e, I_STATUS) | (any_bad ? MG_STATUS_NONFINITE : 0);
        if (status) { g.i[I_STATUS] = status; if (S.status) S.status[env] |= status; }
        /* scores read goals / targets too: use the HBM copy just written */
        double rew = 0.0, eval_score = 0.0;
        if (done) rew = eval_score = env_score(g, T);
        if (IF(e, I_DEBUG_REWARD)) rew = env_debug_reward(g);
        if (reward) reward[env] = (float)rew;
        if (done_out) done_out[env] = done ? 1 : 0;
        if (score_out) score_out[env] = (float)eval_score;
    }
}

__global__ void mg_score_kernel(MgDeviceState S, float *out, int debug_reward)
{
    int env = blockIdx.x * blockDim.x + threadIdx.x;
    if (env >= S.num_envs) return;
    EnvRef e = env_ref(S.dstate, S.istate, env);
    out[env] = (float)(debug_reward ? env_debug_reward(e) : env_score(e, S.tables));
}

/* checkpoint gather / scatter: blob = D_NF doubles then I_NF ints */
__global__ void mg_state_copy_kernel(MgDeviceState S, const int32_t *env_ids, int n, double *dblobs, int *iblobs, int scatter)
{
    int k = blockIdx.x;
    if (k >= n) return;
    int env = env_ids ? env_ids[k] : k;
    EnvRef e = env_ref(S.dstate, S.istate, env);
    for (int f = threadIdx.x; f < D_NF; f += blockDim.x) {
        if (scatter) DF(e, f) = dblobs[(size_t)k * D_NF + f]; else dblobs[(size_t)k * D_NF + f] = DF(e, f);
    }
    for (int f = threadIdx.x; f < I_NF; f += blockDim.x) {
        if (scatter) IF(e, f) = iblobs[(size_t)k * I_NF + f]; else iblobs[(size_t)k * I_NF + f] = IF(e, f);
    }
    if (scatter) {
        __syncthreads();
        if (threadIdx.x == 0) export_poses(e, S.poses + (size_t)env * MG_POSE_STRIDE);
    }
}

/* perm <- envs sorted by min(cached arbiters, 3), ascending and stable: envs without contacts first.  One CTA: every
 * thread owns a contiguous run of envs, counts its four classes, and block-wide exclusive scans place them. */
__global__ void __launch_bounds__(1024) mg_order_kernel(MgDeviceState S, int *perm)
{
    __shared__ int warp_sum[32];
    __shared__ int class_base[5];
    const int B = S.num_envs, t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int per = (B + 1023) / 1024, e0 = t * per, e1 = min(B, e0 + per);
    /* the kernel is one CTA on the step's critical path and its time is the latency of these loads: up to KEYS envs per
     * thread are read once, all loads in flight together, and kept in registers for the scatter pass */
    constexpr int KEYS = 8;
    int key[KEYS];
    const bool in_regs = per <= KEYS;
    if (in_regs) {
#pragma unroll
        for (int j = 0; j < KEYS; j++) key[j] = e0 + j < e1 ? S.istate[mg_field_index(e0 + j, I_NF, I_NARBS)] : 0;
    }
    int cnt[4] = { 0, 0, 0, 0 };
    if (in_regs) {
#pragma unroll
        for (int j = 0; j < KEYS; j++) if (e0 + j < e1) cnt[min(key[j], 3)]++;
    } else {
        for (int env = e0; env < e1; env++) cnt[min(S.istate[mg_field_index(env, I_NF, I_NARBS)], 3)]++;
    }
    int pos[4];
    if (t == 0) class_base[0] = 0;
    for (int c = 0; c < 4; c++) {
        int v = cnt[c], incl = v;
        for (int off = 1; off < 32; off <<= 1) { int o = __shfl_up_sync(0xffffffffu, incl, off); if (lane >= off) incl += o; }
        if (lane == 31) warp_sum[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            int w = warp_sum[lane], wi = w;
            for (int off = 1; off < 32; off <<= 1) { int o = __shfl_up_sync(0xffffffffu, wi, off); if (lane >= off) wi += o; }
            warp_sum[lane] = wi - w;
            if (lane == 31) class_base[c + 1] = class_base[c] + wi;
        }
        __syncthreads();
        pos[c] = class_base[c] + warp_sum[warp] + incl - v;
        __syncthreads();
    }
    if (in_regs) {
#pragma unroll
        for (int j = 0; j < KEYS; j++) if (e0 + j < e1) perm[pos[min(key[j], 3)]++] = e0 + j;
    } else {
        for (int env = e0; env < e1; env++) perm[pos[min(S.istate[mg_field_index(env, I_NF, I_NARBS)], 3)]++] = env;
    }
}

void mg_launch_order(const MgDeviceState &S, int *perm, cudaStream_t st)
{
    mg_order_kernel<<<1, 1024, 0, st>>>(S, perm);
}

void mg_launch_init(const MgDeviceState &S, const mg_scene *scenes_dev, const int32_t *env_ids_dev, int n, cudaStream_t st)
{
    if (n <= 0) return;
    mg_init_kernel<<<(n + 31) / 32, 32, 0, st>>>(S, scenes_dev, env_ids_dev, n);
}
__global__ void mg_step_io_kernel(int n, const int32_t *actions_src, int32_t *actions_dst, const float *reward_src, float *reward_dst,
                                  const uint8_t *done_src, uint8_t *done_dst, const float *score_src, float *score_dst)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    if (actions_dst) actions_dst[k] = actions_src[k];
    if (reward_dst) reward_dst[k] = reward_src[k];
    if (done_dst) done_dst[k] = done_src[k];
    if (score_dst) score_dst[k] = score_src[k];
}
void mg_launch_step_io(int n, const int32_t *actions_src, int32_t *actions_dst, const float *reward_src, float *reward_dst,
                       const uint8_t *done_src, uint8_t *done_dst, const float *score_src, float *score_dst, cudaStream_t st)
{
    mg_step_io_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, actions_src, actions_dst, reward_src, reward_dst, done_src, done_dst, score_src, score_dst);
}
void mg_launch_mark_pending(const MgDeviceState &S, const int32_t *env_ids_dev, int n, cudaStream_t st)
{
    if (n <= 0) return;
    mg_mark_pending_kernel<<<(n + 127) / 128, 128, 0, st>>>(S, env_ids_dev, n);
}
void mg_launch_physics(const MgDeviceState &S, const int32_t *actions, float *reward, uint8_t *done, float *score, int max_blocks, cudaStream_t st)
{
    /* the shared-memory opt-in is a per-device (per-context) function attribute: track it per CUDA ordinal */
    static size_t attr_set_dev[MG_MAX_DEVICES] = { 0 };
    static std::mutex attr_mu;
    int dev_ = 0;
    cudaGetDevice(&dev_);
    std::lock_guard<std::mutex> attr_lock(attr_mu);
    size_t &attr_set = attr_set_dev[dev_ < 0 || dev_ >= MG_MAX_DEVICES ? 0 : dev_];
    if (max_blocks < 0) max_blocks = 0;
    if (max_blocks > MG_MAX_BLOCKS) max_blocks = MG_MAX_BLOCKS;
    /* 16 envs per CTA while two such CTAs fit one SM; larger scenes run 8 envs per CTA so that three (two
     * at the largest block counts) stay resident instead of one.  The shared-memory pool of live arbiters takes
     * whatever is left: the largest size that keeps the CTA shape and the number of resident CTAs */
    const size_t smem_sm = 233472, smem_cta_reserved = 1024;
    auto per_env_bytes = [&](int pool) { return (size_t)phys_smem_doubles(max_blocks, pool) * sizeof(double) + sizeof(PhysSmemFixed); };
    static const int force_shape = getenv("MG_PHYS_ENVS_PER_CTA") ? atoi(getenv("MG_PHYS_ENVS_PER_CTA")) : 0;   /* 8 / 16: timing experiments */
    auto is_wide = [&](int pool) {
        if (force_shape == PHYS_ENVS / 2) return false;
        return 2 * (PHYS_TABLE_BYTES + per_env_bytes(pool) * PHYS_ENVS + smem_cta_reserved) <= smem_sm;
    };
    auto resident = [&](int pool) {
        const int envs_ = is_wide(pool) ? PHYS_ENVS : PHYS_ENVS / 2;
        const int by_smem = (int)(smem_sm / (PHYS_TABLE_BYTES + per_env_bytes(pool) * envs_ + smem_cta_reserved));
        const int by_regs = 512 / (envs_ * PHYS_LANES);            /* __launch_bounds__ of the kernel */
        return by_smem < by_regs ? by_smem : by_regs;
    };
    int pool_n = 0;
    for (int p = MG_ARB_POOL; p > 0; p -= 2)
        if (is_wide(p) == is_wide(0) && resident(p) >= resident(0)) { pool_n = p; break; }
    const size_t per_env = per_env_bytes(pool_n);
    const bool wide = is_wide(pool_n);
    const int envs = wide ? PHYS_ENVS : PHYS_ENVS / 2;
    const size_t smem = PHYS_TABLE_BYTES + per_env * envs;
    const size_t smem_full = PHYS_TABLE_BYTES + per_env_bytes(MG_ARB_POOL) * PHYS_ENVS;
    if (smem_full > attr_set) {
        const int a = (int)smem_full;
        cudaFuncSetAttribute(mg_physics_kernel<PHYS_THREADS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, a);
        cudaFuncSetAttribute(mg_physics_kernel<PHYS_THREADS / 2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, a);
        cudaFuncSetAttribute(mg_physics_kernel<PHYS_THREADS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, a);
        cudaFuncSetAttribute(mg_physics_kernel<PHYS_THREADS / 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, a);
        attr_set = smem_full;
    }
    const int grid = (S.num_envs + envs - 1) / envs;
    /* the PERM instantiations take their envs through S.perm (mg_step in parts) */
#define PHYS_LAUNCH(T_, P_) mg_physics_kernel<T_, P_><<<grid, T_, smem, st>>>(S, actions, reward, done, score, max_blocks, pool_n)
    if (S.perm) {
        if (wide) PHYS_LAUNCH(PHYS_THREADS, true); else PHYS_LAUNCH(PHYS_THREADS / 2, true);
    } else {
        if (wide) PHYS_LAUNCH(PHYS_THREADS, false); else PHYS_LAUNCH(PHYS_THREADS / 2, false);
    }
#undef PHYS_LAUNCH
}
void mg_launch_score(const MgDeviceState &S, float *out, int debug_reward, cudaStream_t st)
{
    mg_score_kernel<<<(S.num_envs + 31) / 32, 32, 0, st>>>(S, out, debug_reward);
}
void mg_launch_state_gather(const MgDeviceState &S, const int32_t *env_ids_dev, int n, double *dblobs, int *iblobs, int scatter, cudaStream_t st)
{
    if (n <= 0) return;
    mg_state_copy_kernel<<<n, 128, 0, st>>>(S, env_ids_dev, n, dblobs, iblobs, scatter);
}
