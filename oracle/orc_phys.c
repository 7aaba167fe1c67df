/*
 * orc_phys.c -- ORACLE (test infrastructure): one Chipmunk-ordered space step,
 * the joints the reference uses, Robot.update and BaseEnv._phys_steps_on_frame.
 *
 * Stands in for Chipmunk2D 7.0.3 (cpSpaceStep.c, cpArbiter.c, cpBody.c,
 * cpPivotJoint.c, cpGearJoint.c, cpRotaryLimitJoint.c, cpSimpleMotor.c,
 * cpDampedRotarySpring.c) as driven by magical/base_env.py:248-255 and
 * magical/entities.py:435-476.  PARITY UNPINNED (see orc.h).
 *
 * Known, documented deviation: Chipmunk discovers colliding pairs in BBTree
 * traversal order; that order is not reproducible without the library, so
 * arbiters are solved in ascending (shape index a, shape index b) order.
 */
#include <math.h>
#include <float.h>
#include <string.h>
#include "orc.h"

static inline vec vsub(vec a, vec b) { vec r = { a.x - b.x, a.y - b.y }; return r; }
static inline vec vadd(vec a, vec b) { vec r = { a.x + b.x, a.y + b.y }; return r; }
static inline vec vmul(vec a, double s) { vec r = { a.x * s, a.y * s }; return r; }
static inline double vdot(vec a, vec b) { return a.x * b.x + a.y * b.y; }
static inline double vcross(vec a, vec b) { return a.x * b.y - a.y * b.x; }
static inline vec vneg(vec a) { vec r = { -a.x, -a.y }; return r; }
static inline vec vperp(vec a) { vec r = { -a.y, a.x }; return r; }
static inline vec vrotate(vec a, vec b) { vec r = { a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x }; return r; }
/* chipmunk.h: cpfmax(a, b) = (a > b) ? a : b, cpfmin(a, b) = (a < b) ? a : b, cpfclamp = cpfmin(cpfmax(f, lo), hi) */
static inline double cp_max(double a, double b) { return a > b ? a : b; }
static inline double cp_min(double a, double b) { return a < b ? a : b; }
static inline double fclamp(double f, double lo, double hi) { return cp_min(cp_max(f, lo), hi); }
static inline vec vclamp(vec v, double len)
{
    /* cpvclamp: (dot > len^2) ? normalize(v) * len : v */
    if (vdot(v, v) > len * len) {
        double l = sqrt(vdot(v, v));
        return vmul(vmul(v, 1.0 / (l + DBL_MIN)), len);
    }
    return v;
}

/* chipmunk_private.h helpers */
static inline vec relative_velocity(const orc_body *a, const orc_body *b, vec r1, vec r2)
{
    vec v1 = vadd(a->v, vmul(vperp(r1), a->w));
    vec v2 = vadd(b->v, vmul(vperp(r2), b->w));
    return vsub(v2, v1);
}
static inline void apply_impulse(orc_body *b, vec j, vec r)
{
    b->v = vadd(b->v, vmul(j, b->m_inv));
    b->w += b->i_inv * vcross(r, j);
}
static inline void apply_impulses(orc_body *a, orc_body *b, vec r1, vec r2, vec j)
{
    apply_impulse(a, vneg(j), r1);
    apply_impulse(b, j, r2);
}
static inline void apply_bias_impulse(orc_body *b, vec j, vec r)
{
    b->v_bias = vadd(b->v_bias, vmul(j, b->m_inv));
    b->w_bias += b->i_inv * vcross(r, j);
}
static inline void apply_bias_impulses(orc_body *a, orc_body *b, vec r1, vec r2, vec j)
{
    apply_bias_impulse(a, vneg(j), r1);
    apply_bias_impulse(b, j, r2);
}
static inline double k_scalar_body(const orc_body *b, vec r, vec n)
{
    double rcn = vcross(r, n);
    return b->m_inv + b->i_inv * rcn * rcn;
}
static inline double k_scalar(const orc_body *a, const orc_body *b, vec r1, vec r2, vec n)
{
    return k_scalar_body(a, r1, n) + k_scalar_body(b, r2, n);
}
static void k_tensor(const orc_body *a, const orc_body *b, vec r1, vec r2, double k[4])
{
    double m_sum = a->m_inv + b->m_inv;
    double k11 = m_sum, k12 = 0.0, k21 = 0.0, k22 = m_sum;
    double a_i_inv = a->i_inv;
    double r1xsq = r1.x * r1.x * a_i_inv, r1ysq = r1.y * r1.y * a_i_inv, r1nxy = -r1.x * r1.y * a_i_inv;
    k11 += r1ysq; k12 += r1nxy; k21 += r1nxy; k22 += r1xsq;
    double b_i_inv = b->i_inv;
    double r2xsq = r2.x * r2.x * b_i_inv, r2ysq = r2.y * r2.y * b_i_inv, r2nxy = -r2.x * r2.y * b_i_inv;
    k11 += r2ysq; k12 += r2nxy; k21 += r2nxy; k22 += r2xsq;
    double det = k11 * k22 - k12 * k21;
    double det_inv = 1.0 / det;
    k[0] = k22 * det_inv; k[1] = -k12 * det_inv; k[2] = -k21 * det_inv; k[3] = k11 * det_inv;
}
static inline double bias_coef(double error_bias, double dt) { return 1.0 - pow(error_bias, dt); }

/* ---------------------------------------------------------------- joints */

static void con_prestep(orc_env *e, orc_constraint *c, double dt)
{
    orc_body *a = &e->bodies[c->a], *b = &e->bodies[c->b];
    switch (c->kind) {
    case ORC_C_PIVOT: { /* cpPivotJoint.c preStep; every body here has cog = 0 */
        c->r1 = vrotate(c->anchor_a, a->rot);
        c->r2 = vrotate(c->anchor_b, b->rot);
        k_tensor(a, b, c->r1, c->r2, c->k);
        vec delta = vsub(vadd(b->p, c->r2), vadd(a->p, c->r1));
        c->bias_v = vclamp(vmul(delta, -bias_coef(c->error_bias, dt) / dt), c->max_bias);
        break;
    }
    case ORC_C_GEAR: { /* cpGearJoint.c preStep */
        double ratio_inv = 1.0 / c->ratio;
        c->isum = 1.0 / (a->i_inv * ratio_inv + c->ratio * b->i_inv);
        c->bias = fclamp(-bias_coef(c->error_bias, dt) * (b->a * c->ratio - a->a - c->phase) / dt,
                         -c->max_bias, c->max_bias);
        break;
    }
    case ORC_C_SPRING: { /* cpDampedRotarySpring.c preStep (default torque function) */
        double moment = a->i_inv + b->i_inv;
        c->isum = 1.0 / moment;
        c->w_coef = 1.0 - exp(-c->damping * dt * moment);
        c->target_wrn = 0.0;
        double j_spring = (((a->a - b->a) - c->rest) * c->stiffness) * dt;
        c->jacc = j_spring;
        a->w -= j_spring * a->i_inv;
        b->w += j_spring * b->i_inv;
        break;
    }
    case ORC_C_LIMIT: { /* cpRotaryLimitJoint.c preStep */
        double dist = b->a - a->a;
        double pdist = 0.0;
        if (dist > c->hi) pdist = c->hi - dist;
        else if (dist < c->lo) pdist = c->lo - dist;
        c->isum = 1.0 / (a->i_inv + b->i_inv);
        c->bias = fclamp(-bias_coef(c->error_bias, dt) * pdist / dt, -c->max_bias, c->max_bias);
        if (!c->bias) c->jacc = 0.0;
        break;
    }
    case ORC_C_MOTOR: /* cpSimpleMotor.c preStep */
        c->isum = 1.0 / (a->i_inv + b->i_inv);
        break;
    }
}

static void con_apply_cached(orc_env *e, orc_constraint *c, double dt_coef)
{
    orc_body *a = &e->bodies[c->a], *b = &e->bodies[c->b];
    switch (c->kind) {
    case ORC_C_PIVOT:
        apply_impulses(a, b, c->r1, c->r2, vmul(c->jacc_v, dt_coef));
        break;
    case ORC_C_GEAR: {
        double j = c->jacc * dt_coef;
        a->w -= j * a->i_inv * (1.0 / c->ratio);
        b->w += j * b->i_inv;
        break;
    }
    case ORC_C_SPRING:
        break;
    case ORC_C_LIMIT:
    case ORC_C_MOTOR: {
        double j = c->jacc * dt_coef;
        a->w -= j * a->i_inv;
        b->w += j * b->i_inv;
        break;
    }
    }
}

static void con_apply_impulse(orc_env *e, orc_constraint *c, double dt)
{
    orc_body *a = &e->bodies[c->a], *b = &e->bodies[c->b];
    switch (c->kind) {
    case ORC_C_PIVOT: {
        vec vr = relative_velocity(a, b, c->r1, c->r2);
        vec d = vsub(c->bias_v, vr);
        vec j = { d.x * c->k[0] + d.y * c->k[1], d.x * c->k[2] + d.y * c->k[3] }; /* cpMat2x2Transform */
        vec j_old = c->jacc_v;
        c->jacc_v = vclamp(vadd(c->jacc_v, j), c->max_force * dt);
        j = vsub(c->jacc_v, j_old);
        apply_impulses(a, b, c->r1, c->r2, j);
        break;
    }
    case ORC_C_GEAR: {
        double wr = b->w * c->ratio - a->w;
        double j_max = c->max_force * dt;
        double j = (c->bias - wr) * c->isum;
        double j_old = c->jacc;
        c->jacc = fclamp(j_old + j, -j_max, j_max);
        j = c->jacc - j_old;
        a->w -= j * a->i_inv * (1.0 / c->ratio);
        b->w += j * b->i_inv;
        break;
    }
    case ORC_C_SPRING: {
        double wrn = a->w - b->w;
        double w_damp = (c->target_wrn - wrn) * c->w_coef;
        c->target_wrn = wrn + w_damp;
        double j_damp = w_damp * c->isum;
        c->jacc += j_damp;
        a->w += j_damp * a->i_inv;
        b->w -= j_damp * b->i_inv;
        break;
    }
    case ORC_C_LIMIT: {
        if (!c->bias) return;
        double wr = b->w - a->w;
        double j_max = c->max_force * dt;
        double j = -(c->bias + wr) * c->isum;
        double j_old = c->jacc;
        if (c->bias < 0.0) c->jacc = fclamp(j_old + j, 0.0, j_max);
        else c->jacc = fclamp(j_old + j, -j_max, 0.0);
        j = c->jacc - j_old;
        a->w -= j * a->i_inv;
        b->w += j * b->i_inv;
        break;
    }
    case ORC_C_MOTOR: {
        double wr = b->w - a->w + c->rate;
        double j_max = c->max_force * dt;
        double j = -wr * c->isum;
        double j_old = c->jacc;
        c->jacc = fclamp(j_old + j, -j_max, j_max);
        j = c->jacc - j_old;
        a->w -= j * a->i_inv;
        b->w += j * b->i_inv;
        break;
    }
    }
}

/* ---------------------------------------------------------------- arbiters */

static orc_arbiter *arb_find(orc_env *e, int sa, int sb)
{
    for (int i = 0; i < e->n_arbs; i++)
        if (e->arbs[i].sa == sa && e->arbs[i].sb == sb) return &e->arbs[i];
    return NULL;
}

/* cpSpaceCollideShapes + cpArbiterUpdate for one candidate pair (i < j by shape index) */
static void collide_pair(orc_env *e, int i, int j)
{
    orc_shape *a = &e->shapes[i], *b = &e->shapes[j];
    /* QueryReject */
    if (!(a->bb_l <= b->bb_r && b->bb_l <= a->bb_r && a->bb_b <= b->bb_t && b->bb_b <= a->bb_t)) return;
    if (a->body == b->body) return;
    if (orc_filter_reject(a, b)) return;
    orc_manifold m;
    if (orc_collide_shapes(a, b, &m) == 0) return;
    int sa = m.sa == 0 ? i : j, sb = m.sa == 0 ? j : i;
    orc_shape *A = &e->shapes[sa], *B = &e->shapes[sb];
    /* sensors and pairs of infinite-mass bodies are processed but never reach the solver */
    if (A->sensor || B->sensor) return;
    if (e->bodies[A->body].m == INFINITY && e->bodies[B->body].m == INFINITY) return;

    orc_arbiter *arb = arb_find(e, sa, sb);
    if (!arb) {
        if (e->n_arbs >= ORC_MAX_ARBS) return;
        arb = &e->arbs[e->n_arbs++];
        memset(arb, 0, sizeof(*arb));
        arb->sa = sa; arb->sb = sb;
        arb->state = ORC_ARB_FIRST; /* cpArbiterInit */
    }
    arb->ba = A->body; arb->bb = B->body;
    orc_contact fresh[2];
    for (int k = 0; k < m.count; k++) {
        orc_contact *con = &fresh[k];
        memset(con, 0, sizeof(*con));
        con->r1 = vsub(m.p1[k], e->bodies[A->body].p);
        con->r2 = vsub(m.p2[k], e->bodies[B->body].p);
        con->hash = m.hash[k];
        con->jn_acc = con->jt_acc = 0.0;
        for (int q = 0; q < arb->count; q++) {
            if (con->hash == arb->con[q].hash) {
                con->jn_acc = arb->con[q].jn_acc;
                con->jt_acc = arb->con[q].jt_acc;
            }
        }
    }
    arb->count = m.count;
    for (int k = 0; k < m.count; k++) arb->con[k] = fresh[k];
    arb->n = m.n;
    arb->e = 0.0;            /* a->e * b->e, all elasticities are 0 */
    arb->u = A->u * B->u;
    if (arb->state == ORC_ARB_CACHED) arb->state = ORC_ARB_FIRST;
    arb->stamp = e->stamp;
    arb->live = 1;
}

/* cpArbiterPreStep */
static void arb_prestep(orc_env *e, orc_arbiter *arb, double dt, double slop, double bias)
{
    orc_body *a = &e->bodies[arb->ba], *b = &e->bodies[arb->bb];
    vec n = arb->n;
    vec body_delta = vsub(b->p, a->p);
    for (int i = 0; i < arb->count; i++) {
        orc_contact *con = &arb->con[i];
        con->n_mass = 1.0 / k_scalar(a, b, con->r1, con->r2, n);
        con->t_mass = 1.0 / k_scalar(a, b, con->r1, con->r2, vperp(n));
        double dist = vdot(vadd(vsub(con->r2, con->r1), body_delta), n);
        con->bias = -bias * cp_min(0.0, dist + slop) / dt;
        con->j_bias = 0.0;
        con->bounce = vdot(relative_velocity(a, b, con->r1, con->r2), n) * arb->e;
    }
}

/* cpArbiterApplyCachedImpulse */
static void arb_apply_cached(orc_env *e, orc_arbiter *arb, double dt_coef)
{
    if (arb->state == ORC_ARB_FIRST) return;
    orc_body *a = &e->bodies[arb->ba], *b = &e->bodies[arb->bb];
    for (int i = 0; i < arb->count; i++) {
        orc_contact *con = &arb->con[i];
        vec jn = { con->jn_acc, con->jt_acc };
        vec j = vrotate(arb->n, jn);
        apply_impulses(a, b, con->r1, con->r2, vmul(j, dt_coef));
    }
}

/* cpArbiterApplyImpulse */
static void arb_apply_impulse(orc_env *e, orc_arbiter *arb)
{
    orc_body *a = &e->bodies[arb->ba], *b = &e->bodies[arb->bb];
    vec n = arb->n;
    double friction = arb->u;
    for (int i = 0; i < arb->count; i++) {
        orc_contact *con = &arb->con[i];
        double n_mass = con->n_mass;
        vec r1 = con->r1, r2 = con->r2;
        vec vb1 = vadd(a->v_bias, vmul(vperp(r1), a->w_bias));
        vec vb2 = vadd(b->v_bias, vmul(vperp(r2), b->w_bias));
        vec vr = relative_velocity(a, b, r1, r2); /* + surface_vr (zero) */
        double vbn = vdot(vsub(vb2, vb1), n);
        double vrn = vdot(vr, n);
        double vrt = vdot(vr, vperp(n));
        double jbn = (con->bias - vbn) * n_mass;
        double jbn_old = con->j_bias;
        con->j_bias = cp_max(jbn_old + jbn, 0.0);
        double jn = -(con->bounce + vrn) * n_mass;
        double jn_old = con->jn_acc;
        con->jn_acc = cp_max(jn_old + jn, 0.0);
        double jt_max = friction * con->jn_acc;
        double jt = -vrt * con->t_mass;
        double jt_old = con->jt_acc;
        con->jt_acc = fclamp(jt_old + jt, -jt_max, jt_max);
        apply_bias_impulses(a, b, r1, r2, vmul(n, con->j_bias - jbn_old));
        vec dj = { con->jn_acc - jn_old, con->jt_acc - jt_old };
        apply_impulses(a, b, r1, r2, vrotate(n, dj));
    }
}

/* ---------------------------------------------------------------- cpSpaceStep */

void orc_space_step(orc_env *e, double dt)
{
    if (dt == 0.0) return;
    e->stamp++;
    double prev_dt = e->curr_dt;
    e->curr_dt = dt;

    /* reset last step's arbiter list */
    for (int i = 0; i < e->n_arbs; i++) {
        if (e->arbs[i].live) e->arbs[i].state = ORC_ARB_NORMAL;
        e->arbs[i].live = 0;
    }

    /* integrate positions: cpBodyUpdatePosition for every dynamic + kinematic body */
    for (int i = 0; i < e->n_bodies; i++) {
        orc_body *b = &e->bodies[i];
        if (b->type == ORC_BODY_STATIC) continue;
        b->p = vadd(b->p, vmul(vadd(b->v, b->v_bias), dt));
        orc_body_set_angle(b, b->a + (b->w + b->w_bias) * dt);
        b->v_bias.x = 0.0; b->v_bias.y = 0.0;
        b->w_bias = 0.0;
    }
    /* refresh shape caches of non-static bodies, then find colliding pairs */
    for (int i = 0; i < e->n_shapes; i++)
        if (e->bodies[e->shapes[i].body].type != ORC_BODY_STATIC) orc_shape_update(e, i);
    for (int i = 0; i < e->n_shapes; i++)
        for (int j = i + 1; j < e->n_shapes; j++) {
            int si = e->bodies[e->shapes[i].body].type == ORC_BODY_STATIC;
            int sj = e->bodies[e->shapes[j].body].type == ORC_BODY_STATIC;
            if (si && sj) continue; /* the static index is never collided with itself */
            collide_pair(e, i, j);
        }

    /* cpSpaceArbiterSetFilter: drop cached arbiters idle for >= collision_persistence stamps */
    {
        int w = 0;
        for (int i = 0; i < e->n_arbs; i++) {
            orc_arbiter *arb = &e->arbs[i];
            int ticks = e->stamp - arb->stamp;
            if (ticks >= 1 && arb->state != ORC_ARB_CACHED) arb->state = ORC_ARB_CACHED;
            if (ticks >= e->collision_persistence) continue;
            if (w != i) e->arbs[w] = *arb;
            w++;
        }
        e->n_arbs = w;
    }

    /* solver order = ascending (sa', sb') of the original shape pair (i < j) */
    int order[ORC_MAX_ARBS], n_live = 0;
    for (int i = 0; i < e->n_arbs; i++) if (e->arbs[i].live) order[n_live++] = i;
    for (int i = 1; i < n_live; i++) { /* insertion sort by (min idx, max idx) */
        int x = order[i], k = i - 1;
        #define KEY(q) ((e->arbs[q].sa < e->arbs[q].sb ? e->arbs[q].sa : e->arbs[q].sb) * 1024 + \
                        (e->arbs[q].sa < e->arbs[q].sb ? e->arbs[q].sb : e->arbs[q].sa))
        while (k >= 0 && KEY(order[k]) > KEY(x)) { order[k + 1] = order[k]; k--; }
        order[k + 1] = x;
        #undef KEY
    }

    double slop = e->collision_slop;
    double bias_c = 1.0 - pow(e->collision_bias, dt);
    for (int i = 0; i < n_live; i++) arb_prestep(e, &e->arbs[order[i]], dt, slop, bias_c);
    for (int i = 0; i < e->n_constraints; i++) con_prestep(e, &e->cons[i], dt);

    /* integrate velocities: cpBodyUpdateVelocity with gravity 0, damping 1, no forces.
     * v = v * 1 + (0 + f * m_inv) * dt is the identity; kinematic bodies are skipped. */

    double dt_coef = (prev_dt == 0.0 ? 0.0 : dt / prev_dt);
    for (int i = 0; i < n_live; i++) arb_apply_cached(e, &e->arbs[order[i]], dt_coef);
    for (int i = 0; i < e->n_constraints; i++) con_apply_cached(e, &e->cons[i], dt_coef);

    for (int it = 0; it < e->iterations; it++) {
        for (int j = 0; j < n_live; j++) arb_apply_impulse(e, &e->arbs[order[j]]);
        for (int j = 0; j < e->n_constraints; j++) con_apply_impulse(e, &e->cons[j], dt);
    }
}

/* ---------------------------------------------------------------- Robot + env step */

/* magical/entities.py:148-190: id = ud + 3*lr + 9*grip */
static void robot_set_action(orc_env *e, int action)
{
    int ud = action % 3, lr = (action / 3) % 3, grip = action / 9;
    const double radius = 0.2;
    e->rel_turn_angle = 0.0;
    e->target_speed = 0.0;
    if (ud == 1) e->target_speed += 4.0 * radius;   /* UP */
    if (ud == 2) e->target_speed -= 3.0 * radius;   /* DOWN */
    if (lr == 1) e->rel_turn_angle += 1.5;          /* LEFT */
    if (lr == 2) e->rel_turn_angle -= 1.5;          /* RIGHT */
    if (grip == 0) e->target_finger_angle = M_PI / 8;   /* OPEN: finger_rot_limit_outer */
    else e->target_finger_angle = -0.0;                 /* CLOSE: -finger_rot_limit_inner */
}

/* Robot.update (entities.py:455-476) */
static void robot_update(orc_env *e)
{
    orc_body *robot = &e->bodies[e->robot_body0 + MG_RB_ROBOT];
    orc_body *control = &e->bodies[e->robot_body0 + MG_RB_CONTROL];
    orc_body_set_angle(control, robot->a + e->rel_turn_angle);
    vec x_vel = { 0.0, e->target_speed };
    control->v = vrotate(robot->rot, x_vel); /* rotation_vector.cpvrotate(x_vel_vector) */
    for (int k = 0; k < 2; k++) {
        double side = k == 0 ? -1.0 : 1.0;
        orc_body *finger = &e->bodies[e->robot_body0 + MG_RB_FINGER_L + k];
        double rel_angle = finger->a - robot->a;
        double angle_error = rel_angle + side * e->target_finger_angle;
        double target_rate = fmax(-1.0, fmin(1.0, angle_error * 10));
        if (fabs(target_rate) < 1e-4) target_rate = 0.0;
        e->cons[e->motor_con[k]].rate = target_rate;
    }
}

/* BaseEnv.step minus render/score (base_env.py:267-288) + _phys_steps_on_frame (:248-255) */
void orc_env_step(orc_env *e, int action)
{
    robot_set_action(e, action);
    const int phys_steps = 10;  /* hard-coded in this fork, base_env.py:249 */
    double spf = 1.0 / 8;       /* fps = 8, benchmarks/__init__.py:435 */
    double dt = spf / phys_steps;
    for (int i = 0; i < phys_steps; i++) {
        robot_update(e);
        orc_space_step(e, dt);
    }
    e->episode_steps += 1;
}

/* space.shape_query(shape): every other shape (sensors included) that the narrow
 * phase reports >= 1 contact with, after the filter test (cpSpaceShapeQuery). */
int orc_shape_query(orc_env *e, int si, int *hits, int max_hits)
{
    int n = 0;
    orc_shape_update(e, si);
    orc_shape *a = &e->shapes[si];
    for (int j = 0; j < e->n_shapes; j++) {
        if (j == si) continue;
        orc_shape *b = &e->shapes[j];
        if (orc_filter_reject(a, b)) continue;
        orc_manifold m;
        if (orc_collide_shapes(a, b, &m) > 0) {
            if (n < max_hits) hits[n] = j;
            n++;
        }
    }
    return n;
}
