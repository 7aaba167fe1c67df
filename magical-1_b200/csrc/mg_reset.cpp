/*
 * mg_reset.cpp -- native host reset engine: BaseEnv.reset() up to the first render for the
 * nine tasks, consuming a numpy-legacy MT19937 stream draw for draw like the reference consumes `self.rng`.
 *
 * Reference: magical/base_env.py:190-246 (reset, phys-var sampling), magical/geom.py:116-384
 * (pm_randomise_pose / pm_randomise_all_poses / randomise_hw / pm_shift_bodies) and the `on_reset` of
 * magical/benchmarks/{move_to_corner,move_to_region,match_regions,make_line,find_dupe,fix_colour,cluster,
 * pick_and_place}.py.  The Python restatement of the same code (magical-1_b200/tasks.py) is the readable
 * specification; tests/test_native_reset.py requires both to produce byte-identical scenes and RNG states.
 *
 * numpy.random.RandomState pieces restated (numpy/random/src/mt19937, legacy-distributions.c,
 * _bounded_integers.pyx): random_sample (53-bit double from two 32-bit words), uniform, randint with the
 * masked-rejection 32-bit path, choice(arange(n)) = randint(0, n), shuffle of a Python list
 * (random_interval, from the back).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>
#include <thread>
#include <vector>
#include "../../include/mg_b200.h"

namespace {

/* ------------------------------------------------------------------ numpy legacy MT19937 stream */
struct Mt {
    uint32_t *key;   /* [624] */
    int32_t *pos;

    uint32_t next32()
    {
        const int N = 624, M = 397;
        if (*pos >= N) {
            int kk;
            uint32_t y;
            for (kk = 0; kk < N - M; kk++) {
                y = (key[kk] & 0x80000000u) | (key[kk + 1] & 0x7fffffffu);
                key[kk] = key[kk + M] ^ (y >> 1) ^ (-(int32_t)(y & 1) & 0x9908b0dfu);
            }
            for (; kk < N - 1; kk++) {
                y = (key[kk] & 0x80000000u) | (key[kk + 1] & 0x7fffffffu);
                key[kk] = key[kk + (M - N)] ^ (y >> 1) ^ (-(int32_t)(y & 1) & 0x9908b0dfu);
            }
            y = (key[N - 1] & 0x80000000u) | (key[0] & 0x7fffffffu);
            key[N - 1] = key[M - 1] ^ (y >> 1) ^ (-(int32_t)(y & 1) & 0x9908b0dfu);
            *pos = 0;
        }
        uint32_t y = key[(*pos)++];
        y ^= (y >> 11);
        y ^= (y << 7) & 0x9d2c5680u;
        y ^= (y << 15) & 0xefc60000u;
        y ^= (y >> 18);
        return y;
    }
    /* random_sample */
    double next_double()
    {
        int32_t a = (int32_t)(next32() >> 5), b = (int32_t)(next32() >> 6);
        return (a * 67108864.0 + b) / 9007199254740992.0;
    }
    /* RandomState.uniform(low, high): low + (high - low) * random_sample */
    double uniform(double lo, double hi)
    {
        double range = hi - lo;
        return lo + range * next_double();
    }
    /* masked rejection on 32-bit words: value in [0, max] (random_interval / buffered_bounded_masked_uint32) */
    uint32_t bounded(uint32_t max)
    {
        if (max == 0) return 0;           /* consumes nothing */
        uint32_t mask = max;
        mask |= mask >> 1; mask |= mask >> 2; mask |= mask >> 4; mask |= mask >> 8; mask |= mask >> 16;
        uint32_t v;
        while ((v = (next32() & mask)) > max) {}
        return v;
    }
    /* RandomState.randint(low, high): high exclusive */
    int randint(int lo, int hi) { return lo + (int)bounded((uint32_t)(hi - 1 - lo)); }
    /* RandomState.shuffle(list) */
    void shuffle(int *x, int n)
    {
        for (int i = n - 1; i >= 1; i--) {
            int j = (int)bounded((uint32_t)i);
            int t = x[i]; x[i] = x[j]; x[j] = t;
        }
    }
};

/* ------------------------------------------------------------------ constants (base_env.py:49-65, geom.py) */
const double ROBOT_RAD = 0.2;
const double SHAPE_RAD = ROBOT_RAD * 0.6;
const double ARENA_L = -1.0, ARENA_R = 1.0, ARENA_B = -1.0, ARENA_T = 1.0;
const double RAND_GOAL_MIN_SIZE = 0.5, RAND_GOAL_MAX_SIZE = 0.8;
const double JITTER_PCT = 0.05;
const double JITTER_POS_BOUND = 1.0 * JITTER_PCT / 2.0;
const double JITTER_ROT_BOUND = JITTER_PCT * M_PI;
const double JITTER_TARGET_BOUND = JITTER_PCT * (RAND_GOAL_MAX_SIZE - RAND_GOAL_MIN_SIZE) / 2;
const double PHYS_DEFAULT[5] = { 3, 1, 4, 1.5, 0.1 };
const double PHYS_LO[5] = { 2.2, 0.7, 2.5, 1.0, 0.07 }, PHYS_HI[5] = { 3.5, 1.5, 4.5, 1.8, 0.15 };

struct Placement {};   /* thrown where the reference raises geom.PlacementError */

inline double py_max(double a, double b) { return b > a ? b : a; }   /* Python max(a, b) */
inline double py_min(double a, double b) { return b < a ? b : a; }   /* Python min(a, b) */

/* pymunk Vec2d.rotated */
inline void rotated(double x, double y, double ang, double *ox, double *oy)
{
    double c = cos(ang), s = sin(ang);
    *ox = x * c - y * s;
    *oy = x * s + y * c;
}

enum { E_ROBOT = 0, E_BLOCK = 1, E_GOAL = 2 };
struct Ent { int kind, idx; };
inline bool same(Ent a, Ent b) { return a.kind == b.kind && (a.kind == E_ROBOT || a.idx == b.idx); }
const Ent ROBOT = { E_ROBOT, 0 };
const double NONE = NAN;   /* "limit is None" */

struct Saved { double v[6][3]; };

struct Builder {
    mg_scene *s;
    Mt *rng;
    bool robot_added;

    Builder(mg_scene *scene, Mt *r, int task, int horizon, bool rand_dynamics) : s(scene), rng(r), robot_added(false)
    {
        memset(s, 0, sizeof(*s));
        s->task = task; s->horizon = horizon;
        /* BaseEnv.reset: phys vars are drawn before on_reset (base_env.py:210-215) */
        for (int i = 0; i < 5; i++) s->phys[i] = rand_dynamics ? rng->uniform(PHYS_LO[i], PHYS_HI[i]) : PHYS_DEFAULT[i];
    }

    /* poses of the six robot bodies right after Robot.setup (entities.py:243-254, 267-277, 313-331) */
    void add_robot(double x, double y, double angle, bool first = false)
    {
        double out[6][3] = { { x, y, angle }, { x, y, angle }, { 0.0, 0.0, angle }, { 0.0, 0.0, angle } };
        for (int k = 0; k < 2; k++) {
            int side = k == 0 ? -1 : 1;
            double delta = side < 0 ? M_PI / 8 : -(M_PI / 8);
            double rx, ry;
            rotated(side * ROBOT_RAD * 0.45, ROBOT_RAD * 0.1, angle, &rx, &ry);
            out[4 + k][0] = x + rx; out[4 + k][1] = y + ry; out[4 + k][2] = angle + delta;
        }
        for (int b = 0; b < 6; b++) for (int k = 0; k < 3; k++) s->robot[b][k] = out[b][k];
        if (first) s->robot_first = 1;
        robot_added = true;
    }
    Ent add_block(int type, int colour, double x, double y, double angle, int role = MG_ROLE_NONE)
    {
        int i = s->n_blocks;
        mg_block *b = &s->blocks[i];
        b->type = type; b->colour = colour; b->role = role;
        b->x = x; b->y = y; b->angle = angle;
        s->n_blocks = i + 1;
        return Ent{ E_BLOCK, i };
    }
    /* GoalRegion(x, y, h, w, colour): (x, y) is the top-left corner (entities.py:783-790) */
    Ent add_goal(double x, double y, double h, double w, int colour, int want_block = -1)
    {
        int i = s->n_goals;
        mg_goal *g = &s->goals[i];
        g->colour = colour; g->want_block = want_block;
        g->h = h; g->w = w;
        g->cx = x + w / 2; g->cy = y - h / 2;
        s->n_goals = i + 1;
        return Ent{ E_GOAL, i };
    }

    void get_pose(Ent e, double *x, double *y, double *a) const
    {
        if (e.kind == E_ROBOT) { *x = s->robot[0][0]; *y = s->robot[0][1]; *a = s->robot[0][2]; }
        else if (e.kind == E_BLOCK) { *x = s->blocks[e.idx].x; *y = s->blocks[e.idx].y; *a = s->blocks[e.idx].angle; }
        else { *x = s->goals[e.idx].cx; *y = s->goals[e.idx].cy; *a = 0.0; }
    }
    /* geom.pm_shift_bodies (geom.py:362-384); NAN = keep */
    void shift(Ent e, bool set_pos, double nx, double ny, bool set_angle, double angle)
    {
        double px, py, pa;
        get_pose(e, &px, &py, &pa);
        if (!set_angle) angle = pa;
        if (!set_pos) { nx = px; ny = py; }
        if (e.kind == E_ROBOT) {
            double nw[6][3];
            for (int b = 0; b < 6; b++) {
                double bx = s->robot[b][0], by = s->robot[b][1], ba = s->robot[b][2], dx, dy;
                rotated(bx - px, by - py, angle - pa, &dx, &dy);
                nw[b][0] = nx + dx; nw[b][1] = ny + dy; nw[b][2] = angle + (ba - pa);
            }
            for (int b = 0; b < 6; b++) for (int k = 0; k < 3; k++) s->robot[b][k] = nw[b][k];
        } else if (e.kind == E_BLOCK) {
            double dx, dy;
            rotated(0.0, 0.0, angle - pa, &dx, &dy);
            s->blocks[e.idx].x = nx + dx; s->blocks[e.idx].y = ny + dy; s->blocks[e.idx].angle = angle + (pa - pa);
        } else {
            s->goals[e.idx].cx = nx; s->goals[e.idx].cy = ny;
        }
    }
    Saved save(Ent e) const
    {
        Saved sv;
        memset(&sv, 0, sizeof(sv));
        if (e.kind == E_ROBOT) { for (int b = 0; b < 6; b++) for (int k = 0; k < 3; k++) sv.v[b][k] = s->robot[b][k]; }
        else get_pose(e, &sv.v[0][0], &sv.v[0][1], &sv.v[0][2]);
        return sv;
    }
    void restore(Ent e, const Saved &sv)
    {
        if (e.kind == E_ROBOT) { for (int b = 0; b < 6; b++) for (int k = 0; k < 3; k++) s->robot[b][k] = sv.v[b][k]; }
        else if (e.kind == E_BLOCK) { s->blocks[e.idx].x = sv.v[0][0]; s->blocks[e.idx].y = sv.v[0][1]; s->blocks[e.idx].angle = sv.v[0][2]; }
        else { s->goals[e.idx].cx = sv.v[0][0]; s->goals[e.idx].cy = sv.v[0][1]; }
    }
    std::vector<Ent> all_entities() const
    {
        std::vector<Ent> v;
        for (int g = 0; g < s->n_goals; g++) v.push_back(Ent{ E_GOAL, g });
        for (int b = 0; b < s->n_blocks; b++) v.push_back(Ent{ E_BLOCK, b });
        if (robot_added) v.push_back(ROBOT);
        return v;
    }
    bool overlaps(Ent e, const std::vector<Ent> &active) const
    {
        int robot_active = 0;
        uint32_t bmask = 0, gmask = 0;
        for (const Ent &a : active) {
            if (same(a, e)) continue;
            if (a.kind == E_ROBOT) robot_active = 1;
            else if (a.kind == E_BLOCK) bmask |= 1u << a.idx;
            else gmask |= 1u << a.idx;
        }
        int code = e.kind == E_ROBOT ? MG_ENT_ROBOT : (e.kind == E_BLOCK ? MG_ENT_BLOCK : MG_ENT_GOAL);
        return mg_host_entity_overlaps(s, code, e.idx, robot_active, bmask, gmask) != 0;
    }
    /* geom.pm_randomise_pose (geom.py:116-264); `active` = entities whose shapes can be hit; limits NAN = None */
    int randomise_pose(Ent e, const std::vector<Ent> &active, bool rand_pos, bool rand_rot, double pos_limit, double rot_limit)
    {
        Saved saved = save(e);
        double ox, oy, oa;
        get_pose(e, &ox, &oy, &oa);
        double x_lo = ARENA_L, x_hi = ARENA_R, y_lo = ARENA_B, y_hi = ARENA_T;
        if (!isnan(pos_limit)) {
            x_lo = py_max(ARENA_L, ox - pos_limit); x_hi = py_min(ARENA_R, ox + pos_limit);
            y_lo = py_max(ARENA_B, oy - pos_limit); y_hi = py_min(ARENA_T, oy + pos_limit);
        }
        double rot_min = -M_PI, rot_max = M_PI;
        if (!isnan(rot_limit)) { rot_min = oa - rot_limit; rot_max = oa + rot_limit; }
        const int max_tries = 10000;
        int n_tries = 0;
        while (n_tries < max_tries) {
            double nx = ox, ny = oy, na = oa;
            if (rand_pos) { nx = rng->uniform(x_lo, x_hi); ny = rng->uniform(y_lo, y_hi); }
            if (rand_rot) na = rng->uniform(rot_min, rot_max);
            shift(e, true, nx, ny, true, na);
            if (!overlaps(e, active)) return n_tries;
            n_tries++;
        }
        restore(e, saved);
        throw Placement();
    }
    /* geom.pm_randomise_all_poses (geom.py:281-341); per-entity limits / flags */
    void randomise_all_poses(const std::vector<Ent> &ents, const std::vector<char> &rand_rot, const std::vector<double> &pos_limits,
                             const std::vector<double> &rot_limits, const std::vector<Ent> &ignore = {}, int max_retries = 10)
    {
        std::vector<Ent> others;
        for (const Ent &a : all_entities()) {
            bool skip = false;
            for (const Ent &x : ents) skip = skip || same(a, x);
            for (const Ent &x : ignore) skip = skip || same(a, x);
            if (!skip) others.push_back(a);
        }
        for (int retry = 0; retry < max_retries; retry++) {
            std::vector<Ent> active = others;
            bool ok = true;
            for (size_t k = 0; k < ents.size(); k++) {
                active.push_back(ents[k]);
                try {
                    randomise_pose(ents[k], active, true, rand_rot[k] != 0, pos_limits[k], rot_limits[k]);
                } catch (Placement &) {
                    if (retry == max_retries - 1) throw;
                    ok = false;
                    break;
                }
            }
            if (ok) break;
        }
    }
    void randomise_all_poses(const std::vector<Ent> &ents, double pos_limit, double rot_limit)
    {
        randomise_all_poses(ents, std::vector<char>(ents.size(), 1), std::vector<double>(ents.size(), pos_limit),
                            std::vector<double>(ents.size(), rot_limit));
    }
};

/* geom.randomise_hw (geom.py:344-359); linf_bound NAN = None */
void randomise_hw(double min_side, double max_side, Mt *rng, double cur_h, double cur_w, double linf_bound, double *h, double *w)
{
    double lo[2] = { min_side, min_side }, hi[2] = { max_side, max_side }, cur[2] = { cur_h, cur_w };
    if (!isnan(linf_bound))
        for (int k = 0; k < 2; k++) { lo[k] = fmax(lo[k], cur[k] - linf_bound); hi[k] = fmin(hi[k], cur[k] + linf_bound); }
    *h = rng->uniform(lo[0], hi[0]);
    *w = rng->uniform(lo[1], hi[1]);
}

/* ------------------------------------------------------------------ the nine tasks */

enum { SQ = MG_SQUARE, PE = MG_PENTAGON, ST = MG_STAR, CI = MG_CIRCLE };
enum { RD = MG_RED, GR = MG_GREEN, BL = MG_BLUE, YE = MG_YELLOW };

/* rng.choice over en.SHAPE_COLOURS / en.SHAPE_TYPES (4 entries) = randint(0, 4) */
inline int choice4(Mt *rng) { return rng->randint(0, 4); }

struct Pose { double x, y, a; };

/* MoveToCornerEnv.on_reset (move_to_corner.py:31-65); kw: rand_shape_colour, rand_shape_type, rand_poses, debug_reward */
void reset_move_to_corner(Builder &b, const int32_t *kw)
{
    Mt *rng = b.rng;
    double rx = rng->next_double(), ry = rng->next_double();     /* rng.rand(2) */
    b.add_robot(rx, ry, 0.55 * M_PI, true);
    int colour = RD, stype = SQ;
    if (kw[0]) colour = choice4(rng);
    if (kw[1]) stype = choice4(rng);
    Ent blk = b.add_block(stype, colour, 0.1, -0.65, 0.13 * M_PI);
    b.s->debug_reward = kw[3] ? 1 : 0;
    if (kw[2]) b.randomise_all_poses({ ROBOT, blk }, JITTER_POS_BOUND, JITTER_ROT_BOUND);
}

/* MoveToRegionEnv.on_reset (move_to_region.py:30-83); kw: rand_poses_minor, rand_poses_full, rand_goal_colour */
void reset_move_to_region(Builder &b, const int32_t *kw)
{
    Mt *rng = b.rng;
    const bool minor = kw[0], full = kw[1];
    double x = -0.62, y = -0.17, h = 0.76, w = 0.75;
    if (minor || full) randomise_hw(RAND_GOAL_MIN_SIZE, RAND_GOAL_MAX_SIZE, rng, h, w, minor ? JITTER_TARGET_BOUND : NONE, &h, &w);
    int colour = kw[2] ? choice4(rng) : BL;
    Ent goal = b.add_goal(x, y, h, w, colour);
    b.add_robot(0.058, 0.53, -2.13);
    if (minor || full) {
        std::vector<double> pos(2, minor ? JITTER_POS_BOUND : NONE);
        std::vector<double> rot = { NONE, minor ? JITTER_ROT_BOUND : NONE };
        b.randomise_all_poses({ goal, ROBOT }, { 0, 1 }, pos, rot);
    }
}

/* MatchRegionsEnv.on_reset (match_regions.py:44-191);
 * kw: rand_target_colour, rand_shape_type, rand_shape_count, rand_layout_minor, rand_layout_full */
void reset_match_regions(Builder &b, const int32_t *kw)
{
    Mt *rng = b.rng;
    const bool minor = kw[3], full = kw[4];
    int target_colour = kw[0] ? choice4(rng) : GR;
    int distractor_colours[3], nd = 0;
    for (int c = 0; c < 4; c++) if (c != target_colour) distractor_colours[nd++] = c;
    double th = 0.7, tw = 0.6, tx = 0.1, ty = 0.7;
    if (minor || full) randomise_hw(RAND_GOAL_MIN_SIZE, RAND_GOAL_MAX_SIZE, rng, th, tw, minor ? JITTER_TARGET_BOUND : NONE, &th, &tw);
    Ent sensor = b.add_goal(tx, ty, th, tw, target_colour);
    const int def_target_types[2] = { ST, SQ };
    const int def_distractor_types[3][2] = { { 0, 0 }, { PE, 0 }, { CI, PE } };
    const Pose def_target_poses[2] = { { 0.8, -0.7, 2.37 }, { -0.68, 0.72, 1.28 } };
    const Pose def_distractor_poses[3][2] = { { { 0, 0, 0 }, { 0, 0, 0 } }, { { -0.05, -0.2, -1.09 }, { 0, 0, 0 } },
                                              { { -0.75, -0.55, 2.78 }, { 0.3, -0.82, -1.15 } } };
    int target_count = 2, distractor_counts[3] = { 0, 1, 2 };
    if (kw[2]) {
        target_count = rng->randint(1, 2 + 1);
        for (int k = 0; k < 3; k++) distractor_counts[k] = rng->randint(0, 2 + 1);
    }
    int target_types[2], distractor_types[3][2];
    if (kw[1]) {
        for (int k = 0; k < target_count; k++) target_types[k] = choice4(rng);
        for (int d = 0; d < 3; d++) for (int k = 0; k < distractor_counts[d]; k++) distractor_types[d][k] = choice4(rng);
    } else {
        for (int k = 0; k < 2; k++) target_types[k] = def_target_types[k];
        for (int d = 0; d < 3; d++) for (int k = 0; k < 2; k++) distractor_types[d][k] = def_distractor_types[d][k];
    }
    std::vector<Ent> shape_ents;
    /* zip(types, poses): without rand_layout_full the default pose lists bound the counts (they have the
     * default lengths, and rand_shape_count implies rand_layout_full) */
    for (int k = 0; k < target_count; k++) {
        Pose p = full ? Pose{ 0, 0, 0 } : def_target_poses[k];
        shape_ents.push_back(b.add_block(target_types[k], target_colour, p.x, p.y, p.a, MG_ROLE_TARGET));
    }
    for (int d = 0; d < 3; d++)
        for (int k = 0; k < distractor_counts[d]; k++) {
            Pose p = full ? Pose{ 0, 0, 0 } : def_distractor_poses[d][k];
            shape_ents.push_back(b.add_block(distractor_types[d][k], distractor_colours[d], p.x, p.y, p.a, MG_ROLE_DISTRACTOR));
        }
    b.add_robot(-0.5, 0.1, -M_PI * 1.2);
    if (minor || full) {
        std::vector<Ent> all = { sensor, ROBOT };
        all.insert(all.end(), shape_ents.begin(), shape_ents.end());
        std::vector<char> rr(all.size(), 1);
        rr[0] = 0;
        b.randomise_all_poses(all, rr, std::vector<double>(all.size(), minor ? JITTER_POS_BOUND : NONE),
                              std::vector<double>(all.size(), minor ? JITTER_ROT_BOUND : NONE));
    }
}

/* MakeLineEnv.on_reset (make_line.py:93-140); kw: rand_colours, rand_shapes, rand_count, rand_layout_minor, rand_layout_full */
void reset_make_line(Builder &b, const int32_t *kw)
{
    Mt *rng = b.rng;
    const bool minor = kw[3], full = kw[4];
    int shapes[4] = { ST, CI, ST, PE }, colours[4] = { BL, YE, RD, GR };
    Pose poses[4] = { { 0.790, -0.820, -0.721 }, { -0.177, 0.383, -1.733 }, { -0.051, -0.128, 2.696 }, { -0.292, -0.745, -0.159 } };
    int n_blocks = 4;
    if (kw[2]) {
        n_blocks = rng->randint(3, 4 + 1);
        for (int k = 1; k < 4; k++) poses[k] = poses[0];
    }
    if (kw[0]) for (int k = 0; k < n_blocks; k++) colours[k] = choice4(rng);
    if (kw[1]) for (int k = 0; k < n_blocks; k++) shapes[k] = choice4(rng);
    std::vector<Ent> ents = { ROBOT };
    for (int k = 0; k < n_blocks; k++) ents.push_back(b.add_block(shapes[k], colours[k], poses[k].x, poses[k].y, poses[k].a));
    b.add_robot(0.702, -0.255, 0.347);
    if (minor || full) b.randomise_all_poses(ents, minor ? JITTER_POS_BOUND : NONE, minor ? JITTER_ROT_BOUND : NONE);
}

/* FindDupeEnv.on_reset (find_dupe.py:72-201); kw: rand_colours, rand_shapes, rand_count, rand_layout_minor, rand_layout_full */
void reset_find_dupe(Builder &b, const int32_t *kw)
{
    Mt *rng = b.rng;
    const bool minor = kw[3], full = kw[4];
    int query_colour = YE, query_shape = PE;
    int out_colours[6] = { GR, RD, RD, YE, BL, YE }, out_shapes[6] = { PE, CI, CI, SQ, ST, PE };
    Pose out_poses[6] = { { -0.066751, 0.7552, -2.9266 }, { -0.05195, 0.31468, 1.5418 }, { 0.57528, -0.46865, -2.2141 },
                          { 0.40594, -0.74977, 0.24582 }, { 0.45254, 0.3681, -1.0834 }, { 0.76849, -0.10652, 0.10028 } };
    int n_out = 6;
    if (kw[2]) n_out = rng->randint(1, 5 + 1) + 1;
    const int n_distractors = n_out - 1;
    /* zip(out_shapes, out_colours, out_poses) stops at the shortest list: a randomised list has n_out entries */
    int n_zip = 6;
    if (kw[0]) {
        query_colour = choice4(rng);
        for (int k = 0; k < n_distractors; k++) out_colours[k] = choice4(rng);
        out_colours[n_distractors] = query_colour;
        n_zip = n_out < n_zip ? n_out : n_zip;
    }
    if (kw[1]) {
        query_shape = choice4(rng);
        for (int k = 0; k < n_distractors; k++) out_shapes[k] = choice4(rng);
        out_shapes[n_distractors] = query_shape;
        n_zip = n_out < n_zip ? n_out : n_zip;
    }
    double gx = -0.72, gy = -0.22, gh = 0.67, gw = 0.72;
    if (minor || full) randomise_hw(RAND_GOAL_MIN_SIZE, RAND_GOAL_MAX_SIZE, rng, gh, gw, minor ? JITTER_TARGET_BOUND : NONE, &gh, &gw);
    Ent sensor = b.add_goal(gx, gy, gh, gw, query_colour);
    if (kw[2]) {
        for (int k = 0; k < 6; k++) out_poses[k] = Pose{ 0, 0, 0 };
        n_zip = n_out < n_zip ? n_out : n_zip;
    }
    std::vector<Ent> outside;
    for (int k = 0; k < n_zip; k++) {
        bool is_target = out_colours[k] == query_colour && out_shapes[k] == query_shape;
        outside.push_back(b.add_block(out_shapes[k], out_colours[k], out_poses[k].x, out_poses[k].y, out_poses[k].a,
                                      is_target ? MG_ROLE_TARGET : MG_ROLE_DISTRACTOR));
    }
    Ent query = b.add_block(query_shape, query_colour, -0.33, -0.49, -0.51, MG_ROLE_TARGET);
    b.add_robot(-0.57, 0.25, 3.83);
    if (minor || full) {
        const double pos_limit = minor ? JITTER_POS_BOUND : NONE, rot_limit = minor ? JITTER_ROT_BOUND : NONE;
        std::vector<Ent> all = { sensor, ROBOT };
        all.insert(all.end(), outside.begin(), outside.end());
        std::vector<char> rr(all.size(), 1);
        rr[0] = 0;
        b.randomise_all_poses(all, rr, std::vector<double>(all.size(), pos_limit), std::vector<double>(all.size(), rot_limit), { query });
        double query_pos_limit = py_max(0, py_min(gh, gw) / 2 - SHAPE_RAD / 2);
        if (minor) query_pos_limit = py_min(JITTER_POS_BOUND, query_pos_limit);
        double sx, sy, sa;
        b.get_pose(sensor, &sx, &sy, &sa);
        b.shift(query, true, sx, sy, false, 0.0);
        std::vector<Ent> active;
        for (const Ent &e : b.all_entities()) if (!same(e, sensor)) active.push_back(e);
        b.randomise_pose(query, active, true, true, query_pos_limit, rot_limit);
    }
}

/* FixColourEnv.on_reset (fix_colour.py:67-191); kw: rand_colours, rand_shapes, rand_count, rand_layout_minor, rand_layout_full */
void reset_fix_colour(Builder &b, const int32_t *kw)
{
    Mt *rng = b.rng;
    const bool minor = kw[3], full = kw[4];
    int block_colours[3] = { GR, GR, BL }, region_colours[3] = { GR, GR, RD }, block_shapes[3] = { PE, SQ, PE };
    Pose block_poses[3] = { { 0.289, 0.030, 0.307 }, { 0.133, -0.561, 1.699 }, { -0.336, 0.000, -1.529 } };
    double region_xyhw[3][4] = { { -0.032, 0.348, 0.427, 0.468 }, { 0.019, -0.391, 0.460, 0.458 }, { -0.681, 0.196, 0.498, 0.418 } };
    int n_regions = 3;
    if (kw[2]) {
        n_regions = rng->randint(2, 3 + 1);
        for (int k = 1; k < 3; k++) { block_poses[k] = block_poses[0]; for (int q = 0; q < 4; q++) region_xyhw[k][q] = region_xyhw[0][q]; }
    }
    if (kw[0]) {
        for (int k = 0; k < n_regions; k++) region_colours[k] = choice4(rng);
        for (int k = 0; k < n_regions; k++) block_colours[k] = region_colours[k];
        int odd_idx = rng->randint(0, n_regions);
        int new_col_idx = rng->randint(0, 4 - 1);
        if (new_col_idx == block_colours[odd_idx]) new_col_idx += 1;
        block_colours[odd_idx] = new_col_idx;
    }
    if (kw[1]) for (int k = 0; k < n_regions; k++) block_shapes[k] = choice4(rng);
    /* the zips below stop at the shortest list: n_regions unless nothing was randomised (then 3) */
    const int n_regions_hw = kw[2] ? n_regions : 3;
    if (minor || full)
        for (int k = 0; k < n_regions_hw; k++)
            randomise_hw(0.4, 0.5, rng, region_xyhw[k][2], region_xyhw[k][3], minor ? JITTER_TARGET_BOUND : NONE,
                         &region_xyhw[k][2], &region_xyhw[k][3]);
    int n = n_regions;
    std::vector<Ent> sensors, blocks;
    for (int k = 0; k < n; k++)
        /* fix_colour.py:127-141: the region keeps its block only when the colours agree */
        sensors.push_back(b.add_goal(region_xyhw[k][0], region_xyhw[k][1], region_xyhw[k][2], region_xyhw[k][3], region_colours[k],
                                     block_colours[k] == region_colours[k] ? k : -1));
    for (int k = 0; k < n; k++) blocks.push_back(b.add_block(block_shapes[k], block_colours[k], block_poses[k].x, block_poses[k].y, block_poses[k].a));
    b.add_robot(0.368, 0.586, 0.718);
    if (minor || full) {
        const double pos_limit = minor ? JITTER_POS_BOUND : NONE, rot_limit = minor ? JITTER_ROT_BOUND : NONE;
        std::vector<Ent> ents = sensors;
        ents.push_back(ROBOT);
        std::vector<char> rr(ents.size(), 0);
        rr[n_regions] = 1;
        b.randomise_all_poses(ents, rr, std::vector<double>(ents.size(), pos_limit), std::vector<double>(ents.size(), rot_limit), blocks);
        for (int k = 0; k < n; k++) {
            double sx, sy, sa;
            b.get_pose(sensors[k], &sx, &sy, &sa);
            b.shift(blocks[k], true, sx, sy, false, 0.0);
        }
        for (int k = 0; k < n; k++) {
            double block_pos_limit = py_max(0, py_min(region_xyhw[k][2], region_xyhw[k][3]) / 2 - SHAPE_RAD);
            if (minor) block_pos_limit = py_min(JITTER_POS_BOUND, block_pos_limit);
            std::vector<Ent> active;
            for (const Ent &e : b.all_entities()) if (!same(e, sensors[k])) active.push_back(e);
            b.randomise_pose(blocks[k], active, true, true, block_pos_limit, rot_limit);
        }
    }
}

/* BaseClusterEnv.on_reset (cluster.py:67-164) with the defaults of ClusterColourEnv (:219-256) and
 * ClusterShapeEnv (:259-297); kw: rand_shape_colour, rand_shape_type, rand_layout_minor, rand_layout_full, rand_shape_count */
void reset_cluster(Builder &b, const int32_t *kw, bool by_type)
{
    Mt *rng = b.rng;
    const bool minor = kw[2], full = kw[3];
    static const int col_c[8] = { BL, BL, BL, GR, GR, RD, YE, YE }, shp_c[8] = { CI, ST, SQ, PE, PE, SQ, ST, PE };
    static const Pose pose_c[8] = { { -0.5147, 0.14149, -0.38871 }, { -0.1347, -0.71414, 1.0533 }, { -0.74247, -0.097592, 1.1571 },
                                    { -0.077363, -0.42964, -0.64379 }, { 0.51978, 0.1853, -1.1762 }, { -0.5278, -0.21642, 2.9356 },
                                    { -0.54039, 0.48292, 0.072818 }, { -0.16761, 0.64303, -2.3255 } };
    static const int col_t[8] = { YE, BL, RD, RD, GR, YE, BL, GR }, shp_t[8] = { SQ, PE, PE, PE, CI, ST, ST, CI };
    static const Pose pose_t[8] = { { -0.414, 0.297, -1.731 }, { 0.068, 0.705, 2.184 }, { 0.821, 0.220, 0.650 }, { -0.461, -0.749, -2.673 },
                                    { 0.867, -0.149, -2.215 }, { -0.785, -0.140, -0.405 }, { -0.305, -0.226, 1.341 }, { 0.758, -0.708, -2.140 } };
    const int *dcol = by_type ? col_t : col_c, *dshp = by_type ? shp_t : shp_c;
    const Pose *dpose = by_type ? pose_t : pose_c;
    const Pose robot = by_type ? Pose{ 0.286, -0.202, -1.878 } : Pose{ 0.71692, -0.34374, 0.83693 };
    int n_shapes = 8;
    Pose poses[MG_MAX_BLOCKS];
    if (kw[4]) {
        n_shapes = rng->randint(7, 10 + 1);
        for (int k = 0; k < n_shapes; k++) poses[k] = Pose{ 0, 0, 0 };
    } else {
        for (int k = 0; k < 8; k++) poses[k] = dpose[k];
    }
    int colours[MG_MAX_BLOCKS], types[MG_MAX_BLOCKS];
    int n_col = 8, n_typ = 8;
    if (kw[0]) {
        colours[0] = RD; colours[1] = GR; colours[2] = BL; colours[3] = YE;
        for (int k = 4; k < n_shapes; k++) colours[k] = choice4(rng);
        n_col = n_shapes > 4 ? n_shapes : 4;
        rng->shuffle(colours, n_col);
    } else {
        for (int k = 0; k < 8; k++) colours[k] = dcol[k];
    }
    if (kw[1]) {
        types[0] = SQ; types[1] = PE; types[2] = ST; types[3] = CI;
        for (int k = 4; k < n_shapes; k++) types[k] = choice4(rng);
        n_typ = n_shapes > 4 ? n_shapes : 4;
        rng->shuffle(types, n_typ);
    } else {
        for (int k = 0; k < 8; k++) types[k] = dshp[k];
    }
    /* zip(poses, colours, shape_types) */
    int n = kw[4] ? n_shapes : 8;
    n = n < n_col ? n : n_col;
    n = n < n_typ ? n : n_typ;
    std::vector<Ent> ents = { ROBOT };
    for (int k = 0; k < n; k++) ents.push_back(b.add_block(types[k], colours[k], poses[k].x, poses[k].y, poses[k].a));
    b.add_robot(robot.x, robot.y, robot.a);
    if (minor || full) b.randomise_all_poses(ents, full ? NONE : JITTER_POS_BOUND, full ? NONE : JITTER_ROT_BOUND);
}

/* PickAndPlaceEnv.on_reset (pick_and_place.py:31-79); kw: rand_shape_colour, rand_shape_type, rand_poses, debug_reward */
void reset_pick_and_place(Builder &b, const int32_t *kw)
{
    Mt *rng = b.rng;
    b.add_robot(0.0, 0.0, 0.55 * M_PI, true);
    std::vector<Ent> ents = { ROBOT };
    int st[3], col[3];
    for (int k = 0; k < 3; k++) {
        col[k] = RD; st[k] = SQ;
        if (kw[0]) col[k] = choice4(rng);
        if (kw[1]) st[k] = choice4(rng);
        ents.push_back(b.add_block(st[k], col[k], 0.1, -0.65, 0.13 * M_PI));
    }
    double tx = rng->next_double() * 2 - 1, ty = rng->next_double() * 2 - 1;     /* rng.rand(2) * 2 - 1 */
    int valid[3], nv = 0;
    for (int k = 0; k < 3; k++) if (st[k] == st[0] && col[k] == col[0]) valid[nv++] = k;
    int target = rng->randint(0, nv);                                           /* rng.choice(np.arange(len(valid))) */
    b.s->target_block = valid[target]; b.s->n_valid_targets = nv;
    b.s->target_pos[0] = tx; b.s->target_pos[1] = ty;
    b.s->debug_reward = kw[3] ? 1 : 0;
    if (kw[2]) b.randomise_all_poses(ents, NONE, NONE);
}

int sample_one(const mg_reset_spec *spec, uint32_t *mt_key, int32_t *mt_pos, mg_scene *out)
{
    if (!spec || !mt_key || !mt_pos || !out || spec->task < 0 || spec->task >= MG_NUM_TASKS) return MG_ERR_ARG;
    Mt rng{ mt_key, mt_pos };
    try {
        Builder b(out, &rng, spec->task, spec->horizon, spec->rand_dynamics != 0);
        switch (spec->task) {
        case MG_TASK_MOVE_TO_CORNER: reset_move_to_corner(b, spec->kw); break;
        case MG_TASK_MOVE_TO_REGION: reset_move_to_region(b, spec->kw); break;
        case MG_TASK_MATCH_REGIONS: reset_match_regions(b, spec->kw); break;
        case MG_TASK_MAKE_LINE: reset_make_line(b, spec->kw); break;
        case MG_TASK_FIND_DUPE: reset_find_dupe(b, spec->kw); break;
        case MG_TASK_FIX_COLOUR: reset_fix_colour(b, spec->kw); break;
        case MG_TASK_CLUSTER_COLOUR: reset_cluster(b, spec->kw, false); break;
        case MG_TASK_CLUSTER_SHAPE: reset_cluster(b, spec->kw, true); break;
        default: reset_pick_and_place(b, spec->kw); break;
        }
    } catch (Placement &) {
        return MG_ERR_PLACEMENT;
    }
    return MG_OK;
}

}  // namespace

extern "C" int mg_host_sample_scene(const mg_reset_spec *spec, uint32_t *mt_key, int32_t *mt_pos, mg_scene *out)
{
    return sample_one(spec, mt_key, mt_pos, out);
}

extern "C" int mg_host_sample_scenes(const mg_reset_spec *specs, int n, uint32_t *mt_keys, int32_t *mt_pos, mg_scene *out,
                                     int32_t *status, int n_threads)
{
    if (!specs || n < 0 || !mt_keys || !mt_pos || !out || !status) return MG_ERR_ARG;
    if (n_threads < 1) n_threads = 1;
    if (n_threads > n) n_threads = n > 0 ? n : 1;
    auto work = [&](int t) {
        for (int i = t; i < n; i += n_threads)
            status[i] = sample_one(&specs[i], mt_keys + (size_t)i * 624, &mt_pos[i], &out[i]);
    };
    if (n_threads == 1) { work(0); return MG_OK; }
    std::vector<std::thread> pool;
    for (int t = 0; t < n_threads; t++) pool.emplace_back(work, t);
    for (auto &th : pool) th.join();
    return MG_OK;
}
