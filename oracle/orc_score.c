/*
 * orc_score.c -- ORACLE (test infrastructure): score_on_end_of_traj of the nine
 * task classes in magical/benchmarks/, and the two debug shaped rewards.
 * Each function cites the reference lines it restates.  The numpy-only scores
 * (MoveToCorner, MakeLine, Cluster*, PickAndPlace) are additionally checked
 * against the reference source itself in tests/test_oracle_scores.py.
 */
#include <math.h>
#include <stdlib.h>
#include "orc.h"

static vec block_pos(const orc_env *e, int bi) { return e->bodies[e->block_body[bi]].p; }
static vec robot_pos(const orc_env *e) { return e->bodies[e->robot_body0].p; }

/* GoalRegion.get_overlapping_ents(com_overlap=True) restricted to blocks
 * (entities.py:803-863): a block counts iff every one of its shapes (a) is
 * reported by space.shape_query(goal_shape) and (b) has its body position inside
 * the goal's AABB (bb.contains_vect). */
static void goal_overlaps(orc_env *e, int gi, int *in_goal /* [n_blocks] */)
{
    int gs = e->goal_shape[gi];
    int hits[ORC_MAX_SHAPES];
    int nh = orc_shape_query(e, gs, hits, ORC_MAX_SHAPES);
    const orc_shape *g = &e->shapes[gs];
    for (int bi = 0; bi < e->scene.n_blocks; bi++) {
        const orc_entity *en = &e->ents[e->block_ent[bi]];
        vec p = block_pos(e, bi);
        int com_in = g->bb_l <= p.x && g->bb_r >= p.x && g->bb_b <= p.y && g->bb_t >= p.y;
        int all = 1;
        for (int s = en->first_shape; s < en->first_shape + en->n_shapes; s++) {
            int found = 0;
            for (int k = 0; k < nh; k++) if (hits[k] == s) found = 1;
            if (!(found && com_in)) all = 0;
        }
        in_goal[bi] = all;
    }
}

/* move_to_corner.py:67-77 */
static double score_move_to_corner(orc_env *e)
{
    vec p = robot_pos(e);
    double dx = -1.0 - p.x, dy = 1.0 - p.y;
    double dist = sqrt(dx * dx + dy * dy);
    double succeed_dist = sqrt(2.0) / 2;
    double furthest_dist = sqrt(2.0);
    double drange = furthest_dist - succeed_dist;
    return fmin(1.0, fmax(0.0, furthest_dist - dist) / drange);
}

/* move_to_region.py:85-94: goal_shape.point_query(robot.position).distance <= 0.
 * cpPolyShapePointQuery on the axis-aligned goal box (radius 0): the distance is
 * <= 0 exactly when the point is inside or on the rectangle. */
static double score_move_to_region(orc_env *e)
{
    const orc_shape *g = &e->shapes[e->goal_shape[0]];
    vec p = robot_pos(e);
    double maxd = -INFINITY;
    for (int i = 0; i < g->count; i++) {
        double d = (p.x - g->tv[i].x) * g->tn[i].x + (p.y - g->tv[i].y) * g->tn[i].y;
        if (d > maxd) maxd = d;
    }
    return maxd <= 0.0 ? 1.0 : 0.0;
}

/* match_regions.py:193-213 */
static double score_match_regions(orc_env *e)
{
    int in_goal[MG_MAX_BLOCKS];
    goal_overlaps(e, 0, in_goal);
    int n_t = 0, n_ot = 0, n_od = 0, n_o = 0;
    for (int i = 0; i < e->scene.n_blocks; i++) {
        int role = e->scene.blocks[i].role;
        if (role == MG_ROLE_TARGET) n_t++;
        if (in_goal[i]) {
            n_o++;
            if (role == MG_ROLE_TARGET) n_ot++;
            if (role == MG_ROLE_DISTRACTOR) n_od++;
        }
    }
    double target_frac_done = (double)n_ot / (double)n_t;
    double contamination = n_o == 0 ? 0.0 : (double)n_od / (double)n_o;
    return target_frac_done * (1 - contamination);
}

static int cmp_double(const void *a, const void *b)
{
    double x = *(const double *)a, y = *(const double *)b;
    return (x > y) - (x < y);
}

/* make_line.longest_line (make_line.py:31-71) */
static int longest_line(const vec *pts, int npts, double inlier_dist, double max_sep)
{
    int best = npts < 1 ? npts : 1;
    for (int i = 0; i < npts - 1; i++)
        for (int j = i + 1; j < npts; j++) {
            vec pi = pts[i];
            vec pj_off = { pts[j].x - pi.x, pts[j].y - pi.y };
            double nrm = sqrt(pj_off.x * pj_off.x + pj_off.y * pj_off.y);
            vec u = { pj_off.x / nrm, pj_off.y / nrm };
            double proj[MG_MAX_BLOCKS], inl[MG_MAX_BLOCKS];
            int n_in = 0;
            for (int k = 0; k < npts; k++) {
                vec off = { pts[k].x - pi.x, pts[k].y - pi.y };
                proj[k] = off.x * u.x + off.y * u.y;
                double rx = off.x - proj[k] * u.x, ry = off.y - proj[k] * u.y;
                double dist = sqrt(rx * rx + ry * ry);
                if (dist <= inlier_dist) inl[n_in++] = proj[k];
            }
            if (n_in <= best) continue;
            qsort(inl, n_in, sizeof(double), cmp_double);
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

/* make_line.py:142-152 */
static double score_make_line(orc_env *e)
{
    const double SHAPE_RAD = 0.2 * 0.6;
    vec pts[MG_MAX_BLOCKS];
    int n = e->scene.n_blocks;
    for (int i = 0; i < n; i++) pts[i] = block_pos(e, i);
    int line_len = longest_line(pts, n, SHAPE_RAD * 1.5, SHAPE_RAD * 3.5);
    int max_line_len = n;
    int min_line_len = max_line_len - 2 > 2 ? max_line_len - 2 : 2;
    int num = line_len - min_line_len > 0 ? line_len - min_line_len : 0;
    return (double)num / (double)(max_line_len - min_line_len);
}

/* find_dupe.py:203-216 */
static double score_find_dupe(orc_env *e)
{
    int in_goal[MG_MAX_BLOCKS];
    goal_overlaps(e, 0, in_goal);
    int n_ot = 0, n_od = 0, n_o = 0;
    for (int i = 0; i < e->scene.n_blocks; i++) {
        if (!in_goal[i]) continue;
        n_o++;
        if (e->scene.blocks[i].role == MG_ROLE_TARGET) n_ot++;
        if (e->scene.blocks[i].role == MG_ROLE_DISTRACTOR) n_od++;
    }
    double have_two = n_ot >= 2 ? 1.0 : 0.0;
    double contamination = n_o == 0 ? 0.0 : (double)n_od / (double)n_o;
    return have_two * (1 - contamination);
}

/* fix_colour.py:193-202 */
static double score_fix_colour(orc_env *e)
{
    for (int g = 0; g < e->scene.n_goals; g++) {
        int in_goal[MG_MAX_BLOCKS];
        goal_overlaps(e, g, in_goal);
        int want = e->scene.goals[g].want_block;
        int n_o = 0, has_want = 0;
        for (int i = 0; i < e->scene.n_blocks; i++)
            if (in_goal[i]) { n_o++; if (i == want) has_want = 1; }
        int ok = want < 0 ? (n_o == 0) : (n_o == 1 && has_want);
        if (!ok) return 0.0;
    }
    return 1.0;
}

/* cluster.py:166-216.  Characteristic values are np.unique() of the colour/type
 * names, i.e. sorted by string: colours blue<green<red<yellow, types
 * circle<pentagon<square<star.  Only the grouping matters for the result. */
static double score_cluster(orc_env *e, int by_type)
{
    static const int colour_rank[4] = { 2, 1, 0, 3 }; /* red, green, blue, yellow -> rank */
    static const int type_rank[4] = { 2, 1, 3, 0 };   /* square, pentagon, star, circle -> rank */
    int n = e->scene.n_blocks;
    int label[MG_MAX_BLOCKS], present[4] = { 0, 0, 0, 0 };
    for (int i = 0; i < n; i++) {
        label[i] = by_type ? type_rank[e->scene.blocks[i].type] : colour_rank[e->scene.blocks[i].colour];
        present[label[i]] = 1;
    }
    int idx_of[4], nvals = 0;
    for (int r = 0; r < 4; r++) idx_of[r] = present[r] ? nvals++ : -1;
    vec centroids[4];
    for (int r = 0; r < 4; r++) {
        if (!present[r]) continue;
        double sx = 0.0, sy = 0.0; int cnt = 0;
        for (int i = 0; i < n; i++) if (label[i] == r) { vec p = block_pos(e, i); sx += p.x; sy += p.y; cnt++; }
        centroids[idx_of[r]].x = sx / cnt; centroids[idx_of[r]].y = sy / cnt;
    }
    const double min_margin = 2.0;
    int n_blocks = 0, n_correct = 0;
    for (int r = 0; r < 4; r++) {
        if (!present[r]) continue;
        int c_idx = idx_of[r];
        for (int i = 0; i < n; i++) {
            if (label[i] != r) continue;
            n_blocks++;
            vec p = block_pos(e, i);
            double true_sse = 0.0, nearest_bad = INFINITY;
            for (int c = 0; c < nvals; c++) {
                double dx = p.x - centroids[c].x, dy = p.y - centroids[c].y;
                double sse = dx * dx + dy * dy;
                if (c == c_idx) true_sse = sse;
                else if (sse < nearest_bad) nearest_bad = sse;
            }
            double margin = min_margin * true_sse;
            n_correct += sqrt(true_sse) < sqrt(nearest_bad) - margin ? 1 : 0;
        }
    }
    double frac_correct = (double)n_correct / (double)(n_blocks > 1 ? n_blocks : 1);
    double thresh = 0.75;
    return fmax(frac_correct - thresh, 0) / (1 - thresh);
}

/* pick_and_place.py:88-102 (every loop iteration evaluates self.target_shape) */
static double score_pick_and_place(orc_env *e)
{
    const double SHAPE_RAD = 0.2 * 0.6;
    double best = -1e9;
    for (int k = 0; k < e->scene.n_valid_targets; k++) {
        vec sp = block_pos(e, e->scene.target_block);
        double dx = e->scene.target_pos[0] - sp.x, dy = e->scene.target_pos[1] - sp.y;
        double dist = sqrt(dx * dx + dy * dy);
        double furthest = sqrt(2.0);
        double drange = furthest - SHAPE_RAD;
        double score = fmin(1.0, fmax(0.0, furthest - dist) / drange);
        best = fmax(best, score);
    }
    return best;
}

double orc_score(orc_env *e)
{
    switch (e->scene.task) {
    case MG_TASK_MOVE_TO_CORNER: return score_move_to_corner(e);
    case MG_TASK_MOVE_TO_REGION: return score_move_to_region(e);
    case MG_TASK_MATCH_REGIONS: return score_match_regions(e);
    case MG_TASK_MAKE_LINE: return score_make_line(e);
    case MG_TASK_FIND_DUPE: return score_find_dupe(e);
    case MG_TASK_FIX_COLOUR: return score_fix_colour(e);
    case MG_TASK_CLUSTER_COLOUR: return score_cluster(e, 0);
    case MG_TASK_CLUSTER_SHAPE: return score_cluster(e, 1);
    case MG_TASK_PICK_AND_PLACE: return score_pick_and_place(e);
    }
    return 0.0;
}

/* move_to_corner.py:86-100 and pick_and_place.py:114-124 */
double orc_debug_reward(orc_env *e)
{
    vec rp = robot_pos(e);
    if (e->scene.task == MG_TASK_MOVE_TO_CORNER) {
        vec sp = block_pos(e, 0);
        double dx = sp.x - 0.0, dy = sp.y - 1.0;
        double shape_to_corner = sqrt(dx * dx + dy * dy);
        double rx = rp.x - sp.x, ry = rp.y - sp.y;
        double robot_to_shape = sqrt(rx * rx + ry * ry);
        double shaping = -shape_to_corner / 5 - fmax(robot_to_shape, 0.2) / 20;
        return shaping + score_move_to_corner(e);
    }
    if (e->scene.task == MG_TASK_PICK_AND_PLACE) {
        const double SHAPE_RAD = 0.2 * 0.6;
        vec sp = block_pos(e, e->scene.target_block);
        double dx = sp.x - e->scene.target_pos[0], dy = sp.y - e->scene.target_pos[1];
        double shape_to_target = sqrt(dx * dx + dy * dy);
        double rx = rp.x - sp.x, ry = rp.y - sp.y;
        double robot_to_shape = sqrt(rx * rx + ry * ry);
        double shaping = -shape_to_target / 5 - fmax(robot_to_shape, SHAPE_RAD) / 10;
        return shaping + score_pick_and_place(e);
    }
    return 0.0;
}
