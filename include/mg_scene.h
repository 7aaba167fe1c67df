/*
 * mg_scene.h -- the initial-state record ("scene") that crosses the drop-in boundary.
 *
 * One mg_scene describes everything a MAGICAL env holds right after
 * `BaseEnv.reset()` has run `on_reset()` (reference: magical/base_env.py:190-246):
 * which entities exist, in which add order, with which poses, colours, shape
 * types and physics variables, plus the few task parameters the score
 * functions need.  It is plain data (no pointers), so the same bytes are read by
 *   - the host reset engine that samples it (magical-1_b200/csrc/host_reset.cpp),
 *   - the CUDA init kernel that expands it into solver/raster state,
 *   - the CPU oracle under oracle/ (test infrastructure only).
 *
 * Entity add order (= draw order = constraint order) is always
 *     arena, goals[0..n_goals), [robot if robot_first], blocks[0..n_blocks), [robot otherwise]
 * which covers every task in the magical/benchmarks task files.
 */
#ifndef MG_SCENE_H
#define MG_SCENE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MG_MAX_BLOCKS 10
#define MG_MAX_GOALS 3
#define MG_NUM_ACTIONS 18
#define MG_RES 384      /* DEFAULT_RES, magical/benchmarks/__init__.py:28 */
#define MG_LORES 96     /* small_res of the LoRes* preprocessors, :272-306 */

/* magical/entities.py:564-570 (SHAPE_TYPES order) */
enum { MG_SQUARE = 0, MG_PENTAGON = 1, MG_STAR = 2, MG_CIRCLE = 3 };
/* magical/entities.py:571-577 (SHAPE_COLOURS order) */
enum { MG_RED = 0, MG_GREEN = 1, MG_BLUE = 2, MG_YELLOW = 3 };

/* one value per task class in magical/benchmarks/ */
enum {
    MG_TASK_MOVE_TO_CORNER = 0,
    MG_TASK_MOVE_TO_REGION = 1,
    MG_TASK_MATCH_REGIONS = 2,
    MG_TASK_MAKE_LINE = 3,
    MG_TASK_FIND_DUPE = 4,
    MG_TASK_FIX_COLOUR = 5,
    MG_TASK_CLUSTER_COLOUR = 6,
    MG_TASK_CLUSTER_SHAPE = 7,
    MG_TASK_PICK_AND_PLACE = 8,
    MG_NUM_TASKS = 9
};

/* robot rigid group, in Robot.setup add order (magical/entities.py:238-351) */
enum {
    MG_RB_ROBOT = 0,
    MG_RB_CONTROL = 1,
    MG_RB_PUPIL_L = 2,
    MG_RB_PUPIL_R = 3,
    MG_RB_FINGER_L = 4,
    MG_RB_FINGER_R = 5,
    MG_RB_COUNT = 6
};

/* block roles used by score functions */
enum { MG_ROLE_NONE = 0, MG_ROLE_TARGET = 1, MG_ROLE_DISTRACTOR = 2 };

typedef struct mg_block {
    int32_t type;   /* MG_SQUARE.. */
    int32_t colour; /* MG_RED.. */
    /* MatchRegions / FindDupe: MG_ROLE_*; Cluster*: unused (label derives from
     * colour or type); others: MG_ROLE_NONE */
    int32_t role;
    int32_t pad_;
    double x, y, angle;
} mg_block;

typedef struct mg_goal {
    int32_t colour;
    /* FixColour (fix_colour.py:127-141): index of the block that must be the
     * region's only occupant, or -1 when the region must end up empty. */
    int32_t want_block;
    double cx, cy; /* goal body position = (x + w/2, y - h/2), entities.py:790 */
    double h, w;
} mg_goal;

typedef struct mg_scene {
    int32_t task;
    int32_t horizon; /* max_episode_steps */
    int32_t n_blocks;
    int32_t n_goals;
    int32_t robot_first;  /* 1: MoveToCorner, PickAndPlace add the robot before the blocks */
    int32_t debug_reward; /* shaped per-step reward (move_to_corner.py:86, pick_and_place.py:114) */
    int32_t target_block; /* PickAndPlace: index of self.target_shape */
    int32_t n_valid_targets; /* PickAndPlace: len(valid_target_shapes) */
    /* PhysicsVariables in declaration order (base_env.py:53-57):
     * robot_pos_joint, robot_rot_joint, robot_finger, shape_trans_joint, shape_rot_joint max forces */
    double phys[5];
    double target_pos[2]; /* PickAndPlace target_position */
    double robot[MG_RB_COUNT][3]; /* x, y, angle of the six robot bodies */
    mg_block blocks[MG_MAX_BLOCKS];
    mg_goal goals[MG_MAX_GOALS];
} mg_scene;

#ifdef __cplusplus
}
#endif
#endif /* MG_SCENE_H */
