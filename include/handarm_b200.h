/*
 * handarm_b200 -- C-ABI of the B200-native task-side hot path of the Isaac Gym
 * hand / hand-arm environments.
 *
 * The reference (maltemosbach/isaacgym-hand-arm) has no FFI for this path: the
 * boundary is the Python VecTask class contract (isaacgymenvs/tasks/base/
 * vec_task.py:390-441 `step`, :378-388 the `pre_physics_step` /
 * `post_physics_step` hooks).  This library is the engine behind a drop-in
 * Python VecTask: each entry point replaces one Python hook of one task family
 * and works directly on the tensors that hook works on.
 *
 * Conventions
 *   - every pointer in a *Buffers struct is a DEVICE pointer, row-major, fp32
 *     unless the field name / comment says otherwise; quaternions are xyzw;
 *   - no allocation, no host synchronisation, stream-ordered, capturable in a
 *     CUDA graph; the library never frees or retains caller memory;
 *   - return value: 0 = ok, <0 = error (see ha_last_error());
 *   - there is NO CPU fallback: without a CUDA device every compute entry
 *     point fails with HA_ERR_CUDA.
 */
#ifndef HANDARM_B200_H
#define HANDARM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HA_ABI_VERSION 6

#define HA_OK 0
#define HA_ERR_INVALID (-1)   /* bad argument / unsupported dimension */
#define HA_ERR_CUDA (-2)      /* CUDA runtime error (message in ha_last_error) */
#define HA_ERR_ALIGN (-3)     /* a pointer violates the alignment contract */

#define HA_MAX_DOFS 48        /* two arms: 2 x (7 + 16) */
#define HA_MAX_TIPS 8
#define HA_MAX_KEYPOINTS 4
#define HA_MAX_ARMS 2
#define HA_ENVS_PER_CTA 32    /* env tile handled by one thread block */
#define HA_STATS_LEN 16

/* opaque CUDA stream handle (cudaStream_t); 0 = legacy default stream */
typedef void* ha_stream_t;

/* ------------------------------------------------------------------------ */
/* AllegroKuka family (single arm: reorientation / regrasping / throw;       */
/* two arms: regrasping / reorientation)                                     */
/* replaces allegro_kuka_base.py:1355-1424 (pre_physics_step incl. reset_idx  */
/* :1246-1353, reset_target_pose :1191-1196, reset_object_pose :1198-1224) and */
/* :1426-1443 (post_physics_step: compute_observations :991-1089,             */
/* compute_full_state :1091-1172, compute_kuka_reward :854-930) and the       */
/* VecTask.step tail vec_task.py:424-432.                                     */
/* ------------------------------------------------------------------------ */

enum { HA_KUKA_REORIENTATION = 0, HA_KUKA_REGRASPING = 1, HA_KUKA_THROW = 2 };

/* Layout of the simulator-owned root / rigid-body state tensors.
 *   HA_STATE_ISAACGYM  what Isaac Gym's tensor API hands out (allegro_kuka_base.py:280-295):
 *                      root [n_envs*actors_per_env, 13], rigid-body [n_envs*n_bodies, 13] (AoS,
 *                      52-byte rows); every needed row costs whole 128-byte DRAM lines.
 *   HA_STATE_SOA       component-major planes for producers that can emit them (the north star's
 *                      "coalesced 128-bit loads of SoA state"): root [actors_per_env][13][n_envs],
 *                      rigid-body [n_bodies][13][n_envs]; needs the regular actor numbering
 *                      (root_regular).  dof_state keeps Isaac Gym's [n_envs, dofs, 2] in both. */
enum { HA_STATE_ISAACGYM = 0, HA_STATE_SOA = 1 };

/* reward terms kept per episode, in the reference's key order
 * (allegro_kuka_base.py:355-368) */
enum {
  HA_KUKA_RAW_FINGERTIP_DELTA = 0, HA_KUKA_RAW_HAND_DELTA, HA_KUKA_RAW_LIFTING, HA_KUKA_RAW_KEYPOINT,
  HA_KUKA_FINGERTIP_DELTA, HA_KUKA_HAND_DELTA, HA_KUKA_LIFTING, HA_KUKA_LIFT_BONUS,
  HA_KUKA_KEYPOINT, HA_KUKA_BONUS, HA_KUKA_KUKA_PENALTY, HA_KUKA_ALLEGRO_PENALTY,
  HA_KUKA_N_EPISODE_TERMS
};
/* per-step terms exported as extras["episode_cumulative"]
 * (allegro_kuka_base.py:912-927), same order as the reference's list */
enum {
  HA_KUKA_CUM_FINGERTIP_DELTA = 0, HA_KUKA_CUM_HAND_DELTA, HA_KUKA_CUM_LIFTING, HA_KUKA_CUM_LIFT_BONUS,
  HA_KUKA_CUM_KEYPOINT, HA_KUKA_CUM_KUKA_PENALTY, HA_KUKA_CUM_ALLEGRO_PENALTY, HA_KUKA_CUM_BONUS,
  HA_KUKA_N_STEP_TERMS
};

/* layout of one env's row of externally supplied random numbers (parity mode);
 * raw U(0,1) / N(0,1) draws, in the order the reference consumes them
 * (SURVEY.md appendix B).  Fast mode generates the same slots with Philox. */
enum {
  HA_KUKA_RND_GOAL_POS = 0,      /* 3: target position, goal-only reset        */
  HA_KUKA_RND_GOAL_ROT = 3,      /* 3: target rotation uvw, goal-only reset    */
  HA_KUKA_RND_RESET_GOAL_POS = 6,/* 3: target position inside reset_idx        */
  HA_KUKA_RND_RESET_GOAL_ROT = 9,/* 3                                          */
  HA_KUKA_RND_OBJ_POS = 12,      /* 3: object position noise                   */
  HA_KUKA_RND_OBJ_ROT = 15,      /* 3: object rotation uvw                     */
  HA_KUKA_RND_FORCE_PROB = 18,   /* 1                                          */
  HA_KUKA_RND_FORCE_TEST = 19,   /* 1: Bernoulli test of the random force      */
  HA_KUKA_RND_FORCE_NORMAL = 20, /* 3: N(0,1)                                  */
  HA_KUKA_RND_SIDE = 23,         /* 2: two-arm coin flips (object side, goal side) */
  HA_KUKA_RND_DOF_POS = 25,      /* n_dofs                                     */
  /* HA_KUKA_RND_DOF_POS + n_dofs : n_dofs DOF velocity draws                   */
};
#define HA_KUKA_RND_STRIDE(n_dofs) (25 + 2 * (n_dofs))

/* two-arm regrasping (allegro_kuka_two_arms.py:1298-1351 + allegro_kuka_two_arms_regrasping.py:75-110): a goal block is
 * target position 3, target side 1, then the object reset that _reset_target performs (side 1, position 3, rotation 3);
 * reset_idx draws a second goal block and one more object block (reset_object_pose is called twice, the last one wins) */
enum {
  HA_KUKA2_RND_GOAL = 0,         /* 11: goal-only reset                                      */
  HA_KUKA2_RND_RESET_GOAL = 11,  /* 11: reset_target_pose inside reset_idx                   */
  HA_KUKA2_RND_RESET_OBJ = 22,   /* 7: reset_object_pose inside reset_idx                    */
  HA_KUKA2_RND_FORCE_PROB = 29,
  HA_KUKA2_RND_FORCE_TEST = 30,
  HA_KUKA2_RND_FORCE_NORMAL = 31, /* 3: N(0,1)                                               */
  HA_KUKA2_RND_DOF_POS = 34      /* n_dofs position draws, then n_dofs velocity draws        */
};
#define HA_KUKA2_RND_STRIDE(n_dofs) (34 + 2 * (n_dofs))
/* two arms keep 8 per-episode terms / 5 per-step terms (allegro_kuka_two_arms.py:345-354, :861-867):
 * raw_fingertip_delta, raw_lifting, raw_keypoint, fingertip_delta, lifting, lift_bonus, keypoint, bonus */
#define HA_KUKA2_N_EPISODE_TERMS 8
#define HA_KUKA2_N_STEP_TERMS 5

typedef struct HaKukaParams {
  int32_t abi_version;
  int32_t n_envs;
  int32_t n_arms;           /* 1 or 2 */
  int32_t n_dofs;           /* n_arms * 23 hand-arm DOFs that the task uses */
  int32_t dofs_per_env;     /* row count per env of dof_state / targets (>= n_dofs) */
  int32_t n_bodies;         /* rigid bodies per env */
  int32_t n_tips;           /* fingertips, all arms (4 or 8) */
  int32_t n_keypoints;      /* 1 or 4 */
  int32_t obs_dim;          /* 117 single arm K=4; see ha_kuka_obs_dim() */
  int32_t subtask;          /* HA_KUKA_* */
  int32_t palm_body[HA_MAX_ARMS];
  int32_t tip_body[HA_MAX_TIPS];
  int32_t object_body;      /* rigid-body handle of the object (object_rb_handles[0]) */
  int32_t max_episode_length;
  int32_t max_consecutive_successes;
  int32_t success_steps;
  int32_t use_fast_rng;     /* 0: consume ext_rand, 1: in-kernel Philox */
  int32_t force_enabled;    /* force_scale > 0 */
  /* root_regular != 0: the actor rows follow Isaac Gym's regular numbering
   * row(env, slot) = env * actors_per_env + slot, so the gather needs no
   * dependent index load; the host verifies this against the index tables */
  int32_t root_regular;
  int32_t actors_per_env;
  int32_t object_slot;
  int32_t goal_slot;
  int32_t state_layout;     /* HA_STATE_* */
  /* test / tuning hook: > 0 forces the CTA count of the persistent single-arm post grid (capped at HA_KUKA_STAT_SLOTS), so a
   * small env count can exercise the many-tiles-per-CTA regime of the headline size; 0: one full wave of resident CTAs */
  int32_t post_grid;

  float palm_offset[3];
  float tip_offset[HA_MAX_TIPS][3];
  float dof_lower[HA_MAX_DOFS];
  float dof_upper[HA_MAX_DOFS];
  float dof_default[HA_MAX_DOFS];
  float dof_reset_noise[HA_MAX_DOFS];   /* 0.1 arm / 0.1 fingers coefficients */
  float dof_range_rcp[HA_MAX_DOFS];     /* fp32(1 / (dof_upper - dof_lower)), see ha_kuka_fill_derived() */
  float clamp_abs_obs;      /* <=0: no clamp; VecTask clip_obs already folded in (min) */
  /* reward */
  float lifting_bonus_threshold, lifting_bonus, lifting_rew_scale;
  float keypoint_rew_scale, distance_delta_rew_scale, hand_delta_rew_scale /* = dist*0 */;
  float bonus_per_step;     /* reach_goal_bonus / success_steps */
  float kuka_actions_penalty_scale, allegro_actions_penalty_scale;
  float fall_z;             /* 0.1 */
  float far_tip_dist;       /* 1.5; reorientation `_extra_reset_rules` */
  /* pre-physics */
  float arm_action_gain;    /* dofSpeedScale * dt (relative control of the 7 arm DOFs) */
  float act_moving_average, one_minus_act_moving_average;
  float force_decay_pow;    /* force_decay ** (dt / force_decay_interval) */
  float force_gain[2];      /* object mass, force_scale (applied in this order) */
  float force_log_c0, force_log_c1; /* log(lo) - log(hi), log(hi) */
  /* reset */
  float target_min[3], target_size[3];      /* target volume */
  float reset_pos_noise[3];
  float reset_dof_vel_noise;
  float two_arm_x_ofs;      /* two arms: +-x offset of the goal side (0.75) */
  float two_arm_obj_x_ofs;  /* two arms: +-x offset of the object side (table_width / 2 - 0.2) */
  float reserved_f[5];
} HaKukaParams;

/* values that change while training (tolerance curriculum); they live in
 * device memory so that captured CUDA graphs see the updates */
typedef struct HaKukaDynamic {
  float keypoint_success_tolerance;  /* success_tolerance * keypoint_scale, rounded to fp32 */
  float true_objective_scale;        /* true_objective = successes * scale + offset */
  float true_objective_offset;
  float pad0;
  uint64_t rng_seed;
  uint64_t rng_step;                 /* one tick per ha_kuka_post_step / ha_kuka_reset_idx (fast RNG mode) */
} HaKukaDynamic;

typedef struct HaKukaBuffers {
  /* ---- simulator-owned state (Isaac Gym layout, AoS) ---- */
  float* root_state;          /* [n_envs*n_actors, 13] */
  const float* rigid_body_state; /* [n_envs, n_bodies, 13] */
  float* dof_state;           /* [n_envs, dofs_per_env, 2] */
  /* ---- index tables ---- */
  const int32_t* object_actor;   /* [n_envs] row of the object in root_state */
  const int32_t* goal_actor;     /* [n_envs] row of the goal object (-1 if none) */
  const int32_t* hand_actor;     /* [n_envs, n_arms] rows of the arm actors */
  /* ---- task constants per env ---- */
  const float* object_init_state;     /* [n_envs, 13] */
  const float* object_init_z;         /* [n_envs] compact copy of object_init_state[:, 2] (nullable: read from the rows) */
  const float* object_keypoint_offsets; /* [n_envs, K, 3] */
  const float* object_scales;         /* [n_envs, 3] */
  const float* finger_rew_coeffs;     /* [n_envs, n_tips] */
  /* ---- task state ---- */
  float* goal_states;         /* [n_envs, 13] */
  float* obs_buf;             /* [n_envs, obs_dim] */
  float* rew_buf;             /* [n_envs] */
  int64_t* reset_buf;         /* [n_envs] */
  int64_t* reset_goal_buf;    /* [n_envs] */
  int64_t* progress_buf;      /* [n_envs] */
  int64_t* randomize_buf;     /* [n_envs] (nullable) */
  uint8_t* timeout_buf;       /* [n_envs] bool */
  float* successes;           /* [n_envs] */
  float* prev_episode_successes;
  float* true_objective;
  float* prev_episode_true_objective;
  int32_t* near_goal_steps;   /* [n_envs] */
  uint8_t* lifted_object;     /* [n_envs] bool */
  float* closest_keypoint_max_dist; /* [n_envs] */
  float* closest_fingertip_dist;    /* [n_envs, n_tips] */
  float* furthest_hand_dist;        /* [n_envs] (single arm only) */
  float* rewards_episode;     /* [HA_KUKA_N_EPISODE_TERMS, n_envs] */
  float* episode_cumulative;  /* [HA_KUKA_N_STEP_TERMS, n_envs] */
  /* ---- control ---- */
  float* prev_targets;        /* [n_envs, dofs_per_env] */
  float* cur_targets;         /* [n_envs, dofs_per_env] */
  float* actions;             /* [n_envs, n_dofs] task-owned copy */
  float* rb_forces;           /* HA_STATE_ISAACGYM: [n_envs, n_bodies, 3] (apply_rigid_body_force_tensors layout);
                                 HA_STATE_SOA: [n_bodies, 3, n_envs] component planes like the state tensors */
  float* random_force_prob;   /* [n_envs] */
  /* ---- outputs for the simulator's *_indexed setters ---- */
  int32_t* root_reset_indices;  /* [2*n_envs] ascending actor rows */
  int32_t* hand_reset_indices;  /* [n_arms*n_envs] ascending actor rows */
  int32_t* reset_counts;        /* [4]: n_root, n_hand, n_reset_envs, n_goal_resets; 8-byte aligned */
  /* ---- reductions ---- */
  float* stats;                 /* [HA_STATS_LEN], see HA_KUKA_STAT_* */
  int64_t* stat_accum;          /* fixed-point statistics scratch.
                                   single arm, and two arms with n_envs <= 32 * HA_KUKA_STAT_SLOTS:
                                   [HA_KUKA_N_PARTIAL_STATS][HA_KUKA_STAT_SLOTS], one slot per CTA of the last
                                   ha_kuka_post_step (plain stores; initialise every slot of row i once with ha_kuka_stat_accum_init(i)), folded into stats[] by the next
                                   ha_kuka_pre_step / ha_kuka_reset_idx or by ha_kuka_stats_finalize;
                                   two arms beyond that size: [HA_STATS_LEN] accumulators, initialise once with ha_kuka_stat_accum_init(i):
                                   0 for sums, INT32_MIN for the two MAX slots (the kernel re-arms them itself) */
  uint32_t* tickets;            /* [8], zero-initialised once by the caller */
  /* ---- dynamic scalars ---- */
  HaKukaDynamic* dynamic;       /* device */
  /* ---- optional: cross-rank statistics exchange inside ha_kuka_post_step (single arm), no extra launch ----
   * xchg: NULL, or a HaStatsExchange (see the end of this header) that lives in DEVICE memory.  The last CTA of the post grid (it
   * owns the fewest tiles) then runs one exchange per launch in one of its warps: it publishes stats[] as the preceding
   * ha_kuka_pre_step folded it (the statistics of the previous post step) to every peer over NVLink and leaves the newest
   * whole-box reduction in xchg_reduced [HA_STATS_LEN] (device).  Same wait-free protocol as ha_stats_exchange. */
  const void* xchg;
  float* xchg_reduced;
} HaKukaBuffers;

/* per-CTA statistic slots of the single-arm post kernel (its persistent grid never exceeds this) */
#define HA_KUKA_STAT_SLOTS 768

/* stats[] of the last ha_kuka_post_step (sums over this GPU's env block).  Two arms: written by the post step itself.
 * Single arm: valid after the next ha_kuka_pre_step / ha_kuka_reset_idx or after ha_kuka_stats_finalize. */
enum {
  HA_KUKA_STAT_N = 0,            /* n_envs */
  HA_KUKA_STAT_SUM_REWARD,
  HA_KUKA_STAT_SUM_RESETS,
  HA_KUKA_STAT_SUM_SUCCESS_EVENTS,  /* is_success this step */
  HA_KUKA_STAT_SUM_PREV_EP_SUCCESSES,
  HA_KUKA_STAT_SUM_TRUE_OBJECTIVE,
  HA_KUKA_STAT_MAX_TRUE_OBJECTIVE,
  HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE, /* -min, so one MAX all-reduce covers both */
  HA_KUKA_STAT_SUM_TIMEOUTS,
  HA_KUKA_STAT_SUM_LIFTED,
  HA_KUKA_N_PARTIAL_STATS,          /* the entries above are per-rank partials (all-reduce SUM / MAX) */
  HA_KUKA_STAT_MEAN_REWARD = HA_KUKA_N_PARTIAL_STATS,   /* derived on device for extras[] */
  HA_KUKA_STAT_MEAN_PREV_EP_SUCCESSES,
  HA_KUKA_STAT_MEAN_TRUE_OBJECTIVE,
  HA_KUKA_STAT_MIN_TRUE_OBJECTIVE,
  HA_KUKA_N_STATS
};

int ha_abi_version(void);
const char* ha_last_error(void);
size_t ha_sizeof_kuka_params(void);
size_t ha_sizeof_kuka_buffers(void);
size_t ha_sizeof_kuka_dynamic(void);

/* obs row width for (n_arms, n_keypoints); allegro_kuka_base.py:185-220,
 * allegro_kuka_two_arms.py:176-211 */
int ha_kuka_obs_dim(int n_arms, int n_keypoints);
/* algorithmic bytes per env-step, pre + post */
size_t ha_kuka_bytes_per_env_step(const HaKukaParams* p);
size_t ha_kuka_post_bytes_per_env_step(const HaKukaParams* p);
size_t ha_kuka_pre_bytes_per_env_step(const HaKukaParams* p);
/* fills the derived fields of HaKukaParams (dof_range_rcp) from the primary ones; call once after
 * dof_lower / dof_upper are set */
void ha_kuka_fill_derived(HaKukaParams* p);
/* initial value of stat_accum[i] (host writes these once before the first post step) */
int64_t ha_kuka_stat_accum_init(int i);

/* pre_physics_step: consume reset / reset_goal masks, reset scatter and
 * re-randomisation, action -> DOF targets, random forces.
 * actions: [n_envs, n_dofs] device, already clamped to +-clip_actions.
 * ext_rand: [n_envs, HA_KUKA_RND_STRIDE(n_dofs)] device, or NULL iff use_fast_rng. */
int ha_kuka_pre_step(const HaKukaParams* p, const HaKukaBuffers* b, const float* actions,
                     const float* ext_rand, ha_stream_t stream);

/* reset_idx outside of a step (VecTask.reset_done, vec_task.py:476-491; allegro_kuka_base.py:1246-1353):
 * resets exactly the envs whose reset_buf is non-zero; no action / force / goal-only processing. */
int ha_kuka_reset_idx(const HaKukaParams* p, const HaKukaBuffers* b, const float* ext_rand, ha_stream_t stream);

/* post_physics_step + VecTask.step tail: progress += 1, gather, keypoints,
 * observations, reward, reset masks, episode accumulators, timeout, stats. */
int ha_kuka_post_step(const HaKukaParams* p, const HaKukaBuffers* b, ha_stream_t stream);

/* compute_observations only (VecTask.reset(), vec_task.py:459-474): fills
 * obs_buf from the current state without touching reward / reset state. */
int ha_kuka_observe(const HaKukaParams* p, const HaKukaBuffers* b, ha_stream_t stream);

/* Folds the per-CTA statistic slots of the last single-arm ha_kuka_post_step into stats[] (the extras means of
 * allegro_kuka_base.py:903-910).  One small launch, idempotent; the next ha_kuka_pre_step / ha_kuka_reset_idx does the
 * same implicitly, so a step loop never needs it.  Two arms: no-op (stats[] is written by the post step itself). */
int ha_kuka_stats_finalize(const HaKukaParams* p, const HaKukaBuffers* b, ha_stream_t stream);


/* ------------------------------------------------------------------------ */
/* In-hand reorientation family: ShadowHand (shadow_hand.py) and AllegroHand   */
/* (allegro_hand.py).  ha_hand_pre_step replaces pre_physics_step              */
/* (shadow_hand.py:670-708 / allegro_hand.py:586-625) incl. reset_idx          */
/* (:604-668 / :524-584) and reset_target_pose (:586-602 / :506-522);          */
/* ha_hand_post_step replaces post_physics_step (:710-715 / :627-632) =        */
/* compute_observations (:437-584 / :406-504, all observation layouts and the  */
/* asymmetric states_buf) + compute_hand_reward (:746-800 / :663-719) and the  */
/* VecTask.step tail (vec_task.py:424-432).                                    */
/* ------------------------------------------------------------------------ */
#define HA_HAND_MAX_DOFS 24
#define HA_HAND_MAX_TIPS 5
#define HA_HAND_MAX_OBS 224

/* observationType (shadow_hand.py:111-116, allegro_hand.py:108-112) */
enum { HA_HAND_OBS_OPENAI = 0, HA_HAND_OBS_FULL_NO_VEL = 1, HA_HAND_OBS_FULL = 2, HA_HAND_OBS_FULL_STATE = 3 };

/* one env's row of externally supplied raw draws (parity mode): U(0,1) everywhere, N(0,1) in the force-normal slots */
enum {
  HA_HAND_RND_GOAL = 0,          /* 4: reset_target_pose of a goal-only reset                 */
  HA_HAND_RND_RESET_GOAL = 4,    /* 4: reset_target_pose inside reset_idx                      */
  HA_HAND_RND_FORCE_PROB = 8,    /* 1                                                          */
  HA_HAND_RND_FORCE_TEST = 9,    /* 1                                                          */
  HA_HAND_RND_FORCE_NORMAL = 10, /* 3: N(0,1)                                                  */
  HA_HAND_RND_RESET = 13         /* 2*n_dofs + 5: rand_floats of reset_idx                     */
};
#define HA_HAND_RND_STRIDE(n_dofs) (18 + 2 * (n_dofs))

typedef struct HaHandParams {
  int32_t abi_version;
  int32_t n_envs;
  int32_t is_shadow;        /* 1: ShadowHand (fingertip states + force sensors observed), 0: AllegroHand */
  int32_t n_dofs;           /* 24 / 16 */
  int32_t dofs_per_env;     /* row width of dof_state / targets (>= n_dofs) */
  int32_t n_actions;        /* 20 / 16 */
  int32_t n_bodies;
  int32_t n_tips;           /* 5 / 0 */
  int32_t obs_type;         /* HA_HAND_OBS_* */
  int32_t obs_dim;          /* see ha_hand_obs_dim() */
  int32_t states_dim;       /* 0, or the full_state width when asymmetric_observations */
  int32_t object_body;
  int32_t max_episode_length;
  int32_t max_consecutive_successes;
  int32_t use_relative_control;
  int32_t ignore_z_rot;     /* objectType == "pen" */
  int32_t use_fast_rng;
  int32_t force_enabled;
  int32_t root_regular;     /* actor rows follow env * actors_per_env + slot */
  int32_t actors_per_env, object_slot, goal_slot;
  int32_t up_axis_idx;
  int32_t actuated_dof[HA_HAND_MAX_DOFS];   /* actuated_dof_indices, n_actions entries */
  int32_t tip_body[HA_HAND_MAX_TIPS];
  float dof_lower[HA_HAND_MAX_DOFS], dof_upper[HA_HAND_MAX_DOFS], dof_default[HA_HAND_MAX_DOFS];
  float dof_range_rcp[HA_HAND_MAX_DOFS];    /* filled by ha_hand_fill_derived() */
  float clip_obs;           /* clipObservations (vec_task.py:432); <= 0: none */
  float vel_obs_scale, force_torque_obs_scale;
  /* reward (compute_hand_reward) */
  float dist_reward_scale, rot_reward_scale, rot_eps, action_penalty_scale, success_tolerance, reach_goal_bonus;
  float fall_dist, fall_penalty, av_factor;
  /* pre-physics */
  float rel_action_gain;    /* dofSpeedScale * dt */
  float act_moving_average, one_minus_act_moving_average;
  float force_decay_pow, force_gain[2], force_log_c0, force_log_c1;
  /* reset */
  float reset_position_noise, reset_dof_pos_noise, reset_dof_vel_noise;
  float goal_displacement[3];
} HaHandParams;

typedef struct HaHandDynamic {
  uint64_t rng_seed;
  uint64_t rng_step;
} HaHandDynamic;

typedef struct HaHandBuffers {
  /* simulator-owned (Isaac Gym layout) */
  float* root_state;              /* [n_envs*actors_per_env, 13] */
  const float* rigid_body_state;  /* [n_envs, n_bodies, 13] */
  float* dof_state;               /* [n_envs, dofs_per_env, 2] */
  const float* dof_force;         /* [n_envs, n_dofs] (full_state / asymmetric; nullable otherwise) */
  const float* force_sensors;     /* [n_envs, 6*n_tips] (ShadowHand full_state / asymmetric; nullable otherwise) */
  const int32_t* object_actor;    /* [n_envs] */
  const int32_t* goal_actor;      /* [n_envs] */
  const int32_t* hand_actor;      /* [n_envs] */
  /* task constants */
  const float* object_init_state; /* [n_envs, 13] */
  const float* goal_init_state;   /* [n_envs, 13] */
  /* task state */
  float* goal_states;             /* [n_envs, 13] */
  float* obs_buf;                 /* [n_envs, obs_dim], clipped to +-clip_obs */
  float* states_buf;              /* [n_envs, states_dim] (nullable iff states_dim == 0) */
  float* rew_buf;
  int64_t* reset_buf;
  int64_t* reset_goal_buf;
  int64_t* progress_buf;
  int64_t* randomize_buf;         /* nullable */
  uint8_t* timeout_buf;
  float* successes;
  float* consecutive_successes;   /* [1] */
  float* prev_targets;            /* [n_envs, dofs_per_env] */
  float* cur_targets;
  float* actions;                 /* [n_envs, n_actions] task-owned copy */
  float* rb_forces;               /* [n_envs, n_bodies, 3] */
  float* random_force_prob;       /* [n_envs] */
  /* outputs for the simulator's *_indexed setters */
  int32_t* root_reset_indices;    /* [2*n_envs] */
  int32_t* hand_reset_indices;    /* [n_envs] */
  int32_t* reset_counts;          /* [4]: n_root, n_hand, n_reset_envs, n_goal_resets */
  /* reductions */
  float* stats;                   /* [HA_STATS_LEN], HA_HAND_STAT_* */
  int64_t* stat_accum;            /* [HA_HAND_N_PARTIAL_STATS][HA_HAND_STAT_SLOTS], zero-initialised once.  n_envs <=
                                     32 * HA_HAND_STAT_SLOTS: one slot of partials per CTA of the last ha_hand_post_step,
                                     folded into stats / consecutive_successes by the next ha_hand_pre_step /
                                     ha_hand_reset_idx or by ha_hand_stats_finalize (whichever comes first; the running
                                     average is applied once per post step).  Larger n_envs: the first HA_STATS_LEN
                                     entries are atomic accumulators and the post launch finishes the statistics itself */
  uint32_t* tickets;              /* [8], zero-initialised once */
  HaHandDynamic* dynamic;
} HaHandBuffers;

#define HA_HAND_STAT_SLOTS 768      /* statistic slots per row of HaHandBuffers.stat_accum (one per CTA of the post kernel) */
enum {
  HA_HAND_STAT_N = 0,
  HA_HAND_STAT_SUM_REWARD,
  HA_HAND_STAT_SUM_RESETS,               /* num_resets of compute_hand_reward */
  HA_HAND_STAT_SUM_FINISHED_SUCCESSES,   /* sum(successes * resets) */
  HA_HAND_STAT_SUM_GOAL_RESETS,
  HA_HAND_STAT_SUM_SUCCESSES,
  HA_HAND_STAT_SUM_TIMEOUTS,
  HA_HAND_N_PARTIAL_STATS,
  HA_HAND_STAT_MEAN_REWARD = HA_HAND_N_PARTIAL_STATS,
  HA_HAND_STAT_CONSECUTIVE_SUCCESSES,
  HA_HAND_N_STATS
};

size_t ha_sizeof_hand_params(void);
size_t ha_sizeof_hand_buffers(void);
int ha_hand_obs_dim(int is_shadow, int obs_type);
void ha_hand_fill_derived(HaHandParams* p);
size_t ha_hand_post_bytes_per_env_step(const HaHandParams* p);
size_t ha_hand_pre_bytes_per_env_step(const HaHandParams* p);
size_t ha_hand_bytes_per_env_step(const HaHandParams* p);
int ha_hand_pre_step(const HaHandParams* p, const HaHandBuffers* b, const float* actions, const float* ext_rand, ha_stream_t stream);
int ha_hand_reset_idx(const HaHandParams* p, const HaHandBuffers* b, const float* ext_rand, ha_stream_t stream);
int ha_hand_post_step(const HaHandParams* p, const HaHandBuffers* b, ha_stream_t stream);
int ha_hand_observe(const HaHandParams* p, const HaHandBuffers* b, ha_stream_t stream);
/* Fold the statistic slots of the last ha_hand_post_step now (stats, consecutive_successes) instead of at the head of the next
 * pre step: what the reference has after compute_hand_reward returns (shadow_hand.py:415-435).  No-op work when nothing is
 * pending or n_envs > 32 * HA_HAND_STAT_SLOTS. */
int ha_hand_stats_finalize(const HaHandParams* p, const HaHandBuffers* b, ha_stream_t stream);

/* =====================================================================================================
 * DeXtreme: AllegroHandDextremeADR (config 5) -- per-step tensor path of
 * isaacgymenvs/tasks/dextreme/allegro_hand_dextreme.py (pre_physics_step :912-949, apply_actions :955-1015 with the ADR
 * apply_action_noise_latency :1430-1441, reset_idx :775-858, reset_target_pose :709-733, apply_random_forces :1017-1030,
 * post_physics_step :1033-1048, compute_observations :606-692 + :1443-1489, compute_poses_wrt_wrist :533-588,
 * get_random_cube_observation :739-771, compute_hand_reward :1598-1663) and the dict-observation step tail of
 * adr_vec_task.py:289-325.  The host side of ADR (apply_randomizations: Isaac Gym property DR, range update, per-env
 * resampling) stays in Python; its per-env results are the adr_* / affine_* / action_latency / cube_pose_refresh_* inputs.
 * ===================================================================================================== */
#define HA_DEX_DOFS 16
#define HA_DEX_TIPS 4
#define HA_DEX_OBS_DIM 285          /* the 20 dict entries of get_num_obs_dict (:168-199), concatenated in declaration order */
#define HA_DEX_MAX_QUEUE 64
#define HA_DEX_N_STEP_TERMS 8       /* dist, rot, action, action_delta, velocity, reach_goal, fall, timeout (:455-464) */

enum {   /* column offsets of the dict entries inside an observation row */
  HA_DEX_OBS_DOF_POS = 0, HA_DEX_OBS_DOF_POS_RANDOMIZED = 16, HA_DEX_OBS_DOF_VEL = 32, HA_DEX_OBS_DOF_FORCE = 48,
  HA_DEX_OBS_OBJECT_VELS = 64, HA_DEX_OBS_LAST_ACTIONS = 70, HA_DEX_OBS_CUBE_RANDOM_PARAMS = 86, HA_DEX_OBS_HAND_RANDOM_PARAMS = 89,
  HA_DEX_OBS_GRAVITY_VEC = 90, HA_DEX_OBS_FT_STATES = 93, HA_DEX_OBS_FT_FORCE_TORQUES = 145, HA_DEX_OBS_RB_FORCES = 169,
  HA_DEX_OBS_ROT_DIST = 172, HA_DEX_OBS_STOCHASTIC_DELAY_PARAMS = 174, HA_DEX_OBS_AFFINE_PARAMS = 178, HA_DEX_OBS_OBJECT_POSE = 256,
  HA_DEX_OBS_GOAL_POSE = 263, HA_DEX_OBS_GOAL_RELATIVE_ROT = 270, HA_DEX_OBS_OBJECT_POSE_CAM_RANDOMIZED = 274,
  HA_DEX_OBS_GOAL_RELATIVE_ROT_CAM_RANDOMIZED = 281
};
enum {   /* pre-step random table [n_envs, HA_DEX_RND_STRIDE]: U(0,1), N(0,1) in the *_WHITE / *_NORMAL slots */
  HA_DEX_RND_GOAL = 0,          /* 4 rand_floats + 3 quaternion uniforms (goal-only reset_target_pose) */
  HA_DEX_RND_RESET_GOAL = 7,    /* the same 7 for reset_target_pose inside reset_idx */
  HA_DEX_RND_RESET = 14,        /* 2*16+5 rand_floats + 3 quaternion uniforms */
  HA_DEX_RND_FORCE_PROB = 54,
  HA_DEX_RND_PROGRESS = 55,     /* progress_buf of an all-env reset: floor(u * max_episode_length) (:848-849) */
  HA_DEX_RND_ACTION_DELAY = 56,
  HA_DEX_RND_ACTION_WHITE = 57, /* 16 normals */
  HA_DEX_RND_FORCE_TEST = 73,
  HA_DEX_RND_FORCE_NORMAL = 74, /* 3 normals */
  HA_DEX_RND_STRIDE = 77
};
enum {   /* post-step random table [n_envs, HA_DEX_POST_RND_STRIDE] */
  HA_DEX_POST_RND_POSE_WHITE = 0,   /* 7 normals */
  HA_DEX_POST_RND_CUBE = 7,         /* 5 rand_floats */
  HA_DEX_POST_RND_CUBE_QUAT = 12,   /* 3 uniforms */
  HA_DEX_POST_RND_CUBE_MASK = 15,
  HA_DEX_POST_RND_DELAY = 16,
  HA_DEX_POST_RND_DOF_WHITE = 17,   /* 16 normals */
  HA_DEX_POST_RND_STRIDE = 33
};
enum {   /* HaDexBuffers.dynamic: int64 [HA_DEX_DYN_LEN], device resident, advanced by the kernels */
  HA_DEX_DYN_RNG_SEED = 0, HA_DEX_DYN_RNG_STEP, HA_DEX_DYN_FRAME /* self.frame */, HA_DEX_DYN_QUEUE_HEAD /* ring slot of the newest action */,
  HA_DEX_DYN_N_APPLY /* envs with apply_reset_buf set when the last pre step ended */, HA_DEX_DYN_APPLY_ACC /* scratch of that count */,
  HA_DEX_DYN_LEN = 8
};

typedef struct HaDexParams {
  int32_t abi_version;
  int32_t n_envs;
  int32_t dofs_per_env;     /* row width of dof_state (>= 16) */
  int32_t n_bodies;
  int32_t palm_body;        /* palm_link (:390) */
  int32_t tip_body[HA_DEX_TIPS];
  int32_t object_body;
  int32_t max_episode_length;
  int32_t max_consecutive_successes;
  int32_t num_success_hold_steps;
  int32_t queue_len;        /* actionLatencyMax + 1 */
  int32_t use_capped_dof_control;
  int32_t apply_random_quat;
  int32_t use_fast_rng;
  int32_t force_enabled;
  int32_t root_regular, actors_per_env, object_slot, goal_slot;
  float dof_lower[HA_DEX_DOFS], dof_upper[HA_DEX_DOFS];
  float dof_range_rcp[HA_DEX_DOFS];   /* filled by ha_dex_fill_derived() */
  float clip_obs;           /* clipObservations (adr_vec_task.py:319); <= 0: none */
  float vel_obs_scale, force_torque_obs_scale;
  float dist_reward_scale, rot_reward_scale, rot_eps, action_penalty_scale, action_delta_penalty_scale, success_tolerance;
  float reach_goal_bonus, fall_dist, fall_penalty, av_factor;
  float act_moving_average, one_minus_act_moving_average;   /* annealed on the host (:896-909) */
  float max_dof_delta;      /* max_dof_radians_per_second * dt * control_freq_inv */
  float force_decay_pow, force_gain[2], force_log_c0, force_log_c1;
  float reset_position_noise, reset_position_noise_z, reset_dof_pos_noise, reset_dof_vel_noise;
  float goal_displacement[3];
  float random_cube_pose_prob;
  int32_t n_view_cols[2];   /* widths of the two concatenated views K_post also emits (0: off), see HaDexBuffers.view_* */
} HaDexParams;

typedef struct HaDexBuffers {
  /* simulator-owned (Isaac Gym layout) */
  float* root_state;                /* [n_envs*actors_per_env, 13] */
  const float* rigid_body_state;    /* [n_envs, n_bodies, 13] */
  float* dof_state;                 /* [n_envs, dofs_per_env, 2] */
  const float* dof_force;           /* [n_envs, 16] */
  const float* force_sensors;       /* [n_envs, 24] */
  const int32_t* object_actor;      /* [n_envs] */
  const int32_t* goal_actor;
  const int32_t* hand_actor;
  /* task constants */
  const float* object_init_state;   /* [n_envs, 13] */
  const float* goal_init_state;     /* [n_envs, 13] */
  /* per-env ADR state sampled by apply_randomizations (:1392-1416): read by the step kernels, written by ha_dex_adr_resample */
  int64_t* action_latency;            /* [n_envs], < queue_len */
  int64_t* cube_pose_refresh_rate;    /* [n_envs], >= 1 */
  int64_t* cube_pose_refresh_offset;  /* [n_envs] */
  const float* adr_action_delay_prob;       /* [n_envs] each: get_adr_tensor(...) */
  const float* adr_affine_action_white;
  const float* adr_affine_cube_pose_white;
  const float* adr_cube_obs_delay_prob;
  const float* adr_affine_dof_pos_white;
  const float* adr_rna_alpha;               /* nullable unless rna_perturbation is passed */
  float* affine_actions_scaling;      /* [n_envs, 16] */
  float* affine_actions_additive;
  float* affine_cube_pose_scaling;    /* [n_envs, 7] */
  float* affine_cube_pose_additive;
  float* affine_dof_pos_scaling;      /* [n_envs, 16] */
  float* affine_dof_pos_additive;
  const float* cube_random_params;          /* [n_envs, 3] */
  const float* hand_random_params;          /* [n_envs, 1] */
  const float* gravity_vec;                 /* [n_envs, 3] */
  /* task state */
  float* goal_states;               /* [n_envs, 13] */
  float* obs_rows;                  /* [n_envs, HA_DEX_OBS_DIM]: every dict entry, clipped to +-clip_obs */
  float* rew_buf;
  int64_t* reset_buf;
  int64_t* reset_goal_buf;
  int64_t* progress_buf;
  int64_t* hold_count_buf;
  int64_t* apply_reset_buf;
  uint8_t* timeout_buf;
  float* successes;
  float* consecutive_successes;     /* [1] */
  float* best_rotation_dist;        /* [n_envs] */
  float* prev_targets;              /* [n_envs, 16] */
  float* cur_targets;
  float* prev_prev_targets;
  float* actions;                   /* [n_envs, 16] task-owned copy */
  float* prev_actions;              /* [n_envs, 16] */
  float* action_queue;              /* [queue_len, n_envs, 16] ring buffer; slot (head - k) mod queue_len = the action of k steps ago */
  float* dof_delta;                 /* [n_envs, 16] */
  float* rb_forces;                 /* [n_envs, n_bodies, 3] */
  float* random_force_prob;         /* [n_envs] */
  float* object_pose;               /* [n_envs, 7]: the object pose gathered by the last compute_observations (:613) */
  float* obs_object_pose;           /* [n_envs, 7] */
  float* obs_object_pose_freq;      /* [n_envs, 7] */
  float* episode_cumulative;        /* [HA_DEX_N_STEP_TERMS, n_envs] */
  /* outputs for the simulator's *_indexed setters */
  int32_t* root_reset_indices;      /* [2*n_envs] */
  int32_t* hand_reset_indices;      /* [n_envs] */
  int32_t* reset_counts;            /* [4]: n_root, n_hand, n_reset_envs, n_randomized */
  /* reductions (HA_HAND_STAT_* layout) */
  float* stats;
  int64_t* stat_accum;
  uint32_t* tickets;                /* [8], zero-initialised once */
  int64_t* dynamic;                 /* [HA_DEX_DYN_LEN] */
  /* optional concatenated views for the rl_games adapter (ComplexObsRLGPUEnv._generate_obs, rlgames_utils.py:341-399: torch.cat of the
     policy inputs -> "obs", of the critic inputs -> "states"): view v is [n_envs, n_view_cols[v]] with column j = obs_rows column
     view_cols[v][j]; written by K_post from the same staged tile, so the two cats per step disappear.  nullable when the width is 0 */
  const int32_t* view_cols[2];
  float* view_rows[2];
} HaDexBuffers;

/* ---- per-env ADR resampling on reset: recycle_envs (adr_vec_task.py:740-758), sample_adr_tensor (:666-730) for every tensorised
 * parameter, and the Dextreme-specific draws of allegro_hand_dextreme.py:1392-1416, for the envs in reset_buf & ~apply_reset_buf (the
 * mask pre_physics_step hands to apply_randomizations, :930-937).  Launch it BEFORE ha_dex_pre_step of the same step.  The ADR range
 * update itself (adr_update: host deques, thresholds) stays on the host and only refreshes this table. ---- */
#define HA_DEX_ADR_MAX_PARAMS 16
enum {
  HA_DEX_ADR_RND_WORKER = 0,
  HA_DEX_ADR_RND_MODE = 1,       /* adr_modes: floor(u * 2 * n_params_total) */
  HA_DEX_ADR_RND_PARAM = 2,      /* one uniform per tensorised parameter */
  HA_DEX_ADR_RND_LATENCY = 18,
  HA_DEX_ADR_RND_RATE = 19,
  HA_DEX_ADR_RND_OFFSET = 20,
  HA_DEX_ADR_RND_AFFINE = 21,    /* 16 + 16 + 7 + 7 + 16 + 16 normals */
  HA_DEX_ADR_RND_STRIDE = 99
};
typedef struct HaDexAdrTable {
  int32_t n_params_total;                       /* num_adr_params: adr_modes ranges over 2 * n_params_total */
  int32_t n_tensor;                             /* tensorised parameters (<= HA_DEX_ADR_MAX_PARAMS), in adr_tensor_values order */
  int32_t param_index[HA_DEX_ADR_MAX_PARAMS];   /* index of each among all ADR parameters */
  float lo[HA_DEX_ADR_MAX_PARAMS];
  float span[HA_DEX_ADR_MAX_PARAMS];            /* float(hi - lo), the difference taken in double like the reference's Python floats */
  float low_value[HA_DEX_ADR_MAX_PARAMS];       /* what a boundary worker evaluating the lower end gets: range[0], or next_limits[0] */
  float high_value[HA_DEX_ADR_MAX_PARAMS];
  float boundary_fraction;                      /* worker_adr_boundary_fraction */
  int32_t idx_action_latency;                   /* positions (0 .. n_tensor-1) of the parameters the Dextreme draws read */
  int32_t idx_cube_pose_refresh_rate;
  int32_t idx_affine[6];                        /* action scaling / additive, cube pose scaling / additive, dof pos scaling / additive */
  int32_t all_envs;                             /* first_randomization: every env is resampled */
} HaDexAdrTable;
typedef struct HaDexAdrBuffers {
  float* values[HA_DEX_ADR_MAX_PARAMS];         /* adr_tensor_values[k], [n_envs] each */
  int64_t* worker_types;                        /* [n_envs] RolloutWorkerModes */
  int64_t* adr_modes;                           /* [n_envs] */
} HaDexAdrBuffers;
size_t ha_sizeof_dex_adr_table(void);
size_t ha_sizeof_dex_adr_buffers(void);
int ha_dex_adr_resample(const HaDexParams* p, const HaDexBuffers* b, const HaDexAdrTable* t, const HaDexAdrBuffers* ab, const float* ext_rand,
                        ha_stream_t stream);

size_t ha_sizeof_dex_params(void);
size_t ha_sizeof_dex_buffers(void);
void ha_dex_fill_derived(HaDexParams* p);
size_t ha_dex_post_bytes_per_env_step(const HaDexParams* p);
size_t ha_dex_pre_bytes_per_env_step(const HaDexParams* p);
size_t ha_dex_bytes_per_env_step(const HaDexParams* p);
/* rna_perturbation: nullable [n_envs, 16], the RNA network's DOF-space perturbation (:864-879) computed by the caller */
int ha_dex_pre_step(const HaDexParams* p, const HaDexBuffers* b, const float* actions, const float* rna_perturbation,
                    const float* ext_rand, ha_stream_t stream);
int ha_dex_post_step(const HaDexParams* p, const HaDexBuffers* b, const float* ext_rand_post, ha_stream_t stream);
int ha_dex_observe(const HaDexParams* p, const HaDexBuffers* b, const float* ext_rand_post, ha_stream_t stream);

/* =====================================================================================================
 * Observation / action noise of domain randomisation: the closures VecTask.apply_randomizations builds (vec_task.py:684-754) and
 * VecTask.step applies to the actions before pre_physics_step (:398-400) and to obs_buf after post_physics_step (:427-429):
 *   gaussian: noise = (corr * var_corr + mu_corr) + N(0,1) * var + mu        uniform: noise = (corr * (hi_c - lo_c) + lo_c) + U(0,1) * (hi - lo) + lo
 *   out = x + noise (additive) | x * noise (scaling);   corr = the persistent N(0,1) tensor drawn on the closure's first use.
 * One streaming pass: x, corr read once, out written once (12 B / element), white noise from in-kernel Philox (or a table in parity mode).
 * ===================================================================================================== */
enum { HA_NOISE_GAUSSIAN = 0, HA_NOISE_UNIFORM = 1 };
enum { HA_NOISE_ADDITIVE = 0, HA_NOISE_SCALING = 1 };
typedef struct HaNoiseParams {
  int32_t abi_version;
  int32_t distribution;   /* HA_NOISE_GAUSSIAN / HA_NOISE_UNIFORM */
  int32_t operation;      /* HA_NOISE_ADDITIVE / HA_NOISE_SCALING */
  int32_t use_fast_rng;
  float corr_scale, corr_shift;     /* var_corr, mu_corr | hi_corr - lo_corr, lo_corr */
  float white_scale, white_shift;   /* var, mu | hi - lo, lo */
  float clip;                       /* > 0: clamp the result to +-clip (vec_task.py:432 follows the observation noise) */
  uint64_t rng_seed, rng_step;
} HaNoiseParams;
size_t ha_sizeof_noise_params(void);
/* x, corr, out: [count] floats (out may alias x); ext_white: nullable [count] raw draws (parity mode) */
int ha_dr_noise_apply(const HaNoiseParams* p, const float* x, const float* corr, float* out, size_t count, const float* ext_white,
                      ha_stream_t stream);


/* =====================================================================================================
 * rl_games observer, episode_cumulative half: RLGPUAlgoObserver.process_infos
 * (isaacgymenvs/utils/rlgames_utils.py:149-172) adds every step's per-env reward terms to running sums and, for each
 * finished env in ascending env order, appends the sums to one deque(maxlen = games_to_track) per term (a Python loop with
 * one .item() sync per finished env and term) and zeroes them; after_print_stats (:192-205) logs mean / min / max of each
 * deque.  ha_observer_step does the per-step half in one launch from the done MASK (no nonzero(), no host sync): running sums
 * += terms; finished envs get their deque position from an order-preserving scan over the grid and write ring slot
 * (appended + rank) mod capacity when the deque would still hold them; ha_observer_summary reduces the ring to
 * [mean, min, max][n_terms] on the device (one small copy to the host at logging time).
 * ===================================================================================================== */
#define HA_OBSERVER_MAX_TERMS 16
#define HA_OBSERVER_MAX_CTAS 592
typedef struct HaObserverBuffers {
  const float* terms;      /* [n_terms, n_envs] this step's terms: the rows of extras["episode_cumulative"] */
  const int64_t* dones;    /* [n_envs] non-zero = episode finished this step (the dones rl_games derives done_indices from) */
  float* cumulative;       /* [n_terms, n_envs] running sums since each env's last finished episode; zero-initialised */
  float* ring;             /* [n_terms, capacity] finished-episode sums, ring buffer == deque(maxlen=capacity) */
  int64_t* state;          /* [4]: episodes appended so far, launch epoch, ticket, unused; zero-initialised */
  int64_t* block_counts;   /* [HA_OBSERVER_MAX_CTAS] per-CTA finished-env counts tagged with the epoch; zero-initialised */
  float* summary;          /* [3, n_terms] mean, min, max over the ring's valid entries (NaN while it is empty) */
} HaObserverBuffers;
size_t ha_sizeof_observer_buffers(void);
int ha_observer_step(int32_t n_envs, int32_t n_terms, int32_t capacity, const HaObserverBuffers* b, ha_stream_t stream);
int ha_observer_summary(int32_t n_terms, int32_t capacity, const HaObserverBuffers* b, ha_stream_t stream);


/* =====================================================================================================
 * Random network adversary of the DeXtreme tasks: RandomNetworkAdversary.forward
 * (isaacgymenvs/utils/rna_util.py:124-150) and its call site get_random_network_adversary_action
 * (allegro_hand_dextreme.py:864-879).  Chain per step: cat(dof_pos, object_pose_wrt_wrist) [N,23] -> fc1 + ReLU [N,512] ->
 * fc1_1, * dropout_masks1 -> fc2 + ReLU [N,1024] -> fc2_1, * dropout_masks2 -> fc3 [N,16*32] -> softmax -> argmax over the 32
 * bins -> gather(discretised_dofs).  The thin first layer (K = 23, no concatenation), the bias / ReLU / mask epilogues and the
 * bias + argmax + gather tail are the kernels below; the four wide layers are fp32 GEMMs the host side issues between them
 * (handarm_b200/rna.py).  The result is the DOF-space perturbation ha_dex_pre_step takes as rna_perturbation (it unscales and
 * blends with rna_alpha, :879-889).
 * ===================================================================================================== */
#define HA_RNA_DOFS 16
#define HA_RNA_BINS 32
#define HA_RNA_IN_DIMS 23
#define HA_RNA_IN_PAD 64          /* fc1's K padded with zeros to one tensor-core k-block */
#define HA_RNA_FEATS1 512
#define HA_RNA_FEATS2 1024
/* out[N,512] = relu(cat(dof_state[:, :, 0], object_pose_wrt_wrist) @ w1^T + b1); dof_state [N,16,2], pose rows of 7 floats at
 * pose_stride floats (7 if packed; 285 for the object_pose band of the DeXtreme observation rows), w1 [512,23] */
int ha_rna_input_layer(int32_t n_envs, const float* dof_state, const float* object_pose_wrt_wrist, int32_t pose_stride, const float* w1, const float* b1,
                       float* out, void* out_hi, void* out_lo, ha_stream_t stream);   /* out fp32, or (out NULL) bf16 hi / lo planes */
/* x[rows, cols] = (x + bias[cols]), then ReLU if relu != 0, then * mask[rows, cols] if mask != NULL; in place */
int ha_rna_bias_act(float* x, const float* bias, const float* mask, int32_t relu, size_t rows, int32_t cols, ha_stream_t stream);
/* out[N,16] = discretised_dofs[d, argmax_bin(logits[n, d, :] + bias3[d, :])]; idx_out: nullable [N,16] chosen bins */
int ha_rna_select(int32_t n_envs, const float* logits, const float* bias3, const float* discretised_dofs, float* out, int32_t* idx_out,
                  ha_stream_t stream);

/* Tensor-core path of the wide layers (csrc/ha_rna_tc.cu): tcgen05 / TMEM / TMA GEMMs on split-bf16 operands.  An fp32 tensor x
 * travels as two bf16 planes hi = bf16(x), lo = bf16(x - hi); a layer accumulates A_hi.W_hi + A_hi.W_lo + A_lo.W_hi in fp32
 * (~1e-5 of an fp32 GEMM).  Planes are [rows, cols] row-major bf16, 16-byte aligned; n % 128 == 0, k % 32 == 0. */
int ha_rna_split_bf16(const float* x, void* hi, void* lo, size_t count, ha_stream_t stream);
/* fc1's A operand without torch.cat: planes [n_envs, HA_RNA_IN_PAD] = split([dof_state[:, :, 0] | pose | 0 ...]); fc1 then runs as
 * ha_rna_layer_tc(n_envs, 512, HA_RNA_IN_PAD, ...) on W1 padded the same way */
int ha_rna_input_planes(int32_t n_envs, const float* dof_state, const float* object_pose_wrt_wrist, int32_t pose_stride, void* out_hi, void* out_lo,
                        ha_stream_t stream);
/* out planes [m, n] = split((A . W^T + bias) [relu] [* mask]); A planes [m, k], W planes [n, k];
 * mask_bits: nullable [m, n / 32] words, bit j of word (row, c) = dropout mask {0, 1} of column 32 c + j */
int ha_rna_layer_tc(int32_t m, int32_t n, int32_t k, const void* a_hi, const void* a_lo, const void* w_hi, const void* w_lo, const float* bias,
                    const uint32_t* mask_bits, int32_t relu, void* out_hi, void* out_lo, ha_stream_t stream);
/* last layer fused with the bin selection: out[m,16] = discretised_dofs[d, argmax_bin(A . W3^T + bias3)], W3 planes [512, k] */
int ha_rna_select_tc(int32_t m, int32_t k, const void* a_hi, const void* a_lo, const void* w_hi, const void* w_lo, const float* bias3,
                     const float* discretised_dofs, float* out, int32_t* idx_out, ha_stream_t stream);


/* =====================================================================================================
 * ADR range update, data half: adr_update (adr_vec_task.py:760-812) extends, per ADR parameter and bound, a
 * deque(maxlen = adr_queue_threshold_length) with the objectives of the boundary workers that finished an episode, each via boolean
 * masks, a gather and a .cpu().numpy().tolist() sync.  ha_adr_queue_step keeps all 2 * n_params queues as device rings: one
 * order-preserving compaction of the finished boundary workers (ascending env index, as the boolean-mask gather lists them) and one
 * launch with a CTA per queue that extends its ring and leaves [mean, length] in summary.  The host reads summary once per step and
 * runs the scalar range logic (:814-870, handarm_b200/tasks/dextreme.py adr_update); clearing a queue = appended[q] = 0.
 * ===================================================================================================== */
#define HA_ADR_WORKER_ROLLOUT 0
#define HA_ADR_WORKER_BOUNDARY 1
typedef struct HaAdrQueueBuffers {
  const int64_t* reset_buf;        /* [n_envs]  rand_env_mask = all_envs | (reset_buf & ~apply_reset_buf) */
  const int64_t* apply_reset_buf;  /* [n_envs], nullable */
  const int64_t* worker_types;     /* [n_envs] RolloutWorkerModes (before this step's recycle_envs) */
  const int64_t* adr_modes;        /* [n_envs] queue index 2n (lower bound of parameter n) / 2n + 1 (upper) */
  const float* objective;          /* [n_envs] adr_objective (successes) */
  float* queues;                   /* [n_queues, capacity] rings */
  int64_t* appended;               /* [n_queues] entries appended since the queue was last cleared; zero-initialised */
  int32_t* list;                   /* [n_envs] scratch: finished boundary workers in env order */
  int64_t* state;                  /* [4] scan state; zero-initialised */
  int64_t* block_counts;           /* [HA_OBSERVER_MAX_CTAS]; zero-initialised */
  double* summary;                 /* [n_queues, 2]: float64 mean (0 if empty), length */
} HaAdrQueueBuffers;
size_t ha_sizeof_adr_queue_buffers(void);
int ha_adr_queue_step(int32_t n_envs, int32_t n_queues, int32_t capacity, int32_t all_envs, const HaAdrQueueBuffers* b, ha_stream_t stream);

/* Allocation-free variant of ha_adr_queue_step.  It has NO counterpart in the reference: it was added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same inputs, outputs and results as ha_adr_queue_step, but the three scratch fields of HaAdrQueueBuffers (list,
 * state, block_counts) are IGNORED (may be NULL): the scratch is a file-scope static __device__ array inside the library
 * (sizeof = 32 + 8 * HA_OBSERVER_MAX_CTAS + 4 * HA_ADR_STATIC_MAX_ENVS bytes, about 261 KB; every GPU that loads the library
 * holds it until the process exits).  The call makes no cudaMalloc / cudaFree / allocating library call: one cudaMemsetAsync
 * of the scratch's scan state (which must start at zero; the array keeps the previous call's contents) and the same two launches,
 * all stream-ordered.  n_envs > HA_ADR_STATIC_MAX_ENVS returns HA_ERR_INVALID with a message naming the limit.
 * All calls on one device share the one array: calls that overlap on the GPU (different streams, or graph branches) race --
 * serialise them, or use ha_adr_queue_step with separate caller-provided scratch. */
#define HA_ADR_STATIC_MAX_ENVS 65536
int ha_adr_queue_step_static(int32_t n_envs, int32_t n_queues, int32_t capacity, int32_t all_envs, const HaAdrQueueBuffers* b, ha_stream_t stream);


/* =====================================================================================================
 * Cross-rank episode statistics over NVLink peer memory.  The env shards of a box never exchange data; the only
 * cross-rank quantities are the <= 16 partial statistics of a step (sums, counts, [max, -min]).  A library all-reduce for 64
 * bytes puts a collective kernel next to the persistent step kernels every few steps (measured: 8 GPUs at 92 % of linear with it,
 * 99.6 % without).  The exchange here is wait-free on both sides and needs no fence: every rank stores its vector into slot
 * (exchange number % HA_XCHG_RING) of every peer's exchange buffer (peer pointers from CUDA IPC / torch symmetric memory) as 16
 * aligned 8-byte words (exchange number << 32 | float bits) -- one store is atomic, so a value never travels without its tag --
 * then head[rank] = number; it reduces the newest exchange every rank claims to have published (min over head[]) if all of its
 * W x 16 words carry that number, otherwise it leaves `reduced` alone and looks again at the next call.  SUM in rank order (the
 * same bits on every rank), MAX for the slots in max_mask.  A late rank delays the exchange number the others reduce, never the
 * others; while a rank is more than HA_XCHG_RING - 2 exchanges behind the others skip (status |= 2); a rank that finds itself
 * lapped rejoins at the fastest rank's exchange number (status |= 1).  One warp runs an exchange: as a launch of its own
 * (ha_stats_exchange) or inside the single-arm post-physics kernel (HaKukaBuffers.xchg, no launch).
 * ===================================================================================================== */
#define HA_XCHG_RING 16
#define HA_XCHG_MAX_RANKS 16
typedef struct HaStatsExchange {
  float** peer_bufs;         /* device array [world]: exchange buffer of every rank (own one included), each
                                ha_stats_exchange_bytes(world) bytes, 8-byte aligned, zero-initialised before the first exchange
                                of any rank */
  int32_t rank, world;
  uint32_t max_mask;         /* bit i: statistic i is MAX-reduced, otherwise SUM */
  int32_t interval;          /* in-kernel path (HaKukaBuffers.xchg) only: exchange on every interval-th post step (<= 1: every one);
                                the same value on every rank */
  int64_t* seq;              /* [4]: exchanges done by this rank, number of the last exchange reduced, post steps seen by the
                                in-kernel path, unused; zero-initialised */
  int32_t* status;           /* [1]: bit 1 (2) = a peer fell a whole ring behind at some point, bit 0 (1) = this rank was lapped
                                and rejoined; zero-initialised */
} HaStatsExchange;
size_t ha_sizeof_stats_exchange(void);
size_t ha_stats_exchange_bytes(int32_t world);
/* stats: [HA_STATS_LEN] of this rank; reduced: [HA_STATS_LEN] whole-box values of the newest complete exchange (left as it is
 * while there is none) */
int ha_stats_exchange(const HaStatsExchange* x, const float* stats, float* reduced, ha_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* HANDARM_B200_H */
