/* piano_step.h - C-ABI of the B200-native batched RoboPianist environment step.
 *
 * The reference has no FFI for this path: it calls MuJoCo through dm_control from Python.
 * Each entry point below names the reference interface it replaces.
 *
 *   pp_create   <- building N envs: VectorizedPianoEnv.__init__
 *                  (reference parallelized_base_v2.py:21-46) / suite.load
 *                  (robopianist/suite/__init__.py:50-93): compiles the model, uploads tables.
 *   pp_reset    <- VectorizedPianoEnv.reset (parallelized_base_v2.py:48-51) ->
 *                  composer Environment.reset -> task.initialize_episode
 *                  (piano_with_shadow_hands.py:169-174, piano.py:145-152).
 *   pp_step     <- VectorizedPianoEnv.step (parallelized_base_v2.py:53-60): for every env,
 *                  CanonicalSpecWrapper action scaling, task.before_step (:176-186),
 *                  n_substeps x [mj_step + Piano.after_substep (piano.py:154-192)],
 *                  task.after_step (:188-204), observation update (:371-412), reward
 *                  (composite_reward.py:46-56), termination (:209-220).
 *   pp_step_host<- the same call with HOST (numpy) buffers, as the reference's callers pass.
 *   pp_get_state / pp_set_state <- physics.bind(...).qpos/qvel / physics.set_state, used by
 *                  parity tests ("identical qpos/qvel/ctrl").
 *   pp_get_diag <- mjData fields (qacc, actuator_force, sensordata, site_xpos, ncon, ...).
 *
 * Conventions: every *_dev pointer is a CUDA device pointer (e.g. torch tensor.data_ptr()),
 * caller-owned, contiguous, 16-byte aligned; `stream` is a cudaStream_t (0 = legacy default).
 * Nothing here synchronises the host except pp_create/pp_destroy/pp_step_host/pp_get_*_host.
 * Return value 0 = ok, negative = error (message via pp_last_error()). One handle per GPU;
 * a handle is not thread-safe. No CPU fallback exists: without a CUDA device pp_create fails.
 */
#ifndef PIANO_STEP_H_
#define PIANO_STEP_H_

#include <stdint.h>

#include "pp_model.h"

#ifdef __cplusplus
extern "C" {
#endif

enum { PP_TASK_SELF_ACTUATED_PIANO = 0, PP_TASK_PIANO_WITH_SHADOW_HANDS = 1 };
enum { PP_REWARD_NEGATIVE_L2 = 0, PP_REWARD_NEGATIVE_XENT = 1 };
enum { PP_OK = 0, PP_ERR_INVALID = -1, PP_ERR_CUDA = -2, PP_ERR_UNSUPPORTED = -3 };

/* reward_terms columns of pp_step (PianoWithShadowHands; insertion order of
 * piano_with_shadow_hands.py:130-144): */
enum { PP_TERM_KEY_PRESS = 0, PP_TERM_SUSTAIN = 1, PP_TERM_ENERGY = 2, PP_TERM_FINGERING = 3,
       PP_TERM_FOREARM = 4, PP_NUM_TERMS = 5 };

typedef struct pp_config {
  int32_t num_envs;
  int32_t device;            /* CUDA ordinal */
  int32_t task;              /* PP_TASK_* */
  int32_t n_substeps;        /* control_timestep / physics_timestep (10; 2 in config 1) */
  int32_t n_steps_lookahead; /* L: goal observation has (L+1)*89 entries */
  int32_t n_frames;          /* T = len(NoteTrajectory.notes) */
  int32_t pmax;              /* max notes per frame in `pairs` */
  int32_t canonical_actions; /* 1: actions are in [-1,1] and are affinely mapped to the spec */
  int32_t has_fingering;     /* 1: fingering obs + fingering reward; 0: OT fingering reward; 2: neither
                              * (piano_with_one_shadow_hand.py:117-118 with disable_fingering_reward) */
  int32_t disable_forearm_reward;
  int32_t wrong_press_termination;
  int32_t reward_type;       /* PP_REWARD_* (self-actuated piano only) */
  int32_t auto_reset;        /* 1: an env whose previous step was LAST resets on the next step */
  int32_t max_contacts;      /* 0 = library default */
  int32_t non_legacy_step;   /* 0: dm_control's default legacy_step=True (mj_step2 ... mj_step1: the task reads positions,
                              * contacts and actuator velocities of the INTEGRATED state); 1: legacy_step=False (plain
                              * mj_step: it reads those of the last substep's start; qpos / qvel are the same either way).
                              * The self-actuated piano task reads nothing position-derived: the flag changes nothing there */
  double energy_penalty_coef;
  const uint8_t *goal;  /* host [T,89]: keys to hold + sustain, per control frame */
  const int16_t *pairs; /* host [T,pmax,2]: (key, finger 0..9), key=-1 padding; may be NULL */
} pp_config;

typedef struct pp_env pp_env;

int pp_create(const pp_model *model, const pp_config *config, pp_env **out);
int pp_destroy(pp_env *env);
int pp_num_envs(const pp_env *env);
int pp_nv(const pp_env *env);
int pp_action_dim(const pp_env *env);
int pp_obs_dim(const pp_env *env);
/* action spec (minimum/maximum of the un-canonicalised action), host arrays [action_dim] */
int pp_action_bounds(const pp_env *env, float *minimum, float *maximum);

/* mask_dev: uint8 [N] (1 = reset this env) or NULL for all. obs_dev may be NULL. */
int pp_reset(pp_env *env, const uint8_t *mask_dev, float *obs_dev, void *stream);
/* actions_dev [N,action_dim] f32; obs_dev [N,obs_dim] f32; reward_dev [N] f32;
 * reward_terms_dev [N,PP_NUM_TERMS] f32 or NULL; done_dev [N] u8 (1 = LAST);
 * discount_dev [N] f32 or NULL. */
int pp_step(pp_env *env, const float *actions_dev, float *obs_dev, float *reward_dev,
            float *reward_terms_dev, uint8_t *done_dev, float *discount_dev, void *stream);
/* Host-buffer variant: H2D of actions, step, D2H of obs/reward/done inside the call; returns
 * after the results are on the host. Page-locked caller buffers (cudaHostAlloc, cudaHostRegister,
 * torch pin_memory) are the DMA endpoints themselves; pageable ones go through pinned staging
 * owned by the handle plus one host memcpy each. Large two-hand batches are stepped in pieces on internal
 * streams so that a piece's D2H overlaps the later pieces' steps (same results); while the schedule tuner is still
 * measuring, and under the divergent schedule, the batch is stepped whole. */
/* Stream contract: the *_host entry points run on the handle's own non-blocking streams and return with the
 * results on the host. They are NOT ordered against work the caller queued earlier through the device entry points
 * (pp_step, pp_reset, pp_set_state, pp_set_hand_offsets, pp_set_qfrc_applied) on the caller's streams: synchronise
 * those streams before mixing the two families on one handle. */
int pp_step_host(pp_env *env, const float *actions, float *obs, float *reward,
                 float *reward_terms, uint8_t *done, float *discount);
int pp_reset_host(pp_env *env, const uint8_t *mask, float *obs);
/* Diagnostic (no reference counterpart): CUDA-event timeline of a host-buffer step. enable != 0 arms the
 * recording for the following pp_step_host calls; ms_out (1 + 2 * 8 floats, may be NULL) receives the times of
 * the LAST call in ms from its start: [0] actions on the device, then per piece [step done, results on the
 * host]; *n_pieces = pieces the batch was stepped in. */
int pp_host_timeline(pp_env *env, int enable, float *ms_out, int *n_pieces);

/* state: qpos/qvel/qacc_warmstart [N,nv] f32, t_idx [N] i32, sustain [N] f32 (device);
 * any pointer may be NULL */
int pp_get_state(pp_env *env, float *qpos_dev, float *qvel_dev, float *warm_dev,
                 int32_t *t_idx_dev, float *sustain_dev, void *stream);
int pp_set_state(pp_env *env, const float *qpos_dev, const float *qvel_dev, const float *warm_dev,
                 const int32_t *t_idx_dev, const float *sustain_dev, void *stream);
/* generalized force added to every substep (tests use it like physics.bind().qfrc_applied) */
int pp_set_qfrc_applied(pp_env *env, const float *qfrc_dev, void *stream);

/* MidiEvaluationWrapper (wrappers/evaluation.py:67-177): the step accumulates, per env, the per-step
 * precision / recall / F1 (sklearn binary, zero_division=1) of key activations against the frame's notes
 * and of the sustain activation against the pedal. episode_metrics_dev [N,6] f32 = means over the steps of
 * the env's LAST FINISHED episode {key P, R, F1, sustain P, R, F1}; episode_count_dev [N] i32 = finished
 * episodes since pp_create. Either pointer may be NULL. */
int pp_get_metrics(pp_env *env, float *episode_metrics_dev, int32_t *episode_count_dev, void *stream);

/* randomize_hand_positions (piano_with_shadow_hands.py:491-499, Entity.shift_pose): y_off_dev [N] f32 is
 * the world-y shift of BOTH hands that env i takes at its NEXT reset (the caller draws it, e.g.
 * uniform(-0.05, 0.05)); it stays in force until the reset after. NULL: zero shift from the next reset on. */
int pp_set_hand_offsets(pp_env *env, const float *y_off_dev, void *stream);

/* MIDI augmentations at episode start (robopianist/suite/variations.py:27-184 applied by
 * piano_with_shadow_hands.py:151-165 `_maybe_change_midi` / `_reset_trajectory`): n_songs note trajectories resident
 * on the device, back to back. goal [sum frames, 89] u8 and pairs [sum frames, pmax, 2] i16 (NULL for an unfingered
 * task) are HOST pointers, song_frames [n_songs] the length of each song; the call synchronises the device and
 * replaces the handle's tables (reset afterwards). pp_set_song_ids: ids_dev [N] int32 (device) = the song each env
 * takes at its NEXT reset (explicit or automatic); ids are clamped to the bank. Two-hands task only. */
int pp_set_song_bank(pp_env *env, const uint8_t *goal, const int16_t *pairs, const int32_t *song_frames,
                     int n_songs, int pmax);
int pp_set_song_ids(pp_env *env, const int32_t *ids_dev, void *stream);

/* Piano.after_substep -> MidiModule.after_substep (robopianist/models/piano/piano.py:154-162,
 * midi_module.py:47-98): the reference turns the key / pedal activations after EVERY physics substep into MIDI
 * note-on / note-off / sustain events. buf_dev: caller-owned uint32 [N, n_substeps, 4], 16-byte aligned, written
 * by every later pp_step / pp_step_host: words 0-2 = activation bits of keys 0-31, 32-63, 64-87, word 3 = pedal.
 * NULL stops the recording. Two-hands task only. */
int pp_set_substep_activations(pp_env *env, uint32_t *buf_dev);

/* diagnostics of the LAST substep of the last pp_step (device pointers, any may be NULL):
 * qacc [N,nv]; actuator_force [N,nu]; actuator_velocity [N,nu]; site_xpos [N,nsite,3];
 * activation [N,88] u8; counters [N,8] i32 = {ncon, nefc, newton iterations, dropped contacts} of the last substep, then
 * sums over the substeps of the last step: {newton iterations, line-search iterations, substeps
 * that took the dense fallback solver, cross-finger contacts (low-rank correction or dense)} */
int pp_get_diag(pp_env *env, float *qacc_dev, float *actuator_force_dev,
                float *actuator_velocity_dev, float *site_xpos_dev, uint8_t *activation_dev,
                int32_t *counters_dev, void *stream);

/* Launch shape of the two-hands step. The step is one warp per env; how many envs share a thread block decides
 * whether the warps of an SM walk through the (large) kernel together:
 *   PP_SCHED_LOCKSTEP   2 envs per block - best when every env does the same work (replaying one action table);
 *   PP_SCHED_DIVERGENT  10 envs per block, envs ordered by the Newton iterations of their previous step so that a
 *                       block's envs need similar work - best when the actions differ per env (training, random
 *                       actions; the reference's loop, parallelized_base_v2.py:116-166, is this case);
 *   PP_SCHED_COMPACTED  every substep as three launches of the same step code - prologue (kinematics .. constraint
 *                       rows), Newton loop, finish (integration of the keys, after the last substep the task
 *                       epilogue) - with the solver state of an env parked in device memory in between (about 30 KB
 *                       per env, allocated at the first such step). The Newton launch is one persistent block per
 *                       SM whose warps take the next env of a list whenever their env's loop is over, the lists
 *                       ordered by cost class (cross-finger contacts = right-hand sides per solve): no warp idles
 *                       while a block-mate iterates, and the block's warps still run the same trip of the loop
 *                       together. Best for batches of thousands of envs with per-env actions (1.2x the divergent shape at 4096 envs, 1.4x at 16384);
 *   PP_SCHED_AUTO       (default) times the three in turn on full-batch steps 4..15 of the handle (CUDA events) and
 *                       keeps the lockstep shape unless another is 3 % faster. Results are identical in every mode:
 *                       an env's result never depends on its block or launch. pp_schedule_in_force = the shape the
 *                       next step uses (PP_SCHED_LOCKSTEP / DIVERGENT / COMPACTED). */
enum { PP_SCHED_AUTO = 0, PP_SCHED_LOCKSTEP = 1, PP_SCHED_DIVERGENT = 2, PP_SCHED_COMPACTED = 3 };
int pp_set_schedule(pp_env *env, int mode);
int pp_get_schedule(const pp_env *env, int *mode, int *block_envs, int *sorted_by_work);
int pp_schedule_in_force(const pp_env *env);

/* number of kernels launched by this handle since creation (for bench's gpu_launches) */
long long pp_launch_count(const pp_env *env);
const char *pp_last_error(void);
const char *pp_version(void);

#ifdef __cplusplus
}
#endif
#endif /* PIANO_STEP_H_ */
