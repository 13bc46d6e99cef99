// pp_env.cuh - kernel argument blocks (plain structs of device pointers) and launchers.
#pragma once
#include "pp_common.cuh"

enum { PP_MODE_STEP = 0, PP_MODE_RESET = 1 };

struct PianoOnlyArgs {
  const DevModel* dm;
  int num_envs, mode, n_substeps, n_frames, lookahead, obs_dim;
  int canonical, reward_type, auto_reset;
  const uint8_t* goal;        // [T,89]
  const uint8_t* reset_mask;  // [N] or null (RESET mode)
  const float* actions;       // [N,89]
  const float* qfrc_applied;  // [N,88] or null
  float *qpos, *qvel;         // [N,88]
  int* t_idx;                 // [N]
  float* sustain;             // [N]
  uint8_t* needs_reset;       // [N]
  float* obs;                 // [N,obs_dim] or null
  float *reward, *reward_terms, *discount;
  uint8_t* done;
  float* qacc;                // diag [N,88] or null
  uint8_t* activation;        // diag [N,88] or null
  EvalAccum eval;             // musical metrics accumulators
};

struct HandsArgs {
  const DevModel* dm;
  int num_envs, mode, n_substeps, n_frames, lookahead, obs_dim, pmax;
  int canonical, has_fingering, disable_forearm_reward, wrong_press_termination, auto_reset;
  int max_contacts;
  float energy_coef;
  const uint8_t* goal;    // [T,89]
  const int16_t* pairs;   // [T,pmax,2]
  const uint8_t* reset_mask;
  const float* actions;   // [N,45]
  const float* act_min;   // [45] action spec
  const float* act_max;   // [45]
  const float* qfrc_applied;  // [N,140] or null
  float* hand_off;                // [N] y shift of both hands in force (null: feature unused)
  const float* hand_off_pending;  // [N] y shift an env takes at its next reset
  float *qpos, *qvel, *warm;  // [N,140]
  int* t_idx;
  float* sustain;
  uint8_t* needs_reset;
  unsigned* stick_bits;  // [N,2] per hand: dofs whose frictionloss row ended the last solve in its quadratic zone
  float* obs;
  float *reward, *reward_terms, *discount;
  uint8_t* done;
  // diagnostics (always allocated by the handle)
  float *qacc, *act_force, *act_vel, *site_xpos;
  uint8_t* activation;
  int* counters;  // [N,8]
  const int* order;  // [N] launch slot -> env (null: identity); envs of similar expected work share a block
  uint4* substep_act;  // [N, n_substeps] or null: key activation bits 0-31 / 32-63 / 64-87 and the pedal after every substep
  EvalAccum eval;  // musical metrics accumulators
  // song bank (MIDI augmentations, suite/variations.py): goal / pairs hold n_songs songs back to back; null = one song
  int* song_id;                // [N] song each env is playing
  const int* song_id_pending;  // [N] song an env takes at its next reset
  const int* song_off;         // [n_songs] first frame of the song inside goal / pairs
  const int* song_T;           // [n_songs] frames of the song
  int n_songs;
  // legacy_step=False (dm_control Physics.step = plain mj_step, no trailing mj_step1): position- and velocity-dependent
  // quantities the task reads after the step are those of the LAST SUBSTEP'S START. stale [N,7,32]: hand qpos / qvel
  // (2 + 2 rows, lane = dof) and key qpos (3 rows) saved there; null = legacy_step=True
  float* stale;
  // split schedule (pp_hands.cu MODE): one substep = prologue kernel, Newton kernel over the list of stepping envs,
  // finish kernel; the solver state of an env waits in sp_regs / sp_smem between them
  int sub_index;            // substep this launch works on
  int sp_chain;             // finish kernel: go on into the next substep's prologue (all but the last substep)
  float* sp_regs;           // [N, PP_SP_NREG, 32] per-lane solver state
  float4* sp_smem;          // [N, sizeof(WarpShared) / 16] the env's shared-memory image
  uint8_t* sp_skip;         // [N] env reset at substep 0 and sits the control step out
  int* sp_list_out;         // prologue: the envs that step, in PP_SP_BUCKETS lists of [N]; sp_count_out[k] = entries of list k
  int* sp_count_out;
  const int* sp_list_in;    // Newton kernel: the same list; *sp_cursor = next entry to take
  const int* sp_count_in;
  int* sp_cursor;
};
constexpr int PP_SP_NREG = 64;
constexpr int PP_SP_BUCKETS = 12;  // lists of the split schedule: 11+, 10, .. 0 cross-front contacts, heaviest first

void pp_launch_piano_only(const PianoOnlyArgs& A, cudaStream_t stream);
// launches the step for envs [env_begin, env_begin + env_count) (env_count < 0: to the end); returns the number of kernels
// block_envs: 2 (envs that do the same work at the same time) or 10 (envs that diverge: see pp_set_schedule)
int pp_launch_hands(const HandsArgs& A, cudaStream_t stream, int env_begin = 0, int env_count = -1, int block_envs = 2);
// order[slot] = env, by decreasing counters[env*8+4] (Newton iterations of the env's previous control step)
void pp_launch_work_order(const int* counters, int* order, int num_envs, cudaStream_t stream);
size_t pp_hands_smem_bytes();
// split schedule (pp_hands_split.cu): one control step as prologue / Newton-trip / finish launches per substep
int pp_launch_hands_split(const HandsArgs& A, cudaStream_t stream, int sms, int* list, int* counts);
size_t pp_hands_split_smem_words();  // float4 words of an env's shared-memory image
