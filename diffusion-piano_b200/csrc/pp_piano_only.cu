// pp_piano_only.cu - SelfActuatedPiano step (BASELINE config 1): 88 decoupled key hinges with
// torque actuators, no hands. One warp per environment, lane l owns keys l, l+32, l+64.
//
// Replaces, per env: SelfActuatedPiano.before_step/after_step/_compute_key_press_reward
// (reference self_actuated_piano.py:143-167,203-208), Piano.apply_action and the actuator
// branch of Piano._update_key_state (piano.py:180-182,208-218), n_substeps x mj_step,
// _update_goal_state (:210-228) and the observation assembly.
#include "pp_env.cuh"
#include "pp_keys.cuh"

__global__ void __launch_bounds__(128) pp_piano_only_step_kernel(PianoOnlyArgs A) {
  const int lane = threadIdx.x & 31;
  const int env = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (env >= A.num_envs) return;
  const DevModel& dm = *A.dm;
  const Options& o = dm.opt;
  const int T = A.n_frames, L = A.lookahead, D = A.obs_dim;
  float* obs = A.obs ? A.obs + (size_t)env * D : nullptr;

  int t_idx = A.t_idx[env];
  const bool do_reset = A.mode == PP_MODE_RESET ? (A.reset_mask ? A.reset_mask[env] != 0 : true)
                                                : (A.auto_reset && A.needs_reset[env]);
  if (A.mode == PP_MODE_RESET && !do_reset) return;

  float q[3], v[3], ctrl_raw[3];
  float sustain = 0.f;
  if (do_reset) {
#pragma unroll
    for (int i = 0; i < 3; i++) { q[i] = 0.f; v[i] = 0.f; ctrl_raw[i] = 0.f; }
    t_idx = 0;
  } else {
    const float* act = A.actions + (size_t)env * 89;
#pragma unroll
    for (int i = 0; i < 3; i++) {
      int k = lane + 32 * i;
      if (k < PP_NKEY) {
        q[i] = A.qpos[(size_t)env * PP_NKEY + k];
        v[i] = A.qvel[(size_t)env * PP_NKEY + k];
        float a = act[k];
        if (A.canonical) a = 0.5f * (a + 1.f) * dm.key_ctrl_hi[k];  // spec minimum is 0
        ctrl_raw[i] = a;
      }
    }
    sustain = act[88];
    if (A.canonical) sustain = 0.5f * (sustain + 1.f);
    // ---- n_substeps of mj_step on the decoupled hinges
    for (int s = 0; s < A.n_substeps; s++) {
#pragma unroll
      for (int i = 0; i < 3; i++) {
        int k = lane + 32 * i;
        if (k >= PP_NKEY) continue;
        float hi = dm.key.hi[k];
        float c = clampf(ctrl_raw[i], 0.f, dm.key_ctrl_hi[k]);  // mj_fwdActuation clamps ctrl
        float applied = A.qfrc_applied ? A.qfrc_applied[(size_t)env * PP_NKEY + k] : 0.f;
        float fs = pp_key_smooth(dm.key, k, q[i], v[i], c + applied);
        LimitRow r = pp_limit_row(o, q[i], v[i], 0.f, hi, dm.key.invw[k]);
        float M = dm.key.M[k];
        float a = pp_key_solve(M, fs, r);
        float fcon = (r.sgn != 0.f && r.sgn * a - r.aref < 0.f) ? -r.D * (r.sgn * a - r.aref) * r.sgn : 0.f;
        if (A.qacc) A.qacc[(size_t)env * PP_NKEY + k] = a;
        pp_scalar_euler(M, dm.key.damping, o.h, fs + fcon, q[i], v[i]);
      }
    }
    t_idx += 1;
  }

  // ---- key state (actuator branch): activation = ctrl >= midpoint of ctrlrange
  bool act_on[3];
  float sq = 0.f, xent = 0.f;
  const bool sustain_on = sustain >= 0.5f;
  const uint8_t* goal_cur = A.goal + (size_t)(t_idx > 0 ? t_idx - 1 : 0) * 89;
#pragma unroll
  for (int i = 0; i < 3; i++) {
    int k = lane + 32 * i;
    act_on[i] = false;
    if (k >= PP_NKEY) continue;
    act_on[i] = ctrl_raw[i] >= 0.5f * dm.key_ctrl_hi[k];
    float p = act_on[i] ? 1.f : 0.f, g = (float)goal_cur[k];
    sq += (p - g) * (p - g);
    xent += g * logf(p + 1e-6f) + (1.f - g) * logf(1.f - p + 1e-6f);
    if (A.activation) A.activation[(size_t)env * PP_NKEY + k] = act_on[i];
  }
  if (lane == 0) {
    float p = sustain_on ? 1.f : 0.f, g = (float)goal_cur[88];
    sq += (p - g) * (p - g);
    xent += g * logf(p + 1e-6f) + (1.f - g) * logf(1.f - p + 1e-6f);
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    sq += __shfl_xor_sync(0xffffffffu, sq, off);
    xent += __shfl_xor_sync(0xffffffffu, xent, off);
  }

  // ---- write back state
#pragma unroll
  for (int i = 0; i < 3; i++) {
    int k = lane + 32 * i;
    if (k < PP_NKEY) {
      A.qpos[(size_t)env * PP_NKEY + k] = q[i];
      A.qvel[(size_t)env * PP_NKEY + k] = v[i];
    }
  }
  const bool done = !do_reset && t_idx == T;
  pp_eval_metrics(A.eval, env, lane, act_on, goal_cur, sustain_on, do_reset, done);
  if (lane == 0) {
    A.t_idx[env] = t_idx;
    A.sustain[env] = sustain;
    A.needs_reset[env] = done ? 1 : 0;
    if (A.mode == PP_MODE_STEP) {
      float rew = do_reset ? 0.f : (A.reward_type == 0 ? -sqrtf(sq) : xent);
      A.reward[env] = rew;
      if (A.reward_terms) {
        for (int j = 0; j < 5; j++) A.reward_terms[(size_t)env * 5 + j] = j == 0 ? rew : 0.f;
      }
      A.done[env] = done ? 1 : 0;
      if (A.discount) A.discount[env] = 1.f;
    }
  }
  // ---- observation: goal (L+1 frames from t_idx; unchanged when t_idx == T), activation, sustain
  if (obs) {
    const int t0 = t_idx == T ? T - 1 : t_idx;  // the goal update is skipped on the last step
    for (int i = lane; i < (L + 1) * 89; i += 32) {
      int f = t0 + i / 89;
      obs[i] = f < T ? (float)A.goal[(size_t)f * 89 + (i % 89)] : 0.f;
    }
    float* o2 = obs + (L + 1) * 89;
#pragma unroll
    for (int i = 0; i < 3; i++) {
      int k = lane + 32 * i;
      if (k < PP_NKEY) o2[k] = act_on[i] ? 1.f : 0.f;
    }
    if (lane == 0) o2[88] = sustain_on ? 1.f : 0.f;
  }
}

void pp_launch_piano_only(const PianoOnlyArgs& A, cudaStream_t stream) {
  const int warps = 4;
  dim3 grid((A.num_envs + warps - 1) / warps), block(32 * warps);
  pp_piano_only_step_kernel<<<grid, block, 0, stream>>>(A);
}
