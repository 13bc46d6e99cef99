// pp_common.cuh - device-side model tables (fp32) and small math helpers shared by the
// step kernels. The tables are the fp32 image of include/pp_model.h, re-indexed for the
// fixed scene structure (88 keys + 2 hands x [25 bodies, 26 dofs]); see DESIGN.md.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define PP_NKEY 88
#define PP_HB 25          // bodies per hand
#define PP_HD 26          // dofs per hand
#define PP_NHB (2 * PP_HB)
#define PP_NHD (2 * PP_HD)
#define PP_NV (PP_NKEY + PP_NHD)
#define PP_MAXGEOM 40     // hand collider geoms (both hands)
#define PP_MAXCHAIN 9     // dofs between a hand body and the world
#define PP_MAXPAIR 256
#define PP_NSITE 10
#define PP_NU_HANDS 44
#define PP_MINVAL 1e-15f

struct KeyTables {
  float M[PP_NKEY];        // hinge inertia incl. armature
  float mghx[PP_NKEY];     // m*g*hx: gravity torque magnitude at q=0
  float hi[PP_NKEY];       // joint range upper bound (lower is 0)
  float invw[PP_NKEY];     // dof_invweight0
  float binvw[PP_NKEY];    // body_invweight0 (translational)
  float pos[PP_NKEY][3];   // body position (geom centre at q=0)
  float half[PP_NKEY][3];  // box half sizes
  float stiffness, springref, damping;
  int8_t white_key[52];    // key number of the i-th white key
  int8_t boundary_black[53];  // key number of the black key on boundary n (between white n-1 and n), -1 none
  float y0, pitch;         // white key i spans y0 + i*pitch .. y0 + i*pitch + width
};

struct HandTables {
  // bodies (index = hand*25 + local body)
  float body_pos[PP_NHB][3], body_quat[PP_NHB][4], ipos[PP_NHB][3], iquat[PP_NHB][4];
  float mass[PP_NHB], inertia[PP_NHB][3], gravcomp[PP_NHB], binvw[PP_NHB];
  // dofs (index = hand*26 + local dof)
  float axis[PP_NHD][3], range[PP_NHD][2], damping[PP_NHD], armature[PP_NHD];
  float frictionloss[PP_NHD], invw[PP_NHD];
  // collider geoms
  int ngeom;
  int8_t geom_body[PP_MAXGEOM];   // hand body index 0..49
  int8_t geom_cls[PP_MAXGEOM];
  int8_t geom_nchain[PP_MAXGEOM];
  int8_t geom_chain[PP_MAXGEOM][PP_MAXCHAIN];  // hand dof indices 0..51, root first
  float geom_pos[PP_MAXGEOM][3], geom_r[PP_MAXGEOM], geom_hl[PP_MAXGEOM];
  int npair;        // fixed hand-hand candidate pairs; [0,npair_same) same hand, the rest rh x lh
  int npair_same;
  int8_t pair_g1[PP_MAXPAIR], pair_g2[PP_MAXPAIR];
  int8_t forearm_geom[2];  // collider index of each hand's forearm capsule
  int8_t dof_act[PP_NHD];  // local actuator (0..21) driving each hand dof
  // sites
  int8_t site_body[PP_NSITE];
  float site_pos[PP_NSITE][3];
  // actuators
  int8_t act_trn0[PP_NU_HANDS], act_trn1[PP_NU_HANDS], act_forcelimited[PP_NU_HANDS];
  float act_kp[PP_NU_HANDS], act_ctrlrange[PP_NU_HANDS][2], act_forcerange[PP_NU_HANDS][2];
};

struct Options {
  float h, g[3], impratio, tolerance, meaninertia;
  int iterations, ls_iterations;
  float limit_solref[2], limit_solimp[5], friction_solref[2], friction_solimp[5];
  float pair_solref[2][2], pair_solimp[2][5], pair_friction[2], pair_margin[2];
  float base_pos[3], base_half[3];
};

struct DevModel {
  Options opt;
  KeyTables key;
  HandTables hand;
  // piano actuators (self-actuated task): ctrlrange hi per key
  float key_ctrl_hi[PP_NKEY];
};

// ------------------------------------------------------------------ math
struct f3 { float x, y, z; };
struct f4 { float w, x, y, z; };
__device__ __forceinline__ f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ f3 ld3(const float* p) { return mk3(p[0], p[1], p[2]); }
__device__ __forceinline__ void st3(float* p, f3 v) { p[0] = v.x; p[1] = v.y; p[2] = v.z; }
__device__ __forceinline__ f3 operator+(f3 a, f3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ f3 operator-(f3 a, f3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ f3 operator*(float s, f3 a) { return mk3(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ float dot(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ f3 cross(f3 a, f3 b) {
  return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ float norm(f3 a) { return sqrtf(dot(a, a)); }
__device__ __forceinline__ f4 qmul(f4 a, f4 b) {
  f4 r;
  r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
  r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
  r.y = a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x;
  r.z = a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w;
  return r;
}
__device__ __forceinline__ f3 qrot(f4 q, f3 v) {  // R(q) v
  f3 u = mk3(q.x, q.y, q.z);
  f3 t = 2.f * cross(u, v);
  return v + q.w * t + cross(u, t);
}
__device__ __forceinline__ f3 qrot_inv(f4 q, f3 v) {
  f3 u = mk3(-q.x, -q.y, -q.z);
  f3 t = 2.f * cross(u, v);
  return v + q.w * t + cross(u, t);
}
__device__ __forceinline__ f4 qnormalize(f4 q) {
  float s = rsqrtf(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
  f4 r; r.w = q.w * s; r.x = q.x * s; r.y = q.y * s; r.z = q.z * s; return r;
}
__device__ __forceinline__ float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

// general power-law branch of getimpedance(), kept out of line (powf is ~200 instructions)
static __device__ __noinline__ float pp_impedance_pow(float x, float mid, float power) {
  return x <= mid ? powf(x, power) / powf(mid, power - 1.f) : 1.f - powf(1.f - x, power) / powf(1.f - mid, power - 1.f);
}
// sin and cos for |x| <= ~2 (joint half angles, key angles): Taylor/Horner, error < 2e-8 for |x| <= 1
__device__ __forceinline__ void pp_sincos_small(float x, float* s, float* c) {
  const float x2 = x * x;
  *s = x * (1.f - x2 * (1.f / 6.f) * (1.f - x2 * (1.f / 20.f) * (1.f - x2 * (1.f / 42.f) * (1.f - x2 * (1.f / 72.f) * (1.f - x2 * (1.f / 110.f) * (1.f - x2 * (1.f / 156.f)))))));
  *c = 1.f - x2 * 0.5f * (1.f - x2 * (1.f / 12.f) * (1.f - x2 * (1.f / 30.f) * (1.f - x2 * (1.f / 56.f) * (1.f - x2 * (1.f / 90.f) * (1.f - x2 * (1.f / 132.f) * (1.f - x2 * (1.f / 182.f)))))));
}
// getimpedance(): MuJoCo's position-dependent impedance d(r) (power-law sigmoid)
__device__ __forceinline__ float pp_impedance(const float* solimp, float pos, float margin) {
  float dmin = clampf(solimp[0], 1e-4f, 0.9999f), dmax = clampf(solimp[1], 1e-4f, 0.9999f);
  float width = fmaxf(solimp[2], PP_MINVAL), mid = clampf(solimp[3], 1e-4f, 0.9999f);
  float power = fmaxf(solimp[4], 1.f);
  if (dmin == dmax) return 0.5f * (dmin + dmax);
  float x = fabsf((pos - margin) / width), y;
  if (x >= 1.f) return dmax;
  if (x <= 0.f) return dmin;
  if (power == 1.f) y = x;
  else if (power == 2.f) y = x <= mid ? x * x / mid : 1.f - (1.f - x) * (1.f - x) / (1.f - mid);  // MuJoCo default
  else y = pp_impedance_pow(x, mid, power);
  return dmin + y * (dmax - dmin);
}
// solref -> (K, B), with the refsafe clamp timeconst >= 2*timestep
__device__ __forceinline__ void pp_kb(const float* solref, const float* solimp, float h, float& K, float& B) {
  float tc = solref[0], dr = solref[1], dmax = clampf(solimp[1], 1e-4f, 0.9999f);
  if (tc > 0.f) {
    tc = fmaxf(tc, 2.f * h);
    K = 1.f / fmaxf(PP_MINVAL, dmax * dmax * tc * tc * dr * dr);
    B = 2.f / fmaxf(PP_MINVAL, dmax * tc);
  } else {
    K = -tc / fmaxf(PP_MINVAL, dmax * dmax);
    B = -dr / fmaxf(PP_MINVAL, dmax);
  }
}
// dm_control.utils.rewards.tolerance(x, bounds=(0,hi), margin, sigmoid="gaussian"),
// value_at_margin = 0.1 (reference piano_with_shadow_hands.py:25 import)
__device__ __forceinline__ float pp_tolerance(float x, float hi, float margin) {
  if (x >= 0.f && x <= hi) return 1.f;
  float d = (x < 0.f ? -x : x - hi) / margin;
  const float scale = 2.1459660262893472f;  // sqrt(-2 ln 0.1)
  float t = d * scale;
  return expf(-0.5f * t * t);
}

// ------------------------------------------------------------------ musical metrics
// MidiEvaluationWrapper (reference wrappers/evaluation.py:67-177): every step, precision / recall / F1
// (sklearn precision_recall_fscore_support, average="binary", zero_division=1) of the 88 key
// activations against the frame's notes, and of the sustain activation against the frame's pedal; the
// episode's metrics are the means over its steps. Per env: running sums [6] = key P, R, F1, sustain P,
// R, F1; the finished episode's means; a count of finished episodes.
struct EvalAccum {
  float* sum;      // [N,6] running sums of the current episode
  int* steps;      // [N]
  float* episode;  // [N,6] means of the last finished episode
  int* count;      // [N] finished episodes
};
__device__ __forceinline__ float pp_prf_div(int num, int den) { return den > 0 ? (float)num / (float)den : 1.f; }
__device__ __forceinline__ void pp_eval_metrics(const EvalAccum& E, int env, int lane, const bool act[3],
                                                const uint8_t* goal_cur, bool sustain_on, bool do_reset, bool done) {
  if (!E.sum) return;
  if (do_reset) {
    if (lane < 6) E.sum[(size_t)env * 6 + lane] = 0.f;
    if (lane == 0) E.steps[env] = 0;
    return;
  }
  int tp = 0, fp = 0, fn = 0;
#pragma unroll
  for (int i = 0; i < 3; i++) {
    const int k = lane + 32 * i;
    const unsigned ba = __ballot_sync(0xffffffffu, k < PP_NKEY && act[i]);
    const unsigned bg = __ballot_sync(0xffffffffu, k < PP_NKEY && goal_cur[k] != 0);
    tp += __popc(ba & bg); fp += __popc(ba & ~bg); fn += __popc(~ba & bg);
  }
  const bool st = goal_cur[88] != 0;
  const int stp = (sustain_on && st) ? 1 : 0, sfp = (sustain_on && !st) ? 1 : 0, sfn = (!sustain_on && st) ? 1 : 0;
  float v = 0.f;
  if (lane == 0) v = pp_prf_div(tp, tp + fp);
  else if (lane == 1) v = pp_prf_div(tp, tp + fn);
  else if (lane == 2) v = pp_prf_div(2 * tp, 2 * tp + fp + fn);
  else if (lane == 3) v = pp_prf_div(stp, stp + sfp);
  else if (lane == 4) v = pp_prf_div(stp, stp + sfn);
  else if (lane == 5) v = pp_prf_div(2 * stp, 2 * stp + sfp + sfn);
  const int n = E.steps[env] + 1;
  __syncwarp();
  if (lane < 6) {
    const float s = E.sum[(size_t)env * 6 + lane] + v;
    E.sum[(size_t)env * 6 + lane] = s;
    if (done) E.episode[(size_t)env * 6 + lane] = s / (float)n;
  }
  if (lane == 0) { E.steps[env] = n; if (done) E.count[env] += 1; }
}
