// pp_keys.cuh - closed-form dynamics of one piano key (a 1-DoF hinge about world +y).
//
// Each key is an independent sub-problem of mj_step unless a finger touches it:
//   M a = spring + damper + gravity + actuator + applied  (+ soft joint-limit force)
// With one scalar unknown the Newton solve of mj_fwdConstraint has a closed form, so the
// GPU path never iterates for untouched keys (the oracle solves the same rows with its
// generic dense Newton). Model numbers: reference piano_constants.py:22-85, piano_mjcf.py:65-100.
#pragma once
#include "pp_common.cuh"

struct LimitRow {
  float sgn;   // +1 lower limit, -1 upper limit, 0 inactive
  float D;     // 1/R
  float aref;  // reference acceleration
};

// joint-limit constraint row of a dof with range [lo,hi], margin 0 (mj_makeConstraint)
__device__ __forceinline__ LimitRow pp_limit_row(const Options& o, float q, float v, float lo, float hi,
                                                 float invw) {
  LimitRow r; r.sgn = 0.f; r.D = 0.f; r.aref = 0.f;
  float dist;
  if (q - lo < 0.f) { r.sgn = 1.f; dist = q - lo; }
  else if (hi - q < 0.f) { r.sgn = -1.f; dist = hi - q; }
  else return r;
  float K, B;
  pp_kb(o.limit_solref, o.limit_solimp, o.h, K, B);
  float imp = pp_impedance(o.limit_solimp, dist, 0.f);
  float R = fmaxf(PP_MINVAL, (1.f - imp) * invw / imp);
  r.D = 1.f / R;
  r.aref = -B * r.sgn * v - K * imp * dist;
  return r;
}

// qfrc_smooth of key k: passive - bias + extra (actuator / applied)
__device__ __forceinline__ float pp_key_smooth(const KeyTables& kt, int k, float q, float v, float extra) {
  float passive = -kt.stiffness * (q - kt.springref) - kt.damping * v;
  float bias = -kt.mghx[k] * cosf(q);
  return passive - bias + extra;
}

// exact minimiser of 1/2 M (a - fs/M)^2 + s(sgn*a - aref) for a single inequality row
__device__ __forceinline__ float pp_key_solve(float M, float fs, const LimitRow& r) {
  float a = fs / M;
  if (r.sgn != 0.f && r.sgn * a - r.aref < 0.f) a = (fs + r.D * r.sgn * r.aref) / (M + r.D);
  return a;
}

// mj_Euler with implicit damping for a scalar dof: (M + h b) a' = f_smooth + f_constraint
__device__ __forceinline__ void pp_scalar_euler(float M, float damping, float h, float ftotal,
                                                float& q, float& v) {
  float a = ftotal / (M + h * damping);
  v += h * a;
  q += h * v;
}

// Piano._update_key_state (reference piano.py:178-192): clip, normalise, threshold
__device__ __forceinline__ void pp_key_state(float q, float hi, float& normalized, bool& active) {
  float s = clampf(q, 0.f, hi);
  normalized = s / hi;
  active = fabsf(s - hi) <= 0.00872665f;
}
