// pp_hands.cu - PianoWithShadowHands step (BASELINE configs 2-5): 88 key hinges + two 26-dof
// Shadow Hands. ONE WARP OWNS ONE ENVIRONMENT for the whole control step (n_substeps x mj_step
// + task epilogue); lane l owns hand dof l of BOTH hands (two independent instruction streams
// per lane) and piano keys l, l+32, l+64.
//
// Replaces, per env and per call (reference file:line):
//   CanonicalSpecWrapper.step affine map                     parallelized_base_v2.py:44
//   PianoWithShadowHands.before_step / apply_action          piano_with_shadow_hands.py:176-186
//   n_substeps x physics.step() (MuJoCo mj_step, third party) + Piano.after_substep
//                                                            piano.py:154-192
//   after_step, rewards, goal/fingering observables, termination
//                                                            piano_with_shadow_hands.py:188-412
// Per substep (names follow mjData; the tests check every stage against the fp64 CPU restatement):
//   kinematics + RNE forward pass (level loop over the fixed tree, warp shuffles)
//   CRBA: composite inertias by subtree sums, M as <=9 ancestor entries per dof (registers)
//   actuator / passive / bias forces -> qfrc_smooth
//   collision: finger capsules vs the <=9 candidate keys under them (y-band lookup) + base,
//              fixed hand-hand capsule pairs
//   constraints: dof frictionloss, joint limits, pyramidal contacts (4 rows each)
//   Newton on the primal cost with exact line search. Untouched keys are solved in closed
//   form (pp_keys.cuh); touched keys are eliminated first (Schur complement), the hand blocks
//   are factorised leaves-first so the kinematic-tree sparsity of M + path cliques is kept
//   (envelope Cholesky in shared memory), dense / coupled envelopes only when a hand-hand
//   contact needs them.
//   mj_Euler with implicit joint damping.
#include "../../include/pp_model.h"
#include "pp_env.cuh"
#include "pp_keys.cuh"
#ifndef PP_T2
#define PP_T2 1
#endif
#ifndef PP_T4
#define PP_T4 1
#endif

namespace {

constexpr int NC = 24;   // contact cap per env
constexpr int KS = 16;   // touched-key slots per env
constexpr int NE = 18;   // jacobian entries per contact (two 9-dof chains, or chain + key)
constexpr unsigned FULL = 0xffffffffu;
constexpr float ZCULL = 0.045f;  // nothing of the piano is above this height

constexpr int LRC = 8;       // contacts that couple two fronts (hand-hand, doubly touched key)
constexpr int LRR = 3 * LRC;   // rank of their Woodbury correction (3 per contact: U V, W = V V^T)
constexpr int ZS = 52 + KS + 1;  // row stride of the solved vectors (odd: conflict-free down a column)

// Shared memory of one warp. Two phases of a substep alias each other: `k` (kinematics ..
// mj_makeConstraint) and `s` (the solver).
struct __align__(16) WarpShared {
  union {
    struct {
      float xpos[52][3];   // joint anchor (= body origin) of every hand dof node, world
      float xaxis[52][3];  // joint axis, world
      float xquat[52][4];  // body orientation of the node
      float S[52][6];      // spatial motion axis about the hand's reference point
      float raw[52][16];   // body spatial inertia (10) + RNE force (6)
      float comp[10][16];  // composite sums of the finger chains
      float c_geo[7][NC];  // contact pos, normal, dist
      float gpos[PP_MAXGEOM][3], gaxis[PP_MAXGEOM][3];  // world centre and axis of every hand collider
    } k;
    struct {
      union {
        struct {
          float Z[LRR + 1][ZS];  // row 0: -grad -> y0; rows 1..m: low-rank columns B -> H_t^-1 B
          float C[LRR][LRR + 1];      // capacitance I + B^T H_t^-1 B | B^T y0
        };
        struct {                      // dense fallback (more than LRC cross-finger contacts)
          float Hd[52 * 53 / 2];      // packed: A(gi <= gk) at gk(gk+1)/2 + gi
          float Kcd[KS][52];
        };
      };
      float Ht[52][9];            // H_t(ancestor j of dof, dof): tree entries only (like Mt)
      float Kc[KS][9];            // H_t(ancestor j on the slot's root path, key slot)
      float hss[KS];              // 1 / H_t(slot, slot)
    } s;
  } a;
  float kq[PP_NKEY], kv[PP_NKEY], kw[PP_NKEY];  // key qpos, qvel, warm-start acceleration
  int8_t key_slot[PP_NKEY];
  int8_t slot_key[KS];
  unsigned slot_mask[KS]; // hand dofs coupled to the slot (a root path of one finger)
  int8_t slot_hand[KS];
  float Mt[52][9];        // M(ancestor j of dof, dof): j < 4 root dofs 0..3, j >= 4 the dof's chain
  float va[52 + KS];      // vector exchange (acceleration / velocity / search direction)
  float dg[52];           // diagonal added to M when the fronts are loaded (row curvatures; Euler: h * damping)
  // contacts, SoA over the contact index
  float c_u[NE * 3][NC];  // per entry: (n, t1, t2) . jacobian column
  uint8_t c_col[NE][NC];  // column of the entry: hand dof 0..51 or 52+slot
  unsigned c_mask[2][NC]; // hand dofs in the entry list, per hand (ascending dof order)
  int8_t c_slot[NC], c_nent[NC], c_lr[NC];
  float c_w[5][NC];       // D * (sum c, mu (c0-c1), mu^2 (c0+c1), mu (c2-c3), mu^2 (c2+c3))
  float c_wf[3][NC];      // 3-vector weight of the rows (forces, or curvature * jp)
  int c_ids[NC];          // h1 | node1<<1 | kind<<6 | (key or h2|node2<<1)<<8
};
static_assert(sizeof(WarpShared) % 16 == 0, "WarpShared is laid out back to back, one per warp");
static_assert(10 * sizeof(WarpShared) + 5 * 1024 <= 228 * 1024, "10 warps (5 blocks of 2) per SM must fit the 228 KB of shared memory");

struct NodeInfo { int parent, depth, nsub, cs; unsigned anc; };
__device__ __forceinline__ NodeInfo node_info(int l) {
  NodeInfo n;
  if (l < 4) { n.cs = 0; n.parent = l > 0 ? l - 1 : 0; n.depth = l; n.nsub = 26 - l; n.anc = (2u << l) - 1u; }
  else {
    const int cs = l >= 21 ? 21 : l >= 16 ? 16 : l >= 12 ? 12 : l >= 8 ? 8 : 4;
    const int ce = cs == 4 ? 7 : cs == 8 ? 11 : cs == 12 ? 15 : cs == 16 ? 20 : 25;
    n.cs = cs; n.parent = l == cs ? 3 : l - 1; n.depth = 4 + l - cs; n.nsub = ce - l + 1;
    n.anc = 0xFu | (((2u << (l - cs)) - 1u) << cs);
  }
  return n;
}
__device__ __forceinline__ int nsub_of(int l) {
  if (l < 4) return 26 - l;
  const int ce = l >= 21 ? 25 : l >= 16 ? 20 : l >= 12 ? 15 : l >= 8 ? 11 : 7;
  return ce - l + 1;
}
__device__ __forceinline__ unsigned anc_of(int l) { return node_info(l).anc; }
__device__ __forceinline__ int tri(int i) { return (i * (i + 1)) >> 1; }
__device__ __forceinline__ int cs_of(int l) { return l >= 21 ? 21 : l >= 16 ? 16 : l >= 12 ? 12 : l >= 8 ? 8 : l >= 4 ? 4 : 0; }
// position of ancestor `a` in the ancestor list of a dof whose chain starts at cs
__device__ __forceinline__ int apos(int a, int cs) { return a < 4 ? a : 4 + a - cs; }
__device__ __forceinline__ f3 shfl3(f3 v, int src) {
  return mk3(__shfl_sync(FULL, v.x, src), __shfl_sync(FULL, v.y, src), __shfl_sync(FULL, v.z, src));
}
__device__ __forceinline__ f4 shfl4(f4 v, int src) {
  f4 r; r.w = __shfl_sync(FULL, v.w, src); r.x = __shfl_sync(FULL, v.x, src);
  r.y = __shfl_sync(FULL, v.y, src); r.z = __shfl_sync(FULL, v.z, src); return r;
}
__device__ __forceinline__ float wsum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
  return v;
}
__device__ __forceinline__ f4 ld4(const float* p) { f4 r; r.w = p[0]; r.x = p[1]; r.y = p[2]; r.z = p[3]; return r; }

// ------------------------------------------------------------------ narrow phase
// Closest-feature rules shared with the fp64 CPU restatement (sphere_box, capsule_box,
// capsule_capsule): normal from geom1 to geom2, dist < 0 is penetration, pos is the midpoint.
struct Hit { bool ok; float dist; f3 n, pos; };
struct Box { f3 p; float cq, sq; f3 h; };  // box rotated by angle q about world +y

__device__ __forceinline__ f3 to_box(const Box& b, f3 r) { return mk3(r.x * b.cq - r.z * b.sq, r.y, r.x * b.sq + r.z * b.cq); }
__device__ __forceinline__ f3 from_box(const Box& b, f3 r) { return mk3(r.x * b.cq + r.z * b.sq, r.y, -r.x * b.sq + r.z * b.cq); }

__device__ __forceinline__ Hit sphere_box(f3 c, float r, const Box& b, float margin) {
  Hit o; o.ok = false; o.dist = 0.f; o.n = mk3(0, 0, 1); o.pos = c;
  const f3 cl = to_box(b, c - b.p);
  const f3 q = mk3(clampf(cl.x, -b.h.x, b.h.x), clampf(cl.y, -b.h.y, b.h.y), clampf(cl.z, -b.h.z, b.h.z));
  bool outside = (q.x != cl.x) || (q.y != cl.y) || (q.z != cl.z);
  f3 nl = q - cl;
  float len = 0.f, dist;
  if (outside) { len = norm(nl); if (len < PP_MINVAL) outside = false; }
  if (outside) {
    dist = len - r;
    if (dist >= margin) return o;
    nl = (1.f / len) * nl;
  } else {
    const float vx = fabsf(cl.x) - b.h.x, vy = fabsf(cl.y) - b.h.y, vz = fabsf(cl.z) - b.h.z;
    int best = 0; float bestv = vx;
    if (vy > bestv) { bestv = vy; best = 1; }
    if (vz > bestv) { bestv = vz; best = 2; }
    const float cb = best == 0 ? cl.x : best == 1 ? cl.y : cl.z;
    const float s = cb < 0.f ? 1.f : -1.f;
    nl = mk3(best == 0 ? s : 0.f, best == 1 ? s : 0.f, best == 2 ? s : 0.f);
    dist = bestv - r;
  }
  o.n = from_box(b, nl);
  o.pos = c + (r + 0.5f * dist) * o.n;
  o.dist = dist; o.ok = true;
  return o;
}

__device__ __forceinline__ float seg_box_dgrad(f3 al, f3 dv, f3 h, float t) {
  const float sx = al.x + t * dv.x, sy = al.y + t * dv.y, sz = al.z + t * dv.z;
  const float ex = sx < -h.x ? sx + h.x : (sx > h.x ? sx - h.x : 0.f);
  const float ey = sy < -h.y ? sy + h.y : (sy > h.y ? sy - h.y : 0.f);
  const float ez = sz < -h.z ? sz + h.z : (sz > h.z ? sz - h.z : 0.f);
  return dv.x * ex + dv.y * ey + dv.z * ez;
}

// capsule / sphere (geom1) against a box: contact at the closest segment point, plus one at the
// far segment end when it also penetrates
__device__ __noinline__ void capsule_box(f3 cp, f3 axis, float r, float hl, const Box& b, float margin,
                                            Hit& h0, Hit& h1) {
  h1.ok = false;
  if (hl == 0.f) { h0 = sphere_box(cp, r, b, margin); return; }
  const f3 a = cp - hl * axis, e = cp + hl * axis;
  const f3 al = to_box(b, a - b.p), bl = to_box(b, e - b.p), dv = bl - al;
  float lo = 0.f, hi = 1.f, glo = seg_box_dgrad(al, dv, b.h, 0.f), ghi = seg_box_dgrad(al, dv, b.h, 1.f), t;
  if (glo >= 0.f) t = 0.f;
  else if (ghi <= 0.f) t = 1.f;
  else {
    const float alv[3] = {al.x, al.y, al.z}, dvv[3] = {dv.x, dv.y, dv.z}, hv[3] = {b.h.x, b.h.y, b.h.z};
#pragma unroll
    for (int i = 0; i < 3; i++) {
      if (fabsf(dvv[i]) <= 1e-12f) continue;
#pragma unroll
      for (int sg = -1; sg <= 1; sg += 2) {
        const float tb = ((float)sg * hv[i] - alv[i]) / dvv[i];
        if (tb > lo && tb < hi) {
          const float g = seg_box_dgrad(al, dv, b.h, tb);
          if (g <= 0.f) { lo = tb; glo = g; } else { hi = tb; ghi = g; }
        }
      }
    }
    t = (ghi - glo > 0.f) ? lo + (hi - lo) * (-glo) / (ghi - glo) : lo;
  }
  h0 = sphere_box(a + (2.f * hl * t) * axis, r, b, margin);
  h1 = sphere_box(t > 0.5f ? a : e, r, b, margin);
}

__device__ __noinline__ Hit capsule_capsule(f3 p1, f3 u, float r1, float h1, f3 p2, f3 v, float r2, float h2,
                                               float margin) {
  Hit o; o.ok = false; o.dist = 0.f; o.n = mk3(0, 0, 1); o.pos = p1;
  const f3 w = p1 - p2;
  const float b = dot(u, v), dd = dot(u, w), e = dot(v, w), den = 1.f - b * b;
  float s, t;
  if (h1 == 0.f) s = 0.f;
  else if (den > 1e-9f) s = clampf((b * e - dd) / den, -h1, h1);
  else s = 0.f;
  t = h2 == 0.f ? 0.f : b * s + e;
  t = clampf(t, -h2, h2);
  if (h1 != 0.f) s = clampf(b * t - dd, -h1, h1);
  const f3 c1 = p1 + s * u, c2 = p2 + t * v;
  f3 n = c2 - c1;
  const float len = norm(n), dist = len - r1 - r2;
  if (dist >= margin) return o;
  n = len < PP_MINVAL ? mk3(0, 0, 1) : (1.f / len) * n;
  o.n = n; o.pos = c1 + (r1 + 0.5f * dist) * n; o.dist = dist; o.ok = true;
  return o;
}

// ------------------------------------------------------------------ constraint rows
// inequality row (limit / contact): active while x < 0
__device__ __forceinline__ void row_ineq(float D, float x, float& f, float& c) {
  if (x < 0.f) { f = -D * x; c = D; } else { f = 0.f; c = 0.f; }
}
// frictionloss row: quadratic inside |x| < floss/D, linear outside. zone: -1, 0, +1
__device__ __forceinline__ void row_fric(float D, float floss, float x, float& f, float& c, int& zone) {
  const float rf = floss / D;
  if (x <= -rf) { f = floss; c = 0.f; zone = -1; }
  else if (x >= rf) { f = -floss; c = 0.f; zone = 1; }
  else { f = -D * x; c = D; zone = 0; }
}

struct LaneCtx {
  int lane;
  NodeInfo ni;
  bool dof;  // lane < 26
};

// One warp-wide ordered append of a contact candidate (deterministic order: call order, then lane)
__device__ __forceinline__ void push_contact(WarpShared& W, bool ok, int& ncon, int& ndrop, int cap, const Hit& h,
                                             int ids) {
  const unsigned b = __ballot_sync(FULL, ok);
  if (b == 0u) return;
  const int idx = ncon + __popc(b & ((1u << (threadIdx.x & 31)) - 1u));
  if (ok && idx < cap) {
    W.a.k.c_geo[0][idx] = h.pos.x; W.a.k.c_geo[1][idx] = h.pos.y; W.a.k.c_geo[2][idx] = h.pos.z;
    W.a.k.c_geo[3][idx] = h.n.x; W.a.k.c_geo[4][idx] = h.n.y; W.a.k.c_geo[5][idx] = h.n.z;
    W.a.k.c_geo[6][idx] = h.dist;
    W.c_ids[idx] = ids;
  }
  const int tot = ncon + __popc(b);
  if (tot > cap) { ndrop += tot - cap; ncon = cap; } else ncon = tot;
}

// sum over contacts of U[entry of my dof] . wf[c]: J^T applied to per-contact 3-vector weights.
// out[h] for the lane's hand dofs, outs for the lane's key slot.
__device__ __forceinline__ void gather_jt(const WarpShared& W, int ncon, int lane, float out[2], float& outs) {
  out[0] = out[1] = 0.f; outs = 0.f;
  const unsigned lt = (1u << lane) - 1u;
  for (int c = 0; c < ncon; c++) {
    const unsigned m0 = W.c_mask[0][c], m1 = W.c_mask[1][c];
    const float w0 = W.c_wf[0][c], w1 = W.c_wf[1][c], w2 = W.c_wf[2][c];
    const int n0 = __popc(m0);
    if ((m0 >> lane) & 1u) {
      const int e = __popc(m0 & lt) * 3;
      out[0] += W.c_u[e][c] * w0 + W.c_u[e + 1][c] * w1 + W.c_u[e + 2][c] * w2;
    }
    if ((m1 >> lane) & 1u) {
      const int e = (n0 + __popc(m1 & lt)) * 3;
      out[1] += W.c_u[e][c] * w0 + W.c_u[e + 1][c] * w1 + W.c_u[e + 2][c] * w2;
    }
    if (W.c_slot[c] == lane) {
      const int e = (n0 + __popc(m1)) * 3;
      outs += W.c_u[e][c] * w0 + W.c_u[e + 1][c] * w1 + W.c_u[e + 2][c] * w2;
    }
  }
}

// U x for contact `lane`: x lives in W.va (hand dofs 0..51, key slots 52..)
__device__ __forceinline__ f3 contact_ux(const WarpShared& W, int c, int nent) {
  f3 r = mk3(0, 0, 0);
  for (int e = 0; e < nent; e++) {
    const float x = W.va[W.c_col[e][c]];
    r.x += W.c_u[3 * e][c] * x; r.y += W.c_u[3 * e + 1][c] * x; r.z += W.c_u[3 * e + 2][c] * x;
  }
  return r;
}

// ------------------------------------------------------------------ dense fallback
// More than LRC cross-finger contacts: the whole 52x52 hand block is factorised densely in shared
// memory (rows eliminated 51 .. 0, lane l owns rows l and 26+l). Rare, kept out of line.
__device__ __forceinline__ int hidx(int gi, int gk) { return tri(gk) + gi; }  // gi <= gk
__device__ __noinline__ void dense_factor_solve(float* H, int lane, bool dof, float b[2]) {
  float dinv[2] = {1.f, 1.f};
  for (int gj = 51; gj >= 0; gj--) {
    float s[2];
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const int g = h * 26 + lane;
      float acc = 0.f;
      if (dof && g <= gj) {
        acc = H[hidx(g, gj)];
        int base = tri(gj + 1);
        for (int k = gj + 1; k < 52; k++) { acc -= H[base + g] * H[base + gj]; base += k + 1; }
      }
      s[h] = acc;
    }
    const int hj = gj >= 26 ? 1 : 0, lj = gj - 26 * hj;
    const float inv = rsqrtf(fmaxf(__shfl_sync(FULL, hj ? s[1] : s[0], lj), 1e-30f));
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const int g = h * 26 + lane;
      if (dof && g <= gj) H[hidx(g, gj)] = s[h] * inv;
      if (g == gj) dinv[h] = inv;
    }
    __syncwarp();
  }
  for (int gj = 51; gj >= 0; gj--) {
    const int hj = gj >= 26 ? 1 : 0, lj = gj - 26 * hj;
    const float y = __shfl_sync(FULL, (hj ? b[1] : b[0]) * (hj ? dinv[1] : dinv[0]), lj);
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const int g = h * 26 + lane;
      if (g == gj) b[h] = y;
      else if (dof && g < gj) b[h] -= H[hidx(g, gj)] * y;
    }
  }
  for (int gi = 0; gi < 52; gi++) {
    const int hi = gi >= 26 ? 1 : 0, li = gi - 26 * hi;
    const float x = __shfl_sync(FULL, (hi ? b[1] : b[0]) * (hi ? dinv[1] : dinv[0]), li);
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const int g = h * 26 + lane;
      if (g == gi) b[h] = x;
      else if (dof && g > gi) b[h] -= H[hidx(gi, g)] * x;
    }
  }
}

// ------------------------------------------------------------------ OT fingering reward
// _compute_ot_fingering_reward (piano_with_shadow_hands.py:333-369): distances between the 10
// fingertips (left hand first) and the K keys to press, scipy.optimize.linear_sum_assignment
// (rectangular shortest-augmenting-path; transposed when K < 10), mean tolerance over the min(10, K)
// assigned pairs. One warp: column c of the cost matrix lives in lane c % 32, slot c / 32.
// `sm` is scratch shared memory (>= 1300 floats); once per control step, kept out of line.
__device__ __noinline__ float ot_fingering_reward(float* sm, int lane, f3 site, const float qk[3],
                                                  const uint8_t* goal_cur, const KeyTables& kt) {
  float* cost = sm;              // [10][88]
  float* kp = sm + 880;          // [88][3] key surface points of the keys to press
  float* tips = kp + 264;        // [10][3]
  float* u = tips + 30;          // [10] row duals
  int* col4row = reinterpret_cast<int*>(u + 10);  // [10]
  int* inSR = col4row + 10;      // [10]
  const unsigned lt = (1u << lane) - 1u;
  int K = 0;
#pragma unroll
  for (int i = 0; i < 3; i++) {
    const int k = lane + 32 * i;
    const bool on = k < PP_NKEY && goal_cur[k] != 0;
    const unsigned b = __ballot_sync(FULL, on);
    if (on) {
      const int idx = K + __popc(b & lt);
      float sq_, cq_;
      pp_sincos_small(qk[i], &sq_, &cq_);
      const float hx = kt.half[k][0];
      kp[3 * idx] = kt.pos[k][0] - hx + hx * cq_ + 0.35f * hx;
      kp[3 * idx + 1] = kt.pos[k][1];
      kp[3 * idx + 2] = kt.pos[k][2] - hx * sq_ + 0.5f * kt.half[k][2];
    }
    K += __popc(b);
  }
  if (K == 0) return 1.f;
  if (lane < PP_NSITE) { const int r = lane >= 5 ? lane - 5 : lane + 5; st3(tips + 3 * r, site); }
  __syncwarp();
  const bool tr = K < 10;  // rows = keys, columns = fingers
  const int n = tr ? K : 10, m = tr ? 10 : K;
  for (int r = 0; r < n; r++)
#pragma unroll
    for (int i = 0; i < 3; i++) {
      const int c = lane + 32 * i;
      if (c < m) cost[r * 88 + c] = norm(ld3(kp + 3 * (tr ? r : c)) - ld3(tips + 3 * (tr ? c : r)));
    }
  if (lane < 10) { u[lane] = 0.f; col4row[lane] = -1; }
  __syncwarp();
  const float INF = 3.0e38f;
  float v[3] = {0.f, 0.f, 0.f};
  int row4col[3] = {-1, -1, -1};
  for (int cur = 0; cur < n; cur++) {
    float shortest[3] = {INF, INF, INF};
    int path[3] = {-1, -1, -1};
    bool inSC[3] = {false, false, false};
    if (lane < 10) inSR[lane] = 0;
    __syncwarp();
    int i = cur, sink = -1;
    float min_val = 0.f;
    while (sink < 0) {
      if (lane == 0) inSR[i] = 1;
      const float ui = u[i];
      float best = INF; int bslot = 0;
#pragma unroll
      for (int k = 0; k < 3; k++) {
        const int c = lane + 32 * k;
        if (c < m && !inSC[k]) {
          const float r = min_val + cost[i * 88 + c] - ui - v[k];
          if (r < shortest[k]) { shortest[k] = r; path[k] = i; }
          // ties prefer a column without a row (as scipy does)
          if (shortest[k] < best || (shortest[k] == best && row4col[k] < 0)) { best = shortest[k]; bslot = k; }
        }
      }
      float gmin = best;
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) gmin = fminf(gmin, __shfl_xor_sync(FULL, gmin, off));
      if (!(gmin < INF)) return 0.f;  // infeasible: cannot happen with finite distances
      const unsigned who = __ballot_sync(FULL, best == gmin);
      const int wl = __ffs(who) - 1;
      const int jslot = __shfl_sync(FULL, bslot, wl);
      const int jrow = __shfl_sync(FULL, bslot == 0 ? row4col[0] : bslot == 1 ? row4col[1] : row4col[2], wl);
      min_val = gmin;
      if (lane == wl) {
        if (jslot == 0) inSC[0] = true; else if (jslot == 1) inSC[1] = true; else inSC[2] = true;
      }
      if (jrow < 0) sink = wl + 32 * jslot; else i = jrow;
      __syncwarp();
    }
    // dual updates; shortest[] goes through shared memory (kp is dead once the cost matrix is built)
#pragma unroll
    for (int k = 0; k < 3; k++) {
      const int c = lane + 32 * k;
      if (c < m) kp[c] = shortest[k];
    }
    __syncwarp();
    if (lane < n && inSR[lane]) {
      if (lane == cur) u[lane] += min_val;
      else u[lane] += min_val - kp[col4row[lane]];
    }
#pragma unroll
    for (int k = 0; k < 3; k++) if (inSC[k]) v[k] -= min_val - shortest[k];
    __syncwarp();
    // augment along the alternating path
    int j = sink;
    for (;;) {
      const int jl = j & 31, js = j >> 5;
      const int pi = __shfl_sync(FULL, js == 0 ? path[0] : js == 1 ? path[1] : path[2], jl);
      if (lane == jl) { if (js == 0) row4col[0] = pi; else if (js == 1) row4col[1] = pi; else row4col[2] = pi; }
      const int prev = col4row[pi];
      __syncwarp();
      if (lane == 0) col4row[pi] = j;
      __syncwarp();
      j = prev;
      if (pi == cur) break;
    }
  }
  float tot = 0.f;
  if (lane < n) tot = pp_tolerance(cost[lane * 88 + col4row[lane]], 0.01f, 0.1f);
  return wsum(tot) / (float)n;
}

// ------------------------------------------------------------------ kinematics
struct Kin {
  f3 pos[2];   // node origin (world)
  f4 quat[2];  // node orientation
  f3 axw[2];   // joint axis (world)
  f3 ref[2];   // per-hand reference point for spatial quantities
  f3 w[2], vl[2], aa[2], al[2];  // spatial velocity / bias acceleration about ref (DYN only)
};

template <bool DYN>
__device__ __forceinline__ void forward_kinematics(const HandTables& ht, const Options& o, const LaneCtx& L,
                                                   const float q[2], const float v[2], float yoff, Kin& K) {
  const int l = L.dof ? L.lane : 25;
  f3 lpos[2], axl[2]; f4 lquat[2];
#pragma unroll
  for (int h = 0; h < 2; h++) {
    const int hb = h * PP_HB + (l >= 2 ? l - 1 : 0);
    lpos[h] = ld3(ht.body_pos[hb]); lquat[h] = ld4(ht.body_quat[hb]);
    axl[h] = ld3(ht.axis[h * PP_HD + l]);
  }
  // forearm (nodes 0, 1): two slides, constant orientation -> closed form
#pragma unroll
  for (int h = 0; h < 2; h++) {
    const f4 qf = ld4(ht.body_quat[h * PP_HB]);
    const f3 a0 = qrot(qf, ld3(ht.axis[h * PP_HD])), a1 = qrot(qf, ld3(ht.axis[h * PP_HD + 1]));
    const float q0 = __shfl_sync(FULL, q[h], 0), q1 = __shfl_sync(FULL, q[h], 1);
    f3 pf = ld3(ht.body_pos[h * PP_HB]) + q0 * a0 + q1 * a1;
    pf.y += yoff;  // Entity.shift_pose of the attachment frame (randomize_hand_positions)
    K.ref[h] = pf + qrot(qf, mk3(0.f, 0.f, 0.25f));
    K.pos[h] = pf; K.quat[h] = qf; K.axw[h] = l == 0 ? a0 : a1;
    if (DYN) {
      const float v0 = __shfl_sync(FULL, v[h], 0), v1 = __shfl_sync(FULL, v[h], 1);
      K.w[h] = mk3(0, 0, 0); K.aa[h] = mk3(0, 0, 0);
      K.vl[h] = l == 0 ? v0 * a0 : v0 * a0 + v1 * a1;
      K.al[h] = mk3(-o.g[0], -o.g[1], -o.g[2]);
    }
  }
  float sh[2], ch[2];
#pragma unroll
  for (int h = 0; h < 2; h++) pp_sincos_small(0.5f * q[h], &sh[h], &ch[h]);
  const int par = L.ni.parent;
  for (int d = 2; d <= 8; d++) {
    const bool mine = L.dof && L.ni.depth == d;
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const f3 pp = shfl3(K.pos[h], par); const f4 qp = shfl4(K.quat[h], par);
      f3 wp, vlp, aap, alp;
      if (DYN) { wp = shfl3(K.w[h], par); vlp = shfl3(K.vl[h], par); aap = shfl3(K.aa[h], par); alp = shfl3(K.al[h], par); }
      if (mine) {
        const f4 qpre = qmul(qp, lquat[h]);
        const f3 pos = pp + qrot(qp, lpos[h]);
        const f3 axw = qrot(qpre, axl[h]);
        f4 qj; qj.w = ch[h]; qj.x = sh[h] * axl[h].x; qj.y = sh[h] * axl[h].y; qj.z = sh[h] * axl[h].z;
        K.quat[h] = qnormalize(qmul(qpre, qj));
        K.pos[h] = pos; K.axw[h] = axw;
        if (DYN) {
          const f3 sl = cross(pos - K.ref[h], axw);
          const float qd = v[h];
          K.aa[h] = aap + qd * cross(wp, axw);
          K.al[h] = alp + qd * (cross(wp, sl) + cross(vlp, axw));
          K.w[h] = wp + qd * axw;
          K.vl[h] = vlp + qd * sl;
        }
      }
    }
  }
}

}  // namespace

// ====================================================================== the step kernel
// ---- split schedule: an env's solver state between the kernels of a substep (see MODE below)
__device__ __forceinline__ void sp_put(float* r, int& k, float x) { r[(k++) * 32] = x; }
__device__ __forceinline__ void sp_put(float* r, int& k, int x) { r[(k++) * 32] = __int_as_float(x); }
__device__ __forceinline__ void sp_put(float* r, int& k, unsigned x) { r[(k++) * 32] = __uint_as_float(x); }
__device__ __forceinline__ void sp_put(float* r, int& k, bool x) { r[(k++) * 32] = x ? 1.f : 0.f; }
__device__ __forceinline__ void sp_get(const float* r, int& k, float& x) { x = r[(k++) * 32]; }
__device__ __forceinline__ void sp_get(const float* r, int& k, int& x) { x = __float_as_int(r[(k++) * 32]); }
__device__ __forceinline__ void sp_get(const float* r, int& k, unsigned& x) { x = __float_as_uint(r[(k++) * 32]); }
__device__ __forceinline__ void sp_get(const float* r, int& k, bool& x) { x = r[(k++) * 32] != 0.f; }
#define SP_PUT(x) sp_put(r_, k_, x);
#define SP_GET(x) sp_get(r_, k_, x);
// what a Newton trip changes (the last one integrates q and v) ...
#define PP_SP_RW(OP) OP(a[0]) OP(a[1]) OP(Ma[0]) OP(Ma[1]) OP(sMa) OP(sa) OP(jar[0]) OP(jar[1]) OP(jar[2]) OP(jar[3]) \
  OP(zone_old) OP(iter) OP(ls_total) OP(euler) OP(dense_mode) OP(q[0]) OP(q[1]) OP(v[0]) OP(v[1])
// ... and what the substep's prologue hands to the loop and to the finish
#define PP_SP_RO(OP) OP(fs[0]) OP(fs[1]) OP(fsk[0]) OP(fsk[1]) OP(fsk[2]) OP(cD) OP(cmu) OP(fD[0]) OP(fD[1]) \
  OP(faref[0]) OP(faref[1]) OP(floss[0]) OP(floss[1]) OP(lsg[0]) OP(lsg[1]) OP(lD[0]) OP(lD[1]) OP(laref[0]) OP(laref[1]) \
  OP(sM) OP(sfs) OP(ssg) OP(sD) OP(saref) OP(ncon) OP(ndrop) OP(nslot) OP(my_nent) OP(my_slot) OP(nlr) OP(nefc) \
  OP(lrb) OP(slot_lr_mask) OP(my_lr) OP(slot_lr) OP(act_force[0]) OP(act_force[1])
#define PP_SP_REGS() float* r_ = A.sp_regs + (size_t)env * PP_SP_NREG * 32 + lane; int k_ = 0
#define PP_SP_SAVE_ALL() do { PP_SP_REGS(); PP_SP_RW(SP_PUT) PP_SP_RO(SP_PUT) } while (0)
#define PP_SP_SAVE_RW() do { PP_SP_REGS(); PP_SP_RW(SP_PUT) } while (0)
#define PP_SP_LOAD_ALL() do { PP_SP_REGS(); PP_SP_RW(SP_GET) PP_SP_RO(SP_GET) } while (0)
// bytes [b0, b1) of the warp's shared memory to / from the env's image (16-byte words, both ends rounded outwards)
template <bool STORE>
__device__ __forceinline__ void sp_copy_smem(const HandsArgs& A, int env, WarpShared& W, int lane, size_t b0, size_t b1) {
  float4* img = A.sp_smem + (size_t)env * (sizeof(WarpShared) / 16);
  float4* sm = reinterpret_cast<float4*>(&W);
  if (STORE) {
    for (int i = (int)(b0 / 16) + lane; i < (int)((b1 + 15) / 16); i += 32) img[i] = sm[i];
  } else {
    // asynchronous copies (LDGSTS): all in flight at once instead of one global latency per 16 bytes of the lane
    const unsigned sbase = (unsigned)__cvta_generic_to_shared(sm);
    for (int i = (int)(b0 / 16) + lane; i < (int)((b1 + 15) / 16); i += 32)
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sbase + 16u * (unsigned)i), "l"(img + i) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
  }
}
// WPB warps (= envs) per block. With WPB > 1 the warps of a block meet at the phase boundaries
// of every substep, so that they fetch the same stretch of the (instruction-cache bound) code
// together; the meeting is skipped for a block in which any env resets (it runs no substeps).
// MODE: 0 = the whole control step in one launch. The split schedule cuts every substep into three launches so that no
// env waits for the slowest Newton loop of its block while every warp of an SM still runs the same phase of the code:
// PP_SPLIT_PRO = substep `A.sub_index` up to the top of the Newton loop (an env that resets does so here, at substep 0,
// and sits the rest of the control step out), PP_SPLIT_NEWTON = the Newton loop, by warps that take the next env of
// the list whenever they are done with one, PP_SPLIT_FIN = the rest of the substep and, after the last one, the task
// epilogue. Same code, same arithmetic: the three kernels are this function with the other phases compiled out.
constexpr int PP_SPLIT_PRO = 1, PP_SPLIT_NEWTON = 2, PP_SPLIT_FIN = 3;
// diagnostic build -DPP_SP_TIMING (tools/trip_cost_model.py): the Newton kernel leaves the cycles of an env's trips in
// counters 1, 3, 5, 6, which the finish then keeps
#ifdef PP_SP_TIMING
constexpr bool kSpTiming = true;
#else
constexpr bool kSpTiming = false;
#endif
template <int WPB, int MODE>
__device__ __forceinline__ void hands_step_env(const HandsArgs& A, WarpShared& W, const int env) {
  const int lane = threadIdx.x & 31;
#ifdef PP_SP_TIMING
  const long long dbg_entry = clock64();
#endif
  const DevModel& dm = *A.dm;
  const Options& o = dm.opt;
  const HandTables& ht = dm.hand;
  const KeyTables& kt = dm.key;
  const int Lk = A.lookahead, OD = A.obs_dim;
  LaneCtx L; L.lane = lane; L.dof = lane < PP_HD; L.ni = node_info(L.dof ? lane : 25);
  const int l = L.dof ? lane : 25;  // safe table index

  int t_idx = A.t_idx[env];
  const bool first_launch = MODE == 0 || (MODE == PP_SPLIT_PRO && A.sub_index == 0);
  const bool do_reset = A.mode == PP_MODE_RESET ? (A.reset_mask ? A.reset_mask[env] != 0 : true)
                                                : (first_launch && A.auto_reset && A.needs_reset[env]);
  bool sits_out = false;
  if (MODE == PP_SPLIT_PRO && A.sub_index == 0) { if (lane == 0) A.sp_skip[env] = do_reset ? 1 : 0; }
  else if (MODE == PP_SPLIT_PRO || MODE == PP_SPLIT_FIN) sits_out = A.sp_skip[env] != 0;
  bool meet = false;
  if (WPB > 1 && MODE != PP_SPLIT_NEWTON) meet = __syncthreads_and(A.mode != PP_MODE_RESET && !do_reset && !sits_out) != 0;
  if (A.mode == PP_MODE_RESET && !do_reset) return;
  if (sits_out) return;
#define PP_MEET() do { if (WPB > 1 && meet) __syncthreads(); } while (0)

  // ---- state: lane l holds hand dof l of both hands and keys l, l+32, l+64
  float q[2] = {0.f, 0.f}, v[2] = {0.f, 0.f}, acc[2] = {0.f, 0.f};
  float qk[3] = {0.f, 0.f, 0.f}, vk[3] = {0.f, 0.f, 0.f}, ak[3] = {0.f, 0.f, 0.f};
  float ctrl[2] = {0.f, 0.f}, sustain = 0.f, act_force[2] = {0.f, 0.f};
  int cnt_ncon = 0, cnt_nefc = 0, cnt_iter = 0, cnt_drop = 0;
  bool stick[2] = {false, false};  // frictionloss row of the dof ended the last solve in its quadratic zone
  int sum_iter = 0, sum_ls = 0, sum_dense = 0, sum_coupled = 0;
  const size_t sbase = (size_t)env * PP_NV;
  const int mc_ = A.max_contacts & 255;
  const bool dbg_all_lr = (A.max_contacts & 256) != 0;  // test hook: every contact through the low-rank path
  const bool dbg_dense = (A.max_contacts & 512) != 0;   // test hook: every substep through the dense factorisation
  const int cap = mc_ > 0 && mc_ < NC ? mc_ : NC;
  const int ngh = ht.ngeom >> 1;
  // _randomize_initial_hand_positions (piano_with_shadow_hands.py:491-499): one y shift of both hands per episode
  float yoff = 0.f;
  if (A.hand_off) {
    if (do_reset) { yoff = A.hand_off_pending[env]; if (lane == 0) A.hand_off[env] = yoff; }
    else yoff = A.hand_off[env];
  }

  if (!do_reset) {
#pragma unroll
    for (int h = 0; h < 2; h++)
      if (L.dof) {
        q[h] = A.qpos[sbase + PP_NKEY + h * PP_HD + lane];
        v[h] = A.qvel[sbase + PP_NKEY + h * PP_HD + lane];
        acc[h] = A.warm[sbase + PP_NKEY + h * PP_HD + lane];
      }
#pragma unroll
    for (int i = 0; i < 3; i++) {
      const int k = lane + 32 * i;
      if (k < PP_NKEY) { qk[i] = A.qpos[sbase + k]; vk[i] = A.qvel[sbase + k]; ak[i] = A.warm[sbase + k]; }
    }
    const float* act = A.actions + (size_t)env * 45;
#pragma unroll
    for (int h = 0; h < 2; h++)
      if (lane < 22) {
        float a = act[h * 22 + lane];
        if (A.canonical) { const float lo = A.act_min[h * 22 + lane], hi = A.act_max[h * 22 + lane]; a = lo + 0.5f * (a + 1.f) * (hi - lo); }
        ctrl[h] = a;
      }
    if (A.stick_bits) {
#pragma unroll
      for (int h = 0; h < 2; h++) stick[h] = (A.stick_bits[(size_t)env * 2 + h] >> lane) & 1u;
    }
    sustain = act[44];
    if (A.canonical) sustain = 0.5f * (sustain + 1.f);

  }
  // ================================================================ substeps (+ the trailing mj_step1)
  const int nsub = do_reset ? 0 : A.n_substeps;
  if (MODE == PP_SPLIT_FIN && A.sub_index > 0) {  // the sums of the control step's earlier substeps
    const int* cn = A.counters + (size_t)env * 8;
    sum_iter = cn[4]; sum_ls = cn[5]; sum_dense = cn[6]; sum_coupled = cn[7];
  }
  {
    bool resume = MODE == PP_SPLIT_NEWTON || MODE == PP_SPLIT_FIN;  // enter the first substep of the launch at its Newton loop
    for (int sub = MODE == 0 ? 0 : A.sub_index;; sub++) {
      // state that crosses the phases of a substep (split schedule: saved / loaded between the kernels)
      float fs[2], fsk[3], cD, cmu, fD[2], faref[2], floss[2], lsg[2], lD[2], laref[2];
      float sM, sfs, sa, ssg, sD, saref, a[2], Ma[2], sMa, jar[4];
      int ncon, ndrop, nslot, my_nent, my_slot, nlr, nefc, zone_old, iter, ls_total;
      unsigned lrb, slot_lr_mask;
      bool my_lr, dense_mode, slot_lr, euler;
      float (*Ht)[9] = W.a.s.Ht;
      const int posl = lane < 4 ? lane : 4 + lane - L.ni.cs;  // own position in own ancestor list
      // front lane identity: lane = replica*16 + hand*8 + chain
      const int frep = lane >> 4, fh = (lane >> 3) & 1, fc = lane & 7;
      const bool fvalid = fc < 5;
      const int fce = fc == 0 ? 7 : fc == 1 ? 11 : fc == 2 ? 15 : fc == 3 ? 20 : 25;
      const int fn = (fc == 3 || fc == 4) ? 5 : 4;
      const int fcs = fce - fn + 1;
      const int fg0 = fh * 26;

      // out = M x for x in W.va: ancestors from registers, descendants from the pristine copy Mt
      auto matvec_M = [&](float out[2]) {
#pragma unroll
        for (int h = 0; h < 2; h++) {
          float s = 0.f;
          if (L.dof) {
#pragma unroll
            for (int j = 0; j < 9; j++) {
              const int an = j < 4 ? j : L.ni.cs + j - 4;
              const bool valid = j < 4 ? (j <= lane) : (lane >= 4 && an <= lane);
              if (valid) s += W.Mt[h * 26 + lane][j] * W.va[h * 26 + an];
            }
            for (int d = lane + 1; d < lane + L.ni.nsub; d++) s += W.Mt[h * 26 + d][posl] * W.va[h * 26 + d];
          }
          out[h] = s;
        }
      };
      // H_t <- M: tree entries only (the diagonal rows are added when the fronts are loaded)
      auto build_M = [&]() {
        if (L.dof) {
#pragma unroll
          for (int h = 0; h < 2; h++) {
#pragma unroll
            for (int j = 0; j < 9; j++) {
              const int an = j < 4 ? j : L.ni.cs + j - 4;
              const bool valid = j < 4 ? (j <= lane) : (lane >= 4 && an <= lane);
              if (valid) Ht[h * 26 + lane][j] = W.Mt[h * 26 + lane][j];
            }
          }
        }
        __syncwarp();
      };
      const float scale = 1.f / (o.meaninertia * (float)PP_NV);
      const int maxit = min(o.iterations, 50);
      if (!resume) {
      PP_MEET();
      // ---- key state to shared memory
#pragma unroll
      for (int i = 0; i < 3; i++) {
        const int k = lane + 32 * i;
        if (k < PP_NKEY) { W.kq[k] = qk[i]; W.kv[k] = vk[i]; W.kw[k] = ak[i]; W.key_slot[k] = -1; }
      }
      if (A.stale && sub == nsub - 1) {
        // legacy_step=False: the task reads positions, contacts and actuator velocities as the last mj_step's forward pass
        // left them, i.e. at the last substep's start: keep that state for the sensors below
        float* st = A.stale + (size_t)env * 7 * 32 + lane;
#pragma unroll
        for (int h = 0; h < 2; h++) { st[h * 32] = q[h]; st[(2 + h) * 32] = v[h]; }
#pragma unroll
        for (int i = 0; i < 3; i++) st[(4 + i) * 32] = qk[i];
      }
      // ---- mj_kinematics + RNE forward pass
      Kin K;
      forward_kinematics<true>(ht, o, L, q, v, yoff, K);
      f3 Sl[2];
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const bool slide = lane < 2;
        Sl[h] = slide ? K.axw[h] : cross(K.pos[h] - K.ref[h], K.axw[h]);
        if (L.dof) {
          const int g = h * 26 + lane;
          st3(W.a.k.xpos[g], K.pos[h]); st3(W.a.k.xaxis[g], K.axw[h]);
          W.a.k.xquat[g][0] = K.quat[h].w; W.a.k.xquat[g][1] = K.quat[h].x; W.a.k.xquat[g][2] = K.quat[h].y; W.a.k.xquat[g][3] = K.quat[h].z;
          const f3 sa = slide ? mk3(0, 0, 0) : K.axw[h];
          W.a.k.S[g][0] = sa.x; W.a.k.S[g][1] = sa.y; W.a.k.S[g][2] = sa.z; W.a.k.S[g][3] = Sl[h].x; W.a.k.S[g][4] = Sl[h].y; W.a.k.S[g][5] = Sl[h].z;
        }
      }
      if (sub == nsub) break;  // trailing mj_step1: positions at the integrated state only
      // ---- body spatial inertia about ref + RNE body force (node 0 is massless)
#pragma unroll
      for (int h = 0; h < 2; h++) {
        float r16[16];
#pragma unroll
        for (int i = 0; i < 16; i++) r16[i] = 0.f;
        if (L.dof && lane >= 1) {
          const int hb = h * PP_HB + lane - 1;
          const float m = ht.mass[hb];
          const f3 c = K.pos[h] + qrot(K.quat[h], ld3(ht.ipos[hb])) - K.ref[h];
          const f4 qi = qmul(K.quat[h], ld4(ht.iquat[hb]));
          const f3 di = ld3(ht.inertia[hb]);
          // rotation matrix columns of the inertial frame
          const f3 ex = qrot(qi, mk3(1, 0, 0)), ey = qrot(qi, mk3(0, 1, 0)), ez = qrot(qi, mk3(0, 0, 1));
          const float cc = dot(c, c);
          r16[0] = m; r16[1] = m * c.x; r16[2] = m * c.y; r16[3] = m * c.z;
          r16[4] = di.x * ex.x * ex.x + di.y * ey.x * ey.x + di.z * ez.x * ez.x + m * (cc - c.x * c.x);  // xx
          r16[5] = di.x * ex.x * ex.y + di.y * ey.x * ey.y + di.z * ez.x * ez.y - m * c.x * c.y;         // xy
          r16[6] = di.x * ex.x * ex.z + di.y * ey.x * ey.z + di.z * ez.x * ez.z - m * c.x * c.z;         // xz
          r16[7] = di.x * ex.y * ex.y + di.y * ey.y * ey.y + di.z * ez.y * ez.y + m * (cc - c.y * c.y);  // yy
          r16[8] = di.x * ex.y * ex.z + di.y * ey.y * ey.z + di.z * ez.y * ez.z - m * c.y * c.z;         // yz
          r16[9] = di.x * ex.z * ex.z + di.y * ey.z * ey.z + di.z * ez.z * ez.z + m * (cc - c.z * c.z);  // zz
          // f = I a_eff + v x* (I v), through the centre of mass (no cancellation)
          const float gc = ht.gravcomp[hb];
          const f3 ale = K.al[h] + gc * mk3(o.g[0], o.g[1], o.g[2]);
          auto icw = [&](f3 x) {  // R diag(I) R^T x
            return (di.x * dot(ex, x)) * ex + (di.y * dot(ey, x)) * ey + (di.z * dot(ez, x)) * ez;
          };
          const f3 f1 = m * (ale + cross(K.aa[h], c));
          const f3 n1 = icw(K.aa[h]) + cross(c, f1);
          const f3 fh = m * (K.vl[h] + cross(K.w[h], c));
          const f3 nh = icw(K.w[h]) + cross(c, fh);
          const f3 nn = n1 + cross(K.w[h], nh) + cross(K.vl[h], fh);
          const f3 ff = f1 + cross(K.w[h], fh);
          r16[10] = nn.x; r16[11] = nn.y; r16[12] = nn.z; r16[13] = ff.x; r16[14] = ff.y; r16[15] = ff.z;
        }
        if (L.dof) {
#pragma unroll
          for (int i = 0; i < 16; i++) W.a.k.raw[h * 26 + lane][i] = r16[i];
        }
      }
      __syncwarp();
      // ---- composite (subtree) sums: chains first, then the 4 root nodes
      float comp[2][16];
#pragma unroll
      for (int h = 0; h < 2; h++) {
#pragma unroll
        for (int i = 0; i < 16; i++) comp[h][i] = 0.f;
        if (L.dof && lane >= 4) {
          for (int k = 0; k < L.ni.nsub; k++)
#pragma unroll
            for (int i = 0; i < 16; i++) comp[h][i] += W.a.k.raw[h * 26 + lane + k][i];
          if (lane == L.ni.cs) {
            const int ci = h * 5 + (lane == 4 ? 0 : lane == 8 ? 1 : lane == 12 ? 2 : lane == 16 ? 3 : 4);
#pragma unroll
            for (int i = 0; i < 16; i++) W.a.k.comp[ci][i] = comp[h][i];
          }
        }
      }
      __syncwarp();
#pragma unroll
      for (int h = 0; h < 2; h++) {
        if (lane < 4) {
          for (int k = lane; k < 4; k++)
#pragma unroll
            for (int i = 0; i < 16; i++) comp[h][i] += W.a.k.raw[h * 26 + k][i];
          for (int k = 0; k < 5; k++)
#pragma unroll
            for (int i = 0; i < 16; i++) comp[h][i] += W.a.k.comp[h * 5 + k][i];
        }
      }
      __syncwarp();  // u.dyn is dead from here on (H aliases it)
      // ---- CRBA row (ancestors of the dof) and RNE bias
      float bias[2];  // M(ancestor j, dof) goes to shared memory (Mt) only: 18 registers less through the solver
#pragma unroll
      for (int h = 0; h < 2; h++) {
        const float* cI = comp[h];
        const f3 sa = lane < 2 ? mk3(0, 0, 0) : K.axw[h], sl = Sl[h];
        const f3 mc = mk3(cI[1], cI[2], cI[3]);
        const f3 Is = mk3(cI[4] * sa.x + cI[5] * sa.y + cI[6] * sa.z, cI[5] * sa.x + cI[7] * sa.y + cI[8] * sa.z,
                          cI[6] * sa.x + cI[8] * sa.y + cI[9] * sa.z);
        const f3 Fn = Is + cross(mc, sl);
        const f3 Ff = cI[0] * sl - cross(mc, sa);
        bias[h] = sa.x * cI[10] + sa.y * cI[11] + sa.z * cI[12] + sl.x * cI[13] + sl.y * cI[14] + sl.z * cI[15];
#pragma unroll
        for (int j = 0; j < 9; j++) {
          const int a = j < 4 ? j : L.ni.cs + j - 4;
          const bool valid = L.dof && (j < 4 ? (j <= lane) : (lane >= 4 && a <= lane));
          float mv = 0.f;
          if (valid) {
            const float* Sa = W.a.k.S[h * 26 + a];
            mv = Sa[0] * Fn.x + Sa[1] * Fn.y + Sa[2] * Fn.z + Sa[3] * Ff.x + Sa[4] * Ff.y + Sa[5] * Ff.z;
            if (a == lane) mv += ht.armature[h * PP_HD + l];
          }
          if (L.dof) W.Mt[h * 26 + lane][j] = mv;
        }
      }
      // ---- mj_fwdActuation (position servos on joints / 2-joint fixed tendons) + qfrc_smooth
#pragma unroll
      for (int h = 0; h < 2; h++) {
        float f = 0.f;
        const int ai = h * 22 + (lane < 22 ? lane : 0);
        const int t0 = ht.act_trn0[ai] - h * 26, t1 = ht.act_trn1[ai] - h * 26;
        float len = __shfl_sync(FULL, q[h], t0);
        const float q1 = __shfl_sync(FULL, q[h], t1 >= 0 ? t1 : 0);
        if (ht.act_trn1[ai] >= 0) len += q1;
        if (lane < 22) {
          const float c = clampf(ctrl[h], ht.act_ctrlrange[ai][0], ht.act_ctrlrange[ai][1]);
          f = ht.act_kp[ai] * c - ht.act_kp[ai] * len;
          if (ht.act_forcelimited[ai]) f = clampf(f, ht.act_forcerange[ai][0], ht.act_forcerange[ai][1]);
          act_force[h] = f;
        }
        const float fa = __shfl_sync(FULL, f, ht.dof_act[h * PP_HD + l]);
        const float damp_h = ht.damping[h * PP_HD + l];
        const float applied = (A.qfrc_applied && L.dof) ? A.qfrc_applied[sbase + PP_NKEY + h * PP_HD + lane] : 0.f;
        fs[h] = L.dof ? (-damp_h * v[h] - bias[h] + fa + applied) : 0.f;
      }
#pragma unroll
      for (int i = 0; i < 3; i++) {
        const int k = lane + 32 * i;
        fsk[i] = 0.f;
        if (k < PP_NKEY) {
          const float applied = A.qfrc_applied ? A.qfrc_applied[sbase + k] : 0.f;
          fsk[i] = pp_key_smooth(kt, k, qk[i], vk[i], applied);
        }
      }

      PP_MEET();
      // ================================================================ mj_collision
      ncon = 0; ndrop = 0;
      {
        f3 gc[2], gax[2]; float gr = 0.f, ghl = 0.f; int gnode = 0; bool finger = false;
        const bool hasg = lane < ngh;
        if (hasg) { gr = ht.geom_r[lane]; ghl = ht.geom_hl[lane]; gnode = ht.geom_body[lane] + 1; if (gnode == 1) gnode = 1; finger = ht.geom_cls[lane] == PP_CLS_RH_FINGER; }
        float ylo[2], yhi[2];
#pragma unroll
        for (int h = 0; h < 2; h++) {
          gc[h] = mk3(0, 0, 1); gax[h] = mk3(0, 0, 1); ylo[h] = 1e9f; yhi[h] = -1e9f;
          if (hasg) {
            const int g = h * 26 + gnode;
            f4 qq; qq.w = W.a.k.xquat[g][0]; qq.x = W.a.k.xquat[g][1]; qq.y = W.a.k.xquat[g][2]; qq.z = W.a.k.xquat[g][3];
            gc[h] = ld3(W.a.k.xpos[g]) + qrot(qq, ld3(ht.geom_pos[h * ngh + lane]));
            gax[h] = qrot(qq, mk3(0, 0, 1));
            ylo[h] = gc[h].y - gr - ghl; yhi[h] = gc[h].y + gr + ghl;
            st3(W.a.k.gpos[h * ngh + lane], gc[h]); st3(W.a.k.gaxis[h * ngh + lane], gax[h]);
          }
        }
        // -- finger colliders against the keys under them and the base
        int i_lo[2], nW[2], ntot[2];
        const float Rb = gr + ghl;
#pragma unroll
        for (int h = 0; h < 2; h++) {
          i_lo[h] = 0; nW[h] = 0; ntot[h] = 0;
          if (hasg && finger && gc[h].z - Rb < ZCULL) {
            int a = (int)floorf((gc[h].y - Rb - kt.y0) / kt.pitch), b = (int)floorf((gc[h].y + Rb - kt.y0) / kt.pitch);
            if (b >= 0 && a <= 51) {
              a = max(a, 0); b = min(b, 51);
              i_lo[h] = a; nW[h] = b - a + 1; ntot[h] = 2 * nW[h] + 2;  // whites, boundaries a..b+1, base
            }
          }
        }
        int tmax = max(ntot[0], ntot[1]);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) tmax = max(tmax, __shfl_xor_sync(FULL, tmax, off));
        const float margin_p = o.pair_margin[PP_PAIRCLS_PIANO];
        // candidate t of the lane's collider: whites i_lo.., the black key on every boundary, the base
        auto candidate = [&](int h, int t) -> int {
          if (t < nW[h]) return kt.white_key[i_lo[h] + t];
          if (t < 2 * nW[h] + 1) return kt.boundary_black[i_lo[h] + t - nW[h]];
          return (t == 2 * nW[h] + 1 && ntot[h] > 0) ? PP_NKEY : -1;
        };
        // conservative cull: bounding sphere of the collider against a box that contains the key at every
        // angle of its range (the far end sinks by <= 2 hx sin(q_max) < 16 mm); the exact test only runs for
        // the survivors, compacted so that the warp iterates max-survivors-per-lane times
        unsigned surv[2] = {0u, 0u};
        const float Rc = Rb + fmaxf(margin_p, 0.f) + 1e-4f;
        for (int t = 0; t < tmax; t++) {
#pragma unroll
          for (int h = 0; h < 2; h++) {
            const int key = candidate(h, t);
            if (key < 0) continue;
            f3 bp, bh;
            if (key < PP_NKEY) {
              bp = ld3(kt.pos[key]); bh = ld3(kt.half[key]);
              bp.z -= 0.007f; bh.z += 0.010f; bh.x += 0.002f;
            } else { bp = ld3(o.base_pos); bh = ld3(o.base_half); }
            const float dx = fmaxf(fabsf(gc[h].x - bp.x) - bh.x, 0.f), dy = fmaxf(fabsf(gc[h].y - bp.y) - bh.y, 0.f);
            const float dz = fmaxf(fabsf(gc[h].z - bp.z) - bh.z, 0.f);
            if (dx * dx + dy * dy + dz * dz < Rc * Rc) surv[h] |= 1u << t;
          }
        }
#pragma unroll
        for (int h = 0; h < 2; h++) {
          while (__any_sync(FULL, surv[h] != 0u)) {
            int key = -1;
            if (surv[h]) { const int t = __ffs(surv[h]) - 1; surv[h] &= surv[h] - 1u; key = candidate(h, t); }
            Hit h0, h1; h0.ok = false; h1.ok = false;
            if (key >= 0) {
              Box b;
              if (key < PP_NKEY) {
                float sq_, cq_;
                pp_sincos_small(W.kq[key], &sq_, &cq_);
                const float hx = kt.half[key][0];
                b.p = mk3(kt.pos[key][0] - hx + hx * cq_, kt.pos[key][1], kt.pos[key][2] - hx * sq_);
                b.cq = cq_; b.sq = sq_; b.h = ld3(kt.half[key]);
              } else {
                b.p = ld3(o.base_pos); b.cq = 1.f; b.sq = 0.f; b.h = ld3(o.base_half);
              }
              capsule_box(gc[h], gax[h], gr, ghl, b, margin_p, h0, h1);
            }
            const int ids = h | (gnode << 1) | (0 << 6) | (key << 8);
            push_contact(W, h0.ok, ncon, ndrop, cap, h0, ids);
            push_contact(W, h1.ok, ncon, ndrop, cap, h1, ids);
          }
        }
        // -- fixed hand-hand pairs; the rh x lh block is skipped while the hands' y extents are apart
        float ymin0 = ylo[0], ymax0 = yhi[0], ymin1 = ylo[1], ymax1 = yhi[1];
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
          ymin0 = fminf(ymin0, __shfl_xor_sync(FULL, ymin0, off)); ymax0 = fmaxf(ymax0, __shfl_xor_sync(FULL, ymax0, off));
          ymin1 = fminf(ymin1, __shfl_xor_sync(FULL, ymin1, off)); ymax1 = fmaxf(ymax1, __shfl_xor_sync(FULL, ymax1, off));
        }
        const bool apart = ymin0 > ymax1 || ymin1 > ymax0;
        const int np = apart ? ht.npair_same : ht.npair;
        __syncwarp();  // gpos / gaxis of every collider are in shared memory
        const float margin_h = o.pair_margin[PP_PAIRCLS_HAND];
        // bounding-sphere pass over the lane's pairs (p = lane + 32 i), then the exact capsule-capsule test
        // on the survivors only, compacted: the warp iterates max-survivors-per-lane times
        unsigned near = 0u;
        for (int i = 0; i * 32 < np; i++) {
          const int p = i * 32 + lane;
          if (p < np) {
            const int g1 = ht.pair_g1[p], g2 = ht.pair_g2[p];
            const f3 dc = ld3(W.a.k.gpos[g1]) - ld3(W.a.k.gpos[g2]);
            const float reach = ht.geom_r[g1] + ht.geom_hl[g1] + ht.geom_r[g2] + ht.geom_hl[g2] + margin_h;
            if (dot(dc, dc) <= reach * reach) near |= 1u << i;
          }
        }
        while (__any_sync(FULL, near != 0u)) {
          Hit hh; hh.ok = false; int ids = 0;
          if (near) {
            const int i = __ffs(near) - 1; near &= near - 1u;
            const int p = i * 32 + lane;
            const int g1 = ht.pair_g1[p], g2 = ht.pair_g2[p];
            const int h1 = g1 >= ngh ? 1 : 0, h2 = g2 >= ngh ? 1 : 0;
            const int n1 = ht.geom_body[g1] - h1 * PP_HB + 1, n2 = ht.geom_body[g2] - h2 * PP_HB + 1;
            hh = capsule_capsule(ld3(W.a.k.gpos[g1]), ld3(W.a.k.gaxis[g1]), ht.geom_r[g1], ht.geom_hl[g1],
                                 ld3(W.a.k.gpos[g2]), ld3(W.a.k.gaxis[g2]), ht.geom_r[g2], ht.geom_hl[g2], margin_h);
            ids = h1 | (n1 << 1) | (1 << 6) | ((h2 | (n2 << 1)) << 8);
          }
          push_contact(W, hh.ok, ncon, ndrop, cap, hh, ids);
        }
      }
      __syncwarp();

      // ================================================================ mj_makeConstraint
      // velocities into the exchange vector (for aref)
#pragma unroll
      for (int h = 0; h < 2; h++) if (L.dof) W.va[h * 26 + lane] = v[h];
      // key slots: contacts on the same key share one slot
      nslot = 0;
      my_nent = 0; cD = 0.f; float caref[4] = {0.f, 0.f, 0.f, 0.f}; my_slot = -1;
      unsigned cm[2] = {0u, 0u};
      int ids = 0, kind = 0, key = -1; nlr = 0;
      lrb = 0u;
      my_lr = false; dense_mode = false;
      if (lane < ncon) { ids = W.c_ids[lane]; kind = (ids >> 6) & 1; key = kind == 0 ? (ids >> 8) : -1; }
      {
        const bool wants = lane < ncon && kind == 0 && key < PP_NKEY && key >= 0;
        const unsigned same = __match_any_sync(FULL, wants ? key : -1 - lane);
        const bool leader = wants && (__ffs(same) - 1) == lane;
        const unsigned lead = __ballot_sync(FULL, leader);
        nslot = __popc(lead);
        if (wants) {
          const int s = __popc(lead & ((1u << (__ffs(same) - 1)) - 1u));
          my_slot = s;
          if (leader && s < KS) { W.slot_key[s] = (int8_t)key; W.key_slot[key] = (int8_t)s; }
        }
        if (nslot > KS) {  // out of slots: the contacts on the overflowing keys are dropped
          const bool dropme = my_slot >= KS;
          ndrop += __popc(__ballot_sync(FULL, dropme));
          if (dropme) { my_slot = -2; }
          nslot = KS;
        }
        // A contact is "in front" when its jacobian lives on one root path (one finger chain + the
        // 4 root dofs) plus a key owned by that finger. Hand-hand contacts and contacts on a key
        // already owned by another finger couple two fronts: they are solved as a low-rank
        // correction while the ranks of their active rows sum to LRR at most (else that trip factorises densely).
        const int n1_ = (ids >> 1) & 31;
        const int my_front = (ids & 1) * 8 + (n1_ >= 21 ? 4 : n1_ >= 16 ? 3 : n1_ >= 12 ? 2 : n1_ >= 8 ? 1 : 0);
        const int owner = __shfl_sync(FULL, my_front, __ffs(same) - 1);
        my_lr = lane < ncon && my_slot != -2 && (kind == 1 || dbg_all_lr || (wants && owner != my_front));
        lrb = __ballot_sync(FULL, my_lr);
        nlr = __popc(lrb);
        dense_mode = false;  // set by the trips that need the dense factorisation (rank of the low-rank part > LRR)
      }
      __syncwarp();
      if (lane < ncon && my_slot != -2) {
        const int c = lane;
        const f3 pos = mk3(W.a.k.c_geo[0][c], W.a.k.c_geo[1][c], W.a.k.c_geo[2][c]);
        const f3 n = mk3(W.a.k.c_geo[3][c], W.a.k.c_geo[4][c], W.a.k.c_geo[5][c]);
        const float dist = W.a.k.c_geo[6][c];
        // mju_makeFrame
        f3 t1 = (n.y < 0.5f && n.y > -0.5f) ? mk3(0, 1, 0) : mk3(0, 0, 1);
        t1 = t1 - dot(n, t1) * n; t1 = (1.f / norm(t1)) * t1;
        const f3 t2 = cross(n, t1);
        const int h1 = ids & 1, n1 = (ids >> 1) & 31;
        int h2 = 0, n2 = 0;
        const int pc = kind == 0 ? PP_PAIRCLS_PIANO : PP_PAIRCLS_HAND;
        float tran = ht.binvw[h1 * PP_HB + n1 - 1];
        unsigned a1 = anc_of(n1), a2 = 0u;
        if (kind == 1) {
          h2 = (ids >> 8) & 1; n2 = (ids >> 9) & 31;
          tran += ht.binvw[h2 * PP_HB + n2 - 1];
          a2 = anc_of(n2);
          if (h1 == h2) { const unsigned sh = a1 & a2; a1 &= ~sh; a2 &= ~sh; cm[h1] = a1 | a2; }
          else { cm[h1] = a1; cm[h2] = a2; }
        } else {
          if (key < PP_NKEY) tran += kt.binvw[key];
          cm[h1] = a1;
        }
        // entries: rh dofs ascending, lh dofs ascending, key slot
        int e = 0; f3 uv = mk3(0, 0, 0);
#pragma unroll
        for (int h = 0; h < 2; h++) {
          unsigned m = cm[h];
          while (m) {
            const int d = __ffs(m) - 1; m &= m - 1;
            const int g = h * 26 + d;
            const f3 ax = ld3(W.a.k.xaxis[g]);
            const f3 col = d < 2 ? ax : cross(ax, pos - ld3(W.a.k.xpos[g]));
            // jac(body2) - jac(body1): chain of geom1 enters with -, chain of geom2 with +
            const bool second = kind == 1 && h == h2 && ((a2 >> d) & 1u) && !(h1 == h2 && ((a1 >> d) & 1u));
            const float sg = second ? 1.f : -1.f;
            const f3 u = mk3(sg * dot(n, col), sg * dot(t1, col), sg * dot(t2, col));
            W.c_u[3 * e][c] = u.x; W.c_u[3 * e + 1][c] = u.y; W.c_u[3 * e + 2][c] = u.z;
            W.c_col[e][c] = (uint8_t)g;
            const float vg = W.va[g];
            uv = uv + vg * u;
            e++;
          }
        }
        if (my_slot >= 0) {
          const float hx = kt.half[key][0];
          const f3 anchor = mk3(kt.pos[key][0] - hx, kt.pos[key][1], kt.pos[key][2]);
          const f3 col = cross(mk3(0, 1, 0), pos - anchor);
          const f3 u = mk3(dot(n, col), dot(t1, col), dot(t2, col));
          W.c_u[3 * e][c] = u.x; W.c_u[3 * e + 1][c] = u.y; W.c_u[3 * e + 2][c] = u.z;
          W.c_col[e][c] = (uint8_t)(52 + my_slot);
          uv = uv + W.kv[key] * u;
          e++;
        }
        my_nent = e;
        W.c_mask[0][c] = cm[0]; W.c_mask[1][c] = cm[1]; W.c_slot[c] = (int8_t)my_slot; W.c_nent[c] = (int8_t)e;
        W.c_lr[c] = my_lr ? 1 : 0;
        // impedance, regularisation, reference acceleration (pyramidal, condim 3)
        const float mu = o.pair_friction[pc], margin = o.pair_margin[pc];
        float Kk, Bb;
        pp_kb(o.pair_solref[pc], o.pair_solimp[pc], o.h, Kk, Bb);
        const float imp = pp_impedance(o.pair_solimp[pc], dist, margin);
        const float diag = tran + mu * mu * tran;
        const float R0 = fmaxf(PP_MINVAL, (1.f - imp) * diag / imp);
        const float R = 2.f * mu * mu / o.impratio * R0;
        cD = 1.f / R;
        const float vr[4] = {uv.x + mu * uv.y, uv.x - mu * uv.y, uv.x + mu * uv.z, uv.x - mu * uv.z};
#pragma unroll
        for (int k = 0; k < 4; k++) caref[k] = -Bb * vr[k] - Kk * imp * (dist - margin);
      } else if (lane < ncon) {
        W.c_mask[0][lane] = 0u; W.c_mask[1][lane] = 0u; W.c_slot[lane] = -1; W.c_nent[lane] = 0; W.c_lr[lane] = 0;
      }
      cmu = (lane < ncon) ? o.pair_friction[kind == 0 ? PP_PAIRCLS_PIANO : PP_PAIRCLS_HAND] : 0.f;
      __syncwarp();
      // slot bookkeeping: the root path (one finger of one hand) that owns each touched key
      slot_lr = false;  // a low-rank contact touches this key: its slot entry is needed in every solved vector
      if (lane < nslot) {
        unsigned u0 = 0u, u1 = 0u;
        for (int c = 0; c < ncon; c++)
          if (W.c_slot[c] == lane) {
            if (W.c_lr[c]) slot_lr = true;
            else { u0 |= W.c_mask[0][c]; u1 |= W.c_mask[1][c]; }
          }
        W.slot_mask[lane] = u0 | u1;  // an in-front contact touches one hand only
        W.slot_hand[lane] = u1 ? 1 : 0;
      }
      slot_lr_mask = __ballot_sync(FULL, slot_lr);

      // per-dof rows: frictionloss (always) and joint limit (when violated)
      nefc = 4 * ncon;
      {
        float Kf, Bf;
        pp_kb(o.friction_solref, o.friction_solimp, o.h, Kf, Bf);
        const float impf = pp_impedance(o.friction_solimp, 0.f, 0.f);
#pragma unroll
        for (int h = 0; h < 2; h++) {
          const int hd = h * PP_HD + l;
          floss[h] = L.dof ? ht.frictionloss[hd] : 0.f;
          const float invw = ht.invw[hd];
          fD[h] = 1.f / fmaxf(PP_MINVAL, (1.f - impf) * invw / impf);
          faref[h] = -Bf * v[h];
          LimitRow r; r.sgn = 0.f; r.D = 0.f; r.aref = 0.f;
          if (L.dof) r = pp_limit_row(o, q[h], v[h], ht.range[hd][0], ht.range[hd][1], invw);
          lsg[h] = r.sgn; lD[h] = r.D; laref[h] = r.aref;
          nefc += __popc(__ballot_sync(FULL, floss[h] > 0.f)) + __popc(__ballot_sync(FULL, r.sgn != 0.f));
        }
      }
      // key-slot scalars (lane s < nslot)
      sM = 1.f; sfs = 0.f; sa = 0.f; ssg = 0.f; sD = 0.f; saref = 0.f; int skey = 0;
      if (lane < nslot) {
        skey = W.slot_key[lane];
        const float kq_ = W.kq[skey], kv_ = W.kv[skey];
        const float applied = A.qfrc_applied ? A.qfrc_applied[sbase + skey] : 0.f;
        sM = kt.M[skey]; sfs = pp_key_smooth(kt, skey, kq_, kv_, applied); sa = W.kw[skey];
        const LimitRow r = pp_limit_row(o, kq_, kv_, 0.f, kt.hi[skey], kt.invw[skey]);
        ssg = r.sgn; sD = r.D; saref = r.aref;
      }
      __syncwarp();  // a.kin is dead from here on (a.sol aliases it)

      PP_MEET();
      // ================================================================ mj_fwdConstraint + mj_Euler
      // Newton on the primal cost; every linear solve goes through the multifrontal factorisation of
      // H_t = M + diagonal rows + in-front contact cliques (one lane per finger-chain front, the
      // 4x4 root front replicated), cross-front contacts enter as a Woodbury correction. The last
      // trip of the loop reuses the same code for mj_Euler's (M + h diag(b)) solve.
      // warm start: the previous substep's acceleration; a dof whose frictionloss row ended in its quadratic
      // zone ("sticking") starts at the row's new reference acceleration instead, which keeps it there.
      // (The minimiser is unique, so the start only changes the number of Newton iterations: -9 % on the
      // bench workload, measured on the fp64 restatement.)
#pragma unroll
      for (int h = 0; h < 2; h++) a[h] = (stick[h] && floss[h] > 0.f) ? faref[h] : acc[h];
      Ma[0] = Ma[1] = 0.f; sMa = 0.f;
      jar[0] = jar[1] = jar[2] = jar[3] = 0.f;
      zone_old = -1;
      // Ma at the warm start
#pragma unroll
      for (int h = 0; h < 2; h++) if (L.dof) W.va[h * 26 + lane] = a[h];
      __syncwarp();
      matvec_M(Ma);
      if (lane < nslot) { W.va[52 + lane] = sa; sMa = sM * sa; }
      __syncwarp();
      if (lane < ncon && my_slot != -2) {
        const f3 ua = contact_ux(W, lane, my_nent);
        jar[0] = ua.x + cmu * ua.y - caref[0]; jar[1] = ua.x - cmu * ua.y - caref[1];
        jar[2] = ua.x + cmu * ua.z - caref[2]; jar[3] = ua.x - cmu * ua.z - caref[3];
      }
      iter = 0; ls_total = 0;
      euler = false;
      if (MODE == PP_SPLIT_PRO || MODE == PP_SPLIT_FIN) {  // the env's solver state waits for the Newton kernel
        PP_SP_SAVE_ALL();
        __syncwarp();
        sp_copy_smem<true>(A, env, W, lane, offsetof(WarpShared, kq), sizeof(WarpShared));
        if (lane == 0) {
          // lists by the cost class of the env's trips: every cross-front contact adds 3 right-hand sides to each solve
          const int k = PP_SP_BUCKETS - 1 - min(nlr, PP_SP_BUCKETS - 1);
          A.sp_list_out[(size_t)k * A.num_envs + atomicAdd(A.sp_count_out + k, 1)] = env;
        }
        return;
      }
      } else {
        PP_SP_LOAD_ALL();
        // the Newton loop needs everything behind the phase union, the finish only the key state and the key -> slot map
        sp_copy_smem<false>(A, env, W, lane, offsetof(WarpShared, kq),
                            MODE == PP_SPLIT_FIN ? offsetof(WarpShared, slot_key) : sizeof(WarpShared));
        __syncwarp();
      }
      const bool con_on = lane < ncon && my_slot != -2;
#ifndef PP_TRIP_MEET
#define PP_TRIP_MEET 1
#endif
      // The envs of a block also meet at the top of every trip (a finished env keeps arriving until all are):
      // with different actions per env the warps of an SM would otherwise spread over the whole loop body and
      // miss the instruction cache on most fetches (ncu, random actions: 51 % hit rate, 57 % of the stall
      // samples "no instruction"; with the meeting the warps fetch the same stretch of code together).
      bool fin = false;
#ifdef PP_SP_TIMING
      long long dbg_t0 = 0, dbg_cyc = 0, dbg_first = 0, dbg_refill = 0; int dbg_redo = 0, dbg_trips = 0;
#endif
      if (MODE != PP_SPLIT_FIN)
      for (;;) {
        // Newton kernel: the block's warps start every trip together (see the kernel below), whatever env they are on
#ifdef PP_SP_TIMING  // diagnostic build (tools/trip_cost_model.py): cycles the env's trips take, without the meetings
        if (MODE == PP_SPLIT_NEWTON) {
          const long long now = clock64();
          if (dbg_t0) { dbg_cyc += now - dbg_t0; if (dbg_trips == 1) dbg_first = now - dbg_t0; }
          else dbg_refill = now - dbg_entry;
        }
#endif
        if (MODE == PP_SPLIT_NEWTON && WPB > 1) __syncthreads_and(0);
#ifdef PP_SP_TIMING
        if (MODE == PP_SPLIT_NEWTON) { dbg_t0 = clock64(); dbg_trips++; }
#endif
        if (PP_TRIP_MEET && WPB > 1 && meet) {
          if (__syncthreads_and(fin)) break;
          if (fin) continue;
        }
        // ---- row forces and curvatures at the current point
        float ff[2] = {0.f, 0.f}, fc_[2] = {0.f, 0.f}, lf[2] = {0.f, 0.f}, lc[2] = {0.f, 0.f};
        int fz[2] = {0, 0};
        float cf[4] = {0.f, 0.f, 0.f, 0.f}, cc[4] = {0.f, 0.f, 0.f, 0.f};
        float sf = 0.f, sc = 0.f;
        float grad[2] = {0.f, 0.f}, sgrad = 0.f;
        if (!euler) {
#pragma unroll
          for (int h = 0; h < 2; h++) {
            if (floss[h] > 0.f) row_fric(fD[h], floss[h], a[h] - faref[h], ff[h], fc_[h], fz[h]);
            if (lsg[h] != 0.f) row_ineq(lD[h], lsg[h] * a[h] - laref[h], lf[h], lc[h]);
          }
          if (ssg != 0.f) row_ineq(sD, ssg * sa - saref, sf, sc);
          if (con_on) {
#pragma unroll
            for (int k = 0; k < 4; k++) row_ineq(cD, jar[k], cf[k], cc[k]);
            const int c = lane;
            W.c_wf[0][c] = cf[0] + cf[1] + cf[2] + cf[3];
            W.c_wf[1][c] = cmu * (cf[0] - cf[1]);
            W.c_wf[2][c] = cmu * (cf[2] - cf[3]);
            W.c_w[0][c] = cc[0] + cc[1] + cc[2] + cc[3];
            W.c_w[1][c] = cmu * (cc[0] - cc[1]);
            W.c_w[2][c] = cmu * cmu * (cc[0] + cc[1]);
            W.c_w[3][c] = cmu * (cc[2] - cc[3]);
            W.c_w[4][c] = cmu * cmu * (cc[2] + cc[3]);
          } else if (lane < ncon) {
            for (int k = 0; k < 3; k++) W.c_wf[k][lane] = 0.f;
            for (int k = 0; k < 5; k++) W.c_w[k][lane] = 0.f;
          }
          __syncwarp();
          // ---- gradient
          float gj[2], gjs;
          gather_jt(W, ncon, lane, gj, gjs);
#pragma unroll
          for (int h = 0; h < 2; h++) grad[h] = L.dof ? (Ma[h] - fs[h] - ff[h] - lsg[h] * lf[h] - gj[h]) : 0.f;
          if (lane < nslot) sgrad = sMa - sfs - ssg * sf - gjs;
        }
        // The multifrontal + Woodbury path takes any number of cross-front contacts as long as the ranks of their active
        // rows' weights (W = V V^T per contact, rank <= 3; 0 when no row is active) sum to LRR at most - decided trip by
        // trip, because a dense trip costs three times a low-rank one of full rank (profiles/r02_trip_cost_model.txt).
        // Only counted here, and only when there are more than LRC such contacts; V itself is built with the right-hand sides.
        int m = 0;
        bool dense_now = dbg_dense && !euler;
        if (!euler && nlr > LRC) {
          int rk = 0;
          if (my_lr) {
            const float w0 = W.c_w[0][lane], w1 = W.c_w[1][lane], w2 = W.c_w[2][lane], w3 = W.c_w[3][lane], w4 = W.c_w[4][lane];
            if (w0 > 0.f) {
              const float l00 = sqrtf(w0), l10 = w1 / l00, l20 = w3 / l00;
              const float d1 = w2 - l10 * l10;
              const float l11 = d1 > 1e-6f * w2 ? sqrtf(d1) : 0.f;
              const float l21 = l11 > 0.f ? -l20 * l10 / l11 : 0.f;
              const float d2 = w4 - l20 * l20 - l21 * l21;
              rk = 1 + (l11 != 0.f ? 1 : 0) + (d2 > 1e-6f * w4 ? 1 : 0);
            }
          }
          dense_now |= (int)wsum((float)rk) > LRR;
        }
        dense_mode |= dense_now;  // diagnostic (counters[6]): the substep took the dense path at least once
        float p[2] = {0.f, 0.f}, sp = 0.f;
        if (dense_now) {
          // ---- dense fallback: full Hessian of the 52 hand dofs, every contact as a clique
          zone_old = -1;  // overwrites the slots' 1 / H_ss: the next multifrontal trip assembles H_t again
          float* Hd = W.a.s.Hd;
          for (int i = lane; i < 52 * 53 / 2; i += 32) Hd[i] = 0.f;
          for (int s = 0; s < nslot; s++) { W.a.s.Kcd[s][lane] = 0.f; if (lane + 32 < 52) W.a.s.Kcd[s][lane + 32] = 0.f; }
          __syncwarp();
          if (L.dof) {
#pragma unroll
            for (int h = 0; h < 2; h++) {
#pragma unroll
              for (int j = 0; j < 9; j++) {
                const int an = j < 4 ? j : L.ni.cs + j - 4;
                const bool valid = j < 4 ? (j <= lane) : (lane >= 4 && an <= lane);
                if (valid) Hd[hidx(h * 26 + an, h * 26 + lane)] = W.Mt[h * 26 + lane][j] + (an == lane ? fc_[h] + lc[h] : 0.f);
              }
            }
          }
          __syncwarp();
          float shss = sM + sc;
          const unsigned lt = (1u << lane) - 1u;
          for (int c = 0; c < ncon; c++) {
            const float w0 = W.c_w[0][c];
            if (w0 == 0.f) continue;
            const float w1 = W.c_w[1][c], w2 = W.c_w[2][c], w3 = W.c_w[3][c], w4 = W.c_w[4][c];
            const unsigned m0 = W.c_mask[0][c], m1 = W.c_mask[1][c];
            const int n0 = __popc(m0), n1_ = __popc(m1), slot = W.c_slot[c];
#pragma unroll
            for (int h = 0; h < 2; h++) {
              const unsigned mh = h ? m1 : m0;
              if (!((mh >> lane) & 1u)) continue;
              const int e0 = (h ? n0 : 0) + __popc(mh & lt);
              const float ux = W.c_u[3 * e0][c], uy = W.c_u[3 * e0 + 1][c], uz = W.c_u[3 * e0 + 2][c];
              const float zx = w0 * ux + w1 * uy + w3 * uz, zy = w1 * ux + w2 * uy, zz = w3 * ux + w4 * uz;
              const int g = h * 26 + lane;
              const int etot = n0 + n1_ + (slot >= 0 ? 1 : 0);
              for (int e = e0; e < etot; e++) {
                const float val = zx * W.c_u[3 * e][c] + zy * W.c_u[3 * e + 1][c] + zz * W.c_u[3 * e + 2][c];
                const int col = W.c_col[e][c];
                if (col >= 52) W.a.s.Kcd[col - 52][g] += val;
                else Hd[hidx(g, col)] += val;
              }
            }
            if (slot == lane) {
              const int e0 = n0 + n1_;
              const float ux = W.c_u[3 * e0][c], uy = W.c_u[3 * e0 + 1][c], uz = W.c_u[3 * e0 + 2][c];
              shss += ux * (w0 * ux + w1 * uy + w3 * uz) + uy * (w1 * ux + w2 * uy) + uz * (w3 * ux + w4 * uz);
            }
          }
          if (lane < nslot) W.a.s.hss[lane] = 1.f / shss;
          __syncwarp();
          float b[2] = {-grad[0], -grad[1]};
          const float sb = -sgrad;
          for (int s = 0; s < nslot; s++) {
            const float ih = W.a.s.hss[s];
            const float bs = __shfl_sync(FULL, sb, s) * ih;
            if (L.dof) {
#pragma unroll
              for (int h = 0; h < 2; h++) {
                const int g = h * 26 + lane;
                const float kg = W.a.s.Kcd[s][g];
                if (kg != 0.f) {
                  b[h] -= kg * bs;
                  const float cf_ = kg * ih;
                  for (int gk = g; gk < 52; gk++) Hd[hidx(g, gk)] -= cf_ * W.a.s.Kcd[s][gk];
                }
              }
            }
          }
          __syncwarp();
          dense_factor_solve(Hd, lane, L.dof, b);
          __syncwarp();
#pragma unroll
          for (int h = 0; h < 2; h++) { p[h] = L.dof ? b[h] : 0.f; if (L.dof) W.va[h * 26 + lane] = p[h]; }
          __syncwarp();
          if (lane < nslot) {
            float d = 0.f;
            for (int g = 0; g < 52; g++) d += W.a.s.Kcd[lane][g] * W.va[g];
            sp = (sb - d) * W.a.s.hss[lane];
            W.va[52 + lane] = sp;
          }
          __syncwarp();
        } else {
        // ---- H_t: M + diagonal rows + in-front contact cliques (Euler: M + h diag(b))
        // diagonal of the dof rows (Euler: h * damping), added to M when the fronts are loaded
        if (L.dof) {
#pragma unroll
          for (int h = 0; h < 2; h++) W.dg[h * 26 + lane] = euler ? o.h * ht.damping[h * PP_HD + lane] : fc_[h] + lc[h];
        }
        // H_t without that diagonal (M + cliques of the in-front contacts - Schur complements of the key
        // slots) only changes when an in-front contact row or a slot's limit row changes zone: kept otherwise
        bool redo = false;
        if (!euler) {
          bool chg = zone_old < 0;  // nothing assembled yet (first trip, or the trip before factorised densely)
          int zone = 0;
          if (con_on && !my_lr) zone = (cc[0] != 0.f ? 1 : 0) | (cc[1] != 0.f ? 2 : 0) | (cc[2] != 0.f ? 4 : 0) | (cc[3] != 0.f ? 8 : 0);
          if (lane < nslot && sc != 0.f) zone |= 16;
          chg |= zone != zone_old;
          redo = __any_sync(FULL, chg);
          zone_old = zone;
        }
        float shss = sM + sc;  // H(slot, slot), lane < nslot
        if (redo) {
#ifdef PP_SP_TIMING
          dbg_redo++;
#endif
          build_M();
          for (int s = 0; s < nslot; s++) {
            if (lane < 9) W.a.s.Kc[s][lane] = 0.f;
          }
          __syncwarp();
          const unsigned lt = (1u << lane) - 1u;
          for (int c = 0; c < ncon; c++) {
            const float w0 = W.c_w[0][c];
            if (w0 == 0.f || W.c_lr[c]) continue;  // no active row / handled by the low-rank path
            const float w1 = W.c_w[1][c], w2 = W.c_w[2][c], w3 = W.c_w[3][c], w4 = W.c_w[4][c];
            const unsigned m0 = W.c_mask[0][c], m1 = W.c_mask[1][c];
            const int n0 = __popc(m0), n1_ = __popc(m1), slot = W.c_slot[c];
#pragma unroll
            for (int h = 0; h < 2; h++) {
              const unsigned mh = h ? m1 : m0;
              if (!((mh >> lane) & 1u)) continue;
              const int e0 = (h ? n0 : 0) + __popc(mh & lt);
              const float ux = W.c_u[3 * e0][c], uy = W.c_u[3 * e0 + 1][c], uz = W.c_u[3 * e0 + 2][c];
              const float zx = w0 * ux + w1 * uy + w3 * uz, zy = w1 * ux + w2 * uy, zz = w3 * ux + w4 * uz;
              const int g = h * 26 + lane;
              const int etot = n0 + n1_ + (slot >= 0 ? 1 : 0);
              for (int e = e0; e < etot; e++) {
                const float val = zx * W.c_u[3 * e][c] + zy * W.c_u[3 * e + 1][c] + zz * W.c_u[3 * e + 2][c];
                const int col = W.c_col[e][c];
                if (col >= 52) W.a.s.Kc[col - 52][posl] += val;
                else Ht[col][posl] += val;
              }
            }
            if (slot == lane) {
              const int e0 = n0 + n1_;
              const float ux = W.c_u[3 * e0][c], uy = W.c_u[3 * e0 + 1][c], uz = W.c_u[3 * e0 + 2][c];
              shss += ux * (w0 * ux + w1 * uy + w3 * uz) + uy * (w1 * ux + w2 * uy) + uz * (w3 * ux + w4 * uz);
            }
          }
          if (lane < nslot) W.a.s.hss[lane] = 1.f / shss;
          __syncwarp();
          // ---- eliminate the key slots from H_t (Schur complement on the owner's root path)
          if (L.dof) {
            for (int s = 0; s < nslot; s++) {
              const int h = W.slot_hand[s];
              const unsigned sm = W.slot_mask[s];
              if (!((sm >> lane) & 1u)) continue;
              const int scs = cs_of(31 - __clz(sm));
              const float cf_ = W.a.s.Kc[s][posl] * W.a.s.hss[s];
              unsigned m = sm & ~lt;
              while (m) { const int d = __ffs(m) - 1; m &= m - 1; Ht[h * 26 + d][posl] -= cf_ * W.a.s.Kc[s][apos(d, scs)]; }
            }
          }
          __syncwarp();
        }
        // ---- multifrontal factorisation. Front (hand fh, chain fc): local variables 0..4 = chain dofs
        // deepest first (a 4-dof chain is padded with an identity pivot), 5..8 = root dofs 3,2,1,0.
        float F[45], Rt[10];
        {
          const float (*Hs)[9] = euler ? W.Mt : Ht;  // Euler: M + h diag(b), no constraint rows
          __syncwarp();
#pragma unroll
          for (int i = 0; i < 9; i++) {
#pragma unroll
            for (int j = 0; j <= i; j++) {
              float val = 0.f;
              if (i < 5) {
                if (fvalid && i < fn) val = Hs[fg0 + fce - j][4 + (fce - i) - fcs] + (i == j ? W.dg[fg0 + fce - i] : 0.f);
                else val = i == j ? 1.f : 0.f;
              } else if (j < 5) {
                if (fvalid && j < fn) val = Hs[fg0 + fce - j][8 - i];
              }
              F[i * (i + 1) / 2 + j] = val;
            }
          }
#pragma unroll
          for (int i = 0; i < 4; i++)
#pragma unroll
            for (int j = 0; j <= i; j++) Rt[i * (i + 1) / 2 + j] = Hs[fg0 + 3 - j][3 - i] + (i == j ? W.dg[fg0 + 3 - i] : 0.f);
          // eliminate the 5 chain pivots
#pragma unroll
          for (int k = 0; k < 5; k++) {
            const float inv = rsqrtf(fmaxf(F[k * (k + 1) / 2 + k], 1e-30f));
            F[k * (k + 1) / 2 + k] = inv;
#pragma unroll
            for (int i = k + 1; i < 9; i++) F[i * (i + 1) / 2 + k] *= inv;
#pragma unroll
            for (int i = k + 1; i < 9; i++)
#pragma unroll
              for (int j = k + 1; j <= i; j++) F[i * (i + 1) / 2 + j] -= F[i * (i + 1) / 2 + k] * F[j * (j + 1) / 2 + k];
          }
          // root front = root block of H_t + the update matrices of the hand's 5 chain fronts
#pragma unroll
          for (int i = 0; i < 4; i++)
#pragma unroll
            for (int j = 0; j <= i; j++) {
              float upd = F[(i + 5) * (i + 6) / 2 + j + 5];
              upd += __shfl_xor_sync(FULL, upd, 1); upd += __shfl_xor_sync(FULL, upd, 2); upd += __shfl_xor_sync(FULL, upd, 4);
              Rt[i * (i + 1) / 2 + j] += upd;
            }
#pragma unroll
          for (int k = 0; k < 4; k++) {
            const float inv = rsqrtf(fmaxf(Rt[k * (k + 1) / 2 + k], 1e-30f));
            Rt[k * (k + 1) / 2 + k] = inv;
#pragma unroll
            for (int i = k + 1; i < 4; i++) Rt[i * (i + 1) / 2 + k] *= inv;
#pragma unroll
            for (int i = k + 1; i < 4; i++)
#pragma unroll
              for (int j = k + 1; j <= i; j++) Rt[i * (i + 1) / 2 + j] -= Rt[i * (i + 1) / 2 + k] * Rt[j * (j + 1) / 2 + k];
          }
        }
        // ---- right-hand sides. Row 0: -grad (Euler: f_smooth + f_constraint = M qacc); rows 1..m:
        // sqrt(D) J_r^T of the active rows of the low-rank contacts.
        float V[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};  // lower 3x3: (00, 10, 20, 11, 21, 22)
        unsigned bal[3] = {0u, 0u, 0u};
        int vrow[3] = {0, 0, 0};
        float (*Z)[ZS] = W.a.s.Z;
        const float sb = -sgrad;
        {
          // W = V V^T (3x3 Cholesky of the active pyramid rows' weight, rank <= 3): columns U v_j
          if (!euler && nlr > 0) {
            if (my_lr) {
              const float w0 = W.c_w[0][lane], w1 = W.c_w[1][lane], w2 = W.c_w[2][lane], w3 = W.c_w[3][lane], w4 = W.c_w[4][lane];
              if (w0 > 0.f) {
                const float l00 = sqrtf(w0), l10 = w1 / l00, l20 = w3 / l00;
                const float d1 = w2 - l10 * l10;
                const float l11 = d1 > 1e-6f * w2 ? sqrtf(d1) : 0.f;
                const float l21 = l11 > 0.f ? -l20 * l10 / l11 : 0.f;
                const float d2 = w4 - l20 * l20 - l21 * l21;
                const float l22 = d2 > 1e-6f * w4 ? sqrtf(d2) : 0.f;
                V[0] = l00; V[1] = l10; V[2] = l20; V[3] = l11; V[4] = l21; V[5] = l22;
              }
            }
            bal[0] = __ballot_sync(FULL, V[0] != 0.f); bal[1] = __ballot_sync(FULL, V[3] != 0.f);
            bal[2] = __ballot_sync(FULL, V[5] != 0.f);
            m = __popc(bal[0]) + __popc(bal[1]) + __popc(bal[2]);
            for (int i = lane; i < m * ZS; i += 32) (&Z[1][0])[i] = 0.f;
          }
#pragma unroll
          for (int h = 0; h < 2; h++) if (L.dof) Z[0][h * 26 + lane] = euler ? Ma[h] : -grad[h];
          if (lane < nslot) Z[0][52 + lane] = euler ? 0.f : sb;
          __syncwarp();
          if (m > 0 && my_lr) {
            const unsigned ltc = (1u << lane) - 1u;
            vrow[0] = __popc(bal[0] & ltc);
            vrow[1] = __popc(bal[0]) + __popc(bal[1] & ltc);
            vrow[2] = __popc(bal[0]) + __popc(bal[1]) + __popc(bal[2] & ltc);
            for (int e = 0; e < my_nent; e++) {
              const float ux = W.c_u[3 * e][lane], uy = W.c_u[3 * e + 1][lane], uz = W.c_u[3 * e + 2][lane];
              const int col = W.c_col[e][lane];
              if (V[0] != 0.f) Z[1 + vrow[0]][col] = V[0] * ux + V[1] * uy + V[2] * uz;
              if (V[3] != 0.f) Z[1 + vrow[1]][col] = V[3] * uy + V[4] * uz;
              if (V[5] != 0.f) Z[1 + vrow[2]][col] = V[5] * uz;
            }
          }
          __syncwarp();
          // slot elimination applied to every right-hand side: b_g -= Kc[s][g] b_s / H_ss
          if (!euler && L.dof) {
            for (int s = 0; s < nslot; s++) {
              const unsigned sm = W.slot_mask[s];
              if (!((sm >> lane) & 1u)) continue;
              const int g = W.slot_hand[s] * 26 + lane;
              const float cf_ = W.a.s.Kc[s][posl] * W.a.s.hss[s];
              const int rmax = ((slot_lr_mask >> s) & 1u) ? m : 0;  // other slots: zero in rows 1..m
              for (int r = 0; r <= rmax; r++) Z[r][g] -= cf_ * Z[r][52 + s];
            }
          }
          __syncwarp();
        }
        // ---- solves: two right-hand sides per pass (one per replica of the fronts)
        for (int r0 = 0; r0 <= m; r0 += 2) {
          const int r = min(r0 + frep, m);  // a replica without work repeats the last row (same values)
          float y[9];
#pragma unroll
          for (int i = 0; i < 5; i++) y[i] = (fvalid && i < fn) ? Z[r][fg0 + fce - i] : 0.f;
#pragma unroll
          for (int i = 5; i < 9; i++) y[i] = 0.f;
          float rr[4];
#pragma unroll
          for (int i = 0; i < 4; i++) rr[i] = Z[r][fg0 + 3 - i];
          __syncwarp();
#pragma unroll
          for (int k = 0; k < 5; k++) {
            y[k] *= F[k * (k + 1) / 2 + k];
#pragma unroll
            for (int i = k + 1; i < 9; i++) y[i] -= F[i * (i + 1) / 2 + k] * y[k];
          }
#pragma unroll
          for (int i = 0; i < 4; i++) {
            float t = y[5 + i];
            t += __shfl_xor_sync(FULL, t, 1); t += __shfl_xor_sync(FULL, t, 2); t += __shfl_xor_sync(FULL, t, 4);
            rr[i] += t;
          }
#pragma unroll
          for (int k = 0; k < 4; k++) {
            rr[k] *= Rt[k * (k + 1) / 2 + k];
#pragma unroll
            for (int i = k + 1; i < 4; i++) rr[i] -= Rt[i * (i + 1) / 2 + k] * rr[k];
          }
#pragma unroll
          for (int k = 3; k >= 0; k--) {
            float t = rr[k];
#pragma unroll
            for (int i = k + 1; i < 4; i++) t -= Rt[i * (i + 1) / 2 + k] * rr[i];
            rr[k] = t * Rt[k * (k + 1) / 2 + k];
          }
#pragma unroll
          for (int i = 0; i < 4; i++) y[5 + i] = rr[i];
#pragma unroll
          for (int k = 4; k >= 0; k--) {
            float t = y[k];
#pragma unroll
            for (int i = k + 1; i < 9; i++) t -= F[i * (i + 1) / 2 + k] * y[i];
            y[k] = t * F[k * (k + 1) / 2 + k];
          }
          if (fvalid && (frep == 0 || r0 + 1 <= m)) {
#pragma unroll
            for (int i = 0; i < 5; i++) if (i < fn) Z[r][fg0 + fce - i] = y[i];
            if (fc == 0) {
#pragma unroll
              for (int i = 0; i < 4; i++) Z[r][fg0 + 3 - i] = rr[i];
            }
          }
          __syncwarp();
        }
        if (euler) {
#pragma unroll
          for (int h = 0; h < 2; h++)
            if (L.dof) { const float x = Z[0][h * 26 + lane]; v[h] += o.h * x; q[h] += o.h * v[h]; }
          __syncwarp();
          if (PP_TRIP_MEET && WPB > 1 && meet) { fin = true; continue; }
          break;
        }
        // ---- back-substitute the key slots touched by low-rank contacts in every solved vector
        int bs_h = 0, bs_cs = 0; unsigned bs_m = 0u;
        if (lane < nslot) {
          bs_h = W.slot_hand[lane]; bs_m = W.slot_mask[lane];
          bs_cs = bs_m ? cs_of(31 - __clz(bs_m)) : 0;
          if (slot_lr) {
            for (int r = 0; r <= m; r++) {
              float d = 0.f;
              unsigned mm = bs_m;
              while (mm) { const int dd = __ffs(mm) - 1; mm &= mm - 1; d += W.a.s.Kc[lane][apos(dd, bs_cs)] * Z[r][bs_h * 26 + dd]; }
              Z[r][52 + lane] = (Z[r][52 + lane] - d) * W.a.s.hss[lane];
            }
          }
        }
        __syncwarp();
        // ---- Woodbury: p = y0 - Z (I + B^T Z)^-1 B^T y0
        float (*C)[LRR + 1] = W.a.s.C;
        if (m > 0) {
          // capacitance, lane = solved vector (0: y0 -> right-hand side, 1 + r: column r)
          unsigned lm = lrb;
          while (lm) {
            const int c = __ffs(lm) - 1; lm &= lm - 1;
            const float v0 = __shfl_sync(FULL, V[0], c), v1 = __shfl_sync(FULL, V[1], c), v2 = __shfl_sync(FULL, V[2], c);
            const float v3 = __shfl_sync(FULL, V[3], c), v4 = __shfl_sync(FULL, V[4], c), v5 = __shfl_sync(FULL, V[5], c);
            if (v0 == 0.f) continue;
            const unsigned ltc = (1u << c) - 1u;
            const int q0 = __popc(bal[0] & ltc), q1 = __popc(bal[0]) + __popc(bal[1] & ltc);
            const int q2 = __popc(bal[0]) + __popc(bal[1]) + __popc(bal[2] & ltc);
            if (lane <= m) {
              const int ne = W.c_nent[c];
              f3 uz = mk3(0, 0, 0);
              for (int e = 0; e < ne; e++) {
                const float x = Z[lane][W.c_col[e][c]];
                uz.x += W.c_u[3 * e][c] * x; uz.y += W.c_u[3 * e + 1][c] * x; uz.z += W.c_u[3 * e + 2][c] * x;
              }
              const int col = lane == 0 ? m : lane - 1;
              C[q0][col] = v0 * uz.x + v1 * uz.y + v2 * uz.z + (col == q0 ? 1.f : 0.f);
              if (v3 != 0.f) C[q1][col] = v3 * uz.y + v4 * uz.z + (col == q1 ? 1.f : 0.f);
              if (v5 != 0.f) C[q2][col] = v5 * uz.z + (col == q2 ? 1.f : 0.f);
            }
          }
          __syncwarp();
          // Cholesky of the SPD capacitance (lane = row) and solve C t = B^T y0
          for (int j = 0; j < m; j++) {
            float s = 0.f;
            if (lane >= j && lane < m) {
              s = C[lane][j];
#if PP_T2
#pragma unroll 4
#endif
              for (int k = 0; k < j; k++) s -= C[lane][k] * C[j][k];
            }
            const float inv = rsqrtf(fmaxf(__shfl_sync(FULL, s, j), 1e-30f));
            if (lane >= j && lane < m) C[lane][j] = lane == j ? inv : s * inv;
            __syncwarp();
          }
          float t = lane < m ? C[lane][m] : 0.f;
          for (int j = 0; j < m; j++) {  // forward
            const float yj = __shfl_sync(FULL, t, j) * C[j][j];
            if (lane == j) t = yj; else if (lane > j && lane < m) t -= C[lane][j] * yj;
          }
          for (int j = m - 1; j >= 0; j--) {  // backward
            const float xj = __shfl_sync(FULL, t, j) * C[j][j];
            if (lane == j) t = xj; else if (lane < j) t -= C[j][lane] * xj;
          }
          if (lane < m) C[lane][m] = t;
          __syncwarp();
        }
#pragma unroll
        for (int h = 0; h < 2; h++)
          if (L.dof) {
            float s = Z[0][h * 26 + lane];
            for (int r = 0; r < m; r++) s -= Z[1 + r][h * 26 + lane] * C[r][m];
            p[h] = s; W.va[h * 26 + lane] = s;
          }
        __syncwarp();
        if (lane < nslot) {
          if (slot_lr) {
            float s = Z[0][52 + lane];
            for (int r = 0; r < m; r++) s -= Z[1 + r][52 + lane] * C[r][m];
            sp = s;
          } else {  // the slot's row of H p = -grad, from the final hand part of p
            float d = 0.f;
            unsigned mm = bs_m;
            while (mm) { const int dd = __ffs(mm) - 1; mm &= mm - 1; d += W.a.s.Kc[lane][apos(dd, bs_cs)] * W.va[bs_h * 26 + dd]; }
            sp = (Z[0][52 + lane] - d) * W.a.s.hss[lane];
          }
          W.va[52 + lane] = sp;
        }
        __syncwarp();
        }
        // ---- J p for every row
        float jp[4] = {0.f, 0.f, 0.f, 0.f};
        if (con_on) {
          const f3 up = contact_ux(W, lane, my_nent);
          jp[0] = up.x + cmu * up.y; jp[1] = up.x - cmu * up.y; jp[2] = up.x + cmu * up.z; jp[3] = up.x - cmu * up.z;
        }
        __syncwarp();
        // ---- M p from the pristine tree entries
        float Mp[2], sMp = 0.f;
        matvec_M(Mp);
        if (lane < nslot) sMp = sM * sp;
        // ---- exact line search on phi(alpha) = cost(a + alpha p)
        float g1 = grad[0] * p[0] + grad[1] * p[1] + sgrad * sp;
        float g2 = p[0] * Mp[0] + p[1] * Mp[1] + sp * sMp;
        float gauss1 = (Ma[0] - fs[0]) * p[0] + (Ma[1] - fs[1]) * p[1] + (sMa - sfs) * sp;
#if PP_T4
        // phi''(0) = p^T H p from the curvatures already in hand: the first Newton step on phi starts from
        // alpha = 0 without evaluating the rows there (phi'(0) = g1)
        float dd0 = (fc_[0] + lc[0]) * p[0] * p[0] + (fc_[1] + lc[1]) * p[1] * p[1] + sc * sp * sp;
#pragma unroll
        for (int k = 0; k < 4; k++) dd0 += cc[k] * jp[k] * jp[k];
        dd0 = wsum(dd0);
#endif
        g1 = wsum(g1); g2 = wsum(g2); gauss1 = wsum(gauss1);
        if (!(g1 < 0.f) || 0.5f * scale * (-g1) < 1e-10f) { euler = true; continue; }  // converged
        float alpha = 0.f, lo = 0.f, hi = -1.f;
        bool changed = false;
        int ls0 = 0;
#if PP_T4
        if (dd0 + g2 > 0.f) { alpha = -g1 / (dd0 + g2); ls0 = 1; ls_total++; }
#endif
        for (int ls = ls0; ls < 20; ls++) {
          float dphi = 0.f, ddphi = 0.f;
          bool ch = false;
#pragma unroll
          for (int h = 0; h < 2; h++) {
            float f_, c_; int z_;
            if (floss[h] > 0.f) {
              row_fric(fD[h], floss[h], a[h] - faref[h] + alpha * p[h], f_, c_, z_);
              dphi -= f_ * p[h]; ddphi += c_ * p[h] * p[h]; ch |= z_ != fz[h];
            }
            if (lsg[h] != 0.f) {
              const float jp_ = lsg[h] * p[h];
              row_ineq(lD[h], lsg[h] * a[h] - laref[h] + alpha * jp_, f_, c_);
              dphi -= f_ * jp_; ddphi += c_ * jp_ * jp_; ch |= (c_ != lc[h]);
            }
          }
          if (ssg != 0.f) {
            float f_, c_;
            const float jp_ = ssg * sp;
            row_ineq(sD, ssg * sa - saref + alpha * jp_, f_, c_);
            dphi -= f_ * jp_; ddphi += c_ * jp_ * jp_; ch |= (c_ != sc);
          }
          if (con_on) {
#pragma unroll
            for (int k = 0; k < 4; k++) {
              float f_, c_;
              row_ineq(cD, jar[k] + alpha * jp[k], f_, c_);
              dphi -= f_ * jp[k]; ddphi += c_ * jp[k] * jp[k]; ch |= (c_ != cc[k]);
            }
          }
          dphi = wsum(dphi) + gauss1 + alpha * g2;
          ddphi = wsum(ddphi) + g2;
          changed = __any_sync(FULL, ch);
          if (fabsf(dphi) <= 1e-6f * fabsf(g1) || !(ddphi > 0.f)) break;
          if (dphi < 0.f) lo = alpha; else hi = alpha;
          float next = alpha - dphi / ddphi;
          if (hi >= 0.f && (next <= lo || next >= hi)) next = 0.5f * (lo + hi);
          if (next == alpha) break;
          alpha = next;
          ls_total++;
        }
        // ---- move
#pragma unroll
        for (int h = 0; h < 2; h++) { a[h] += alpha * p[h]; Ma[h] += alpha * Mp[h]; }
        sa += alpha * sp; sMa += alpha * sMp;
#pragma unroll
        for (int k = 0; k < 4; k++) jar[k] += alpha * jp[k];
        iter++;
        // piecewise-quadratic cost: a full Newton step that crosses no breakpoint is exact. This also holds
        // with low-rank rows: over a whole config-2 episode the decrement test of the following trip confirmed
        // convergence in 1360 of 1365 such steps (and the other 5 were marginal), at the price of a full
        // factorisation each; single-substep parity against the fp64 restatement is unchanged without it.
        // (Measured in round 2: also stopping after a step whose improvement 0.5 alpha |phi'(0)| / (meaninertia nv)
        // is below opt.tolerance - MuJoCo's own rule - changes no iteration count on configs 2 and 3: this test
        // already catches those steps.)
        if ((!changed && fabsf(alpha - 1.f) < 1e-3f) || iter >= maxit) euler = true;
      }
      if (MODE == PP_SPLIT_NEWTON) {  // the loop is over: the finish kernel takes the env from here
        PP_SP_SAVE_RW();
#ifdef PP_SP_TIMING
        {
          const long long last = clock64() - dbg_t0;  // the mj_Euler trip and the hand-over
          dbg_cyc += last;
          int* cn = A.counters + (size_t)env * 8;
          if (lane == 0) { cn[1] = (int)(dbg_cyc >> 4); cn[3] = (int)(dbg_first >> 4); cn[5] = (int)(last >> 4); cn[6] = (int)(dbg_refill >> 4); }
        }
#endif
        return;
      }
      // ---- solution: hand dofs, touched keys (slots), untouched keys in closed form
#pragma unroll
      for (int h = 0; h < 2; h++) {
        acc[h] = a[h];
        stick[h] = floss[h] > 0.f && fabsf(a[h] - faref[h]) < floss[h] / fD[h];
      }
      if (lane < nslot) W.va[52 + lane] = sa;
      __syncwarp();
#pragma unroll
      for (int i = 0; i < 3; i++) {
        const int k = lane + 32 * i;
        if (k >= PP_NKEY) continue;
        const float Mk = kt.M[k];
        const int slot = W.key_slot[k];
        float ftot;
        if (slot >= 0 && slot < KS) {
          ak[i] = W.va[52 + slot];
          ftot = Mk * ak[i];
        } else {
          const LimitRow r = pp_limit_row(o, qk[i], vk[i], 0.f, kt.hi[k], kt.invw[k]);
          const float aa_ = pp_key_solve(Mk, fsk[i], r);
          const float fcon = (r.sgn != 0.f && r.sgn * aa_ - r.aref < 0.f) ? -r.D * (r.sgn * aa_ - r.aref) * r.sgn : 0.f;
          ak[i] = aa_;
          ftot = fsk[i] + fcon;
          nefc += 0;
        }
        pp_scalar_euler(Mk, kt.damping, o.h, ftot, qk[i], vk[i]);
      }
      __syncwarp();
      if (A.substep_act) {
        // Piano.after_substep -> MidiModule.after_substep (piano.py:154-162, midi_module.py:47-98): the key and
        // pedal activations after EVERY substep, as bit masks; the host turns their changes into note events
        unsigned bits[3];
#pragma unroll
        for (int i = 0; i < 3; i++) {
          const int k = lane + 32 * i;
          float nrm = 0.f; bool on = false;
          if (k < PP_NKEY) pp_key_state(qk[i], kt.hi[k], nrm, on);
          bits[i] = __ballot_sync(FULL, on);
        }
        if (lane == 0) A.substep_act[(size_t)env * nsub + sub] = make_uint4(bits[0], bits[1], bits[2], sustain >= 0.5f ? 1u : 0u);
      }
      cnt_ncon = ncon; cnt_iter = iter; cnt_drop = ndrop;
      sum_iter += iter; sum_ls += ls_total; sum_dense += dense_mode ? 1 : 0; sum_coupled += nlr;
      {  // key limit rows count towards nefc (diagnostic)
        int nk = 0;
#pragma unroll
        for (int i = 0; i < 3; i++) { const int k = lane + 32 * i; if (k < PP_NKEY && (W.kq[k] < 0.f || W.kq[k] > kt.hi[k])) nk++; }
        cnt_nefc = nefc + (int)wsum((float)nk);
      }
      resume = false;
      if (MODE == PP_SPLIT_FIN && sub + 1 < nsub) {
        // between two substeps of the split schedule the env lives in the handle's state arrays
#pragma unroll
        for (int h = 0; h < 2; h++) {
          if (L.dof) {
            A.qpos[sbase + PP_NKEY + h * PP_HD + lane] = q[h];
            A.qvel[sbase + PP_NKEY + h * PP_HD + lane] = v[h];
            A.warm[sbase + PP_NKEY + h * PP_HD + lane] = acc[h];
          }
          if (A.stick_bits) {
            const unsigned b = __ballot_sync(FULL, stick[h]);
            if (lane == 0) A.stick_bits[(size_t)env * 2 + h] = b;
          }
        }
#pragma unroll
        for (int i = 0; i < 3; i++) {
          const int k = lane + 32 * i;
          if (k < PP_NKEY) { A.qpos[sbase + k] = qk[i]; A.qvel[sbase + k] = vk[i]; A.warm[sbase + k] = ak[i]; }
        }
        if (lane == 0) {
          int* cn = A.counters + (size_t)env * 8;
          cn[0] = cnt_ncon; cn[2] = cnt_iter; cn[4] = sum_iter; cn[7] = sum_coupled;
          if (!kSpTiming) { cn[1] = cnt_nefc; cn[3] = cnt_drop; cn[5] = sum_ls; cn[6] = sum_dense; }
        }
        if (!A.sp_chain) return;  // chained: straight on into the next substep's prologue, which ends this launch
      }
    }  // substeps
  }
  const bool stale = A.stale && nsub > 0;
  if (stale) {  // body frames of the last substep's start over those of the integrated state
    float qs[2];
#pragma unroll
    for (int h = 0; h < 2; h++) qs[h] = A.stale[((size_t)env * 7 + h) * 32 + lane];
    Kin K;
    forward_kinematics<false>(ht, o, L, qs, qs, yoff, K);
    __syncwarp();
#pragma unroll
    for (int h = 0; h < 2; h++) {
      if (L.dof) {
        const int g = h * 26 + lane;
        st3(W.a.k.xpos[g], K.pos[h]);
        W.a.k.xquat[g][0] = K.quat[h].w; W.a.k.xquat[g][1] = K.quat[h].x; W.a.k.xquat[g][2] = K.quat[h].y; W.a.k.xquat[g][3] = K.quat[h].z;
      }
    }
  }
  t_idx = do_reset ? 0 : t_idx + 1;

  // ================================================================== sensors at the integrated state
  __syncwarp();
  // fingertip sites: lanes 0..9
  f3 site = mk3(0, 0, 0);
  if (lane < PP_NSITE) {
    const int hb = ht.site_body[lane], h = hb >= PP_HB ? 1 : 0, g = h * 26 + (hb - h * PP_HB) + 1;
    f4 qq; qq.w = W.a.k.xquat[g][0]; qq.x = W.a.k.xquat[g][1]; qq.y = W.a.k.xquat[g][2]; qq.z = W.a.k.xquat[g][3];
    site = ld3(W.a.k.xpos[g]) + qrot(qq, ld3(ht.site_pos[lane]));
    st3(A.site_xpos + ((size_t)env * PP_NSITE + lane) * 3, site);
  }
  // forearm-forearm contact (forearm reward): the fixed pair of the two forearm capsules
  bool forearm_hit = false;
  if (ht.forearm_geom[0] >= 0 && ht.forearm_geom[1] >= 0) {
    const int g1 = ht.forearm_geom[0], g2 = ht.forearm_geom[1];
    const int a1 = ht.geom_body[g1] + 1, a2 = 26 + ht.geom_body[g2] - PP_HB + 1;
    f4 q1; q1.w = W.a.k.xquat[a1][0]; q1.x = W.a.k.xquat[a1][1]; q1.y = W.a.k.xquat[a1][2]; q1.z = W.a.k.xquat[a1][3];
    f4 q2; q2.w = W.a.k.xquat[a2][0]; q2.x = W.a.k.xquat[a2][1]; q2.y = W.a.k.xquat[a2][2]; q2.z = W.a.k.xquat[a2][3];
    const f3 c1 = ld3(W.a.k.xpos[a1]) + qrot(q1, ld3(ht.geom_pos[g1]));
    const f3 c2 = ld3(W.a.k.xpos[a2]) + qrot(q2, ld3(ht.geom_pos[g2]));
    const Hit hh = capsule_capsule(c1, qrot(q1, mk3(0, 0, 1)), ht.geom_r[g1], ht.geom_hl[g1], c2,
                                   qrot(q2, mk3(0, 0, 1)), ht.geom_r[g2], ht.geom_hl[g2],
                                   o.pair_margin[PP_PAIRCLS_HAND]);
    forearm_hit = hh.ok && hh.dist <= 0.f;
  }
  // actuatorvel sensor at the integrated velocity (refreshed by the trailing mj_step1)
  float act_vel[2] = {0.f, 0.f};
#pragma unroll
  for (int h = 0; h < 2; h++) {
    const int ai = h * 22 + (lane < 22 ? lane : 0);
    const int t0 = ht.act_trn0[ai] - h * 26, t1 = ht.act_trn1[ai] - h * 26;
    const float vs = stale ? A.stale[((size_t)env * 7 + 2 + h) * 32 + lane] : v[h];  // legacy_step=False: mj_fwdVelocity of the last mj_step
    float vv = __shfl_sync(FULL, vs, t0);
    const float v1 = __shfl_sync(FULL, vs, t1 >= 0 ? t1 : 0);
    if (ht.act_trn1[ai] >= 0) vv += v1;
    act_vel[h] = lane < 22 ? vv : 0.f;
  }

  // ================================================================== task epilogue
  // Piano._update_key_state (piano.py:178-192)
  float normalized[3]; bool active[3];
#pragma unroll
  for (int i = 0; i < 3; i++) {
    const int k = lane + 32 * i;
    normalized[i] = 0.f; active[i] = false;
    if (k < PP_NKEY) pp_key_state(qk[i], kt.hi[k], normalized[i], active[i]);
  }
  // the env's song (looked up here, after the substeps: nothing before needs it): one of the bank's (MIDI augmentations: an env takes its pending song at a reset), or the only one
  int T = A.n_frames, song_off = 0;  // frames of the env's song, its first frame inside A.goal / A.pairs
  if (A.song_id) {
    int sid = do_reset ? A.song_id_pending[env] : A.song_id[env];
    sid = min(max(sid, 0), A.n_songs - 1);
    if (do_reset && lane == 0) A.song_id[env] = sid;
    song_off = A.song_off[sid];
    T = A.song_T[sid];
  }
  const bool sustain_on = sustain >= 0.5f;
  const int t_cur = t_idx > 0 ? t_idx - 1 : 0;  // frame just played (goal_current)
  const uint8_t* goal_cur = A.goal + (size_t)(song_off + t_cur) * 89;
  // key_press reward (piano_with_shadow_hands.py:279-298)
  float kp_sum = 0.f; int kp_on = 0; bool wrong = false;
#pragma unroll
  for (int i = 0; i < 3; i++) {
    const int k = lane + 32 * i;
    if (k >= PP_NKEY) continue;
    if (goal_cur[k]) { kp_on++; kp_sum += pp_tolerance(1.f - normalized[i], 0.05f, 0.5f); }
    else if (active[i]) wrong = true;
  }
  kp_sum = wsum(kp_sum);
  const int n_on = (int)(wsum((float)kp_on) + 0.5f);
  const bool failure = __any_sync(FULL, wrong);
  float energy0 = wsum(fabsf(act_force[0]) * fabsf(act_vel[0]));
  float energy1 = wsum(fabsf(act_force[1]) * fabsf(act_vel[1]));
  // fingering reward (:300-331): fingertip site to key surface point, world offsets
  float fing = 0.f;
  float qkp[3];  // key angles behind the key geoms' world positions (legacy_step=False: those of the last mj_step's kinematics)
#pragma unroll
  for (int i = 0; i < 3; i++) qkp[i] = stale ? A.stale[((size_t)env * 7 + 4 + i) * 32 + lane] : qk[i];
  if (A.has_fingering == 1) {
    float tot = 0.f; int cnt = 0;
    for (int p0 = 0; p0 < A.pmax; p0 += 32) {
      int key = -1, finger = 0;
      if (p0 + lane < A.pmax && !do_reset) {
        const int16_t* pr = A.pairs + ((size_t)(song_off + t_cur) * A.pmax + p0 + lane) * 2;
        key = pr[0]; finger = pr[1];
      }
      const bool has = key >= 0;
      // rh fingers 0..4 -> sites 0..4, lh fingers 5..9 -> sites 5..9; -1 wraps like python (rh site 4)
      const int sidx = has ? (finger < 0 ? 4 : min(finger, 9)) : 0;
      const f3 tip = shfl3(site, sidx);
      const int kk = has ? key : 0;
      // the key angle lives in the owner lane's registers
      const float q0 = __shfl_sync(FULL, qkp[0], kk & 31), q1 = __shfl_sync(FULL, qkp[1], kk & 31),
                  q2 = __shfl_sync(FULL, qkp[2], kk & 31);
      float val = 0.f;
      if (has) {
        const float kqv = (kk >> 5) == 0 ? q0 : (kk >> 5) == 1 ? q1 : q2;
        float sq_, cq_;
        pp_sincos_small(kqv, &sq_, &cq_);
        const float hx = kt.half[kk][0];
        // key geom centre (rotated about the hinge) + world-axis offsets (0.35 size_x, 0, 0.5 size_z)
        const f3 kp = mk3(kt.pos[kk][0] - hx + hx * cq_ + 0.35f * hx, kt.pos[kk][1],
                          kt.pos[kk][2] - hx * sq_ + 0.5f * kt.half[kk][2]);
        val = pp_tolerance(norm(kp - tip), 0.01f, 0.1f);
      }
      tot += wsum(val);
      cnt += __popc(__ballot_sync(FULL, has));
    }
    fing = cnt > 0 ? tot / (float)cnt : 0.f;
  } else if (A.has_fingering == 0) {
    // unfingered MIDI: optimal-transport fingering reward. The kinematics scratch is dead by now.
    __syncwarp();
    fing = do_reset ? 0.f : ot_fingering_reward(&W.a.s.Z[0][0], lane, site, qkp, goal_cur, kt);
  }

  const bool should_terminate = !do_reset && t_idx == T;
  bool done = should_terminate;
  float discount = 1.f;
  if (!do_reset && A.wrong_press_termination && failure && !should_terminate) { done = true; discount = 0.f; }

  pp_eval_metrics(A.eval, env, lane, active, goal_cur, sustain_on, do_reset, done);
  // ---- write back state
#pragma unroll
  for (int h = 0; h < 2; h++)
    if (L.dof) {
      A.qpos[sbase + PP_NKEY + h * PP_HD + lane] = q[h];
      A.qvel[sbase + PP_NKEY + h * PP_HD + lane] = v[h];
      A.warm[sbase + PP_NKEY + h * PP_HD + lane] = acc[h];
      A.qacc[sbase + PP_NKEY + h * PP_HD + lane] = acc[h];
    }
#pragma unroll
  for (int i = 0; i < 3; i++) {
    const int k = lane + 32 * i;
    if (k < PP_NKEY) {
      A.qpos[sbase + k] = qk[i]; A.qvel[sbase + k] = vk[i]; A.warm[sbase + k] = ak[i]; A.qacc[sbase + k] = ak[i];
      A.activation[(size_t)env * PP_NKEY + k] = active[i] ? 1 : 0;
    }
  }
#pragma unroll
  for (int h = 0; h < 2; h++)
    if (lane < 22) {
      A.act_force[(size_t)env * PP_NU_HANDS + h * 22 + lane] = act_force[h];
      A.act_vel[(size_t)env * PP_NU_HANDS + h * 22 + lane] = act_vel[h];
    }
  if (A.stick_bits) {
#pragma unroll
    for (int h = 0; h < 2; h++) {
      const unsigned b = __ballot_sync(FULL, stick[h] && !do_reset);
      if (lane == 0) A.stick_bits[(size_t)env * 2 + h] = b;
    }
  }
  if (lane == 0) {
    A.t_idx[env] = t_idx;
    A.sustain[env] = sustain;
    A.needs_reset[env] = done ? 1 : 0;
    int* cn = A.counters + (size_t)env * 8;
    cn[0] = cnt_ncon; cn[2] = cnt_iter; cn[4] = sum_iter; cn[7] = sum_coupled;
    if (!kSpTiming || MODE == 0) { cn[1] = cnt_nefc; cn[3] = cnt_drop; cn[5] = sum_ls; cn[6] = sum_dense; }
    if (A.mode == PP_MODE_STEP) {
      float terms[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
      if (!do_reset) {
        // CompositeReward.compute: sum in insertion order (composite_reward.py:46-56)
        terms[0] = (n_on > 0 ? 0.5f * (kp_sum / (float)n_on) : 0.f) + 0.5f * (failure ? 0.f : 1.f);
        terms[1] = pp_tolerance((float)goal_cur[88] - (sustain_on ? 1.f : 0.f), 0.05f, 0.5f);
        terms[2] = 0.f - A.energy_coef * energy0 - A.energy_coef * energy1;
        terms[3] = fing;
        terms[4] = A.disable_forearm_reward ? 0.f : (forearm_hit ? 0.f : 0.5f);
      }
      float rew = 0.f;
      for (int j = 0; j < 5; j++) rew += terms[j];
      A.reward[env] = rew;
      if (A.reward_terms) for (int j = 0; j < 5; j++) A.reward_terms[(size_t)env * 5 + j] = terms[j];
      A.done[env] = done ? 1 : 0;
      if (A.discount) A.discount[env] = discount;
    }
  }
  // ---- observation: goal | fingering | piano/state | sustain_state | rh joints | lh joints
  if (A.obs) {
    float* obs = A.obs + (size_t)env * OD;
    const int t0 = t_idx == T ? T - 1 : t_idx;  // both updates are skipped after the last frame
    const int ng = (Lk + 1) * 89;
    for (int i = lane; i < ng; i += 32) {
      const int f = t0 + i / 89;
      obs[i] = f < T ? (float)A.goal[(size_t)(song_off + f) * 89 + (i % 89)] : 0.f;
    }
    int off = ng;
    if (A.has_fingering == 1) {
      // (2,5) flags of the fingers used in frame t0 (:391-412)
      unsigned used = 0u;
      for (int p = 0; p < A.pmax; p++) {
        const int16_t* pr = A.pairs + ((size_t)(song_off + t0) * A.pmax + p) * 2;
        if (pr[0] >= 0) {
          const int f = pr[1];
          const int slot = f < 5 ? (f < 0 ? f + 5 : f) : 5 + min(f - 5, 4);  // python negative index wraps
          used |= 1u << slot;
        }
      }
      if (lane < 10) obs[off + lane] = (used >> lane) & 1u ? 1.f : 0.f;
      off += 10;
    }
#pragma unroll
    for (int i = 0; i < 3; i++) {
      const int k = lane + 32 * i;
      if (k < PP_NKEY) obs[off + k] = normalized[i];
    }
    if (lane == 0) obs[off + 88] = sustain;
    off += 89;
    // hand.joints order: the 24 menagerie joints, then forearm_tx, forearm_ty
#pragma unroll
    for (int h = 0; h < 2; h++)
      if (L.dof) obs[off + h * 26 + (lane >= 2 ? lane - 2 : 24 + lane)] = q[h];
  }
}

template <int WPB, int MODE = 0>
__global__ void __launch_bounds__(32 * WPB, 10 / WPB) pp_hands_step_kernel(HandsArgs A, int env_base) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  WarpShared& W = reinterpret_cast<WarpShared*>(smem_raw)[threadIdx.x >> 5];
  if (MODE == PP_SPLIT_NEWTON) {
    // Persistent warps: each takes the next env of the substep's list whenever its env's loop is over, so no warp
    // waits for the slowest env of a block. The block's warps still meet at the top of every trip - left to
    // themselves they spread over the loop body and wait for instructions (measured: 59 % of the stall samples) - but
    // a warp spends the meeting's round on its own env's next trip, not idle. A warp that finds the list empty
    // keeps voting until all have.
    // The lists are taken one after the other, heaviest class first: the warps of a block then mostly hold envs of
    // one class, whose trips take similar times.
    int cnt[PP_SP_BUCKETS], n = 0;
#pragma unroll
    for (int k = 0; k < PP_SP_BUCKETS; k++) { cnt[k] = A.sp_count_in[k]; n += cnt[k]; }
    for (;;) {
      int i = 0;
      if ((threadIdx.x & 31) == 0) i = atomicAdd(A.sp_cursor, 1);
      i = __shfl_sync(FULL, i, 0);
      if (i >= n) break;
      int k = 0;
#pragma unroll
      for (int b = 0; b < PP_SP_BUCKETS - 1; b++) if (k == b && i >= cnt[b]) { i -= cnt[b]; k = b + 1; }
      hands_step_env<WPB, MODE>(A, W, A.sp_list_in[(size_t)k * A.num_envs + i]);
      __syncwarp();
    }
    if (WPB > 1) while (!__syncthreads_and(1)) {}
  } else {
    const int slot = env_base + blockIdx.x * WPB + (threadIdx.x >> 5);
    if (MODE != 0 && slot >= A.num_envs) {  // split launches round the grid up: the spare warps of the last block
      if (WPB > 1) __syncthreads_and(0);    // answer the entry vote (the block then runs without meetings) and leave
      return;
    }
    hands_step_env<WPB, MODE>(A, W, A.order ? A.order[slot] : slot);
  }
}

// The shared-memory attribute is per device and per kernel: set on every device this process launches on
// (a second handle on another GPU of the same process must not inherit the first one's "configured").
template <int WPB, int MODE = 0>
static void configure_kernel() {
  static bool configured[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    cudaFuncSetAttribute(pp_hands_step_kernel<WPB, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                         (int)(sizeof(WarpShared) * WPB));
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
}
template <int WPB, int MODE = 0>
static int launch_wpb(const HandsArgs& A, cudaStream_t stream, int env_begin, int env_count) {
  const size_t smem = sizeof(WarpShared);
  configure_kernel<WPB, MODE>();
  configure_kernel<1, MODE>();
  const int nblk = env_count / WPB, tail = env_count - nblk * WPB;
  if (nblk) pp_hands_step_kernel<WPB, MODE><<<nblk, 32 * WPB, smem * WPB, stream>>>(A, env_begin);
  if (WPB > 1 && tail) pp_hands_step_kernel<1, MODE><<<tail, 32, smem, stream>>>(A, env_begin + nblk * WPB);
  return (nblk ? 1 : 0) + (WPB > 1 && tail ? 1 : 0);
}
#ifdef PP_HANDS_SPLIT_TU
#ifndef PP_SP_NEWTON_WPB
#define PP_SP_NEWTON_WPB 10
#endif
#ifndef PP_SP_EDGE_WPB
#define PP_SP_EDGE_WPB 10
#endif
// One control step of the split schedule: per substep the Newton launch (one block per SM slot, warps pull envs from the
// lists) and, between two of them, one launch that finishes a substep and runs the next one's prologue (which appends
// every stepping env to its list); the first prologue and the last finish (with the task epilogue) stand alone. counts: [n_substeps,
// PP_SP_BUCKETS + 1] ints (envs per list, cursor), zeroed here; list: [PP_SP_BUCKETS, num_envs].
int pp_launch_hands_split(const HandsArgs& A0, cudaStream_t stream, int sms, int* list, int* counts) {
  constexpr int WN = PP_SP_NEWTON_WPB, WE = PP_SP_EDGE_WPB, CS = PP_SP_BUCKETS + 1;
  HandsArgs A = A0;
  const int N = A.num_envs, nsub = A.n_substeps;
  cudaMemsetAsync(counts, 0, sizeof(int) * (size_t)nsub * CS, stream);
  configure_kernel<WN, PP_SPLIT_NEWTON>();
  configure_kernel<WE, PP_SPLIT_PRO>();
  configure_kernel<WE, PP_SPLIT_FIN>();
  const size_t smem = sizeof(WarpShared);
  const int grid_n = std::min((N + WN - 1) / WN, sms * (10 / WN)), grid_e = (N + WE - 1) / WE;
  A.sp_list_out = list; A.sp_list_in = list;
  for (int s = 0; s < nsub; s++) {
    // substep s's prologue: on its own first, then on the tail of the launch that finishes substep s - 1
    A.sub_index = s ? s - 1 : 0; A.sp_chain = s ? 1 : 0;
    A.sp_count_out = counts + CS * s;
    if (s == 0) pp_hands_step_kernel<WE, PP_SPLIT_PRO><<<grid_e, 32 * WE, smem * WE, stream>>>(A, 0);
    else pp_hands_step_kernel<WE, PP_SPLIT_FIN><<<grid_e, 32 * WE, smem * WE, stream>>>(A, 0);
    A.sub_index = s;
    A.sp_count_in = counts + CS * s; A.sp_cursor = counts + CS * s + PP_SP_BUCKETS;
    pp_hands_step_kernel<WN, PP_SPLIT_NEWTON><<<grid_n, 32 * WN, smem * WN, stream>>>(A, 0);
  }
  A.sub_index = nsub - 1; A.sp_chain = 0;
  pp_hands_step_kernel<WE, PP_SPLIT_FIN><<<grid_e, 32 * WE, smem * WE, stream>>>(A, 0);
  return 2 * nsub + 1;
}
#else
#ifndef PP_WPB
#define PP_WPB 2
#endif
// envs per block, measured on B200 (profiles/):
//   every env doing the same work (config 2 replay, env-steps/s): 2 -> 1 175 k, 5 -> 893 k, 10 -> 885 k
//   different actions per env (config 3, B = 16384, ms per step):  2 -> 45.8, 5 -> 33.4, 10 -> 31.8
// so both shapes are built and pp_set_schedule (or its auto-tuner) picks one per handle.
constexpr int WARPS_PER_BLOCK = PP_WPB;
#ifndef PP_WPB_DIVERGENT
#define PP_WPB_DIVERGENT 10
#endif
constexpr int WARPS_PER_BLOCK_DIVERGENT = PP_WPB_DIVERGENT;
int pp_launch_hands(const HandsArgs& A, cudaStream_t stream, int env_begin, int env_count, int block_envs) {
  if (env_count < 0) env_count = A.num_envs - env_begin;
  if (block_envs > WARPS_PER_BLOCK)
    return launch_wpb<WARPS_PER_BLOCK_DIVERGENT>(A, stream, env_begin, env_count);
  return launch_wpb<WARPS_PER_BLOCK>(A, stream, env_begin, env_count);
}

// Counting sort of the envs by the Newton iterations of their previous control step, heaviest first: the envs of
// a block then need similar numbers of trips (less waiting at the meetings) and the longest blocks start first.
// The order only changes which block computes an env, never its result. One block; N up to a few 100 k.
__global__ void __launch_bounds__(1024) pp_work_order_kernel(const int* __restrict__ counters, int* __restrict__ order,
                                                            int n) {
  __shared__ int hist[256], base[256];
  for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0;
  __syncthreads();
  for (int e = threadIdx.x; e < n; e += blockDim.x) atomicAdd(&hist[min(max(counters[(size_t)e * 8 + 4], 0), 255)], 1);
  __syncthreads();
  if (threadIdx.x == 0) {
    int acc = 0;
    for (int k = 255; k >= 0; k--) { base[k] = acc; acc += hist[k]; }
  }
  __syncthreads();
  for (int e = threadIdx.x; e < n; e += blockDim.x) {
    const int k = min(max(counters[(size_t)e * 8 + 4], 0), 255);
    order[atomicAdd(&base[k], 1)] = e;
  }
}
void pp_launch_work_order(const int* counters, int* order, int num_envs, cudaStream_t stream) {
  pp_work_order_kernel<<<1, 1024, 0, stream>>>(counters, order, num_envs);
}
size_t pp_hands_smem_bytes() { return sizeof(WarpShared); }
size_t pp_hands_split_smem_words() { return sizeof(WarpShared) / 16; }
#endif  // PP_HANDS_SPLIT_TU
