/* piano_oracle.c - see piano_oracle.h. TEST INFRASTRUCTURE ONLY ("parity unpinned" vs mj_step).
 *
 * Written for clarity: dense matrices, world-origin spatial algebra, no tree sparsity.
 * Stage names follow MuJoCo's mjData fields so each can be diffed against the CUDA path.
 */
#include "piano_oracle.h"

#include <math.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define MINVAL 1e-15
#define MINIMP 0.0001
#define MAXIMP 0.9999

/* ------------------------------------------------------------------ small math */
static void v3set(double *r, double a, double b, double c) { r[0] = a; r[1] = b; r[2] = c; }
static void v3cpy(double *r, const double *a) { r[0] = a[0]; r[1] = a[1]; r[2] = a[2]; }
static void v3add(double *r, const double *a, const double *b) { r[0] = a[0] + b[0]; r[1] = a[1] + b[1]; r[2] = a[2] + b[2]; }
static void v3sub(double *r, const double *a, const double *b) { r[0] = a[0] - b[0]; r[1] = a[1] - b[1]; r[2] = a[2] - b[2]; }
static void v3axpy(double *r, double s, const double *a) { r[0] += s * a[0]; r[1] += s * a[1]; r[2] += s * a[2]; }
static double v3dot(const double *a, const double *b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static void v3cross(double *r, const double *a, const double *b) {
  double x = a[1] * b[2] - a[2] * b[1], y = a[2] * b[0] - a[0] * b[2], z = a[0] * b[1] - a[1] * b[0];
  r[0] = x; r[1] = y; r[2] = z;
}
static double v3norm(const double *a) { return sqrt(v3dot(a, a)); }

static void qmulq(double *r, const double *a, const double *b) {
  double w = a[0] * b[0] - a[1] * b[1] - a[2] * b[2] - a[3] * b[3];
  double x = a[0] * b[1] + a[1] * b[0] + a[2] * b[3] - a[3] * b[2];
  double y = a[0] * b[2] - a[1] * b[3] + a[2] * b[0] + a[3] * b[1];
  double z = a[0] * b[3] + a[1] * b[2] - a[2] * b[1] + a[3] * b[0];
  r[0] = w; r[1] = x; r[2] = y; r[3] = z;
}
static void q2mat(double *m, const double *q) {
  double w = q[0], x = q[1], y = q[2], z = q[3];
  m[0] = 1 - 2 * (y * y + z * z); m[1] = 2 * (x * y - w * z); m[2] = 2 * (x * z + w * y);
  m[3] = 2 * (x * y + w * z); m[4] = 1 - 2 * (x * x + z * z); m[5] = 2 * (y * z - w * x);
  m[6] = 2 * (x * z - w * y); m[7] = 2 * (y * z + w * x); m[8] = 1 - 2 * (x * x + y * y);
}
static void m3v(double *r, const double *m, const double *v) {
  double x = m[0] * v[0] + m[1] * v[1] + m[2] * v[2];
  double y = m[3] * v[0] + m[4] * v[1] + m[5] * v[2];
  double z = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
static void m3tv(double *r, const double *m, const double *v) { /* m^T v */
  double x = m[0] * v[0] + m[3] * v[1] + m[6] * v[2];
  double y = m[1] * v[0] + m[4] * v[1] + m[7] * v[2];
  double z = m[2] * v[0] + m[5] * v[1] + m[8] * v[2];
  r[0] = x; r[1] = y; r[2] = z;
}
static void qnormalize(double *q) {
  double n = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  for (int i = 0; i < 4; i++) q[i] /= n;
}

/* ------------------------------------------------------------------ data */
static double *zalloc(size_t n) { return (double *)calloc(n ? n : 1, sizeof(double)); }

po_data *po_data_create(const pp_model *m) {
  po_data *d = (po_data *)calloc(1, sizeof(po_data));
  int nv = m->nv, nb = m->nbody, ng = m->ngeom;
  d->nv = nv; d->nbody = nb; d->ngeom = ng; d->nsite = m->nsite; d->nu = m->nu;
  d->qpos = zalloc(nv); d->qvel = zalloc(nv); d->qacc_warmstart = zalloc(nv);
  d->ctrl = zalloc(m->nu); d->qfrc_applied = zalloc(nv);
  d->xpos = zalloc(3 * nb); d->xquat = zalloc(4 * nb); d->xmat = zalloc(9 * nb);
  d->xipos = zalloc(3 * nb); d->ximat = zalloc(9 * nb);
  d->xanchor = zalloc(3 * nv); d->xaxis = zalloc(3 * nv);
  d->geom_xpos = zalloc(3 * ng); d->geom_xmat = zalloc(9 * ng); d->site_xpos = zalloc(3 * m->nsite);
  d->qM = zalloc((size_t)nv * nv);
  d->qfrc_bias = zalloc(nv); d->qfrc_passive = zalloc(nv);
  d->actuator_force = zalloc(m->nu); d->actuator_velocity = zalloc(m->nu);
  d->qfrc_actuator = zalloc(nv); d->qfrc_smooth = zalloc(nv); d->qacc_smooth = zalloc(nv);
  d->qacc = zalloc(nv); d->qfrc_constraint = zalloc(nv);
  po_reset(m, d);
  return d;
}

void po_data_free(po_data *d) {
  if (!d) return;
  double *p[] = {d->qpos, d->qvel, d->qacc_warmstart, d->ctrl, d->qfrc_applied, d->xpos, d->xquat,
                 d->xmat, d->xipos, d->ximat, d->xanchor, d->xaxis, d->geom_xpos, d->geom_xmat,
                 d->site_xpos, d->qM, d->qfrc_bias, d->qfrc_passive, d->actuator_force,
                 d->actuator_velocity, d->qfrc_actuator, d->qfrc_smooth, d->qacc_smooth, d->qacc,
                 d->qfrc_constraint};
  for (size_t i = 0; i < sizeof(p) / sizeof(p[0]); i++) free(p[i]);
  free(d->ws); /* one block: see ws_get */
  free(d);
}

void po_reset(const pp_model *m, po_data *d) { /* mj_resetData */
  for (int i = 0; i < m->nv; i++) {
    d->qpos[i] = m->qpos0[i]; d->qvel[i] = 0; d->qacc_warmstart[i] = 0; d->qfrc_applied[i] = 0;
    d->qacc[i] = 0;
  }
  for (int i = 0; i < m->nu; i++) { d->ctrl[i] = 0; d->actuator_force[i] = 0; }
  d->ncon = d->nefc = d->ncon_dropped = 0;
}

/* ------------------------------------------------------------------ mj_kinematics */
static void kinematics(const pp_model *m, po_data *d) {
  v3set(d->xpos, 0, 0, 0);
  d->xquat[0] = 1; d->xquat[1] = d->xquat[2] = d->xquat[3] = 0;
  q2mat(d->xmat, d->xquat);
  for (int b = 1; b < m->nbody; b++) {
    int p = m->body_parent[b];
    double pos[3], quat[4], mat[9], t[3];
    m3v(t, d->xmat + 9 * p, m->body_pos + 3 * b);
    v3add(pos, d->xpos + 3 * p, t);
    qmulq(quat, d->xquat + 4 * p, m->body_quat + 4 * b);
    for (int k = 0; k < m->body_dofnum[b]; k++) {
      int j = m->body_dofadr[b] + k;
      double q = d->qpos[j] - m->qpos0[j];
      q2mat(mat, quat);
      m3v(t, mat, m->jnt_pos + 3 * j);
      v3add(d->xanchor + 3 * j, pos, t);
      m3v(d->xaxis + 3 * j, mat, m->jnt_axis + 3 * j);
      if (m->jnt_type[j] == PP_JNT_SLIDE) {
        v3axpy(pos, q, d->xaxis + 3 * j);
      } else { /* hinge: rotate about the local axis, then keep the anchor fixed */
        double s = sin(0.5 * q), ql[4] = {cos(0.5 * q), s * m->jnt_axis[3 * j], s * m->jnt_axis[3 * j + 1], s * m->jnt_axis[3 * j + 2]};
        double nq[4];
        qmulq(nq, quat, ql);
        memcpy(quat, nq, sizeof(nq));
        q2mat(mat, quat);
        m3v(t, mat, m->jnt_pos + 3 * j);
        v3sub(pos, d->xanchor + 3 * j, t);
      }
    }
    qnormalize(quat);
    v3cpy(d->xpos + 3 * b, pos);
    memcpy(d->xquat + 4 * b, quat, sizeof(quat));
    q2mat(d->xmat + 9 * b, quat);
    m3v(t, d->xmat + 9 * b, m->body_ipos + 3 * b);
    v3add(d->xipos + 3 * b, pos, t);
    double iq[4];
    qmulq(iq, quat, m->body_iquat + 4 * b);
    q2mat(d->ximat + 9 * b, iq);
  }
  for (int g = 0; g < m->ngeom; g++) {
    int b = m->geom_body[g];
    double t[3], gq[4];
    m3v(t, d->xmat + 9 * b, m->geom_pos + 3 * g);
    v3add(d->geom_xpos + 3 * g, d->xpos + 3 * b, t);
    qmulq(gq, d->xquat + 4 * b, m->geom_quat + 4 * g);
    q2mat(d->geom_xmat + 9 * g, gq);
  }
  for (int s = 0; s < m->nsite; s++) {
    int b = m->site_body[s];
    double t[3];
    m3v(t, d->xmat + 9 * b, m->site_pos + 3 * s);
    v3add(d->site_xpos + 3 * s, d->xpos + 3 * b, t);
  }
}

/* ------------------------------------------------------------------ spatial algebra
 * Everything is expressed about the WORLD ORIGIN: motion (w; vO), force (nO; f). */
static void dof_motion(const pp_model *m, const po_data *d, int j, double *S) {
  const double *ax = d->xaxis + 3 * j;
  if (m->jnt_type[j] == PP_JNT_HINGE) {
    v3cpy(S, ax);
    v3cross(S + 3, d->xanchor + 3 * j, ax);
  } else {
    v3set(S, 0, 0, 0);
    v3cpy(S + 3, ax);
  }
}
static void cross_motion(double *r, const double *v, const double *s) {
  double a[3], b[3];
  v3cross(r, v, s);
  v3cross(a, v, s + 3);
  v3cross(b, v + 3, s);
  v3add(r + 3, a, b);
}
static void cross_force(double *r, const double *v, const double *f) {
  double a[3], b[3];
  v3cross(a, v, f);
  v3cross(b, v + 3, f + 3);
  v3add(r, a, b);
  v3cross(r + 3, v, f + 3);
}
static void body_inertia6(const pp_model *m, const po_data *d, int b, double *I6) {
  double mass = m->body_mass[b];
  const double *R = d->ximat + 9 * b, *c = d->xipos + 3 * b, *di = m->body_inertia + 3 * b;
  double Ic[9];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++)
      Ic[3 * i + j] = R[3 * i] * di[0] * R[3 * j] + R[3 * i + 1] * di[1] * R[3 * j + 1] + R[3 * i + 2] * di[2] * R[3 * j + 2];
  double C[9] = {0, -c[2], c[1], c[2], 0, -c[0], -c[1], c[0], 0};
  memset(I6, 0, 36 * sizeof(double));
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      double cc = 0;
      for (int k = 0; k < 3; k++) cc += C[3 * i + k] * C[3 * k + j];
      I6[6 * i + j] = Ic[3 * i + j] - mass * cc;
      I6[6 * i + 3 + j] = mass * C[3 * i + j];
      I6[6 * (3 + i) + j] = mass * C[3 * j + i];
    }
  for (int i = 0; i < 3; i++) I6[6 * (3 + i) + 3 + i] = mass;
}
static void m6v(double *r, const double *M, const double *v) {
  for (int i = 0; i < 6; i++) {
    double s = 0;
    for (int k = 0; k < 6; k++) s += M[6 * i + k] * v[k];
    r[i] = s;
  }
}
static double v6dot(const double *a, const double *b) {
  double s = 0;
  for (int i = 0; i < 6; i++) s += a[i] * b[i];
  return s;
}

/* mj_crb: composite rigid body algorithm -> dense qM (armature included) */
static void crb_buf(const pp_model *m, po_data *d, double *Ic) {
  int nb = m->nbody, nv = m->nv;
  for (int b = 0; b < nb; b++) body_inertia6(m, d, b, Ic + 36 * b);
  for (int b = nb - 1; b > 0; b--) {
    int p = m->body_parent[b];
    for (int k = 0; k < 36; k++) Ic[36 * p + k] += Ic[36 * b + k];
  }
  memset(d->qM, 0, sizeof(double) * nv * nv);
  for (int i = 0; i < nv; i++) {
    double S[6], F[6], Sj[6];
    dof_motion(m, d, i, S);
    m6v(F, Ic + 36 * m->jnt_body[i], S);
    d->qM[i * nv + i] = v6dot(S, F) + m->dof_armature[i];
    for (int j = m->dof_parent[i]; j >= 0; j = m->dof_parent[j]) {
      dof_motion(m, d, j, Sj);
      d->qM[i * nv + j] = d->qM[j * nv + i] = v6dot(Sj, F);
    }
  }
}
static void crb(const pp_model *m, po_data *d) {
  double *Ic = (double *)malloc(sizeof(double) * 36 * m->nbody);
  crb_buf(m, d, Ic);
  free(Ic);
}

/* mj_rne(flg_acc=0): Coriolis, centrifugal and gravity forces */
static void rne_bias_buf(const pp_model *m, po_data *d, double *vel, double *acc, double *frc) {
  int nb = m->nbody;
  memset(vel, 0, sizeof(double) * 6 * nb); memset(acc, 0, sizeof(double) * 6 * nb); memset(frc, 0, sizeof(double) * 6 * nb);
  for (int k = 0; k < 3; k++) acc[3 + k] = -m->gravity[k];
  for (int b = 1; b < nb; b++) {
    int p = m->body_parent[b];
    double v[6], a[6], S[6], Sd[6], I6[36], h[6], t[6];
    memcpy(v, vel + 6 * p, sizeof(v));
    memcpy(a, acc + 6 * p, sizeof(a));
    for (int k = 0; k < m->body_dofnum[b]; k++) {
      int j = m->body_dofadr[b] + k;
      dof_motion(m, d, j, S);
      cross_motion(Sd, v, S);
      for (int i = 0; i < 6; i++) { a[i] += Sd[i] * d->qvel[j]; v[i] += S[i] * d->qvel[j]; }
    }
    memcpy(vel + 6 * b, v, sizeof(v));
    memcpy(acc + 6 * b, a, sizeof(a));
    body_inertia6(m, d, b, I6);
    m6v(frc + 6 * b, I6, a);
    m6v(h, I6, v);
    cross_force(t, v, h);
    for (int i = 0; i < 6; i++) frc[6 * b + i] += t[i];
  }
  for (int b = nb - 1; b > 0; b--) {
    double S[6];
    for (int k = 0; k < m->body_dofnum[b]; k++) {
      int j = m->body_dofadr[b] + k;
      dof_motion(m, d, j, S);
      d->qfrc_bias[j] = v6dot(S, frc + 6 * b);
    }
    int p = m->body_parent[b];
    for (int i = 0; i < 6; i++) frc[6 * p + i] += frc[6 * b + i];
  }
}
static void rne_bias(const pp_model *m, po_data *d) {
  int nb = m->nbody;
  double *buf = (double *)malloc(sizeof(double) * 18 * nb);
  rne_bias_buf(m, d, buf, buf + 6 * nb, buf + 12 * nb);
  free(buf);
}

/* translational Jacobian row block: jac[3][nv] of a world point attached to body b */
static void jac_point(const pp_model *m, const po_data *d, int b, const double *point, double *jac) {
  int nv = m->nv;
  memset(jac, 0, sizeof(double) * 3 * nv);
  while (b > 0) {
    for (int k = 0; k < m->body_dofnum[b]; k++) {
      int j = m->body_dofadr[b] + k;
      double col[3];
      if (m->jnt_type[j] == PP_JNT_HINGE) {
        double r[3];
        v3sub(r, point, d->xanchor + 3 * j);
        v3cross(col, d->xaxis + 3 * j, r);
      } else {
        v3cpy(col, d->xaxis + 3 * j);
      }
      for (int i = 0; i < 3; i++) jac[i * nv + j] = col[i];
    }
    b = m->body_parent[b];
  }
}

/* mj_passive: joint springs, dampers, gravity compensation */
static void passive_buf(const pp_model *m, po_data *d, double *jac) {
  int nv = m->nv;
  for (int j = 0; j < nv; j++)
    d->qfrc_passive[j] = -m->jnt_stiffness[j] * (d->qpos[j] - m->jnt_springref[j]) - m->dof_damping[j] * d->qvel[j];
  for (int b = 1; b < m->nbody; b++) {
    if (m->body_gravcomp[b] == 0) continue;
    double f[3];
    for (int i = 0; i < 3; i++) f[i] = -m->gravity[i] * m->body_mass[b] * m->body_gravcomp[b];
    jac_point(m, d, b, d->xipos + 3 * b, jac);
    for (int j = 0; j < nv; j++) d->qfrc_passive[j] += jac[j] * f[0] + jac[nv + j] * f[1] + jac[2 * nv + j] * f[2];
  }
}
static void passive(const pp_model *m, po_data *d) {
  double *jac = (double *)malloc(sizeof(double) * 3 * m->nv);
  passive_buf(m, d, jac);
  free(jac);
}

/* mj_fwdActuation for position / fixed-gain actuators on joints or 2-joint fixed tendons */
static void actuation(const pp_model *m, po_data *d) {
  memset(d->qfrc_actuator, 0, sizeof(double) * m->nv);
  for (int i = 0; i < m->nu; i++) {
    int j0 = m->act_trn0[i], j1 = m->act_trn1[i];
    double c = d->ctrl[i], lo = m->act_ctrlrange[2 * i], hi = m->act_ctrlrange[2 * i + 1];
    c = c < lo ? lo : (c > hi ? hi : c);
    double len = d->qpos[j0] + (j1 >= 0 ? d->qpos[j1] : 0.0);
    double f = m->act_gain_only[i] ? m->act_kp[i] * c : m->act_kp[i] * c - m->act_kp[i] * len;
    if (m->act_forcelimited[i]) {
      double flo = m->act_forcerange[2 * i], fhi = m->act_forcerange[2 * i + 1];
      f = f < flo ? flo : (f > fhi ? fhi : f);
    }
    d->actuator_force[i] = f;
    d->qfrc_actuator[j0] += f;
    if (j1 >= 0) d->qfrc_actuator[j1] += f;
  }
}

static void actuator_velocity(const pp_model *m, po_data *d) {
  for (int i = 0; i < m->nu; i++) {
    int j0 = m->act_trn0[i], j1 = m->act_trn1[i];
    d->actuator_velocity[i] = d->qvel[j0] + (j1 >= 0 ? d->qvel[j1] : 0.0);
  }
}

/* dense Cholesky A = L L^T in place (lower); returns 0 on success */
static int chol_factor(double *A, int n) {
  for (int j = 0; j < n; j++) {
    double s = A[j * n + j];
    for (int k = 0; k < j; k++) s -= A[j * n + k] * A[j * n + k];
    if (s <= 0) return -1;
    s = sqrt(s);
    A[j * n + j] = s;
    for (int i = j + 1; i < n; i++) {
      double t = A[i * n + j];
      for (int k = 0; k < j; k++) t -= A[i * n + k] * A[j * n + k];
      A[i * n + j] = t / s;
    }
  }
  return 0;
}
static void chol_solve(const double *L, int n, double *x) {
  for (int i = 0; i < n; i++) {
    double s = x[i];
    for (int k = 0; k < i; k++) s -= L[i * n + k] * x[k];
    x[i] = s / L[i * n + i];
  }
  for (int i = n - 1; i >= 0; i--) {
    double s = x[i];
    for (int k = i + 1; k < n; k++) s -= L[k * n + i] * x[k];
    x[i] = s / L[i * n + i];
  }
}

/* ------------------------------------------------------------------ collision
 * MuJoCo's own narrow-phase routines (mjc_CapsuleBox etc.) are not available; these are
 * restated closest-feature tests with the contact conventions of mjContact: frame[0..2]
 * is the normal pointing from geom1 to geom2, dist < 0 is penetration, pos is the midpoint
 * between the two surfaces. The CUDA path implements the same rules. */
static void make_frame(double *frame) { /* mju_makeFrame with no y hint */
  double *x = frame, *y = frame + 3, *z = frame + 6;
  v3set(y, 0, 0, 0);
  if (x[1] < 0.5 && x[1] > -0.5) y[1] = 1; else y[2] = 1;
  double dt = v3dot(x, y);
  v3axpy(y, -dt, x);
  double n = v3norm(y);
  for (int i = 0; i < 3; i++) y[i] /= n;
  v3cross(z, x, y);
}

static void add_contact(po_data *d, int g1, int g2, double dist, const double *pos, const double *normal) {
  if (d->ncon >= PO_MAXCON) { d->ncon_dropped++; return; }
  po_contact *c = &d->contact[d->ncon++];
  c->geom1 = g1; c->geom2 = g2; c->dist = dist;
  v3cpy(c->pos, pos);
  v3cpy(c->frame, normal);
  make_frame(c->frame);
}

/* sphere (geom1) against box (geom2); c is the sphere centre in world coordinates */
static void sphere_box(po_data *d, int g1, int g2, const double *c, double r, const double *bpos,
                       const double *bmat, const double *h, double margin) {
  double rel[3], cl[3], q[3], nl[3], n[3], pos[3], dist;
  v3sub(rel, c, bpos);
  m3tv(cl, bmat, rel);
  int outside = 0;
  for (int i = 0; i < 3; i++) {
    q[i] = cl[i] < -h[i] ? -h[i] : (cl[i] > h[i] ? h[i] : cl[i]);
    if (q[i] != cl[i]) outside = 1;
  }
  double len = 0;
  if (outside) {
    v3sub(nl, q, cl);
    len = v3norm(nl);
    if (len < MINVAL) outside = 0;
  }
  if (outside) {
    dist = len - r;
    if (dist >= margin) return;
    for (int i = 0; i < 3; i++) nl[i] /= len;
  } else {
    int best = 0;
    double bestv = fabs(cl[0]) - h[0];
    for (int i = 1; i < 3; i++) {
      double v = fabs(cl[i]) - h[i];
      if (v > bestv) { bestv = v; best = i; }
    }
    v3set(nl, 0, 0, 0);
    nl[best] = cl[best] < 0 ? 1.0 : -1.0;
    dist = bestv - r; /* bestv = -(penetration of the centre) */
  }
  m3v(n, bmat, nl);
  v3cpy(pos, c);
  v3axpy(pos, r + 0.5 * dist, n);
  add_contact(d, g1, g2, dist, pos, n);
}

static double seg_box_dgrad(const double *al, const double *dv, const double *h, double t) {
  double g = 0;
  for (int i = 0; i < 3; i++) {
    double s = al[i] + t * dv[i];
    double e = s < -h[i] ? s + h[i] : (s > h[i] ? s - h[i] : 0.0);
    g += dv[i] * e;
  }
  return g;
}

/* capsule/sphere (geom1) against box (geom2): contact at the closest segment point, plus one
 * at the far segment end when it also penetrates (flat-lying capsule) */
static void capsule_box(const pp_model *m, po_data *d, int g1, int g2, double margin) {
  const double *cp = d->geom_xpos + 3 * g1, *cm = d->geom_xmat + 9 * g1;
  const double *bp = d->geom_xpos + 3 * g2, *bm = d->geom_xmat + 9 * g2, *h = m->geom_size + 3 * g2;
  double r = m->geom_size[3 * g1];
  double hl = m->geom_type[g1] == PP_GEOM_CAPSULE ? m->geom_size[3 * g1 + 1] : 0.0;
  double axis[3] = {cm[2], cm[5], cm[8]};
  if (hl == 0.0) { sphere_box(d, g1, g2, cp, r, bp, bm, h, margin); return; }
  double a[3], b[3], rel[3], al[3], bl[3], dv[3];
  v3cpy(a, cp); v3axpy(a, -hl, axis);
  v3cpy(b, cp); v3axpy(b, hl, axis);
  v3sub(rel, a, bp); m3tv(al, bm, rel);
  v3sub(rel, b, bp); m3tv(bl, bm, rel);
  v3sub(dv, bl, al);
  double lo = 0, hi = 1, glo = seg_box_dgrad(al, dv, h, 0), ghi = seg_box_dgrad(al, dv, h, 1), t;
  if (glo >= 0) t = 0;
  else if (ghi <= 0) t = 1;
  else {
    for (int i = 0; i < 3; i++) {
      if (fabs(dv[i]) <= 1e-12) continue;
      for (int sgn = -1; sgn <= 1; sgn += 2) {
        double tb = (sgn * h[i] - al[i]) / dv[i];
        if (tb > lo && tb < hi) {
          double g = seg_box_dgrad(al, dv, h, tb);
          if (g <= 0) { lo = tb; glo = g; } else { hi = tb; ghi = g; }
        }
      }
    }
    t = (ghi - glo > 0) ? lo + (hi - lo) * (-glo) / (ghi - glo) : lo;
  }
  double c[3];
  v3cpy(c, a); v3axpy(c, 2 * hl * t, axis);
  sphere_box(d, g1, g2, c, r, bp, bm, h, margin);
  sphere_box(d, g1, g2, t > 0.5 ? a : b, r, bp, bm, h, margin);
}

/* capsule/sphere against capsule/sphere: closest points of the two axis segments */
static void capsule_capsule(const pp_model *m, po_data *d, int g1, int g2, double margin) {
  const double *p1 = d->geom_xpos + 3 * g1, *m1 = d->geom_xmat + 9 * g1;
  const double *p2 = d->geom_xpos + 3 * g2, *m2 = d->geom_xmat + 9 * g2;
  double r1 = m->geom_size[3 * g1], r2 = m->geom_size[3 * g2];
  double h1 = m->geom_type[g1] == PP_GEOM_CAPSULE ? m->geom_size[3 * g1 + 1] : 0.0;
  double h2 = m->geom_type[g2] == PP_GEOM_CAPSULE ? m->geom_size[3 * g2 + 1] : 0.0;
  double u[3] = {m1[2], m1[5], m1[8]}, v[3] = {m2[2], m2[5], m2[8]}, w[3];
  v3sub(w, p1, p2);
  /* minimise |w + s u - t v|, s in [-h1,h1], t in [-h2,h2] (u, v unit) */
  double b = v3dot(u, v), dd = v3dot(u, w), e = v3dot(v, w), den = 1 - b * b, s, t;
  if (h1 == 0) s = 0;
  else if (den > 1e-9) { s = (b * e - dd) / den; s = s < -h1 ? -h1 : (s > h1 ? h1 : s); }
  else s = 0;
  t = h2 == 0 ? 0 : b * s + e;
  if (t < -h2) t = -h2; else if (t > h2) t = h2;
  if (h1 != 0) { s = b * t - dd; s = s < -h1 ? -h1 : (s > h1 ? h1 : s); }
  double c1[3], c2[3], n[3], pos[3];
  v3cpy(c1, p1); v3axpy(c1, s, u);
  v3cpy(c2, p2); v3axpy(c2, t, v);
  v3sub(n, c2, c1);
  double len = v3norm(n), dist = len - r1 - r2;
  if (dist >= margin) return;
  if (len < MINVAL) v3set(n, 0, 0, 1); else for (int i = 0; i < 3; i++) n[i] /= len;
  v3cpy(pos, c1); v3axpy(pos, r1 + 0.5 * dist, n);
  add_contact(d, g1, g2, dist, pos, n);
}

static double bound_radius(const pp_model *m, int g) {
  const double *s = m->geom_size + 3 * g;
  if (m->geom_type[g] == PP_GEOM_BOX) return sqrt(s[0] * s[0] + s[1] * s[1] + s[2] * s[2]);
  if (m->geom_type[g] == PP_GEOM_CAPSULE) return s[0] + s[1];
  return s[0];
}

static void collision(const pp_model *m, po_data *d) {
  d->ncon = 0; d->ncon_dropped = 0;
  if (!m->has_hands) return;
  for (int g = 0; g < m->ngeom; g++) {
    if (m->geom_cls[g] != PP_CLS_RH_FINGER && m->geom_cls[g] != PP_CLS_LH_FINGER) continue;
    for (int k = 0; k < m->ngeom; k++) {
      if (m->geom_cls[k] != PP_CLS_KEY && m->geom_cls[k] != PP_CLS_BASE) continue;
      double rel[3];
      v3sub(rel, d->geom_xpos + 3 * g, d->geom_xpos + 3 * k);
      double margin = m->pair_margin[PP_PAIRCLS_PIANO];
      if (v3norm(rel) > bound_radius(m, g) + bound_radius(m, k) + margin) continue;
      capsule_box(m, d, g, k, margin);
    }
  }
  for (int p = 0; p < m->npair; p++) {
    int g1 = m->pair_geom1[p], g2 = m->pair_geom2[p];
    double rel[3], margin = m->pair_margin[PP_PAIRCLS_HAND];
    v3sub(rel, d->geom_xpos + 3 * g1, d->geom_xpos + 3 * g2);
    if (v3norm(rel) > bound_radius(m, g1) + bound_radius(m, g2) + margin) continue;
    capsule_capsule(m, d, g1, g2, margin);
  }
}

void po_forward_position(const pp_model *m, po_data *d, int with_collision) {
  kinematics(m, d);
  if (with_collision) collision(m, d);
}

/* ------------------------------------------------------------------ constraints
 * Row order as mj_makeConstraint: dof frictionloss, joint limits, contacts (pyramidal). */
enum { ROW_FRICTION = 0, ROW_INEQ = 1 };

typedef struct {
  int nefc, nv, cap;
  double *J;     /* nefc x nv dense */
  double *D, *aref, *floss;
  int *type;
} efc_t;

static void efc_reserve(efc_t *e, int nv, int cap) {
  e->nv = nv; e->cap = cap; e->nefc = 0;
  e->J = (double *)calloc((size_t)cap * nv, sizeof(double));
  e->D = (double *)calloc(cap, sizeof(double));
  e->aref = (double *)calloc(cap, sizeof(double));
  e->floss = (double *)calloc(cap, sizeof(double));
  e->type = (int *)calloc(cap, sizeof(int));
}
static void efc_free(efc_t *e) { free(e->J); free(e->D); free(e->aref); free(e->floss); free(e->type); }

/* getimpedance(): position-dependent impedance d(r) */
static double impedance(const double *solimp, double pos, double margin) {
  double dmin = solimp[0], dmax = solimp[1], width = solimp[2], mid = solimp[3], power = solimp[4];
  dmin = fmin(fmax(dmin, MINIMP), MAXIMP); dmax = fmin(fmax(dmax, MINIMP), MAXIMP);
  width = fmax(width, MINVAL); mid = fmin(fmax(mid, MINIMP), MAXIMP); power = fmax(power, 1.0);
  if (dmin == dmax) return 0.5 * (dmin + dmax);
  double x = fabs((pos - margin) / width), y;
  if (x >= 1) return dmax;
  if (x <= 0) return dmin;
  if (power == 1) y = x;
  else if (x <= mid) y = pow(x, power) / pow(mid, power - 1);
  else y = 1 - pow(1 - x, power) / pow(1 - mid, power - 1);
  return dmin + y * (dmax - dmin);
}

/* solref -> (K, B) with the refsafe clamp timeconst >= 2*timestep */
static void stiffness_damping(const pp_model *m, const double *solref, const double *solimp, double *K, double *B) {
  double tc = solref[0], dr = solref[1], dmax = fmin(fmax(solimp[1], MINIMP), MAXIMP);
  if (tc > 0) {
    tc = fmax(tc, 2 * m->timestep);
    *K = 1 / fmax(MINVAL, dmax * dmax * tc * tc * dr * dr);
    *B = 2 / fmax(MINVAL, dmax * tc);
  } else {
    *K = -tc / fmax(MINVAL, dmax * dmax);
    *B = -dr / fmax(MINVAL, dmax);
  }
}

static double row_vel(const efc_t *e, int r, const double *qvel) {
  double s = 0;
  for (int j = 0; j < e->nv; j++) s += e->J[(size_t)r * e->nv + j] * qvel[j];
  return s;
}

static void make_constraint(const pp_model *m, po_data *d, efc_t *e) {
  int nv = m->nv;
  efc_reserve(e, nv, 3 * nv + 4 * d->ncon + 1);
  double K, B;
  /* frictionloss */
  stiffness_damping(m, m->friction_solref, m->friction_solimp, &K, &B);
  for (int j = 0; j < nv; j++) {
    if (m->dof_frictionloss[j] <= 0) continue;
    int r = e->nefc++;
    e->J[(size_t)r * nv + j] = 1;
    double imp = impedance(m->friction_solimp, 0, 0);
    double R = fmax(MINVAL, (1 - imp) * m->dof_invweight0[j] / imp);
    e->D[r] = 1 / R; e->type[r] = ROW_FRICTION; e->floss[r] = m->dof_frictionloss[j];
    e->aref[r] = -B * d->qvel[j];
  }
  /* joint limits (margin 0) */
  stiffness_damping(m, m->limit_solref, m->limit_solimp, &K, &B);
  for (int j = 0; j < nv; j++) {
    if (!m->jnt_limited[j]) continue;
    for (int side = 0; side < 2; side++) {
      double sgn = side == 0 ? 1.0 : -1.0;
      double dist = side == 0 ? d->qpos[j] - m->jnt_range[2 * j] : m->jnt_range[2 * j + 1] - d->qpos[j];
      if (dist >= 0) continue;
      int r = e->nefc++;
      e->J[(size_t)r * nv + j] = sgn;
      double imp = impedance(m->limit_solimp, dist, 0);
      double R = fmax(MINVAL, (1 - imp) * m->dof_invweight0[j] / imp);
      e->D[r] = 1 / R; e->type[r] = ROW_INEQ;
      e->aref[r] = -B * sgn * d->qvel[j] - K * imp * dist;
    }
  }
  /* contacts: condim 3, pyramidal cone -> 4 rows each */
  double *j1 = (double *)malloc(sizeof(double) * 3 * nv), *j2 = (double *)malloc(sizeof(double) * 3 * nv);
  for (int c = 0; c < d->ncon; c++) {
    const po_contact *con = &d->contact[c];
    int cls1 = m->geom_cls[con->geom1], cls2 = m->geom_cls[con->geom2];
    int pc = (cls1 <= PP_CLS_BASE || cls2 <= PP_CLS_BASE) ? PP_PAIRCLS_PIANO : PP_PAIRCLS_HAND;
    const double *solref = m->pair_solref[pc], *solimp = m->pair_solimp[pc];
    double mu = m->pair_friction[pc], margin = m->pair_margin[pc];
    int b1 = m->geom_body[con->geom1], b2 = m->geom_body[con->geom2];
    jac_point(m, d, b1, con->pos, j1);
    jac_point(m, d, b2, con->pos, j2);
    stiffness_damping(m, solref, solimp, &K, &B);
    double imp = impedance(solimp, con->dist, margin);
    double tran = m->body_invweight0[2 * b1] + m->body_invweight0[2 * b2];
    double diag = tran + mu * mu * tran;
    double R0 = fmax(MINVAL, (1 - imp) * diag / imp);
    double mureg = mu * sqrt(1 / m->impratio);
    double R = 2 * mureg * mureg * R0;
    for (int k = 0; k < 4; k++) {
      int r = e->nefc++;
      const double *n = con->frame, *t = con->frame + 3 * (1 + k / 2);
      double sg = (k % 2 == 0) ? mu : -mu;
      for (int j = 0; j < nv; j++) {
        double dn = 0, dt = 0;
        for (int i = 0; i < 3; i++) {
          double dj = j2[i * nv + j] - j1[i * nv + j];
          dn += n[i] * dj; dt += t[i] * dj;
        }
        e->J[(size_t)r * nv + j] = dn + sg * dt;
      }
      e->D[r] = 1 / R; e->type[r] = ROW_INEQ;
      e->aref[r] = -B * row_vel(e, r, d->qvel) - K * imp * (con->dist - margin);
    }
  }
  free(j1); free(j2);
  d->nefc = e->nefc;
}

/* per-row cost s(x), force f = -ds/dx, and curvature (D if in the quadratic zone else 0) */
static void row_eval(const efc_t *e, int r, double x, double *cost, double *force, double *curv) {
  double D = e->D[r];
  if (e->type[r] == ROW_INEQ) {
    if (x < 0) { *cost = 0.5 * D * x * x; *force = -D * x; *curv = D; }
    else { *cost = 0; *force = 0; *curv = 0; }
  } else {
    double f = e->floss[r], rf = f / D;
    if (x <= -rf) { *cost = -0.5 * rf * f - f * x; *force = f; *curv = 0; }
    else if (x >= rf) { *cost = -0.5 * rf * f + f * x; *force = -f; *curv = 0; }
    else { *cost = 0.5 * D * x * x; *force = -D * x; *curv = D; }
  }
}

static void matvec(const double *A, int n, const double *x, double *y) {
  for (int i = 0; i < n; i++) {
    double s = 0;
    for (int j = 0; j < n; j++) s += A[(size_t)i * n + j] * x[j];
    y[i] = s;
  }
}

/* total cost at acceleration a; optionally gradient and row forces */
static double total_cost(const pp_model *m, const po_data *d, const efc_t *e, const double *a,
                         double *grad, double *force, double *curv) {
  int nv = m->nv;
  double *Ma = (double *)malloc(sizeof(double) * nv);
  matvec(d->qM, nv, a, Ma);
  double cost = 0;
  for (int j = 0; j < nv; j++) cost += 0.5 * (Ma[j] - d->qfrc_smooth[j]) * (a[j] - d->qacc_smooth[j]);
  if (grad) for (int j = 0; j < nv; j++) grad[j] = Ma[j] - d->qfrc_smooth[j];
  for (int r = 0; r < e->nefc; r++) {
    double x = -e->aref[r], c, f, cu;
    const double *Jr = e->J + (size_t)r * nv;
    for (int j = 0; j < nv; j++) x += Jr[j] * a[j];
    row_eval(e, r, x, &c, &f, &cu);
    cost += c;
    if (force) force[r] = f;
    if (curv) curv[r] = cu;
    if (grad) for (int j = 0; j < nv; j++) grad[j] -= Jr[j] * f;
  }
  free(Ma);
  return cost;
}

/* mj_solNewton: primal Newton with exact line search */
static void solve_newton(const pp_model *m, po_data *d, efc_t *e) {
  int nv = m->nv, ne = e->nefc;
  double *a = d->qacc;
  /* warm start: the better of qacc_warmstart and qacc_smooth */
  double cw = total_cost(m, d, e, d->qacc_warmstart, NULL, NULL, NULL);
  double cs = total_cost(m, d, e, d->qacc_smooth, NULL, NULL, NULL);
  memcpy(a, cw < cs ? d->qacc_warmstart : d->qacc_smooth, sizeof(double) * nv);
  /* experiment (PO_WARM_STICK): dofs whose frictionloss row was in its quadratic zone at the previous
   * solution start at the row's new reference acceleration */
  static int warm_stick = -1;
  static unsigned char stick_prev[512];
  if (warm_stick < 0) { warm_stick = getenv("PO_WARM_STICK") ? atoi(getenv("PO_WARM_STICK")) : 0; memset(stick_prev, 0, sizeof(stick_prev)); }
  if (warm_stick) {
    double *t = (double *)malloc(sizeof(double) * nv);
    memcpy(t, a, sizeof(double) * nv);
    for (int r = 0; r < ne; r++) {
      if (e->type[r] != ROW_FRICTION) continue;
      int j = -1;
      for (int k = 0; k < nv; k++) if (e->J[(size_t)r * nv + k] != 0) { j = k; break; }
      if (j >= 0 && stick_prev[j]) t[j] = e->aref[r];
    }
    double ct = total_cost(m, d, e, t, NULL, NULL, NULL);
    double c0 = cw < cs ? cw : cs;
    if (warm_stick == 2 || ct < c0) memcpy(a, t, sizeof(double) * nv);
    free(t);
  }
  d->solver_iter = 0;
  if (ne == 0) { memcpy(a, d->qacc_smooth, sizeof(double) * nv); memset(d->qfrc_constraint, 0, sizeof(double) * nv); return; }
  double *grad = (double *)malloc(sizeof(double) * nv), *p = (double *)malloc(sizeof(double) * nv);
  double *Mp = (double *)malloc(sizeof(double) * nv), *H = (double *)malloc(sizeof(double) * nv * nv);
  double *force = (double *)malloc(sizeof(double) * ne), *curv = (double *)malloc(sizeof(double) * ne);
  double *jar = (double *)malloc(sizeof(double) * ne), *jp = (double *)malloc(sizeof(double) * ne);
  double scale = 1 / (m->meaninertia * (nv > 1 ? nv : 1));
  double cost = total_cost(m, d, e, a, grad, force, curv);
  for (int iter = 0; iter < m->iterations; iter++) {
    /* H = M + J^T diag(curv) J ; p = -H^-1 grad */
    memcpy(H, d->qM, sizeof(double) * nv * nv);
    for (int r = 0; r < ne; r++) {
      if (curv[r] == 0) continue;
      const double *Jr = e->J + (size_t)r * nv;
      for (int i = 0; i < nv; i++) {
        if (Jr[i] == 0) continue;
        double s = curv[r] * Jr[i];
        for (int j = 0; j < nv; j++) H[(size_t)i * nv + j] += s * Jr[j];
      }
    }
    chol_factor(H, nv);
    for (int j = 0; j < nv; j++) p[j] = -grad[j];
    chol_solve(H, nv, p);
    /* exact line search on phi(alpha) = cost(a + alpha p) */
    matvec(d->qM, nv, p, Mp);
    double g1 = 0, g2 = 0;
    for (int j = 0; j < nv; j++) { g1 += grad[j] * p[j]; g2 += p[j] * Mp[j]; }
    /* note: grad includes constraint forces; rebuild the Gauss-only slope */
    double gauss1 = 0;
    {
      double *Ma = (double *)malloc(sizeof(double) * nv);
      matvec(d->qM, nv, a, Ma);
      for (int j = 0; j < nv; j++) gauss1 += (Ma[j] - d->qfrc_smooth[j]) * p[j];
      free(Ma);
    }
    for (int r = 0; r < ne; r++) {
      const double *Jr = e->J + (size_t)r * nv;
      double x = -e->aref[r], y = 0;
      for (int j = 0; j < nv; j++) { x += Jr[j] * a[j]; y += Jr[j] * p[j]; }
      jar[r] = x; jp[r] = y;
    }
    double alpha = 0, lo = 0, hi = -1, d0 = g1;
    for (int ls = 0; ls < m->ls_iterations; ls++) {
      double dphi = gauss1 + alpha * g2, ddphi = g2;
      for (int r = 0; r < ne; r++) {
        double c, f, cu;
        row_eval(e, r, jar[r] + alpha * jp[r], &c, &f, &cu);
        dphi -= f * jp[r];
        ddphi += cu * jp[r] * jp[r];
      }
      if (fabs(dphi) <= 1e-14 * fabs(d0) || ddphi <= 0) break;
      if (dphi < 0) lo = alpha; else hi = alpha;
      double next = alpha - dphi / ddphi;
      if (hi >= 0 && (next <= lo || next >= hi)) next = 0.5 * (lo + hi);
      if (next == alpha) break;
      alpha = next;
    }
    static int fullstep = -1;
    if (fullstep < 0) fullstep = getenv("PO_NEWTON_FULLSTEP") ? atoi(getenv("PO_NEWTON_FULLSTEP")) : 0;
    if (fullstep) {
      double *t = (double *)malloc(sizeof(double) * nv);
      for (int j = 0; j < nv; j++) t[j] = a[j] + p[j];
      double c1 = total_cost(m, d, e, t, NULL, NULL, NULL);
      for (int j = 0; j < nv; j++) t[j] = a[j] + alpha * p[j];
      double ca = total_cost(m, d, e, t, NULL, NULL, NULL);
      if (fullstep == 1 ? (c1 < cost) : (c1 < cost - 0.5 * (cost - ca))) alpha = 1;
      free(t);
    }
    for (int j = 0; j < nv; j++) a[j] += alpha * p[j];
    double *oldcurv = NULL;
    static int dbg = -1;
    if (dbg < 0) dbg = getenv("PO_DEBUG_NEWTON") != NULL;
    if (dbg) { oldcurv = (double *)malloc(sizeof(double) * ne); memcpy(oldcurv, curv, sizeof(double) * ne); }
    double newcost = total_cost(m, d, e, a, grad, force, curv);
    if (dbg) {
      int nfr = 0, nin = 0;
      for (int r = 0; r < ne; r++) if ((oldcurv[r] != 0) != (curv[r] != 0)) { if (e->type[r] == ROW_INEQ) nin++; else nfr++; }
      fprintf(stderr, "  newton it %d alpha %.6f improvement %.3e changed fric %d ineq %d (nefc %d)\n", iter, alpha,
              scale * (cost - newcost), nfr, nin, ne);
      free(oldcurv);
    }
    double improvement = scale * (cost - newcost), gnorm = 0;
    cost = newcost;
    for (int j = 0; j < nv; j++) gnorm += grad[j] * grad[j];
    gnorm = scale * sqrt(gnorm);
    d->solver_iter = iter + 1;
    if (improvement < m->tolerance || gnorm < m->tolerance) break;
  }
  d->solver_cost = cost;
  if (warm_stick) {
    memset(stick_prev, 0, sizeof(stick_prev));
    for (int r = 0; r < ne; r++) {
      if (e->type[r] != ROW_FRICTION) continue;
      for (int k = 0; k < nv; k++) if (e->J[(size_t)r * nv + k] != 0) { stick_prev[k] = curv[r] != 0; break; }
    }
  }
  memset(d->qfrc_constraint, 0, sizeof(double) * nv);
  for (int r = 0; r < ne; r++)
    for (int j = 0; j < nv; j++) d->qfrc_constraint[j] += e->J[(size_t)r * nv + j] * force[r];
  free(grad); free(p); free(Mp); free(H); free(force); free(curv); free(jar); free(jp);
}

/* ------------------------------------------------------------------ mj_step */
void po_step(const pp_model *m, po_data *d) {
  int nv = m->nv;
  /* mj_fwdPosition */
  kinematics(m, d);
  crb(m, d);
  collision(m, d);
  efc_t e;
  make_constraint(m, d, &e);
  /* mj_fwdVelocity */
  rne_bias(m, d);
  passive(m, d);
  /* mj_fwdActuation, mj_fwdAcceleration */
  actuation(m, d);
  double *L = (double *)malloc(sizeof(double) * nv * nv);
  memcpy(L, d->qM, sizeof(double) * nv * nv);
  chol_factor(L, nv);
  for (int j = 0; j < nv; j++) {
    d->qfrc_smooth[j] = d->qfrc_passive[j] - d->qfrc_bias[j] + d->qfrc_actuator[j] + d->qfrc_applied[j];
    d->qacc_smooth[j] = d->qfrc_smooth[j];
  }
  chol_solve(L, nv, d->qacc_smooth);
  /* mj_fwdConstraint */
  solve_newton(m, d, &e);
  efc_free(&e);
  /* mj_Euler with implicit joint damping: (M + h diag(b)) qacc' = qfrc_smooth + qfrc_constraint */
  double h = m->timestep;
  memcpy(L, d->qM, sizeof(double) * nv * nv);
  double *x = (double *)malloc(sizeof(double) * nv);
  for (int j = 0; j < nv; j++) {
    L[(size_t)j * nv + j] += h * m->dof_damping[j];
    x[j] = d->qfrc_smooth[j] + d->qfrc_constraint[j];
  }
  chol_factor(L, nv);
  chol_solve(L, nv, x);
  for (int j = 0; j < nv; j++) {
    d->qvel[j] += h * x[j];
    d->qpos[j] += h * d->qvel[j];
    d->qacc_warmstart[j] = d->qacc[j];
  }
  free(L); free(x);
  actuator_velocity(m, d); /* actuatorvel sensor as refreshed by the trailing mj_step1 */
}

/* threads of po_step_batch; 0 = OpenMP's default (OMP_NUM_THREADS, which torchrun sets to 1) */
static int po_threads = 0;
void po_set_num_threads(int n) { po_threads = n > 0 ? n : 0; }

void po_step_batch(const pp_model *m, po_data **d, int nenv, int nsub) {
#ifdef _OPENMP
  if (po_threads > 0) omp_set_num_threads(po_threads);
#endif
#pragma omp parallel for schedule(dynamic, 1)
  for (int i = 0; i < nenv; i++)
    for (int s = 0; s < nsub; s++) po_step(m, d[i]);
}


/* ====================================================================== po_step_sparse
 * The same pipeline with this model's structure exploited and no allocation inside the step (piano_oracle.h).
 * Dof sets: T = dofs of kinematic trees with more than one dof (the two hands; fixed per model), C = T plus the
 * 1-dof joints that share a constraint row with another dof in this substep (keys under a finger). M and the
 * Newton Hessian are dense over T / C and diagonal elsewhere. */
typedef struct {
  int nv, cap_rows, cap_nnz, nt, nc, nrows, nnz;
  int *tlist, *tidx;            /* T and dof -> position in T (-1: isolated 1-dof joint) */
  int *clist, *cidx;            /* C and dof -> position in C (-1: scalar) */
  int *rowptr, *col, *type;     /* constraint rows, CSR */
  double *val, *D, *aref, *floss, *force, *curv, *jar, *jp;
  double *H, *LM, *Hd, *grad, *p, *Mp, *Ma, *a, *t1, *t2, *j1, *j2;
  double *Ic, *vel, *acc, *frc; /* crb / rne scratch */
} po_ws;

static po_ws *ws_get(const pp_model *m, po_data *d) {
  if (d->ws) return (po_ws *)d->ws;
  int nv = m->nv, nb = m->nbody;
  int cap_rows = 3 * nv + 4 * PO_MAXCON + 1, cap_nnz = 3 * nv + 4 * PO_MAXCON * 64;
  size_t ni = (size_t)4 * nv + (cap_rows + 1) + cap_nnz + cap_rows;
  size_t nd = (size_t)cap_nnz + 7 * (size_t)cap_rows + 2 * (size_t)nv * nv + 8 * (size_t)nv + 6 * (size_t)nv +
              (size_t)36 * nb + 18 * (size_t)nb;
  char *blk = (char *)calloc(1, sizeof(po_ws) + ni * sizeof(int) + 8 + nd * sizeof(double));
  po_ws *w = (po_ws *)blk;
  int *ip = (int *)(blk + sizeof(po_ws));
  w->nv = nv; w->cap_rows = cap_rows; w->cap_nnz = cap_nnz;
  w->tlist = ip; ip += nv; w->tidx = ip; ip += nv; w->clist = ip; ip += nv; w->cidx = ip; ip += nv;
  w->rowptr = ip; ip += cap_rows + 1; w->col = ip; ip += cap_nnz; w->type = ip; ip += cap_rows;
  double *dp = (double *)(((uintptr_t)ip + 7) & ~(uintptr_t)7);
  w->val = dp; dp += cap_nnz;
  w->D = dp; dp += cap_rows; w->aref = dp; dp += cap_rows; w->floss = dp; dp += cap_rows; w->force = dp; dp += cap_rows;
  w->curv = dp; dp += cap_rows; w->jar = dp; dp += cap_rows; w->jp = dp; dp += cap_rows;
  w->H = dp; dp += (size_t)nv * nv; w->LM = dp; dp += (size_t)nv * nv;
  w->Hd = dp; dp += nv; w->grad = dp; dp += nv; w->p = dp; dp += nv; w->Mp = dp; dp += nv; w->Ma = dp; dp += nv;
  w->a = dp; dp += nv; w->t1 = dp; dp += nv; w->t2 = dp; dp += nv;
  w->j1 = dp; dp += 3 * nv; w->j2 = dp; dp += 3 * nv;
  w->Ic = dp; dp += 36 * nb; w->vel = dp; dp += 6 * nb; w->acc = dp; dp += 6 * nb; w->frc = dp; dp += 6 * nb;
  /* T: a dof with a parent dof, or that is some dof's parent */
  for (int j = 0; j < nv; j++) w->tidx[j] = -1;
  for (int j = 0; j < nv; j++)
    if (m->dof_parent[j] >= 0) { w->tidx[j] = 0; w->tidx[m->dof_parent[j]] = 0; }
  for (int j = 0; j < nv; j++)
    if (w->tidx[j] == 0) { w->tidx[j] = w->nt; w->tlist[w->nt++] = j; }
  d->ws = w;
  return w;
}

/* y = M x with M dense over T and diagonal elsewhere */
static void ws_matvecM(const po_ws *w, const po_data *d, const double *x, double *y) {
  int nv = w->nv;
  for (int j = 0; j < nv; j++) if (w->tidx[j] < 0) y[j] = d->qM[(size_t)j * nv + j] * x[j];
  for (int a = 0; a < w->nt; a++) {
    int i = w->tlist[a];
    const double *Mi = d->qM + (size_t)i * nv;
    double s = 0;
    for (int b = 0; b < w->nt; b++) s += Mi[w->tlist[b]] * x[w->tlist[b]];
    y[i] = s;
  }
}

/* x <- (M + diag(extra))^-1 x : Cholesky over T, scalars elsewhere (extra may be NULL) */
static void ws_solveM(po_ws *w, const po_data *d, const double *extra, double *x) {
  int nv = w->nv, nt = w->nt;
  for (int j = 0; j < nv; j++)
    if (w->tidx[j] < 0) x[j] /= d->qM[(size_t)j * nv + j] + (extra ? extra[j] : 0.0);
  for (int a = 0; a < nt; a++) {
    int i = w->tlist[a];
    for (int b = 0; b < nt; b++) w->LM[a * nt + b] = d->qM[(size_t)i * nv + w->tlist[b]];
    if (extra) w->LM[a * nt + a] += extra[i];
    w->t1[a] = x[i];
  }
  chol_factor(w->LM, nt);
  chol_solve(w->LM, nt, w->t1);
  for (int a = 0; a < nt; a++) x[w->tlist[a]] = w->t1[a];
}

static int ws_new_row(po_ws *w, int type, double D, double floss) {
  int r = w->nrows++;
  w->rowptr[r] = w->nnz;
  w->type[r] = type; w->D[r] = D; w->floss[r] = floss;
  return r;
}
static void ws_push(po_ws *w, int col, double val) { w->col[w->nnz] = col; w->val[w->nnz] = val; w->nnz++; }
static double ws_rowdot(const po_ws *w, int r, const double *x) {
  double s = 0;
  for (int k = w->rowptr[r]; k < w->rowptr[r + 1]; k++) s += w->val[k] * x[w->col[k]];
  return s;
}
static void ws_row_eval(const po_ws *w, int r, double x, double *cost, double *force, double *curv) {
  double D = w->D[r];
  if (w->type[r] == ROW_INEQ) {
    if (x < 0) { *cost = 0.5 * D * x * x; *force = -D * x; *curv = D; }
    else { *cost = 0; *force = 0; *curv = 0; }
  } else {
    double f = w->floss[r], rf = f / D;
    if (x <= -rf) { *cost = -0.5 * rf * f - f * x; *force = f; *curv = 0; }
    else if (x >= rf) { *cost = -0.5 * rf * f + f * x; *force = -f; *curv = 0; }
    else { *cost = 0.5 * D * x * x; *force = -D * x; *curv = D; }
  }
}

/* mj_makeConstraint with sparse rows (same rows, same order, same numbers as make_constraint) */
static void ws_make_constraint(const pp_model *m, po_data *d, po_ws *w) {
  int nv = m->nv;
  w->nrows = 0; w->nnz = 0;
  double K, B;
  stiffness_damping(m, m->friction_solref, m->friction_solimp, &K, &B);
  double impf = impedance(m->friction_solimp, 0, 0);
  for (int j = 0; j < nv; j++) {
    if (m->dof_frictionloss[j] <= 0) continue;
    double R = fmax(MINVAL, (1 - impf) * m->dof_invweight0[j] / impf);
    int r = ws_new_row(w, ROW_FRICTION, 1 / R, m->dof_frictionloss[j]);
    ws_push(w, j, 1.0);
    w->aref[r] = -B * d->qvel[j];
  }
  stiffness_damping(m, m->limit_solref, m->limit_solimp, &K, &B);
  for (int j = 0; j < nv; j++) {
    if (!m->jnt_limited[j]) continue;
    for (int side = 0; side < 2; side++) {
      double sgn = side == 0 ? 1.0 : -1.0;
      double dist = side == 0 ? d->qpos[j] - m->jnt_range[2 * j] : m->jnt_range[2 * j + 1] - d->qpos[j];
      if (dist >= 0) continue;
      double imp = impedance(m->limit_solimp, dist, 0);
      double R = fmax(MINVAL, (1 - imp) * m->dof_invweight0[j] / imp);
      int r = ws_new_row(w, ROW_INEQ, 1 / R, 0);
      ws_push(w, j, sgn);
      w->aref[r] = -B * sgn * d->qvel[j] - K * imp * dist;
    }
  }
  for (int c = 0; c < d->ncon; c++) {
    const po_contact *con = &d->contact[c];
    int cls1 = m->geom_cls[con->geom1], cls2 = m->geom_cls[con->geom2];
    int pc = (cls1 <= PP_CLS_BASE || cls2 <= PP_CLS_BASE) ? PP_PAIRCLS_PIANO : PP_PAIRCLS_HAND;
    const double *solref = m->pair_solref[pc], *solimp = m->pair_solimp[pc];
    double mu = m->pair_friction[pc], margin = m->pair_margin[pc];
    int b1 = m->geom_body[con->geom1], b2 = m->geom_body[con->geom2];
    /* columns: the dofs above either body; w->t1 marks them */
    int ncol = 0, cols[128];
    for (int side = 0; side < 2; side++)
      for (int b = side ? b2 : b1; b > 0; b = m->body_parent[b])
        for (int k = 0; k < m->body_dofnum[b]; k++) {
          int j = m->body_dofadr[b] + k, seen = 0;
          for (int q = 0; q < ncol; q++) if (cols[q] == j) { seen = 1; break; }
          if (!seen && ncol < 128) cols[ncol++] = j;
        }
    /* ascending dof order, like a dense row read left to right */
    for (int a = 1; a < ncol; a++) { int v = cols[a], b = a - 1; while (b >= 0 && cols[b] > v) { cols[b + 1] = cols[b]; b--; } cols[b + 1] = v; }
    jac_point(m, d, b1, con->pos, w->j1);
    jac_point(m, d, b2, con->pos, w->j2);
    stiffness_damping(m, solref, solimp, &K, &B);
    double imp = impedance(solimp, con->dist, margin);
    double tran = m->body_invweight0[2 * b1] + m->body_invweight0[2 * b2];
    double diag = tran + mu * mu * tran;
    double R0 = fmax(MINVAL, (1 - imp) * diag / imp);
    double mureg = mu * sqrt(1 / m->impratio);
    double R = 2 * mureg * mureg * R0;
    for (int k = 0; k < 4; k++) {
      int r = ws_new_row(w, ROW_INEQ, 1 / R, 0);
      const double *n = con->frame, *t = con->frame + 3 * (1 + k / 2);
      double sg = (k % 2 == 0) ? mu : -mu;
      for (int q = 0; q < ncol; q++) {
        int j = cols[q];
        double dn = 0, dt = 0;
        for (int i = 0; i < 3; i++) {
          double dj = w->j2[i * nv + j] - w->j1[i * nv + j];
          dn += n[i] * dj; dt += t[i] * dj;
        }
        double v = dn + sg * dt;
        if (v != 0) ws_push(w, j, v);
      }
      w->rowptr[r + 1] = w->nnz;
      w->aref[r] = -B * ws_rowdot(w, r, d->qvel) - K * imp * (con->dist - margin);
    }
  }
  w->rowptr[w->nrows] = w->nnz;
  d->nefc = w->nrows;
  /* C = T + isolated dofs that share a row with another dof */
  w->nc = 0;
  for (int j = 0; j < nv; j++) w->cidx[j] = -1;
  for (int a = 0; a < w->nt; a++) { w->cidx[w->tlist[a]] = w->nc; w->clist[w->nc++] = w->tlist[a]; }
  for (int r = 0; r < w->nrows; r++) {
    if (w->rowptr[r + 1] - w->rowptr[r] < 2) continue;
    for (int k = w->rowptr[r]; k < w->rowptr[r + 1]; k++) {
      int j = w->col[k];
      if (w->cidx[j] < 0) { w->cidx[j] = w->nc; w->clist[w->nc++] = j; }
    }
  }
}

static double ws_total_cost(const pp_model *m, const po_data *d, po_ws *w, const double *a, double *grad,
                            double *force, double *curv) {
  int nv = m->nv;
  ws_matvecM(w, d, a, w->Ma);
  double cost = 0;
  for (int j = 0; j < nv; j++) cost += 0.5 * (w->Ma[j] - d->qfrc_smooth[j]) * (a[j] - d->qacc_smooth[j]);
  if (grad) for (int j = 0; j < nv; j++) grad[j] = w->Ma[j] - d->qfrc_smooth[j];
  for (int r = 0; r < w->nrows; r++) {
    double x = ws_rowdot(w, r, a) - w->aref[r], c, f, cu;
    ws_row_eval(w, r, x, &c, &f, &cu);
    cost += c;
    if (force) force[r] = f;
    if (curv) curv[r] = cu;
    if (grad) for (int k = w->rowptr[r]; k < w->rowptr[r + 1]; k++) grad[w->col[k]] -= w->val[k] * f;
  }
  return cost;
}

/* mj_solNewton, same iterates as solve_newton: the linear solve is dense over C, scalar elsewhere */
static void ws_solve_newton(const pp_model *m, po_data *d, po_ws *w) {
  int nv = m->nv, ne = w->nrows;
  double *a = d->qacc;
  double cw = ws_total_cost(m, d, w, d->qacc_warmstart, NULL, NULL, NULL);
  double cs = ws_total_cost(m, d, w, d->qacc_smooth, NULL, NULL, NULL);
  memcpy(a, cw < cs ? d->qacc_warmstart : d->qacc_smooth, sizeof(double) * nv);
  d->solver_iter = 0;
  if (ne == 0) { memcpy(a, d->qacc_smooth, sizeof(double) * nv); memset(d->qfrc_constraint, 0, sizeof(double) * nv); return; }
  double *grad = w->grad, *p = w->p, *Mp = w->Mp, *H = w->H;
  double scale = 1 / (m->meaninertia * (nv > 1 ? nv : 1));
  double cost = ws_total_cost(m, d, w, a, grad, w->force, w->curv);
  for (int iter = 0; iter < m->iterations; iter++) {
    int nc = w->nc;
    for (int j = 0; j < nv; j++) w->Hd[j] = d->qM[(size_t)j * nv + j];
    for (int x = 0; x < nc; x++) {
      const double *Mi = d->qM + (size_t)w->clist[x] * nv;
      for (int y = 0; y < nc; y++) H[x * nc + y] = Mi[w->clist[y]];
    }
    for (int r = 0; r < ne; r++) {
      double cu = w->curv[r];
      if (cu == 0) continue;
      int k0 = w->rowptr[r], k1 = w->rowptr[r + 1];
      if (k1 - k0 == 1 && w->cidx[w->col[k0]] < 0) { w->Hd[w->col[k0]] += cu * w->val[k0] * w->val[k0]; continue; }
      for (int ka = k0; ka < k1; ka++) {
        double s = cu * w->val[ka];
        double *Hx = H + (size_t)w->cidx[w->col[ka]] * nc;
        for (int kb = k0; kb < k1; kb++) Hx[w->cidx[w->col[kb]]] += s * w->val[kb];
      }
    }
    chol_factor(H, nc);
    for (int x = 0; x < nc; x++) w->t1[x] = -grad[w->clist[x]];
    chol_solve(H, nc, w->t1);
    for (int j = 0; j < nv; j++) if (w->cidx[j] < 0) p[j] = -grad[j] / w->Hd[j];
    for (int x = 0; x < nc; x++) p[w->clist[x]] = w->t1[x];
    ws_matvecM(w, d, p, Mp);
    double g1 = 0, g2 = 0, gauss1 = 0;
    ws_matvecM(w, d, a, w->Ma);
    for (int j = 0; j < nv; j++) { g1 += grad[j] * p[j]; g2 += p[j] * Mp[j]; gauss1 += (w->Ma[j] - d->qfrc_smooth[j]) * p[j]; }
    for (int r = 0; r < ne; r++) { w->jar[r] = ws_rowdot(w, r, a) - w->aref[r]; w->jp[r] = ws_rowdot(w, r, p); }
    double alpha = 0, lo = 0, hi = -1, d0 = g1;
    for (int ls = 0; ls < m->ls_iterations; ls++) {
      double dphi = gauss1 + alpha * g2, ddphi = g2;
      for (int r = 0; r < ne; r++) {
        double c, f, cu;
        ws_row_eval(w, r, w->jar[r] + alpha * w->jp[r], &c, &f, &cu);
        dphi -= f * w->jp[r];
        ddphi += cu * w->jp[r] * w->jp[r];
      }
      if (fabs(dphi) <= 1e-14 * fabs(d0) || ddphi <= 0) break;
      if (dphi < 0) lo = alpha; else hi = alpha;
      double next = alpha - dphi / ddphi;
      if (hi >= 0 && (next <= lo || next >= hi)) next = 0.5 * (lo + hi);
      if (next == alpha) break;
      alpha = next;
    }
    for (int j = 0; j < nv; j++) a[j] += alpha * p[j];
    double newcost = ws_total_cost(m, d, w, a, grad, w->force, w->curv);
    double improvement = scale * (cost - newcost), gnorm = 0;
    cost = newcost;
    for (int j = 0; j < nv; j++) gnorm += grad[j] * grad[j];
    gnorm = scale * sqrt(gnorm);
    d->solver_iter = iter + 1;
    if (improvement < m->tolerance || gnorm < m->tolerance) break;
  }
  d->solver_cost = cost;
  memset(d->qfrc_constraint, 0, sizeof(double) * nv);
  for (int r = 0; r < ne; r++)
    for (int k = w->rowptr[r]; k < w->rowptr[r + 1]; k++) d->qfrc_constraint[w->col[k]] += w->val[k] * w->force[r];
}


void po_step_sparse(const pp_model *m, po_data *d) {
  int nv = m->nv;
  po_ws *w = ws_get(m, d);
  /* mj_fwdPosition */
  kinematics(m, d);
  crb_buf(m, d, w->Ic);
  collision(m, d);
  ws_make_constraint(m, d, w);
  /* mj_fwdVelocity, mj_fwdActuation */
  rne_bias_buf(m, d, w->vel, w->acc, w->frc);
  passive_buf(m, d, w->j1);
  actuation(m, d);
  /* mj_fwdAcceleration */
  for (int j = 0; j < nv; j++) {
    d->qfrc_smooth[j] = d->qfrc_passive[j] - d->qfrc_bias[j] + d->qfrc_actuator[j] + d->qfrc_applied[j];
    d->qacc_smooth[j] = d->qfrc_smooth[j];
  }
  ws_solveM(w, d, NULL, d->qacc_smooth);
  /* mj_fwdConstraint */
  ws_solve_newton(m, d, w);
  /* mj_Euler with implicit joint damping */
  double h = m->timestep;
  for (int j = 0; j < nv; j++) { w->t2[j] = h * m->dof_damping[j]; w->a[j] = d->qfrc_smooth[j] + d->qfrc_constraint[j]; }
  ws_solveM(w, d, w->t2, w->a);
  for (int j = 0; j < nv; j++) {
    d->qvel[j] += h * w->a[j];
    d->qpos[j] += h * d->qvel[j];
    d->qacc_warmstart[j] = d->qacc[j];
  }
  actuator_velocity(m, d);
}

void po_step_batch_sparse(const pp_model *m, po_data **d, int nenv, int nsub) {
#ifdef _OPENMP
  if (po_threads > 0) omp_set_num_threads(po_threads);
#endif
#pragma omp parallel for schedule(dynamic, 1)
  for (int i = 0; i < nenv; i++)
    for (int s = 0; s < nsub; s++) po_step_sparse(m, d[i]);
}
