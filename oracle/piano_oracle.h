/* piano_oracle.h - CPU fp64 ORACLE for the batched piano/hand step. TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load this; the product (diffusion-piano_b200/) must never import, link or call it.
 *
 * What it restates: one MuJoCo `mj_step` (as driven by dm_control's legacy step:
 * mj_step2 then mj_step1) for the fixed scene the reference builds - the call the
 * reference makes at `physics.step()` inside composer.Environment.step for the envs
 * created at reference parallelized_base_v2.py:39-43 / robopianist/suite/__init__.py:87-93.
 * The arithmetic lives in the third-party package mujoco>=3.1.1 (reference setup.py:39),
 * NOT in /root/reference, so this file restates MuJoCo's published pipeline
 * (engine_forward.c: fwdPosition/fwdVelocity/fwdActuation/fwdAcceleration/fwdConstraint,
 * engine_core_constraint.c: makeConstraint/makeImpedance, engine_solver.c: Newton,
 * mj_Euler with implicit joint damping) over the tables in include/pp_model.h.
 *
 * PARITY STATUS: "parity unpinned" versus real mj_step - MuJoCo, dm_control and the
 * menagerie Shadow Hand XML are absent from this container and from /root/reference, and
 * the reference's tests hold no numeric physics vectors (piano_with_shadow_hands_test.py:274).
 * Pinned pieces: the piano tables (golden layout from the reference's own constants), the
 * closed-form single-key dynamics, and the reference-test behaviours in tests/.
 */
#ifndef PIANO_ORACLE_H_
#define PIANO_ORACLE_H_

#include "../include/pp_model.h"

#ifdef __cplusplus
extern "C" {
#endif

#define PO_MAXCON 256

typedef struct po_contact {
  int geom1, geom2;
  double dist;
  double pos[3];
  double frame[9]; /* rows: normal (geom1 -> geom2), tangent1, tangent2 */
} po_contact;

/* Everything one step computes, kept for stage-by-stage diffing. Sizes use the model's
 * nbody/nv/ngeom/nsite/nu; buffers are allocated by po_data_create. */
typedef struct po_data {
  int nv, nbody, ngeom, nsite, nu;
  /* state (in/out) */
  double *qpos, *qvel, *qacc_warmstart, *ctrl, *qfrc_applied;
  /* position stage */
  double *xpos, *xquat, *xmat, *xipos, *ximat, *xanchor, *xaxis;
  double *geom_xpos, *geom_xmat, *site_xpos;
  double *qM; /* dense nv*nv */
  /* velocity / force stage */
  double *qfrc_bias, *qfrc_passive, *actuator_force, *actuator_velocity, *qfrc_actuator;
  double *qfrc_smooth, *qacc_smooth;
  /* constraints */
  int ncon, nefc, ncon_dropped;
  po_contact contact[PO_MAXCON];
  double *qacc, *qfrc_constraint;
  int solver_iter;
  double solver_cost;
  void *ws; /* workspace of po_step_sparse (allocated at its first call, freed by po_data_free) */
} po_data;

po_data *po_data_create(const pp_model *m);
void po_data_free(po_data *d);
void po_reset(const pp_model *m, po_data *d);
/* kinematics + collision at the current qpos (what the trailing mj_step1 refreshes) */
void po_forward_position(const pp_model *m, po_data *d, int with_collision);
/* one full mj_step: forward dynamics at (qpos,qvel,ctrl), constraint solve, Euler */
void po_step(const pp_model *m, po_data *d);
/* The same mj_step with the structure of THIS model exploited, as an engine written for speed would (MuJoCo's own
 * data structures are sparse): constraint rows as sparse lists, the Newton Hessian and the mass matrix as one dense
 * block over the coupled dofs (kinematic trees with more than one dof + the 1-dof joints a contact touches) plus
 * scalars for every other 1-dof joint, no allocation inside the step. Same Newton iterates as po_step, so the
 * results agree to rounding (tests/test_scene_and_oracle.py); it exists to be the honest timed CPU baseline. */
void po_step_sparse(const pp_model *m, po_data *d);
/* n env-independent steps over a batch, OpenMP over envs; for the CPU baseline */
void po_step_batch(const pp_model *m, po_data **d, int nenv, int nsub);
void po_step_batch_sparse(const pp_model *m, po_data **d, int nenv, int nsub);
void po_set_num_threads(int n); /* threads used by po_step_batch (0 = OpenMP default) */

#ifdef __cplusplus
}
#endif
#endif
