/* pp_model.h - flat compiled-model tables handed to pp_create() (and to the CPU oracle).
 *
 * The reference has no such struct: its model lives inside MuJoCo's mjModel, produced by
 * dm_control from the MJCF the reference builds (reference parallelized_base_v2.py:39-43,
 * robopianist/models/piano/piano_mjcf.py:25, robopianist/models/hands/shadow_hand.py:91).
 * This is the subset of mjModel the hot path reads (SURVEY.md §8a "model constants"), with
 * MuJoCo's field names. All joints of this scene are 1-DoF, so joint index == dof index
 * == qpos index. Arrays are host pointers, caller-owned, only read during pp_create().
 */
#ifndef PP_MODEL_H_
#define PP_MODEL_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { PP_GEOM_SPHERE = 2, PP_GEOM_CAPSULE = 3, PP_GEOM_BOX = 6 };
enum { PP_JNT_SLIDE = 2, PP_JNT_HINGE = 3 };
/* geom classes */
enum { PP_CLS_KEY = 0, PP_CLS_BASE = 1, PP_CLS_RH_FINGER = 2, PP_CLS_LH_FINGER = 3,
       PP_CLS_RH_FOREARM = 4, PP_CLS_LH_FOREARM = 5 };
/* contact parameter classes: 0 = hand-vs-piano geom, 1 = hand-vs-hand geom */
enum { PP_PAIRCLS_PIANO = 0, PP_PAIRCLS_HAND = 1 };

typedef struct pp_model {
  int32_t nbody, nv, ngeom, nsite, nu, npair, nkey, has_hands;
  /* mjOption subset */
  double timestep;
  double gravity[3];
  double impratio;
  double tolerance;      /* Newton termination (scaled), MuJoCo default 1e-8 */
  int32_t iterations;    /* Newton iteration cap, MuJoCo default 100 */
  int32_t ls_iterations; /* line-search iteration cap, MuJoCo default 50 */
  double meaninertia;    /* mjModel.stat.meaninertia: mean diagonal of M(qpos0) */
  /* bodies [nbody] (body 0 = world) */
  const int32_t *body_parent, *body_dofadr, *body_dofnum;
  const double *body_pos;        /* [nbody,3] */
  const double *body_quat;       /* [nbody,4] */
  const double *body_ipos;       /* [nbody,3] */
  const double *body_iquat;      /* [nbody,4] */
  const double *body_mass;       /* [nbody]   */
  const double *body_inertia;    /* [nbody,3] diagonal in the inertial frame */
  const double *body_gravcomp;   /* [nbody]   */
  const double *body_invweight0; /* [nbody,2] */
  /* joints == dofs [nv] */
  const int32_t *jnt_type, *jnt_body, *jnt_limited, *dof_parent;
  const double *jnt_pos;   /* [nv,3] anchor in body frame */
  const double *jnt_axis;  /* [nv,3] unit axis in body frame */
  const double *jnt_range; /* [nv,2] */
  const double *jnt_stiffness, *jnt_springref, *qpos0;
  const double *dof_damping, *dof_armature, *dof_frictionloss, *dof_invweight0;
  double limit_solref[2], limit_solimp[5];       /* jnt_solref / jnt_solimp (uniform) */
  double friction_solref[2], friction_solimp[5]; /* dof_solref / dof_solimp (uniform) */
  /* collision geoms [ngeom] */
  const int32_t *geom_type, *geom_body, *geom_cls;
  const double *geom_pos;  /* [ngeom,3] */
  const double *geom_quat; /* [ngeom,4] */
  const double *geom_size; /* [ngeom,3] */
  /* mixed contact parameters per contact class */
  double pair_solref[2][2], pair_solimp[2][5], pair_friction[2], pair_margin[2];
  /* fixed candidate list for hand-hand geom pairs [npair] */
  const int32_t *pair_geom1, *pair_geom2;
  /* fingertip sites [nsite]: rh th,ff,mf,rf,lf then lh th,ff,mf,rf,lf */
  const int32_t *site_body;
  const double *site_pos; /* [nsite,3] */
  /* actuators [nu]: force = kp*(ctrl - length), or kp*ctrl when gain_only */
  const int32_t *act_trn0, *act_trn1, *act_gain_only, *act_forcelimited;
  const double *act_kp;
  const double *act_ctrlrange;  /* [nu,2] (ctrl is always clamped) */
  const double *act_forcerange; /* [nu,2] */
} pp_model;

#ifdef __cplusplus
}
#endif
#endif /* PP_MODEL_H_ */
