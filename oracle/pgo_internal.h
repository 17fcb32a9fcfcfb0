/* oracle/pgo_internal.h -- private links between the oracle's translation units (TEST INFRASTRUCTURE ONLY, see pgo.h). */
#ifndef PGO_INTERNAL_H
#define PGO_INTERNAL_H
#include "pgo.h"

/* one joint-space solver row of the arm (joint limit or motor), see pgo_arm_substep */
typedef struct { int dof; double sign, rhs, lo, hi, applied, dinv, resp[PGO_MAXD]; } pgo__row;

/* phases of pgo_arm_substep, exposed so that the world step (pgo_world.c) can interleave contact rows exactly
 * where btMultiBodyConstraintSolver::solveSingleIteration does */
void pgo__arm_predict(pgo_arm* a);
int pgo__arm_rows(pgo_arm* a, pgo__row* rows);
double pgo__arm_sweep(const pgo_arm* a, pgo__row* rows, int nr, int it, double* dv);
void pgo__arm_finish(pgo_arm* a, const double* dv);
void pgo__arm_response(const pgo_arm* a, const double* tau, double* out); /* M^-1 tau (valid after predict) */
int pgo__arm_solver_iters(const pgo_arm* a);
double pgo__arm_residual_thr(const pgo_arm* a);
double pgo__arm_dt(const pgo_arm* a);
const double* pgo__arm_qd(const pgo_arm* a);
/* world-frame translational Jacobian (3 x ndof, row-major) of the point p rigidly attached to `link`, current q */
void pgo__arm_point_jacobian(const pgo_arm* a, int link, const double p[3], double* J);

/* primitive `prim` of the arm vs an ORIENTED box (centre bpos, rotation bR row-major box->world, half extents):
 * signed distance, closest points pa (robot) / pb (box) and unit direction n robot -> box, all in world frame */
double pgo__prim_obb(const pgo_arm* a, int prim, const double bpos[3], const double bR[9], const double half[3],
                     double pa[3], double pb[3], double n[3]);
int pgo__prim_link(const pgo_arm* a, int prim);
/* world placement of a primitive at the current q: centre, unit axis, radius, half length, type (0 sphere, 1 cylinder) */
void pgo__prim_world(const pgo_arm* a, int prim, double c[3], double ax[3], double* r, double* h, int* type);
#endif
