/* oracle/pgo_dyn.c -- CPU restatement of the PyBullet calls behind panda_gym/pybullet.py for the Panda arm:
 * stepSimulation (btMultiBody ABA + joint motors / joint limits + PGS + semi-implicit Euler),
 * getLinkState / getJointState, resetJointState, setJointMotorControlArray, calculateInverseKinematics.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/pgo.h).  Algorithms follow SURVEY.md App. A.1-A.5, A.7
 * ([UPSTREAM-RECALL] of pybullet 3.2.6 / Bullet 3); pinned by the KATs of test/pybullet_test.py.
 */
#define _USE_MATH_DEFINES
#define _GNU_SOURCE
#include "pgo.h"
#include "pgo_internal.h"
#include "panda_model_data.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

typedef double m3[9];
typedef double m6[36];

struct pgo_arm {
  int nl, nd, np;
  const pgo_link_t* L;
  const pgo_prim_t* P;
  int dof_of_link[PGO_MAXL];
  int link_of_dof[PGO_MAXD];
  double base_pos[3];
  double lin_damp, ang_damp, dt, g[3];
  int solver_iters;
  double residual_thr;
  double max_coord_vel;
  double q[PGO_MAXD], qd[PGO_MAXD];
  /* joint positions behind btMultiBody's m_cachedWorldTransform, i.e. what getLinkState() reports without
   * computeForwardKinematics: forwardKinematics() runs at the START of each stepSimulation, so link poses
   * lag the joint state by one substep.  Pinned by test/pybullet_test.py:139-170 (orientation/velocity KATs
   * match q_19, not q_20).  resetJointState refreshes it. */
  double q_fk[PGO_MAXD];
  int motor_mode[PGO_MAXD];
  double motor_target[PGO_MAXD], motor_maximp[PGO_MAXD];
  double ik_damping, ik_residual;
  int ik_iters;
  /* kinematics cache (valid for the current q) */
  double Rw[PGO_MAXL][9], ow[PGO_MAXL][3], comw[PGO_MAXL][3];
  /* ABA cache for impulse responses */
  double X[PGO_MAXL][36], S[PGO_MAXL][6], U[PGO_MAXL][6], Dinv[PGO_MAXL];
};

/* ------------------------------------------------------------------ small linear algebra */
static void m3mul(const double* a, const double* b, double* c) {
  double t[9];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) t[3 * i + j] = a[3 * i] * b[j] + a[3 * i + 1] * b[3 + j] + a[3 * i + 2] * b[6 + j];
  memcpy(c, t, sizeof t);
}
static void m3vec(const double* a, const double* v, double* o) {
  double t[3];
  for (int i = 0; i < 3; i++) t[i] = a[3 * i] * v[0] + a[3 * i + 1] * v[1] + a[3 * i + 2] * v[2];
  o[0] = t[0]; o[1] = t[1]; o[2] = t[2];
}
static void m3Tvec(const double* a, const double* v, double* o) {
  double t[3];
  for (int i = 0; i < 3; i++) t[i] = a[i] * v[0] + a[3 + i] * v[1] + a[6 + i] * v[2];
  o[0] = t[0]; o[1] = t[1]; o[2] = t[2];
}
static void cross(const double* a, const double* b, double* o) {
  double t[3] = {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]};
  o[0] = t[0]; o[1] = t[1]; o[2] = t[2];
}
static double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static double norm3(const double* a) { return sqrt(dot3(a, a)); }
static void axis_angle_R(const double* ax, double th, double* R) {
  double c = cos(th), s = sin(th), t = 1 - c, x = ax[0], y = ax[1], z = ax[2];
  R[0] = t * x * x + c;     R[1] = t * x * y - s * z; R[2] = t * x * z + s * y;
  R[3] = t * x * y + s * z; R[4] = t * y * y + c;     R[5] = t * y * z - s * x;
  R[6] = t * x * z - s * y; R[7] = t * y * z + s * x; R[8] = t * z * z + c;
}
static void R_to_quat_xyzw(const double* R, double* q) {
  double tr = R[0] + R[4] + R[8];
  if (tr > 0) {
    double s = sqrt(tr + 1.0) * 2;
    q[3] = 0.25 * s; q[0] = (R[7] - R[5]) / s; q[1] = (R[2] - R[6]) / s; q[2] = (R[3] - R[1]) / s;
  } else if (R[0] > R[4] && R[0] > R[8]) {
    double s = sqrt(1.0 + R[0] - R[4] - R[8]) * 2;
    q[3] = (R[7] - R[5]) / s; q[0] = 0.25 * s; q[1] = (R[1] + R[3]) / s; q[2] = (R[2] + R[6]) / s;
  } else if (R[4] > R[8]) {
    double s = sqrt(1.0 + R[4] - R[0] - R[8]) * 2;
    q[3] = (R[2] - R[6]) / s; q[0] = (R[1] + R[3]) / s; q[1] = 0.25 * s; q[2] = (R[5] + R[7]) / s;
  } else {
    double s = sqrt(1.0 + R[8] - R[0] - R[4]) * 2;
    q[3] = (R[3] - R[1]) / s; q[0] = (R[2] + R[6]) / s; q[1] = (R[5] + R[7]) / s; q[2] = 0.25 * s;
  }
}

/* ------------------------------------------------------------------ construction */
pgo_arm* pgo_arm_create(int model, const double base_pos[3]) {
  pgo_arm* a = (pgo_arm*)calloc(1, sizeof *a);
  if (model == 0) { a->L = PGO_CUSTOM0_LINKS; a->nl = PGO_CUSTOM0_NLINKS; a->P = PGO_CUSTOM0_PRIMS; a->np = PGO_CUSTOM0_NPRIMS; }
  else            { a->L = PGO_STOCK_LINKS;   a->nl = PGO_STOCK_NLINKS;   a->P = PGO_STOCK_PRIMS;   a->np = PGO_STOCK_NPRIMS; }
  for (int i = 0; i < a->nl; i++) {
    a->dof_of_link[i] = -1;
    if (a->L[i].jtype != 0) { a->dof_of_link[i] = a->nd; a->link_of_dof[a->nd] = i; a->nd++; }
  }
  for (int k = 0; k < 3; k++) a->base_pos[k] = base_pos ? base_pos[k] : 0.0;
  a->lin_damp = 0.04; a->ang_damp = 0.04;  /* btMultiBody defaults (App. A.3) */
  a->dt = 1.0 / 500.0;                      /* pybullet.py:50 */
  a->g[0] = 0; a->g[1] = 0; a->g[2] = -9.81; /* pybullet.py:54 */
  a->solver_iters = 50; a->residual_thr = 1e-7; /* PyBullet defaults (App. A.1) */
  a->max_coord_vel = 100.0;
  a->ik_damping = 0.5; a->ik_iters = 20; a->ik_residual = 1e-4;
  for (int d = 0; d < a->nd; d++) { /* createJointMotors: velocity motor, target 0, max impulse 1 */
    a->motor_mode[d] = 0; a->motor_target[d] = 0; a->motor_maximp[d] = 1.0;
  }
  return a;
}
void pgo_arm_destroy(pgo_arm* a) { free(a); }
int pgo_arm_ndof(const pgo_arm* a) { return a->nd; }
int pgo_arm_nlinks(const pgo_arm* a) { return a->nl; }
void pgo_arm_set_damping(pgo_arm* a, double lin, double ang) { a->lin_damp = lin; a->ang_damp = ang; }
void pgo_arm_set_solver(pgo_arm* a, int iters, double thr) { a->solver_iters = iters; a->residual_thr = thr; }
void pgo_arm_set_ik(pgo_arm* a, double damping, int iters, double thr) { a->ik_damping = damping; a->ik_iters = iters; a->ik_residual = thr; }
void pgo_arm_reset_joint(pgo_arm* a, int d, double q) { a->q[d] = q; a->qd[d] = 0; memcpy(a->q_fk, a->q, sizeof a->q); }
void pgo_arm_get_q(const pgo_arm* a, double* q, double* qd) {
  for (int d = 0; d < a->nd; d++) { if (q) q[d] = a->q[d]; if (qd) qd[d] = a->qd[d]; }
}
void pgo_arm_set_q(pgo_arm* a, const double* q, const double* qd) {
  for (int d = 0; d < a->nd; d++) { if (q) a->q[d] = q[d]; if (qd) a->qd[d] = qd[d]; }
  memcpy(a->q_fk, a->q, sizeof a->q);
}
void pgo_arm_set_motor(pgo_arm* a, int d, int mode, double target, double max_force) {
  a->motor_mode[d] = mode; a->motor_target[d] = target; a->motor_maximp[d] = max_force * a->dt;
}

/* ------------------------------------------------------------------ forward kinematics */
static void joint_E(const pgo_arm* a, int i, const double* q, double* E) { /* child->parent rotation */
  const pgo_link_t* l = &a->L[i];
  if (l->jtype == 1) { double Rq[9]; axis_angle_R(l->axis, q[a->dof_of_link[i]], Rq); m3mul(l->org_R, Rq, E); }
  else memcpy(E, l->org_R, sizeof(m3));
}
static void fk_q(const pgo_arm* a, const double* q, double Rw[][9], double ow[][3], double comw[][3]) {
  for (int i = 0; i < a->nl; i++) {
    const pgo_link_t* l = &a->L[i];
    double E[9], t[3];
    joint_E(a, i, q, E);
    int p = l->parent;
    if (p < 0) {
      memcpy(Rw[i], E, sizeof(m3));
      for (int k = 0; k < 3; k++) ow[i][k] = a->base_pos[k] + l->org_xyz[k];
    } else {
      m3mul(Rw[p], E, Rw[i]);
      m3vec(Rw[p], l->org_xyz, t);
      for (int k = 0; k < 3; k++) ow[i][k] = ow[p][k] + t[k];
    }
    if (l->jtype == 2) { m3vec(Rw[i], l->axis, t); for (int k = 0; k < 3; k++) ow[i][k] += t[k] * q[a->dof_of_link[i]]; }
    m3vec(Rw[i], l->com, t);
    for (int k = 0; k < 3; k++) comw[i][k] = ow[i][k] + t[k];
  }
}
static void fk(pgo_arm* a) { fk_q(a, a->q, a->Rw, a->ow, a->comw); }

/* world angular velocity of each link and linear velocity of its COM */
static void link_vel(const pgo_arm* a, double Rw[][9], double ow[][3], double comw[][3], double w[][3], double vc[][3]) {
  double vo[PGO_MAXL][3];
  for (int i = 0; i < a->nl; i++) {
    const pgo_link_t* l = &a->L[i];
    int p = l->parent;
    double wp[3] = {0, 0, 0}, vp[3] = {0, 0, 0}, d[3], t[3];
    if (p >= 0) {
      for (int k = 0; k < 3; k++) { wp[k] = w[p][k]; d[k] = ow[i][k] - ow[p][k]; }
      cross(wp, d, t);
      for (int k = 0; k < 3; k++) vp[k] = vo[p][k] + t[k];
    }
    for (int k = 0; k < 3; k++) { w[i][k] = wp[k]; vo[i][k] = vp[k]; }
    if (l->jtype != 0) {
      double axw[3];
      m3vec(Rw[i], l->axis, axw);
      double qd = a->qd[a->dof_of_link[i]];
      if (l->jtype == 1) for (int k = 0; k < 3; k++) w[i][k] += axw[k] * qd;
      else               for (int k = 0; k < 3; k++) vo[i][k] += axw[k] * qd;
    }
    for (int k = 0; k < 3; k++) d[k] = comw[i][k] - ow[i][k];
    cross(w[i], d, t);
    for (int k = 0; k < 3; k++) vc[i][k] = vo[i][k] + t[k];
  }
}

void pgo_arm_link_state(const pgo_arm* a0, int link, double pos[3], double quat[4], double lin[3], double ang[3]) {
  pgo_arm* a = (pgo_arm*)a0;
  fk(a);
  double w[PGO_MAXL][3], vc[PGO_MAXL][3];
  link_vel(a, a->Rw, a->ow, a->comw, w, vc);
  /* pose from the cached (one-substep-old) transforms; velocities = compTreeLinkVelocities (link-local,
   * current q/qd) rotated to world by the CACHED basis */
  double Rc[PGO_MAXL][9], oc[PGO_MAXL][3], cc[PGO_MAXL][3], vl[3], wl[3];
  fk_q(a, a->q_fk, Rc, oc, cc);
  m3Tvec(a->Rw[link], vc[link], vl);
  m3Tvec(a->Rw[link], w[link], wl);
  if (lin) m3vec(Rc[link], vl, lin);
  if (ang) m3vec(Rc[link], wl, ang);
  for (int k = 0; k < 3; k++) if (pos) pos[k] = cc[link][k];
  if (quat) R_to_quat_xyzw(Rc[link], quat); /* inertial rpy = 0: COM frame shares the link orientation */
}
void pgo_arm_link_frame(const pgo_arm* a0, int link, double R[9], double p[3]) {
  /* CURRENT-q frame of the URDF link (what the colliders / getClosestPoints see) */
  pgo_arm* a = (pgo_arm*)a0;
  fk(a);
  memcpy(R, a->Rw[link], sizeof(m3));
  for (int k = 0; k < 3; k++) p[k] = a->ow[link][k];
}

void pgo_arm_jacobian(const pgo_arm* a0, int link, double* Jt, double* Jr) {
  pgo_arm* a = (pgo_arm*)a0;
  fk(a);
  for (int k = 0; k < 3 * a->nd; k++) { Jt[k] = 0; if (Jr) Jr[k] = 0; }
  for (int i = link; i >= 0; i = a->L[i].parent) {
    const pgo_link_t* l = &a->L[i];
    if (l->jtype == 0) continue;
    int d = a->dof_of_link[i];
    double axw[3], r[3], c[3];
    m3vec(a->Rw[i], l->axis, axw);
    if (l->jtype == 1) {
      for (int k = 0; k < 3; k++) r[k] = a->comw[link][k] - a->ow[i][k];
      cross(axw, r, c);
      for (int k = 0; k < 3; k++) { Jt[k * a->nd + d] = c[k]; if (Jr) Jr[k * a->nd + d] = axw[k]; }
    } else {
      for (int k = 0; k < 3; k++) Jt[k * a->nd + d] = axw[k];
    }
  }
}

void pgo__arm_point_jacobian(const pgo_arm* a0, int link, const double p[3], double* J) {
  pgo_arm* a = (pgo_arm*)a0;
  fk(a);
  for (int k = 0; k < 3 * a->nd; k++) J[k] = 0;
  for (int i = link; i >= 0; i = a->L[i].parent) {
    const pgo_link_t* l = &a->L[i];
    if (l->jtype == 0) continue;
    int d = a->dof_of_link[i];
    double axw[3], r[3], c[3];
    m3vec(a->Rw[i], l->axis, axw);
    if (l->jtype == 1) {
      for (int k = 0; k < 3; k++) r[k] = p[k] - a->ow[i][k];
      cross(axw, r, c);
      for (int k = 0; k < 3; k++) J[k * a->nd + d] = c[k];
    } else {
      for (int k = 0; k < 3; k++) J[k * a->nd + d] = axw[k];
    }
  }
}

/* Panda.get_manipulability (panda.py:314-317): rtb.models.Panda().manipulability(q, axes="trans") = Yoshikawa index
 * sqrt(det(Jv Jv^T)), Jv = 3 x 7 translational Jacobian of roboticstoolbox's URDF Panda at its end-effector link
 * panda_link8 (the flange: link7 frame + (0, 0, 0.107); the toolbox declares panda_hand and below as the gripper, so
 * ee_links = [panda_link8], no tool transform).  The toolbox model is franka_description's chain, whose joint origins
 * equal the loaded URDF's (checked in tests against Franka's published DH table); the index does not depend on the
 * base offset.  PARITY UNPINNED: roboticstoolbox is not installable here and the reference holds no vector for it. */
double pgo_arm_manipulability(const pgo_arm* a0) {
  pgo_arm* a = (pgo_arm*)a0;
  fk(a);
  const int flange = 7; /* panda_link8 in both model tables */
  double J[3 * PGO_MAXD], A[3][3];
  pgo__arm_point_jacobian(a0, flange, a->ow[flange], J);
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) {
      double v = 0;
      for (int d = 0; d < 7; d++) v += J[r * a->nd + d] * J[c * a->nd + d];
      A[r][c] = v;
    }
  double det = A[0][0] * (A[1][1] * A[2][2] - A[1][2] * A[2][1]) - A[0][1] * (A[1][0] * A[2][2] - A[1][2] * A[2][0]) +
               A[0][2] * (A[1][0] * A[2][1] - A[1][1] * A[2][0]);
  return det > 0 ? sqrt(det) : 0.0;
}

/* ------------------------------------------------------------------ articulated-body algorithm
 * Spatial vectors are (angular; linear) in the link's COM frame (Bullet's btMultiBody convention).  */
static void build_X(const double* R, const double* rp, double* X) { /* motion transform parent->child */
  /* v_c = R v_p - R [rp]x w_p ; rp = parent COM -> child COM in the parent frame */
  double K[9] = {0, -rp[2], rp[1], rp[2], 0, -rp[0], -rp[1], rp[0], 0}, RK[9];
  m3mul(R, K, RK);
  memset(X, 0, sizeof(m6));
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      X[6 * i + j] = R[3 * i + j];
      X[6 * (i + 3) + j + 3] = R[3 * i + j];
      X[6 * (i + 3) + j] = -RK[3 * i + j];
    }
}
static void m6vec(const double* A, const double* v, double* o) {
  double t[6];
  for (int i = 0; i < 6; i++) { t[i] = 0; for (int j = 0; j < 6; j++) t[i] += A[6 * i + j] * v[j]; }
  memcpy(o, t, sizeof t);
}
static void m6Tvec(const double* A, const double* v, double* o) {
  double t[6];
  for (int i = 0; i < 6; i++) { t[i] = 0; for (int j = 0; j < 6; j++) t[i] += A[6 * j + i] * v[j]; }
  memcpy(o, t, sizeof t);
}
static double dot6(const double* a, const double* b) { double s = 0; for (int i = 0; i < 6; i++) s += a[i] * b[i]; return s; }

/* forward dynamics qdd = ABA(q, qd, tau=0, gravity, damping); fills the impulse-response cache */
static void aba(pgo_arm* a, double* qdd) {
  int n = a->nl;
  static const double zero6[6] = {0};
  double v[PGO_MAXL][6], c[PGO_MAXL][6], pA[PGO_MAXL][6], IA[PGO_MAXL][36], u[PGO_MAXL], acc[PGO_MAXL][6];
  fk(a);
  for (int i = 0; i < n; i++) {
    const pgo_link_t* l = &a->L[i];
    int p = l->parent;
    double E[9], R[9], rp[3];
    joint_E(a, i, a->q, E);
    for (int r = 0; r < 3; r++) for (int s = 0; s < 3; s++) R[3 * r + s] = E[3 * s + r];
    /* parent COM -> child COM, parent frame */
    double cc[3]; m3vec(E, l->com, cc);
    for (int k = 0; k < 3; k++) rp[k] = l->org_xyz[k] - (p >= 0 ? a->L[p].com[k] : 0.0) + cc[k];
    if (l->jtype == 2) { double ax[3]; m3vec(E, l->axis, ax); for (int k = 0; k < 3; k++) rp[k] += ax[k] * a->q[a->dof_of_link[i]]; }
    build_X(R, rp, a->X[i]);
    /* motion subspace in the COM frame */
    double* S = a->S[i];
    memset(S, 0, 6 * sizeof(double));
    if (l->jtype == 1) { for (int k = 0; k < 3; k++) S[k] = l->axis[k]; cross(l->axis, l->com, S + 3); }
    else if (l->jtype == 2) { for (int k = 0; k < 3; k++) S[3 + k] = l->axis[k]; }
    m6vec(a->X[i], p >= 0 ? v[p] : zero6, v[i]);
    memset(c[i], 0, sizeof c[i]);
    if (l->jtype != 0) {
      double qd = a->qd[a->dof_of_link[i]], vJ[6], t1[3], t2[3];
      for (int k = 0; k < 6; k++) vJ[k] = S[k] * qd;
      for (int k = 0; k < 6; k++) v[i][k] += vJ[k];
      cross(v[i], vJ, c[i]);                       /* w x wJ */
      cross(v[i] + 3, vJ, t1); cross(v[i], vJ + 3, t2); /* v x wJ + w x vJ */
      for (int k = 0; k < 3; k++) c[i][3 + k] = t1[k] + t2[k];
    }
    /* rigid-body inertia and bias force */
    memset(IA[i], 0, sizeof(m6));
    for (int k = 0; k < 3; k++) { IA[i][6 * k + k] = l->inertia[k]; IA[i][6 * (k + 3) + k + 3] = l->mass; }
    double Iw[3], gl[3], t[3];
    for (int k = 0; k < 3; k++) Iw[k] = l->inertia[k] * v[i][k];
    cross(v[i], Iw, pA[i]);                         /* gyroscopic w x Iw */
    cross(v[i], v[i] + 3, t);
    for (int k = 0; k < 3; k++) pA[i][3 + k] = l->mass * t[k]; /* m w x v */
    double wn = norm3(v[i]), vn = norm3(v[i] + 3);
    for (int k = 0; k < 3; k++) {                   /* btMultiBody per-link damping */
      pA[i][k] += Iw[k] * (a->ang_damp + a->ang_damp * wn);
      pA[i][3 + k] += l->mass * v[i][3 + k] * (a->lin_damp + a->lin_damp * vn);
    }
    m3Tvec(a->Rw[i], a->g, gl);                     /* gravity as an external force at the COM */
    for (int k = 0; k < 3; k++) pA[i][3 + k] -= l->mass * gl[k];
  }
  for (int i = n - 1; i >= 0; i--) {
    const pgo_link_t* l = &a->L[i];
    double Ia[36], pa[6];
    memcpy(Ia, IA[i], sizeof Ia);
    memcpy(pa, pA[i], sizeof pa);
    if (l->jtype != 0) {
      m6vec(IA[i], a->S[i], a->U[i]);
      double D = dot6(a->S[i], a->U[i]);
      a->Dinv[i] = 1.0 / D;
      u[i] = -dot6(a->S[i], pA[i]);
      for (int r = 0; r < 6; r++) for (int s = 0; s < 6; s++) Ia[6 * r + s] -= a->U[i][r] * a->U[i][s] * a->Dinv[i];
      double Ic[6]; m6vec(Ia, c[i], Ic);
      for (int k = 0; k < 6; k++) pa[k] += Ic[k] + a->U[i][k] * u[i] * a->Dinv[i];
    }
    int p = l->parent;
    if (p >= 0) {
      double T[36], XT[36];
      for (int r = 0; r < 6; r++) for (int s = 0; s < 6; s++) XT[6 * r + s] = a->X[i][6 * s + r];
      /* IA_p += X^T Ia X */
      for (int r = 0; r < 6; r++) for (int s = 0; s < 6; s++) { double z = 0; for (int k = 0; k < 6; k++) z += Ia[6 * r + k] * a->X[i][6 * k + s]; T[6 * r + s] = z; }
      for (int r = 0; r < 6; r++) for (int s = 0; s < 6; s++) { double z = 0; for (int k = 0; k < 6; k++) z += XT[6 * r + k] * T[6 * k + s]; IA[p][6 * r + s] += z; }
      double pp[6]; m6Tvec(a->X[i], pa, pp);
      for (int k = 0; k < 6; k++) pA[p][k] += pp[k];
    }
  }
  for (int i = 0; i < n; i++) {
    const pgo_link_t* l = &a->L[i];
    int p = l->parent;
    double ap[6];
    m6vec(a->X[i], p >= 0 ? acc[p] : zero6, ap);
    for (int k = 0; k < 6; k++) ap[k] += c[i][k];
    if (l->jtype != 0) {
      double dd = (u[i] - dot6(a->U[i], ap)) * a->Dinv[i];
      qdd[a->dof_of_link[i]] = dd;
      for (int k = 0; k < 6; k++) ap[k] += a->S[i][k] * dd;
    }
    memcpy(acc[i], ap, sizeof ap);
  }
}

/* response of the joint accelerations to a joint-space force (== M^-1 tau), using the ABA cache
 * (btMultiBody::calcAccelerationDeltasMultiDof) */
static void aba_delta(const pgo_arm* a, const double* tau, double* out) {
  int n = a->nl;
  static const double zero6[6] = {0};
  double pA[PGO_MAXL][6], u[PGO_MAXL], acc[PGO_MAXL][6];
  memset(pA, 0, sizeof pA);
  for (int i = n - 1; i >= 0; i--) {
    const pgo_link_t* l = &a->L[i];
    double pa[6];
    memcpy(pa, pA[i], sizeof pa);
    if (l->jtype != 0) {
      u[i] = tau[a->dof_of_link[i]] - dot6(a->S[i], pA[i]);
      for (int k = 0; k < 6; k++) pa[k] += a->U[i][k] * u[i] * a->Dinv[i];
    }
    if (l->parent >= 0) { double pp[6]; m6Tvec(a->X[i], pa, pp); for (int k = 0; k < 6; k++) pA[l->parent][k] += pp[k]; }
  }
  for (int i = 0; i < n; i++) {
    const pgo_link_t* l = &a->L[i];
    double ap[6];
    m6vec(a->X[i], l->parent >= 0 ? acc[l->parent] : zero6, ap);
    if (l->jtype != 0) {
      double dd = (u[i] - dot6(a->U[i], ap)) * a->Dinv[i];
      out[a->dof_of_link[i]] = dd;
      for (int k = 0; k < 6; k++) ap[k] += a->S[i][k] * dd;
    }
    memcpy(acc[i], ap, sizeof ap);
  }
}

void pgo_arm_minv(pgo_arm* a, double* minv) {
  double qdd[PGO_MAXD], save[PGO_MAXD];
  memcpy(save, a->qd, sizeof save);
  aba(a, qdd);
  for (int j = 0; j < a->nd; j++) {
    double tau[PGO_MAXD] = {0}, col[PGO_MAXD];
    tau[j] = 1.0;
    aba_delta(a, tau, col);
    for (int i = 0; i < a->nd; i++) minv[i * a->nd + j] = col[i];
  }
}

/* ------------------------------------------------------------------ one stepSimulation() */
typedef pgo__row row_t; /* pgo_internal.h */

/* 1. forward dynamics with gravity/damping, velocity update (solveExternalForces) */
void pgo__arm_predict(pgo_arm* a) {
  int nd = a->nd;
  double qdd[PGO_MAXD];
  memcpy(a->q_fk, a->q, sizeof a->q); /* btMultiBodyDynamicsWorld::solveConstraints -> forwardKinematics() */
  aba(a, qdd);
  for (int d = 0; d < nd; d++) {
    a->qd[d] += qdd[d] * a->dt;
    if (a->qd[d] > a->max_coord_vel) a->qd[d] = a->max_coord_vel;
    if (a->qd[d] < -a->max_coord_vel) a->qd[d] = -a->max_coord_vel;
  }
}

/* 2. constraint rows: joint limits were added to the world at URDF import, motors afterwards */
int pgo__arm_rows(pgo_arm* a, row_t* rows) {
  int nd = a->nd;
  int nr = 0;
  const double erp = 0.2; /* solverInfo.m_erp (joint-limit rows) */
  for (int d = 0; d < nd; d++) {
    const pgo_link_t* l = &a->L[a->link_of_dof[d]];
    if (!l->has_limit) continue;
    for (int side = 0; side < 2; side++) {
      double pen = side == 0 ? a->q[d] - l->lower : l->upper - a->q[d];
      if (pen > 0) continue; /* btMultiBodyJointLimitConstraint: row only when at/over the limit */
      row_t* r = &rows[nr++];
      double tau[PGO_MAXD] = {0};
      r->dof = d; r->sign = side == 0 ? 1.0 : -1.0;
      tau[d] = r->sign;
      aba_delta(a, tau, r->resp);
      double denom = r->sign * r->resp[d];
      r->dinv = denom > 1.1920929e-07 ? 1.0 / denom : 0.0;
      double rel_vel = r->sign * a->qd[d];
      r->rhs = (-pen * erp / a->dt - rel_vel) * r->dinv;
      r->lo = 0; r->hi = 100.0; r->applied = 0; /* btMultiBodyConstraint m_maxAppliedImpulse = 100 */
    }
  }
  for (int d = 0; d < nd; d++) {
    if (a->motor_maximp[d] == 0.0) continue;
    row_t* r = &rows[nr++];
    double tau[PGO_MAXD] = {0};
    r->dof = d; r->sign = 1.0;
    tau[d] = 1.0;
    aba_delta(a, tau, r->resp);
    double denom = r->resp[d];
    r->dinv = denom > 1.1920929e-07 ? 1.0 / denom : 0.0;
    /* btMultiBodyJointMotor: rhs velocity = kp*erp*(q* - q)/dt + v + kd*(v* - v), erp 1 */
    double v = a->qd[d], want;
    if (a->motor_mode[d] == 1) want = 0.1 * (a->motor_target[d] - a->q[d]) / a->dt + v + 1.0 * (0.0 - v);
    else                       want = 0.0 + v + 1.0 * (a->motor_target[d] - v);
    r->rhs = (want - v) * r->dinv;
    r->lo = -a->motor_maximp[d]; r->hi = a->motor_maximp[d]; r->applied = 0;
  }
  return nr;
}

/* one sweep over the limit + motor rows (btMultiBodyConstraintSolver::solveSingleIteration, non-contact
 * part): forward on odd iterations, backward on even.  returns the largest squared residual */
double pgo__arm_sweep(const pgo_arm* a, row_t* rows, int nr, int it, double* dv) {
  int nd = a->nd;
  double lsq = 0;
  for (int j = 0; j < nr; j++) {
    row_t* r = &rows[(it & 1) ? j : nr - 1 - j];
    double delta = r->rhs - r->sign * dv[r->dof] * r->dinv;
    double sum = r->applied + delta;
    if (sum < r->lo) { delta = r->lo - r->applied; r->applied = r->lo; }
    else if (sum > r->hi) { delta = r->hi - r->applied; r->applied = r->hi; }
    else r->applied = sum;
    for (int d = 0; d < nd; d++) dv[d] += r->resp[d] * delta;
    double res = r->dinv != 0.0 ? delta / r->dinv : 0.0;
    if (res * res > lsq) lsq = res * res;
  }
  return lsq;
}

/* 4. apply the solver's velocity change, integrate positions (stepPositionsMultiDof) */
void pgo__arm_finish(pgo_arm* a, const double* dv) {
  int nd = a->nd;
  for (int d = 0; d < nd; d++) {
    a->qd[d] += dv[d];
    if (a->qd[d] > a->max_coord_vel) a->qd[d] = a->max_coord_vel;
    if (a->qd[d] < -a->max_coord_vel) a->qd[d] = -a->max_coord_vel;
  }
  for (int d = 0; d < nd; d++) a->q[d] += a->qd[d] * a->dt;
}
void pgo__arm_response(const pgo_arm* a, const double* tau, double* out) { aba_delta(a, tau, out); }
int pgo__arm_solver_iters(const pgo_arm* a) { return a->solver_iters; }
double pgo__arm_residual_thr(const pgo_arm* a) { return a->residual_thr; }
double pgo__arm_dt(const pgo_arm* a) { return a->dt; }
const double* pgo__arm_qd(const pgo_arm* a) { return a->qd; }

int pgo_arm_substep(pgo_arm* a) {
  pgo__arm_predict(a);
  row_t rows[3 * PGO_MAXD];
  int nr = pgo__arm_rows(a, rows);
  /* 3. projected Gauss-Seidel (btMultiBodyConstraintSolver::solveSingleIteration) */
  double dv[PGO_MAXD] = {0};
  int it;
  for (it = 0; it < a->solver_iters; it++) {
    double lsq = pgo__arm_sweep(a, rows, nr, it, dv);
    if (lsq <= a->residual_thr || it >= a->solver_iters - 1) { it++; break; }
  }
  pgo__arm_finish(a, dv);
  return it;
}

/* ------------------------------------------------------------------ calculateInverseKinematics */
static void solve_spd(int n, double* A, double* b) { /* Gaussian elimination with partial pivoting, in place */
  for (int c = 0; c < n; c++) {
    int piv = c;
    for (int r = c + 1; r < n; r++) if (fabs(A[r * n + c]) > fabs(A[piv * n + c])) piv = r;
    if (piv != c) { for (int k = 0; k < n; k++) { double t = A[c * n + k]; A[c * n + k] = A[piv * n + k]; A[piv * n + k] = t; } double t = b[c]; b[c] = b[piv]; b[piv] = t; }
    for (int r = c + 1; r < n; r++) {
      double f = A[r * n + c] / A[c * n + c];
      for (int k = c; k < n; k++) A[r * n + k] -= f * A[c * n + k];
      b[r] -= f * b[c];
    }
  }
  for (int r = n - 1; r >= 0; r--) { double s = b[r]; for (int k = r + 1; k < n; k++) s -= A[r * n + k] * b[k]; b[r] = s / A[r * n + r]; }
}

void pgo_arm_ik(const pgo_arm* a0, int link, const double pos[3], const double* quat, double* out) {
  pgo_arm tmp = *a0; /* IK iterates a private copy of q (does not move the robot) */
  pgo_arm* a = &tmp;
  int nd = a->nd;
  double diff = 1e30;
  for (int it = 0; it < a->ik_iters && diff > a->ik_residual; it++) {
    double Jt[3 * PGO_MAXD], Jr[3 * PGO_MAXD], e[6] = {0};
    pgo_arm_jacobian(a, link, Jt, Jr);
    /* PyBullet's IK drives the URDF LINK FRAME origin, not the COM ("its link coordinate, not center of
     * mass coordinate"); pinned by the IK KAT (test/pybullet_test.py:254-266). Shift the COM Jacobian. */
    {
      double r[3], c[3];
      for (int k = 0; k < 3; k++) r[k] = a->ow[link][k] - a->comw[link][k];
      for (int d = 0; d < nd; d++) {
        double w[3] = {Jr[d], Jr[nd + d], Jr[2 * nd + d]};
        cross(w, r, c);
        for (int k = 0; k < 3; k++) Jt[k * nd + d] += c[k];
      }
    }
    for (int k = 0; k < 3; k++) e[k] = pos[k] - a->ow[link][k];
    diff = norm3(e);
    int m = 3;
    if (quat) {
      m = 6;
      double qs[4]; R_to_quat_xyzw(a->Rw[link], qs);
      /* deltaQ = endQ * startQ^-1 */
      double qn = sqrt(quat[0] * quat[0] + quat[1] * quat[1] + quat[2] * quat[2] + quat[3] * quat[3]);
      double ax = quat[0] / qn, ay = quat[1] / qn, az = quat[2] / qn, aw = quat[3] / qn; /* target is normalised */
      double bx = -qs[0], by = -qs[1], bz = -qs[2], bw = qs[3];
      double dq[4] = {aw * bx + ax * bw + ay * bz - az * by, aw * by + ay * bw + az * bx - ax * bz,
                      aw * bz + az * bw + ax * by - ay * bx, aw * bw - ax * bx - ay * by - az * bz};
      double w = dq[3] > 1 ? 1 : (dq[3] < -1 ? -1 : dq[3]);
      double angle = 2 * acos(w), s2 = 1 - w * w, axis[3] = {1, 0, 0};
      if (s2 >= 10 * 2.220446049250313e-16) { double s = 1 / sqrt(s2); axis[0] = dq[0] * s; axis[1] = dq[1] * s; axis[2] = dq[2] * s; }
      if (angle > M_PI) angle -= 2 * M_PI;
      double an = norm3(axis);
      for (int k = 0; k < 3; k++) e[3 + k] = angle * axis[k] / an;
    }
    /* dtheta = (J^T J + diag(damping))^-1 J^T e  (Jacobian::CalcDeltaThetasDLS2), clamp 45 deg */
    double A[PGO_MAXD * PGO_MAXD], b[PGO_MAXD];
    for (int i = 0; i < nd; i++) {
      for (int j = 0; j < nd; j++) {
        double s = 0;
        for (int k = 0; k < 3; k++) s += Jt[k * nd + i] * Jt[k * nd + j];
        if (m == 6) for (int k = 0; k < 3; k++) s += Jr[k * nd + i] * Jr[k * nd + j];
        A[i * nd + j] = s + (i == j ? a->ik_damping : 0.0);
      }
      double s = 0;
      for (int k = 0; k < 3; k++) s += Jt[k * nd + i] * e[k];
      if (m == 6) for (int k = 0; k < 3; k++) s += Jr[k * nd + i] * e[3 + k];
      b[i] = s;
    }
    solve_spd(nd, A, b);
    double mx = 0;
    for (int i = 0; i < nd; i++) if (fabs(b[i]) > mx) mx = fabs(b[i]);
    const double max_dls = 45.0 * M_PI / 180.0;
    if (mx > max_dls) for (int i = 0; i < nd; i++) b[i] *= max_dls / mx;
    for (int i = 0; i < nd; i++) a->q[i] += b[i];
  }
  for (int i = 0; i < nd; i++) out[i] = a->q[i];
}

/* accessors for the geometry module */
/* roboticstoolbox 1.1.0 ERobot.ik_lm_chan(Tep, q0, ilimit, slimit, tol, reject_jl=True, we=1, lambda=1.0) as Panda.rtb_ik calls it
 * (panda.py:220-224): Tep = fkine(q_seed) with its translation replaced by `target`, Chan's Levenberg-Marquardt step
 *   e = angle_axis(Te, Tep) (translation error; rotation error of Tep.R Te.R^T as angle * axis),  E = e.e / 2,
 *   q += (J^T J + lambda E I)^-1 J^T e,  J = jacob0 at the toolbox model's end-effector link panda_link8,
 * stop when E < tol, wrap q into (-pi, pi], reject a solution outside the TOOLBOX model's joint limits (franka_description's,
 * not the simulated URDF's) and restart from a random configuration, at most slimit searches of ilimit iterations.
 * The toolbox's C solver draws restart configurations q_i = lo_i + rand() / RAND_MAX * (hi_i - lo_i) from the C library's rand(),
 * never seeded: every SubprocVecEnv worker process therefore walks glibc's srand(1) stream (pgo_crand below, checked against libc
 * in the tests), one state per environment for the life of the process, NOT reset with the episode seed.  About one target in
 * six of the goal shells needs a restart (the first descent ends outside the joint limits), so this stream shapes the start poses.
 * Frames are the robot's own (base at the origin: the toolbox model has no base offset).
 * PARITY UNPINNED: roboticstoolbox is not installable here and the reference holds no vector for this solver [UPSTREAM-RECALL].
 * returns 0 = solved in the first search, 1 = solved after >= 1 restart, 2 = no solution (q_out = last iterate). */
static const double RTB_QLIM_LO[7] = {-2.8973, -1.7628, -2.8973, -3.0718, -2.8973, -0.0175, -2.8973};
static const double RTB_QLIM_HI[7] = {2.8973, 1.7628, 2.8973, -0.0698, 2.8973, 3.7525, 2.8973};
/* glibc rand(): TYPE_3 additive feedback generator r[k] = r[k-31] + r[k-3] after srandom(seed)'s 310 warm-up draws */
void pgo_crand_seed(pgo_crand* c, uint32_t seed) {
  int32_t word = (int32_t)seed;
  c->r[0] = seed;
  for (int i = 1; i < 31; i++) {
    const int32_t hi = word / 127773, lo = word % 127773;
    word = 16807 * lo - 2836 * hi;
    if (word < 0) word += 2147483647;
    c->r[i] = (uint32_t)word;
  }
  c->f = 3; c->b = 0;
  for (int i = 0; i < 310; i++) pgo_crand_next(c);
}
int pgo_crand_next(pgo_crand* c) {
  c->r[c->f] += c->r[c->b];
  const int res = (int)(c->r[c->f] >> 1);
  c->f = (c->f + 1) % 31; c->b = (c->b + 1) % 31;
  return res;
}
int pgo_arm_ik_lm_chan(const pgo_arm* a, const double target[3], const double* q_seed, int ilimit, int slimit, double tol,
                       double lambda, pgo_crand* restart, double* q_out, int* iters_out) {
  const int flange = 7, n = 7;
  double Rw[PGO_MAXL][9], ow[PGO_MAXL][3], comw[PGO_MAXL][3], q[PGO_MAXD] = {0}, Rt[9];
  memcpy(q, q_seed, n * sizeof(double));
  fk_q(a, q, Rw, ow, comw);
  memcpy(Rt, Rw[flange], sizeof Rt);
  int total = 0, solved = 0, search = 1;
  for (; search <= slimit; search++) {
    for (int it = 1; it <= ilimit; it++, total++) {
      fk_q(a, q, Rw, ow, comw);
      double e[6], R[9];
      for (int k = 0; k < 3; k++) e[k] = target[k] - (ow[flange][k] - a->base_pos[k]);
      for (int i = 0; i < 3; i++) for (int j = 0; j < 3; j++) { double s = 0; for (int k = 0; k < 3; k++) s += Rt[3 * i + k] * Rw[flange][3 * j + k]; R[3 * i + j] = s; }
      double li[3] = {R[7] - R[5], R[2] - R[6], R[3] - R[1]}, ln = norm3(li), tr = R[0] + R[4] + R[8];
      if (ln < 1e-6) { for (int k = 0; k < 3; k++) e[3 + k] = tr > 0 ? 0.0 : M_PI / 2 * (R[4 * k] + 1); }
      else { double ang = atan2(ln, tr - 1); for (int k = 0; k < 3; k++) e[3 + k] = ang * li[k] / ln; }
      double E = 0; for (int k = 0; k < 6; k++) E += 0.5 * e[k] * e[k];
      if (E < tol) {
        for (int d = 0; d < n; d++) q[d] = fmod(q[d] + M_PI, 2 * M_PI) - M_PI;
        solved = 1;
        for (int d = 0; d < n; d++) if (q[d] < RTB_QLIM_LO[d] || q[d] > RTB_QLIM_HI[d]) solved = 0;
        break;
      }
      double J[6][7];
      for (int d = 0; d < n; d++) {
        int l = a->link_of_dof[d];
        double axw[3], r[3], c[3];
        m3vec(Rw[l], a->L[l].axis, axw);
        for (int k = 0; k < 3; k++) r[k] = ow[flange][k] - ow[l][k];
        cross(axw, r, c);
        for (int k = 0; k < 3; k++) { J[k][d] = c[k]; J[3 + k][d] = axw[k]; }
      }
      double A[49], g[7];
      for (int i = 0; i < n; i++) {
        for (int j = 0; j < n; j++) { double s = 0; for (int k = 0; k < 6; k++) s += J[k][i] * J[k][j]; A[i * n + j] = s + (i == j ? lambda * E : 0.0); }
        double s = 0; for (int k = 0; k < 6; k++) s += J[k][i] * e[k];
        g[i] = s;
      }
      solve_spd(n, A, g);
      for (int d = 0; d < n; d++) q[d] += g[d];
    }
    if (solved) break;
    for (int d = 0; d < n; d++) { /* random restart inside the toolbox limits */
      const double u = (double)pgo_crand_next(restart) / 2147483647.0;
      q[d] = RTB_QLIM_LO[d] + u * (RTB_QLIM_HI[d] - RTB_QLIM_LO[d]);
    }
  }
  memcpy(q_out, q, n * sizeof(double));
  if (iters_out) *iters_out = total;
  return solved ? (search > 1 ? 1 : 0) : 2;
}

int pgo__arm_prims(const pgo_arm* a, const pgo_prim_t** P) { *P = a->P; return a->np; }
void pgo__arm_fk(pgo_arm* a) { fk(a); }
const double* pgo__arm_Rw(const pgo_arm* a, int link) { return a->Rw[link]; }
const double* pgo__arm_ow(const pgo_arm* a, int link) { return a->ow[link]; }
const double* pgo__arm_base(const pgo_arm* a) { return a->base_pos; }
const pgo_link_t* pgo__arm_links(const pgo_arm* a) { return a->L; }
