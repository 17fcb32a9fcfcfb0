/* oracle/pgo_world.c -- CPU restatement of one PyBullet stepSimulation() for the Push / PickAndPlace world:
 * the Panda arm (pgo_dyn.c) + one free rigid box ("object", push.py:33-39 / pick_and_place.py:38-44) + the static
 * table and plane (pybullet.py:759-815), coupled by frictional contacts.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/pgo.h).
 *
 * PARITY UNPINNED: the reference has no test, golden vector or artefact for Push / PickAndPlace (and cannot even
 * instantiate them: Task.is_truncated is abstract), pybullet is not installable here.  What is restated
 * is Bullet's published formulation [UPSTREAM-RECALL, btMultiBodyConstraintSolver]:
 *   - createMultiBody bodies are btMultiBody bases: free-body update = ABA on a single body (gravity, velocity
 *     damping m v (k + k|v|), I w (k + k|w|), k = 0.04), world-frame velocities, semi-implicit Euler, quaternion
 *     exponential map for the orientation (btMultiBody::stepPositionsMultiDof);
 *   - contact rows (setupMultiBodyContactConstraint): normal row rhs = -(rel_vel + d/dt) for a speculative contact
 *     (d > 0), -(rel_vel) - d*erp/dt for a penetrating one (erp2 = 0.2), impulse >= 0; two friction rows per contact
 *     along btPlaneSpace1(normal), |impulse| <= mu_A*mu_B * normal impulse (pyramid); contacts exist for d < 0.02
 *     (contact breaking threshold);
 *   - solveSingleIteration order: joint-limit + motor rows (alternating sweep), then all normal rows, then all
 *     friction rows; 50 iterations, residual early-out 1e-7.
 * What is simplified (documented deviation): the contact GENERATOR.  Bullet keeps persistent 4-point
 * manifolds (SAT box-box, GJK/EPA for hulls) and warm-starts impulses; here contacts are rebuilt every substep
 * from closed-form features (box vertices vs the table-top / plane face, one closest point per robot primitive vs the
 * box, lowest point of a robot primitive vs the table-top half-space), the 4 closest per pair are kept, no warm
 * start, no spinning friction (panda.py:71-72 sets 0.001).
 */
#define _GNU_SOURCE
#include "pgo.h"
#include "pgo_internal.h"
#include <math.h>
#include <string.h>

#define NC_PAIR 4            /* contacts kept per body pair (size of a btPersistentManifold) */
#define NC_MAX (3 * NC_PAIR)
#define BREAKING 0.02        /* gContactBreakingThreshold */
#define ERP2 0.2             /* contactERP */
#define BODY_DAMP 0.04       /* btMultiBody linear / angular damping */
#define GRAVITY 9.81         /* pybullet.py:54 */
#define PRIM_FIRST 21        /* collision primitives of panda_link7, link8, hand, ee (model table order) */
#define PRIM_LAST 32
#define EE_LINK_IDX 9        /* panda_ee: lateral friction 1.0 (panda.py:69-70, fingers_indices[0] = 9) */

void pgo_world_default(pgo_world* w) {
  memset(w, 0, sizeof *w);
  for (int k = 0; k < 3; k++) w->half[k] = 0.02;           /* object_size 0.04 (push.py:22) */
  w->mass = 1.0;
  w->table_pos[0] = -0.3; w->table_pos[2] = -0.2;           /* create_table(1.1, 0.7, 0.4, x_offset=-0.3) */
  w->table_half[0] = 0.55; w->table_half[1] = 0.35; w->table_half[2] = 0.2;
  w->plane_pos[2] = -0.41;                                   /* create_plane(z_offset=-0.4) */
  w->plane_half[0] = 3.0; w->plane_half[1] = 3.0; w->plane_half[2] = 0.01;
  w->mu_object = 0.5; w->mu_static = 0.5; w->mu_link = 0.5; w->mu_ee = 1.0; /* Bullet default 0.5; panda.py:69 */
}

static void quat_to_R(const double* q, double* R) { /* xyzw -> row-major body->world */
  double x = q[0], y = q[1], z = q[2], w = q[3];
  R[0] = 1 - 2 * (y * y + z * z); R[1] = 2 * (x * y - z * w); R[2] = 2 * (x * z + y * w);
  R[3] = 2 * (x * y + z * w); R[4] = 1 - 2 * (x * x + z * z); R[5] = 2 * (y * z - x * w);
  R[6] = 2 * (x * z - y * w); R[7] = 2 * (y * z + x * w); R[8] = 1 - 2 * (x * x + y * y);
}
static void cross3(const double* a, const double* b, double* o) {
  o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}
static double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

/* btPlaneSpace1 */
static void plane_space(const double* n, double* p, double* q) {
  if (fabs(n[2]) > 0.7071067811865475244) {
    double a = n[1] * n[1] + n[2] * n[2], k = 1.0 / sqrt(a);
    p[0] = 0; p[1] = -n[2] * k; p[2] = n[1] * k;
    q[0] = a * k; q[1] = -n[0] * p[2]; q[2] = n[0] * p[1];
  } else {
    double a = n[0] * n[0] + n[1] * n[1], k = 1.0 / sqrt(a);
    p[0] = -n[1] * k; p[1] = n[0] * k; p[2] = 0;
    q[0] = -n[2] * p[1]; q[1] = n[2] * p[0]; q[2] = a * k;
  }
}

/* one contact: body A is the object (kind 0) or a robot link (kind 1, 2); body B is static (kind 0, 2) or the
 * object (kind 1).  n points from B towards A. */
typedef pgo_contact contact_t; /* pgo.h */
typedef struct {
  int arm, body;               /* which bodies the row touches */
  double Ja[PGO_MAXD], ra[PGO_MAXD], Jb[6], rb[6];
  double rhs, dinv, applied, lo, hi;
} crow_t;

static void keep_closest(contact_t* list, int* n, const contact_t* c) { /* sorted insert, capacity NC_PAIR, stable */
  int pos = *n;
  while (pos > 0 && c->d < list[pos - 1].d) pos--;
  if (pos >= NC_PAIR) return;
  int last = *n < NC_PAIR ? *n : NC_PAIR - 1;
  for (int i = last; i > pos; i--) list[i] = list[i - 1];
  list[pos] = *c;
  if (*n < NC_PAIR) (*n)++;
}

static int over_rect(const double* p, const double* pos, const double* half) {
  return fabs(p[0] - pos[0]) <= half[0] && fabs(p[1] - pos[1]) <= half[1];
}

int pgo_world_contacts(const pgo_arm* a, const pgo_body* b, const pgo_world* w, pgo_contact* out) {
  contact_t ct[NC_PAIR], cr[NC_PAIR], cs[NC_PAIR];
  int nt = 0, nr = 0, ns = 0;
  double R[9];
  quat_to_R(b->quat, R);
  const double top = w->table_pos[2] + w->table_half[2], ptop = w->plane_pos[2] + w->plane_half[2];
  /* object vs table top (vertices over the table) / plane (vertices beside it) */
  for (int v = 0; v < 8; v++) {
    double l[3] = {(v & 1) ? w->half[0] : -w->half[0], (v & 2) ? w->half[1] : -w->half[1], (v & 4) ? w->half[2] : -w->half[2]}, p[3];
    for (int i = 0; i < 3; i++) p[i] = b->pos[i] + R[3 * i] * l[0] + R[3 * i + 1] * l[1] + R[3 * i + 2] * l[2];
    double z = over_rect(p, w->table_pos, w->table_half) ? top : ptop;
    contact_t c; c.kind = 0; c.link = -1; c.d = p[2] - z; c.mu = w->mu_object * w->mu_static;
    if (!(c.d < BREAKING)) continue;
    c.n[0] = 0; c.n[1] = 0; c.n[2] = 1;
    memcpy(c.pa, p, sizeof p); c.pb[0] = p[0]; c.pb[1] = p[1]; c.pb[2] = z;
    keep_closest(ct, &nt, &c);
  }
  /* robot primitives (link7 .. panda_ee) vs object and vs the table-top half-space */
  for (int k = PRIM_FIRST; k <= PRIM_LAST; k++) {
    contact_t c; double dir[3];
    int link = pgo__prim_link(a, k);
    double mu_l = link == EE_LINK_IDX ? w->mu_ee : w->mu_link;
    c.link = link;
    double cc[3], ax[3], r, h; int type;
    pgo__prim_world(a, k, cc, ax, &r, &h, &type);
    c.kind = 1; c.mu = mu_l * w->mu_object;
    c.d = pgo__prim_obb(a, k, b->pos, R, w->half, c.pa, c.pb, dir);
    for (int i = 0; i < 3; i++) c.n[i] = -dir[i];
    int done = 0;
    if (type == 1 && c.d < BREAKING + 0.05) {
      /* cylinder against a box FACE: the closest point of a generator parallel to the face is not unique (it would
       * be picked by rounding noise), so the contact is spread like a manifold: both ends of the generator facing
       * the box, clipped to the face, each with its own distance */
      double dl[3], cl[3], al[3];
      for (int i = 0; i < 3; i++) {
        dl[i] = R[i] * dir[0] + R[3 + i] * dir[1] + R[6 + i] * dir[2];
        al[i] = R[i] * ax[0] + R[3 + i] * ax[1] + R[6 + i] * ax[2];
        cl[i] = R[i] * (cc[0] - b->pos[0]) + R[3 + i] * (cc[1] - b->pos[1]) + R[6 + i] * (cc[2] - b->pos[2]);
      }
      int kf = fabs(dl[0]) > fabs(dl[1]) ? (fabs(dl[0]) > fabs(dl[2]) ? 0 : 2) : (fabs(dl[1]) > fabs(dl[2]) ? 1 : 2);
      if (fabs(dl[kf]) > 0.999) {
        double sg = dl[kf] > 0 ? 1.0 : -1.0, u[3] = {0, 0, 0}, rr = r + 0.001;
        u[kf] = sg;
        double ua = u[0] * al[0] + u[1] * al[1] + u[2] * al[2], L;
        for (int i = 0; i < 3; i++) u[i] -= ua * al[i];
        L = sqrt(dot3(u, u));
        if (L > 1e-6) {
          double g0[3], t0 = -h, t1 = h;
          for (int i = 0; i < 3; i++) g0[i] = cl[i] + rr * u[i] / L;
          for (int j = 0; j < 3; j++) {
            if (j == kf) continue;
            /* |g0_j + t al_j| <= half_j */
            if (fabs(al[j]) < 1e-9) { if (fabs(g0[j]) > w->half[j]) t1 = t0 - 1.0; continue; }
            double ta = (-w->half[j] - g0[j]) / al[j], tb = (w->half[j] - g0[j]) / al[j];
            if (ta > tb) { double t = ta; ta = tb; tb = t; }
            if (ta > t0) t0 = ta;
            if (tb < t1) t1 = tb;
          }
          if (t1 >= t0) {
            done = 1;
            for (int end = 0; end < 2; end++) {
              if (end == 1 && t1 - t0 < 1e-6) break;
              double t = end == 0 ? t0 : t1, gl[3], nl[3] = {0, 0, 0};
              for (int i = 0; i < 3; i++) gl[i] = g0[i] + t * al[i];
              nl[kf] = -sg;
              c.d = -sg * gl[kf] - w->half[kf];
              for (int i = 0; i < 3; i++) {
                c.pa[i] = b->pos[i] + R[3 * i] * gl[0] + R[3 * i + 1] * gl[1] + R[3 * i + 2] * gl[2];
                c.n[i] = R[3 * i] * nl[0] + R[3 * i + 1] * nl[1] + R[3 * i + 2] * nl[2];
              }
              for (int i = 0; i < 3; i++) c.pb[i] = c.pa[i] - c.d * c.n[i];
              if (c.d < BREAKING) keep_closest(cr, &nr, &c);
            }
          }
        }
      }
    }
    if (!done && c.d < BREAKING) keep_closest(cr, &nr, &c);
    /* table top: lowest point(s) of the primitive; a cylinder contributes both ends of its lowest generator */
    c.kind = 2; c.mu = mu_l * w->mu_static;
    c.n[0] = 0; c.n[1] = 0; c.n[2] = 1;
    for (int end = 0; end < 2; end++) {
      if (type == 0) { if (end) break; c.pa[0] = cc[0]; c.pa[1] = cc[1]; c.pa[2] = cc[2] - r; }
      else {
        double rr = r + 0.001; /* hull margin of a URDF cylinder */
        double t = end == 0 ? -h : h;
        double u[3] = {-ax[2] * ax[0], -ax[2] * ax[1], 1.0 - ax[2] * ax[2]}, L = sqrt(dot3(u, u));
        for (int i = 0; i < 3; i++) c.pa[i] = cc[i] + t * ax[i] - (L > 1e-6 ? rr * u[i] / L : 0.0);
      }
      if (!over_rect(c.pa, w->table_pos, w->table_half)) continue;
      c.d = c.pa[2] - top;
      c.pb[0] = c.pa[0]; c.pb[1] = c.pa[1]; c.pb[2] = top;
      if (c.d < BREAKING) keep_closest(cs, &ns, &c);
    }
  }
  int n = 0;
  for (int i = 0; i < nt; i++) out[n++] = ct[i];
  for (int i = 0; i < nr; i++) out[n++] = cr[i];
  for (int i = 0; i < ns; i++) out[n++] = cs[i];
  return n;
}

/* ------------------------------------------------------------------ rows */
static void body_inertia(const pgo_world* w, double* I) {
  /* btBoxShape::calculateLocalInertia: m/12 * (ly^2 + lz^2, ...) on the full extents */
  double lx = 2 * w->half[0], ly = 2 * w->half[1], lz = 2 * w->half[2];
  I[0] = w->mass / 12.0 * (ly * ly + lz * lz); I[1] = w->mass / 12.0 * (lx * lx + lz * lz); I[2] = w->mass / 12.0 * (lx * lx + ly * ly);
}
static void apply_diag_world(const double* R, const double* diag, const double* x, double* out) { /* R diag R^T x */
  double l[3];
  for (int i = 0; i < 3; i++) l[i] = (R[i] * x[0] + R[3 + i] * x[1] + R[6 + i] * x[2]) * diag[i];
  for (int i = 0; i < 3; i++) out[i] = R[3 * i] * l[0] + R[3 * i + 1] * l[1] + R[3 * i + 2] * l[2];
}

/* fill one solver row for contact c along direction t (normal or friction axis) */
static void setup_row(crow_t* r, const pgo_arm* a, const pgo_body* b, const pgo_world* w, const double* R, const contact_t* c,
                      const double* t, int friction) {
  int nd = pgo_arm_ndof(a);
  memset(r, 0, sizeof *r);
  double denom = 0, rel_vel = 0;
  if (c->kind != 0) { /* body A = robot link: joint-space Jacobian of the contact point, impulse +t */
    double J[3 * PGO_MAXD];
    r->arm = 1;
    pgo__arm_point_jacobian(a, c->link, c->pa, J);
    for (int d = 0; d < nd; d++) r->Ja[d] = J[d] * t[0] + J[nd + d] * t[1] + J[2 * nd + d] * t[2];
    pgo__arm_response(a, r->Ja, r->ra);
    const double* qd = pgo__arm_qd(a);
    for (int d = 0; d < nd; d++) { denom += r->Ja[d] * r->ra[d]; rel_vel += r->Ja[d] * qd[d]; }
  }
  if (c->kind != 2) { /* the object: body A for kind 0 (impulse +t at pa), body B for kind 1 (impulse -t at pb) */
    double s = c->kind == 0 ? 1.0 : -1.0, rel[3], tq[3], Iinv[3], I[3];
    const double* p = c->kind == 0 ? c->pa : c->pb;
    r->body = 1;
    for (int k = 0; k < 3; k++) rel[k] = p[k] - b->pos[k];
    cross3(rel, t, tq);
    for (int k = 0; k < 3; k++) { r->Jb[k] = s * t[k]; r->Jb[3 + k] = s * tq[k]; }
    body_inertia(w, I);
    for (int k = 0; k < 3; k++) { Iinv[k] = 1.0 / I[k]; r->rb[k] = r->Jb[k] / w->mass; }
    apply_diag_world(R, Iinv, r->Jb + 3, r->rb + 3);
    for (int k = 0; k < 6; k++) denom += r->Jb[k] * r->rb[k];
    rel_vel += dot3(r->Jb, b->lin) + dot3(r->Jb + 3, b->ang);
  }
  r->dinv = denom > 1.1920929e-07 ? 1.0 / denom : 0.0;
  const double dt = pgo__arm_dt(a);
  if (friction) { r->rhs = -rel_vel * r->dinv; r->lo = -c->mu; r->hi = c->mu; }
  else {
    double pos_err = 0, vel_err = -rel_vel; /* restitution 0 */
    if (c->d > 0) vel_err -= c->d / dt; else pos_err = -c->d * ERP2 / dt;
    r->rhs = (pos_err + vel_err) * r->dinv; r->lo = 0; r->hi = 1e10;
  }
}

static double solve_row(crow_t* r, double lo, double hi, int nd, double* dva, double* dvb) {
  double jv = 0;
  if (r->arm) for (int d = 0; d < nd; d++) jv += r->Ja[d] * dva[d];
  if (r->body) for (int k = 0; k < 6; k++) jv += r->Jb[k] * dvb[k];
  double delta = r->rhs - jv * r->dinv;
  double sum = r->applied + delta;
  if (sum < lo) { delta = lo - r->applied; r->applied = lo; }
  else if (sum > hi) { delta = hi - r->applied; r->applied = hi; }
  else r->applied = sum;
  if (r->arm) for (int d = 0; d < nd; d++) dva[d] += r->ra[d] * delta;
  if (r->body) for (int k = 0; k < 6; k++) dvb[k] += r->rb[k] * delta;
  return r->dinv != 0.0 ? delta / r->dinv : 0.0;
}

/* ------------------------------------------------------------------ one stepSimulation() of arm + object */
int pgo_world_substep(pgo_arm* a, pgo_body* b, const pgo_world* w) {
  const int nd = pgo_arm_ndof(a);
  const double dt = pgo__arm_dt(a);
  double R[9];
  quat_to_R(b->quat, R);
  /* collision detection sees the poses at the start of the step */
  contact_t ct[NC_MAX];
  const int nc = pgo_world_contacts(a, b, w, ct);
  /* 1. unconstrained velocities: arm (ABA), object (single-body ABA: gravity, damping, gyroscopic term) */
  pgo__arm_predict(a);
  {
    double I[3], Iinv[3], Iw[3], gy[3], tq[3], al[3];
    body_inertia(w, I);
    for (int k = 0; k < 3; k++) Iinv[k] = 1.0 / I[k];
    apply_diag_world(R, I, b->ang, Iw);
    cross3(b->ang, Iw, gy);
    double vn = sqrt(dot3(b->lin, b->lin)), wn = sqrt(dot3(b->ang, b->ang));
    for (int k = 0; k < 3; k++) tq[k] = -gy[k] - Iw[k] * (BODY_DAMP + BODY_DAMP * wn);
    apply_diag_world(R, Iinv, tq, al);
    for (int k = 0; k < 3; k++) {
      b->lin[k] += dt * (-(b->lin[k]) * (BODY_DAMP + BODY_DAMP * vn) - (k == 2 ? GRAVITY : 0.0));
      b->ang[k] += dt * al[k];
    }
  }
  /* 2. rows */
  pgo__row rows[3 * PGO_MAXD];
  const int nr = pgo__arm_rows(a, rows);
  crow_t nrow[NC_MAX], frow[2 * NC_MAX];
  for (int i = 0; i < nc; i++) {
    double t1[3], t2[3];
    setup_row(&nrow[i], a, b, w, R, &ct[i], ct[i].n, 0);
    plane_space(ct[i].n, t1, t2);
    setup_row(&frow[2 * i], a, b, w, R, &ct[i], t1, 1);
    setup_row(&frow[2 * i + 1], a, b, w, R, &ct[i], t2, 1);
  }
  /* 3. PGS: [limits, motors] alternating, then normals, then frictions */
  double dva[PGO_MAXD] = {0}, dvb[6] = {0};
  const int iters = pgo__arm_solver_iters(a);
  int it;
  for (it = 0; it < iters; it++) {
    double lsq = pgo__arm_sweep(a, rows, nr, it, dva);
    for (int i = 0; i < nc; i++) { double res = solve_row(&nrow[i], 0.0, 1e10, nd, dva, dvb); if (res * res > lsq) lsq = res * res; }
    for (int i = 0; i < 2 * nc; i++) {
      double tot = nrow[i / 2].applied;
      if (tot > 0) {
        double res = solve_row(&frow[i], -ct[i / 2].mu * tot, ct[i / 2].mu * tot, nd, dva, dvb);
        if (res * res > lsq) lsq = res * res;
      }
    }
    if (lsq <= pgo__arm_residual_thr(a) || it >= iters - 1) { it++; break; }
  }
  /* 4. apply + integrate */
  pgo__arm_finish(a, dva);
  for (int k = 0; k < 3; k++) { b->lin[k] += dvb[k]; b->ang[k] += dvb[3 + k]; }
  for (int k = 0; k < 3; k++) b->pos[k] += b->lin[k] * dt;
  {
    /* btMultiBody::stepPositionsMultiDof, base body: q <- exp(w dt) * q (world-frame w) */
    double fa = sqrt(dot3(b->ang, b->ang)), ax[3], s;
    if (fa * dt > 0.78539816339744830962) fa = 0.78539816339744830962 / dt; /* ANGULAR_MOTION_THRESHOLD */
    if (fa < 0.001) s = 0.5 * dt - dt * dt * dt * 0.020833333333 * fa * fa;
    else s = sin(0.5 * fa * dt) / fa;
    for (int k = 0; k < 3; k++) ax[k] = b->ang[k] * s;
    double cw = cos(fa * dt * 0.5);
    const double* q = b->quat;
    double nq[4] = {cw * q[0] + ax[0] * q[3] + ax[1] * q[2] - ax[2] * q[1],
                    cw * q[1] + ax[1] * q[3] + ax[2] * q[0] - ax[0] * q[2],
                    cw * q[2] + ax[2] * q[3] + ax[0] * q[1] - ax[1] * q[0],
                    cw * q[3] - ax[0] * q[0] - ax[1] * q[1] - ax[2] * q[2]};
    double L = sqrt(nq[0] * nq[0] + nq[1] * nq[1] + nq[2] * nq[2] + nq[3] * nq[3]);
    for (int k = 0; k < 4; k++) b->quat[k] = nq[k] / L;
  }
  return it;
}

/* btQuaternion::getEulerZYX as pybullet.getEulerFromQuaternion reports it: (roll, pitch, yaw) */
void pgo_quat_to_euler(const double q[4], double rpy[3]) {
  double x = q[0], y = q[1], z = q[2], w = q[3];
  double sqx = x * x, sqy = y * y, sqz = z * z, squ = w * w;
  double sarg = -2.0 * (x * z - w * y);
  if (sarg <= -0.99999) { rpy[0] = 0; rpy[1] = -0.5 * M_PI; rpy[2] = 2 * atan2(x, -y); }
  else if (sarg >= 0.99999) { rpy[0] = 0; rpy[1] = 0.5 * M_PI; rpy[2] = 2 * atan2(-x, y); }
  else {
    rpy[0] = atan2(2 * (y * z + w * x), squ - sqx - sqy + sqz);
    rpy[1] = asin(sarg);
    rpy[2] = atan2(2 * (x * y + w * z), squ + sqx - sqy - sqz);
  }
}
