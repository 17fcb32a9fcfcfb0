/* oracle/pgo_geom.c -- CPU restatement of PyBullet getClosestPoints for the shape pairs the reference uses
 * (reach_ao.py:896-900, 1169-1181; pyb_utils-fork CollisionDetector): robot link primitives
 * (spheres + cylinders of franka_panda_custom_0) vs sphere / axis-aligned box bodies.
 *
 * TEST INFRASTRUCTURE ONLY (see oracle/pgo.h).  PARITY UNPINNED: no reference test touches this path; the
 * shape semantics follow SURVEY.md App. A.2/A.6 ([UPSTREAM-RECALL] of Bullet 3):
 *   - sphere: exact.
 *   - URDF cylinder: convex hull (32-gon) carrying the default collision margin 0.001 on top of its
 *     vertices.  Restated as the exact solid cylinder (r, L) inflated by 0.001 (deviation <= r(1-cos(pi/32))
 *     = 0.29 mm for r = 0.06, inside the hull's own polygonal error).
 *   - box: core shrunk by margin m = min(0.04, 0.1*min half extent) (btBoxShape::setSafeMargin), inflated
 *     back by m => rounded edges of radius m.  sphere-box is Bullet's analytic btSphereBoxCollisionAlgorithm;
 *     cylinder-box is GJK on the two cores minus both margins.
 *   - deep penetration (cores overlap) reports -(sum of margins): only the sign is consumed by the path.
 */
#include "pgo.h"
#include "pgo_internal.h"
#include "panda_model_data.h"
#include <math.h>
#include <string.h>

#define HULL_MARGIN 0.001

int pgo__arm_prims(const pgo_arm* a, const pgo_prim_t** P);
void pgo__arm_fk(pgo_arm* a);
const double* pgo__arm_Rw(const pgo_arm* a, int link);
const double* pgo__arm_ow(const pgo_arm* a, int link);
const double* pgo__arm_base(const pgo_arm* a);

static double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static void sub3(const double* a, const double* b, double* o) { o[0] = a[0] - b[0]; o[1] = a[1] - b[1]; o[2] = a[2] - b[2]; }
static void cross3(const double* a, const double* b, double* o) {
  double t[3] = {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]};
  o[0] = t[0]; o[1] = t[1]; o[2] = t[2];
}

double pgo_obstacle_margin(const pgo_obstacle* ob) { return ob->type == 2 ? 0.001 : pgo_box_margin(ob->size); }
double pgo_box_margin(const double half[3]) {
  double mn = half[0] < half[1] ? half[0] : half[1];
  if (half[2] < mn) mn = half[2];
  return 0.1 * mn < 0.04 ? 0.1 * mn : 0.04;
}

/* ------------------------------------------------------------------ point vs solid cylinder */
static double point_cylinder(const double p[3], const double c[3], const double a[3], double r, double h, double on_cyl[3]) {
  double d[3], rho[3];
  sub3(p, c, d);
  double t = dot3(d, a);
  for (int k = 0; k < 3; k++) rho[k] = d[k] - t * a[k];
  double rn = sqrt(dot3(rho, rho));
  double tc = t > h ? h : (t < -h ? -h : t);
  double dt = fabs(t) - h, dr = rn - r;
  double rc = rn > r ? r : rn;
  for (int k = 0; k < 3; k++) on_cyl[k] = c[k] + tc * a[k] + (rn > 1e-300 ? rho[k] / rn * rc : 0.0);
  if (dt <= 0 && dr <= 0) { /* inside: minimum translation to the surface */
    if (-dt < -dr) { for (int k = 0; k < 3; k++) on_cyl[k] = c[k] + (t >= 0 ? h : -h) * a[k] + rho[k]; return dt; }
    for (int k = 0; k < 3; k++) on_cyl[k] = c[k] + t * a[k] + (rn > 1e-300 ? rho[k] / rn * r : 0.0);
    return dr;
  }
  double ddt = dt > 0 ? dt : 0, ddr = dr > 0 ? dr : 0;
  return sqrt(ddt * ddt + ddr * ddr);
}

/* ------------------------------------------------------------------ point vs core box (Bullet sphere-box) */
static double point_corebox(const double p[3], const double b[3], const double e[3], double on_box[3], double n[3]) {
  double rel[3], cl[3];
  sub3(p, b, rel);
  for (int k = 0; k < 3; k++) cl[k] = rel[k] > e[k] ? e[k] : (rel[k] < -e[k] ? -e[k] : rel[k]);
  double v[3];
  sub3(rel, cl, v);
  double d2 = dot3(v, v);
  if (d2 <= 1.1920929e-07) { /* centre inside the core: getSpherePenetration */
    double best = 1e300; int bk = 0; double bs = 1;
    for (int k = 0; k < 3; k++) {
      double dp = e[k] - rel[k], dm = e[k] + rel[k];
      if (dp < best) { best = dp; bk = k; bs = 1; }
      if (dm < best) { best = dm; bk = k; bs = -1; }
    }
    n[0] = n[1] = n[2] = 0; n[bk] = bs;
    for (int k = 0; k < 3; k++) on_box[k] = b[k] + rel[k];
    on_box[bk] = b[bk] + bs * e[bk];
    return -best;
  }
  double d = sqrt(d2);
  for (int k = 0; k < 3; k++) { n[k] = v[k] / d; on_box[k] = b[k] + cl[k]; }
  return d;
}

/* ------------------------------------------------------------------ GJK distance, cylinder core vs box core */
typedef struct { double c[3], a[3], r, h, b[3], e[3]; } gjk_pair;
static void support(const gjk_pair* g, const double d[3], double sa[3], double sb[3]) {
  /* support of the cylinder along -d ... caller passes the direction for A; B gets the opposite */
  double da = dot3(d, g->a), rad[3];
  for (int k = 0; k < 3; k++) rad[k] = d[k] - da * g->a[k];
  double rn = sqrt(dot3(rad, rad));
  for (int k = 0; k < 3; k++) sa[k] = g->c[k] + (da >= 0 ? g->h : -g->h) * g->a[k] + (rn > 1e-14 ? g->r * rad[k] / rn : 0.0);
  for (int k = 0; k < 3; k++) sb[k] = g->b[k] + (-d[k] >= 0 ? g->e[k] : -g->e[k]);
}

/* closest point to the origin on a simplex of n points W; returns reduced simplex + barycentric weights */
/* Ericson, closest point on triangle to the origin; idx = which vertices stay */
static void closest_triangle(const double* a, const double* b, const double* c, double* l, int* mask) {
  double ab[3], ac[3], ap[3], bp[3], cp[3];
  sub3(b, a, ab); sub3(c, a, ac);
  for (int k = 0; k < 3; k++) { ap[k] = -a[k]; bp[k] = -b[k]; cp[k] = -c[k]; }
  double d1 = dot3(ab, ap), d2 = dot3(ac, ap);
  if (d1 <= 0 && d2 <= 0) { l[0] = 1; l[1] = l[2] = 0; *mask = 1; return; }
  double d3 = dot3(ab, bp), d4 = dot3(ac, bp);
  if (d3 >= 0 && d4 <= d3) { l[1] = 1; l[0] = l[2] = 0; *mask = 2; return; }
  double vc = d1 * d4 - d3 * d2;
  if (vc <= 0 && d1 >= 0 && d3 <= 0) { double v = d1 / (d1 - d3); l[0] = 1 - v; l[1] = v; l[2] = 0; *mask = 3; return; }
  double d5 = dot3(ab, cp), d6 = dot3(ac, cp);
  if (d6 >= 0 && d5 <= d6) { l[2] = 1; l[0] = l[1] = 0; *mask = 4; return; }
  double vb = d5 * d2 - d1 * d6;
  if (vb <= 0 && d2 >= 0 && d6 <= 0) { double w = d2 / (d2 - d6); l[0] = 1 - w; l[2] = w; l[1] = 0; *mask = 5; return; }
  double va = d3 * d6 - d5 * d4;
  if (va <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0) { double w = (d4 - d3) / ((d4 - d3) + (d5 - d6)); l[1] = 1 - w; l[2] = w; l[0] = 0; *mask = 6; return; }
  double den = 1.0 / (va + vb + vc);
  l[1] = vb * den; l[2] = vc * den; l[0] = 1 - l[1] - l[2]; *mask = 7;
}

double pgo_gjk_cyl_box(const double c[3], const double a[3], double r, double h, const double b[3], const double e[3], double pa[3], double pb[3]) {
  gjk_pair g;
  memcpy(g.c, c, 24); memcpy(g.a, a, 24); g.r = r; g.h = h; memcpy(g.b, b, 24); memcpy(g.e, e, 24);
  double W[4][3], A[4][3], B[4][3], lam[4] = {1, 0, 0, 0};
  int n = 0;
  double v[3];
  sub3(c, b, v);
  if (dot3(v, v) < 1e-24) { v[0] = 1; v[1] = 0; v[2] = 0; }
  for (int it = 0; it < 100; it++) {
    double d[3] = {-v[0], -v[1], -v[2]}, sa[3], sb[3], w[3];
    support(&g, d, sa, sb);
    sub3(sa, sb, w);
    double vv = dot3(v, v);
    if (n > 0 && vv - dot3(v, w) <= 1e-14 * vv + 1e-22) break; /* no progress possible: v is the closest point */
    memcpy(W[n], w, 24); memcpy(A[n], sa, 24); memcpy(B[n], sb, 24); n++;
    if (n == 1) { lam[0] = 1; }
    else if (n == 2) {
      double ab[3]; sub3(W[1], W[0], ab);
      double t = -dot3(W[0], ab), den = dot3(ab, ab);
      if (t <= 0) { n = 1; lam[0] = 1; }
      else if (t >= den) { memcpy(W[0], W[1], 24); memcpy(A[0], A[1], 24); memcpy(B[0], B[1], 24); n = 1; lam[0] = 1; }
      else { lam[1] = t / den; lam[0] = 1 - lam[1]; }
    } else if (n == 3) {
      double l[3]; int mask;
      closest_triangle(W[0], W[1], W[2], l, &mask);
      int m = 0;
      for (int k = 0; k < 3; k++) if (mask & (1 << k)) { if (m != k) { memcpy(W[m], W[k], 24); memcpy(A[m], A[k], 24); memcpy(B[m], B[k], 24); } lam[m] = l[k]; m++; }
      n = m;
    } else { /* tetrahedron: origin inside => overlap; else closest face */
      static const int F[4][3] = {{0, 1, 2}, {0, 1, 3}, {0, 2, 3}, {1, 2, 3}};
      static const int O[4] = {3, 2, 1, 0};
      double best = 1e300, bl[3]; int bmask = 0, bf = -1, inside = 1;
      for (int f = 0; f < 4; f++) {
        const double *p0 = W[F[f][0]], *p1 = W[F[f][1]], *p2 = W[F[f][2]], *po = W[O[f]];
        double e1[3], e2[3], nn[3], t0[3], t1[3];
        sub3(p1, p0, e1); sub3(p2, p0, e2); cross3(e1, e2, nn);
        sub3(po, p0, t0);
        for (int k = 0; k < 3; k++) t1[k] = -p0[k];
        double so = dot3(nn, t0), sp = dot3(nn, t1);
        /* the origin is inside only if it lies strictly on the opposite-vertex side of every face of a
         * NON-degenerate tetrahedron; the closest face is searched over all four regardless */
        if (!(so * sp > 0) || fabs(so) <= 1e-10 * sqrt(dot3(nn, nn) * dot3(t0, t0))) inside = 0;
        double l[3]; int mask;
        closest_triangle(p0, p1, p2, l, &mask);
        double q[3] = {0, 0, 0};
        for (int k = 0; k < 3; k++) for (int x = 0; x < 3; x++) q[x] += l[k] * W[F[f][k]][x];
        double d2 = dot3(q, q);
        if (d2 < best) { best = d2; bf = f; bmask = mask; memcpy(bl, l, sizeof bl); }
      }
      if (inside) { /* cores overlap */
        for (int k = 0; k < 3; k++) { pa[k] = c[k]; pb[k] = c[k]; }
        return 0.0;
      }
      double TW[3][3], TA[3][3], TB[3][3];
      for (int k = 0; k < 3; k++) { memcpy(TW[k], W[F[bf][k]], 24); memcpy(TA[k], A[F[bf][k]], 24); memcpy(TB[k], B[F[bf][k]], 24); }
      int m = 0;
      for (int k = 0; k < 3; k++) if (bmask & (1 << k)) { memcpy(W[m], TW[k], 24); memcpy(A[m], TA[k], 24); memcpy(B[m], TB[k], 24); lam[m] = bl[k]; m++; }
      n = m;
    }
    v[0] = v[1] = v[2] = 0;
    for (int k = 0; k < n; k++) for (int x = 0; x < 3; x++) v[x] += lam[k] * W[k][x];
    if (dot3(v, v) < 1e-24) { /* touching / overlapping cores */
      for (int x = 0; x < 3; x++) { pa[x] = 0; for (int k = 0; k < n; k++) pa[x] += lam[k] * A[k][x]; pb[x] = pa[x]; }
      return 0.0;
    }
  }
  for (int x = 0; x < 3; x++) { pa[x] = 0; pb[x] = 0; for (int k = 0; k < n; k++) { pa[x] += lam[k] * A[k][x]; pb[x] += lam[k] * B[k][x]; } }
  return sqrt(dot3(v, v));
}

/* ------------------------------------------------------------------ primitive vs obstacle */
static void prim_world(const pgo_arm* a, const pgo_prim_t* p, double c[3], double ax[3]) {
  if (p->link < 0) { /* base link: identity orientation at the base position */
    const double* bp = pgo__arm_base(a);
    for (int k = 0; k < 3; k++) { c[k] = bp[k] + p->xyz[k]; ax[k] = p->axis[k]; }
    return;
  }
  const double* R = pgo__arm_Rw(a, p->link);
  const double* o = pgo__arm_ow(a, p->link);
  for (int i = 0; i < 3; i++) {
    c[i] = o[i] + R[3 * i] * p->xyz[0] + R[3 * i + 1] * p->xyz[1] + R[3 * i + 2] * p->xyz[2];
    ax[i] = R[3 * i] * p->axis[0] + R[3 * i + 1] * p->axis[1] + R[3 * i + 2] * p->axis[2];
  }
}

static double prim_obstacle(const double c[3], const double ax[3], const pgo_prim_t* p, const pgo_obstacle* ob, double pa[3], double pb[3]) {
  double n[3];
  if (p->type == 0) { /* robot sphere */
    if (ob->type == 0) { /* sphere-sphere */
      double d[3]; sub3(ob->pos, c, d);
      double L = sqrt(dot3(d, d));
      for (int k = 0; k < 3; k++) { n[k] = L > 1.1920929e-07 ? d[k] / L : (k == 0); pa[k] = c[k] + n[k] * p->radius; pb[k] = ob->pos[k] - n[k] * ob->size[0]; }
      return L - p->radius - ob->size[0];
    }
    double m = pgo_obstacle_margin(ob), e[3] = {ob->size[0] - m, ob->size[1] - m, ob->size[2] - m}, onb[3];
    double d = point_corebox(c, ob->pos, e, onb, n);
    for (int k = 0; k < 3; k++) { pb[k] = onb[k] + n[k] * m; pa[k] = c[k] - n[k] * p->radius; }
    return d - m - p->radius;
  }
  /* robot cylinder (hull + margin) */
  if (ob->type == 0) {
    double onc[3];
    double d = point_cylinder(ob->pos, c, ax, p->radius, p->half_len, onc);
    double v[3]; sub3(ob->pos, onc, v);
    double L = sqrt(dot3(v, v));
    if (d < 0) { for (int k = 0; k < 3; k++) v[k] = -v[k]; }
    for (int k = 0; k < 3; k++) { n[k] = L > 1e-300 ? v[k] / L : (k == 0); pa[k] = onc[k] + n[k] * HULL_MARGIN; pb[k] = ob->pos[k] - n[k] * ob->size[0]; }
    return d - HULL_MARGIN - ob->size[0];
  }
  double m = pgo_obstacle_margin(ob), e[3] = {ob->size[0] - m, ob->size[1] - m, ob->size[2] - m}, qa[3], qb[3];
  double d = pgo_gjk_cyl_box(c, ax, p->radius, p->half_len, ob->pos, e, qa, qb);
  double v[3]; sub3(qb, qa, v);
  double L = sqrt(dot3(v, v));
  for (int k = 0; k < 3; k++) { n[k] = L > 1e-300 ? v[k] / L : 0.0; pa[k] = qa[k] + n[k] * HULL_MARGIN; pb[k] = qb[k] - n[k] * m; }
  return d - HULL_MARGIN - m;
}

double pgo_prim_obstacle_distance(const pgo_arm* a, int prim, const pgo_obstacle* ob, double pa[3], double pb[3]) {
  const pgo_prim_t* P; pgo__arm_prims(a, &P);
  pgo__arm_fk((pgo_arm*)a);
  double c[3], ax[3], ta[3], tb[3];
  prim_world(a, &P[prim], c, ax);
  return prim_obstacle(c, ax, &P[prim], ob, pa ? pa : ta, pb ? pb : tb);
}

double pgo_link_obstacle_distance(const pgo_arm* a, int link, const pgo_obstacle* ob, double pa[3], double pb[3]) {
  const pgo_prim_t* P; int np = pgo__arm_prims(a, &P);
  pgo__arm_fk((pgo_arm*)a);
  double best = 1e300;
  for (int i = 0; i < np; i++) {
    if (P[i].link != link) continue;
    double c[3], ax[3], ta[3], tb[3];
    prim_world(a, &P[i], c, ax);
    double d = prim_obstacle(c, ax, &P[i], ob, ta, tb);
    if (d < best) { best = d; for (int k = 0; k < 3; k++) { if (pa) pa[k] = ta[k]; if (pb) pb[k] = tb[k]; } }
  }
  return best;
}

/* ------------------------------------------------------------------ primitive vs oriented box (contact pass of
 * the Push / PickAndPlace world, pgo_world.c): evaluated in the box frame with the routines above */
int pgo__prim_link(const pgo_arm* a, int prim) { const pgo_prim_t* P; pgo__arm_prims(a, &P); return P[prim].link; }
double pgo__prim_obb(const pgo_arm* a, int prim, const double bpos[3], const double bR[9], const double half[3],
                     double pa[3], double pb[3], double n[3]) {
  const pgo_prim_t* P; pgo__arm_prims(a, &P);
  pgo__arm_fk((pgo_arm*)a);
  double c[3], ax[3], cl[3], al[3], la[3], lb[3];
  prim_world(a, &P[prim], c, ax);
  for (int i = 0; i < 3; i++) { /* world -> box frame: R^T (x - bpos) */
    cl[i] = 0; al[i] = 0;
    for (int k = 0; k < 3; k++) { cl[i] += bR[3 * k + i] * (c[k] - bpos[k]); al[i] += bR[3 * k + i] * ax[k]; }
  }
  pgo_obstacle ob; ob.type = 1;
  for (int k = 0; k < 3; k++) { ob.pos[k] = 0.0; ob.size[k] = half[k]; }
  double d = prim_obstacle(cl, al, &P[prim], &ob, la, lb);
  double v[3] = {lb[0] - la[0], lb[1] - la[1], lb[2] - la[2]};
  double L = sqrt(dot3(v, v));
  double nl[3];
  if (L > 1e-9) { double s = d >= 0 ? 1.0 / L : -1.0 / L; for (int k = 0; k < 3; k++) nl[k] = v[k] * s; }
  else { /* deep overlap (cores intersect): push along the line of centres */
    double cn = sqrt(dot3(cl, cl));
    for (int k = 0; k < 3; k++) nl[k] = cn > 1e-12 ? -cl[k] / cn : (k == 2 ? -1.0 : 0.0);
  }
  for (int i = 0; i < 3; i++) {
    pa[i] = bpos[i]; pb[i] = bpos[i]; n[i] = 0;
    for (int k = 0; k < 3; k++) { pa[i] += bR[3 * i + k] * la[k]; pb[i] += bR[3 * i + k] * lb[k]; n[i] += bR[3 * i + k] * nl[k]; }
  }
  return d;
}
void pgo__prim_world(const pgo_arm* a, int prim, double c[3], double ax[3], double* r, double* h, int* type) {
  const pgo_prim_t* P; pgo__arm_prims(a, &P);
  pgo__arm_fk((pgo_arm*)a);
  prim_world(a, &P[prim], c, ax);
  *r = P[prim].radius; *h = P[prim].half_len; *type = P[prim].type;
}
