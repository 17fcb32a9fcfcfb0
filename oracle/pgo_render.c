/* oracle/pgo_render.c -- TEST INFRASTRUCTURE ONLY (see pgo.h): fp64 restatement of the headless render("rgb_array").
 *
 * The reference renders through PyBullet's OpenGL path (panda_gym/pybullet.py:117-180: computeViewMatrixFromYawPitchRoll
 * with upAxisIndex = 2, computeProjectionMatrixFOV(60, width / height, 0.1, 100), getCameraImage); no graphics stack
 * exists on the target machines ("unavailable on this box"), so the image itself is unpinned against the reference.  What
 * IS restated from the reference is the camera (Bullet's yaw / pitch / roll convention) and the scene content: the plane
 * and table boxes (pybullet.py:759-815), the robot's collision primitives, the obstacles and their colours
 * (reach_ao.py:819-860), the Push object (push.py:33-39) and the translucent goal marker (reach.py:36-41,
 * reach_ao.py:276-282, push.py:40-47).  Same mathematics as panda_gym_b200/csrc/pg_render.cuh, in double precision. */
#include <math.h>
#include <string.h>
#include "pgo_internal.h"
#include "panda_model_data.h"

int pgo__arm_prims(const pgo_arm* a, const pgo_prim_t** P);

typedef struct { int type, rot; double c[3], a[3], r, h, e[3], rgb[3]; } rprim;
typedef struct { double t, n[3]; } rhit;

static double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

static int hit_sphere(const double* o, const double* d, const rprim* p, double tmin, rhit* h) {
  double oc[3] = {o[0] - p->c[0], o[1] - p->c[1], o[2] - p->c[2]};
  double b = dot3(oc, d), cc = dot3(oc, oc) - p->r * p->r, disc = b * b - cc;
  if (disc < 0) return 0;
  double sq = sqrt(disc), t = -b - sq;
  if (t < tmin) t = -b + sq;
  if (t < tmin || t >= h->t) return 0;
  h->t = t;
  for (int k = 0; k < 3; k++) h->n[k] = (oc[k] + t * d[k]) / p->r;
  return 1;
}

static int hit_cylinder(const double* o, const double* d, const rprim* p, double tmin, rhit* h) {
  double oc[3] = {o[0] - p->c[0], o[1] - p->c[1], o[2] - p->c[2]};
  double dpar = dot3(d, p->a), opar = dot3(oc, p->a), dperp[3], operp[3];
  for (int k = 0; k < 3; k++) { dperp[k] = d[k] - dpar * p->a[k]; operp[k] = oc[k] - opar * p->a[k]; }
  double A = dot3(dperp, dperp), B = dot3(operp, dperp), C = dot3(operp, operp) - p->r * p->r;
  int any = 0;
  if (A > 1e-12) {
    double disc = B * B - A * C;
    if (disc >= 0) {
      double t = (-B - sqrt(disc)) / A, z = opar + t * dpar;
      if (t >= tmin && t < h->t && fabs(z) <= p->h) {
        h->t = t; any = 1;
        for (int k = 0; k < 3; k++) h->n[k] = (operp[k] + t * dperp[k]) / p->r;
      }
    }
  }
  if (fabs(dpar) > 1e-12)
    for (int s = 0; s < 2; s++) {
      double sg = s ? 1.0 : -1.0, t = (sg * p->h - opar) / dpar, q[3];
      for (int k = 0; k < 3; k++) q[k] = operp[k] + t * dperp[k];
      if (t >= tmin && t < h->t && dot3(q, q) <= p->r * p->r) {
        h->t = t; any = 1;
        for (int k = 0; k < 3; k++) h->n[k] = sg * p->a[k];
      }
    }
  return any;
}

/* R = rows of the box -> world rotation */
static int hit_box(const double* o, const double* d, const rprim* p, const double R[9], double tmin, rhit* h) {
  double ow[3] = {o[0] - p->c[0], o[1] - p->c[1], o[2] - p->c[2]}, ol[3], dl[3];
  for (int i = 0; i < 3; i++) {
    if (p->rot) { ol[i] = R[i] * ow[0] + R[3 + i] * ow[1] + R[6 + i] * ow[2]; dl[i] = R[i] * d[0] + R[3 + i] * d[1] + R[6 + i] * d[2]; }
    else { ol[i] = ow[i]; dl[i] = d[i]; }
  }
  double tn = -3.0e38, tf = 3.0e38, sn = 0, sf = 0;
  int an = 0, af = 0;
  for (int k = 0; k < 3; k++) {
    if (fabs(dl[k]) < 1e-12) { if (fabs(ol[k]) > p->e[k]) return 0; continue; }
    double inv = 1.0 / dl[k], t1 = (-p->e[k] - ol[k]) * inv, t2 = (p->e[k] - ol[k]) * inv, s1 = -1, s2 = 1;
    if (t1 > t2) { double t = t1; t1 = t2; t2 = t; s1 = 1; s2 = -1; }
    if (t1 > tn) { tn = t1; an = k; sn = s1; }
    if (t2 < tf) { tf = t2; af = k; sf = s2; }
  }
  if (tn > tf) return 0;
  double t = tn, sg = sn; int ax = an;
  if (t < tmin) { t = tf; ax = af; sg = sf; }
  if (t < tmin || t >= h->t) return 0;
  h->t = t;
  double nl[3] = {0, 0, 0}; nl[ax] = sg;
  for (int i = 0; i < 3; i++) h->n[i] = p->rot ? R[3 * i] * nl[0] + R[3 * i + 1] * nl[1] + R[3 * i + 2] * nl[2] : nl[i];
  return 1;
}

static int hit_prim(const double* o, const double* d, const rprim* p, const double R[9], double tmin, rhit* h) {
  if (p->type == 0) return hit_sphere(o, d, p, tmin, h);
  if (p->type == 1) return hit_cylinder(o, d, p, tmin, h);
  return hit_box(o, d, p, R, tmin, h);
}

static void shade(const double* rgb, const double* n, double* out) {
  const double L[3] = {0.40824829046386307, -0.40824829046386307, 0.81649658092772615}; /* normalize(1, -1, 2) */
  double l = dot3(n, L);
  if (l < 0) l = 0;
  double k = 0.35 + 0.65 * l;
  for (int c = 0; c < 3; c++) out[c] = rgb[c] * k;
}

static void set_prim(rprim* p, int type, const double* c, const double* a, double r, double h, const double* e, double cr, double cg, double cb, int rot) {
  memset(p, 0, sizeof *p);
  p->type = type; p->rot = rot; p->r = r; p->h = h; p->rgb[0] = cr; p->rgb[1] = cg; p->rgb[2] = cb;
  for (int k = 0; k < 3; k++) { p->c[k] = c[k]; if (a) p->a[k] = a[k]; if (e) p->e[k] = e[k]; }
}

void pgo_render(pgo_env* env, int width, int height, const double target[3], double distance, double yaw_deg, double pitch_deg,
                double roll_deg, unsigned char* rgba) {
  /* ---- camera: eye = target + Rz(yaw) Ry(roll) Rx(pitch) (0, -distance, 0), up = third column of that rotation */
  const double D2R = 0.017453292519943295;
  const double y = yaw_deg * D2R, pt = pitch_deg * D2R, r = roll_deg * D2R;
  const double cy = cos(y), sy = sin(y), cp = cos(pt), sp = sin(pt), cr = cos(r), sr = sin(r);
  const double Rm[3][3] = {{cy * cr, cy * sr * sp - sy * cp, cy * sr * cp + sy * sp},
                           {sy * cr, sy * sr * sp + cy * cp, sy * sr * cp - cy * sp},
                           {-sr, cr * sp, cr * cp}};
  double eye[3], up[3], f[3], s[3], u[3], fl = 0, sl = 0;
  for (int k = 0; k < 3; k++) { eye[k] = target[k] - Rm[k][1] * distance; up[k] = Rm[k][2]; }
  for (int k = 0; k < 3; k++) { f[k] = target[k] - eye[k]; fl += f[k] * f[k]; }
  fl = sqrt(fl);
  for (int k = 0; k < 3; k++) f[k] /= fl;
  s[0] = f[1] * up[2] - f[2] * up[1]; s[1] = f[2] * up[0] - f[0] * up[2]; s[2] = f[0] * up[1] - f[1] * up[0];
  for (int k = 0; k < 3; k++) sl += s[k] * s[k];
  sl = sqrt(sl);
  for (int k = 0; k < 3; k++) s[k] /= sl;
  u[0] = s[1] * f[2] - s[2] * f[1]; u[1] = s[2] * f[0] - s[0] * f[2]; u[2] = s[0] * f[1] - s[1] * f[0];
  const double thf = tan(0.5 * 60.0 * D2R), aspect = (double)width / (double)height, near_plane = 0.1, far_plane = 100.0;

  /* ---- scene */
  rprim prims[8 + 36 + PGO_MAXOBS], goal;
  int np = 0;
  const double z3[3] = {0, 0, 0};
  pgo_env_config cfg; pgo_env_get_config(env, &cfg);
  { const double c[3] = {0, 0, -0.41}, e[3] = {3, 3, 0.01}; set_prim(&prims[np++], 2, c, z3, 0, 0, e, 0.15, 0.15, 0.15, 0); }
  { double c[3], e[3]; pgo_env_get_table(env, c, e); set_prim(&prims[np++], 2, c, z3, 0, 0, e, 0.3, 0.3, 0.3, 0); }
  pgo_arm* arm = pgo_env_arm(env);
  const pgo_prim_t* P; const int narm = pgo__arm_prims(arm, &P);
  for (int i = 0; i < narm; i++) {
    double c[3], a[3], rr, hh; int type;
    pgo__prim_world(arm, i, c, a, &rr, &hh, &type);
    set_prim(&prims[np++], type, c, a, rr, hh, z3, 1, 1, 1, 0);
  }
  double Rb[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  if (cfg.task == 1) {
    pgo_obstacle ob[PGO_MAXOBS];
    const int nobs = pgo_env_get_obstacles(env, ob);
    for (int j = 0; j < nobs; j++) {
      if (ob[j].type < 0 || ob[j].pos[2] < -50.0) continue;
      if (ob[j].type == 0) set_prim(&prims[np++], 0, ob[j].pos, z3, ob[j].size[0], 0, z3, 1, 0, 0, 0);
      else if (ob[j].type == 1) set_prim(&prims[np++], 2, ob[j].pos, z3, 0, 0, ob[j].size, 1, 0.5, 0, 0);
      else set_prim(&prims[np++], 2, ob[j].pos, z3, 0, 0, ob[j].size, 0.6, 0.6, 0.6, 0);  /* benchmark-scene box */
    }
  }
  double g[3]; pgo_env_get_goal(env, g);
  if (cfg.task >= 2) {
    const pgo_body* b = pgo_env_object(env);
    const pgo_world* w = pgo_env_world(env);
    set_prim(&prims[np++], 2, b->pos, z3, 0, 0, w->half, 0.1, 0.9, 0.1, 1);
    const double x = b->quat[0], yy = b->quat[1], z = b->quat[2], ww = b->quat[3];
    Rb[0] = 1 - 2 * (yy * yy + z * z); Rb[1] = 2 * (x * yy - z * ww); Rb[2] = 2 * (x * z + yy * ww);
    Rb[3] = 2 * (x * yy + z * ww); Rb[4] = 1 - 2 * (x * x + z * z); Rb[5] = 2 * (yy * z - x * ww);
    Rb[6] = 2 * (x * z - yy * ww); Rb[7] = 2 * (yy * z + x * ww); Rb[8] = 1 - 2 * (x * x + yy * yy);
    set_prim(&goal, 2, g, z3, 0, 0, w->half, 0.1, 0.9, 0.1, 0);
  } else set_prim(&goal, 0, g, z3, 0.02, 0, z3, 0.1, 0.9, 0.1, 0);
  const double alpha = cfg.task == 1 ? 0.8 : 0.3;
  const double I9[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};

  for (int py = 0; py < height; py++)
    for (int px = 0; px < width; px++) {
      const double sx = (2.0 * (px + 0.5) / width - 1.0) * aspect * thf, sy2 = (1.0 - 2.0 * (py + 0.5) / height) * thf;
      double d[3], dl = 0;
      for (int k = 0; k < 3; k++) { d[k] = f[k] + sx * s[k] + sy2 * u[k]; dl += d[k] * d[k]; }
      dl = sqrt(dl);
      for (int k = 0; k < 3; k++) d[k] /= dl;
      const double tmin = near_plane * dl, tmax = far_plane * dl;
      rhit h; h.t = tmax; h.n[0] = 0; h.n[1] = 0; h.n[2] = 1;
      int best = -1;
      for (int k = 0; k < np; k++)
        if (hit_prim(eye, d, &prims[k], Rb, tmin, &h)) best = k;
      double col[3] = {0.7, 0.7, 0.8};
      if (best >= 0) shade(prims[best].rgb, h.n, col);
      rhit hg; hg.t = h.t; hg.n[0] = 0; hg.n[1] = 0; hg.n[2] = 1;
      if (hit_prim(eye, d, &goal, I9, tmin, &hg)) {
        double gc[3];
        shade(goal.rgb, hg.n, gc);
        for (int c = 0; c < 3; c++) col[c] = alpha * gc[c] + (1.0 - alpha) * col[c];
      }
      unsigned char* o = rgba + ((size_t)py * width + px) * 4;
      for (int c = 0; c < 3; c++) { double v = col[c] < 0 ? 0 : (col[c] > 1 ? 1 : col[c]); o[c] = (unsigned char)(v * 255.0 + 0.5); }
      o[3] = 255;
    }
}
