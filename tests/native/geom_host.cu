// Host build of the product's closed-form cylinder-box code (panda_gym_b200/csrc/pg_geom.cuh is __host__ __device__) so that
// the CPU test suite can compare it with the oracle's independent fp64 GJK.  Test infrastructure; built by tests/test_geom_closed_form.py.
#include "../../panda_gym_b200/csrc/pg_geom.cuh"

extern "C" {
// out: [0] lb, [1] exact flag, [2] s, [3..5] u, [6] ub (distance from the box's closest point to the solid cylinder)
void geom_cyl_box_bound_f64(const double* c, const double* a, double r, double h, const double* e, double* out) {
  using namespace pg;
  CylBoxBoundT<double> b = cyl_box_bound<double>(v3t<double>(c[0], c[1], c[2]), v3t<double>(a[0], a[1], a[2]), r, h, v3t<double>(e[0], e[1], e[2]));
  D3 on;
  out[0] = b.lb; out[1] = b.exact ? 1.0 : 0.0; out[2] = b.s; out[3] = b.u.x; out[4] = b.u.y; out[5] = b.u.z;
  out[6] = point_cylinder<double>(b.pb, v3t<double>(c[0], c[1], c[2]), v3t<double>(a[0], a[1], a[2]), r, h, on);
}
void geom_cyl_box_bound_f32(const double* c, const double* a, double r, double h, const double* e, double* out) {
  using namespace pg;
  CylBoxBoundT<float> b = cyl_box_bound<float>(v3((float)c[0], (float)c[1], (float)c[2]), v3((float)a[0], (float)a[1], (float)a[2]), (float)r, (float)h,
                                               v3((float)e[0], (float)e[1], (float)e[2]));
  out[0] = b.lb; out[1] = b.exact ? 1.0 : 0.0; out[2] = b.s; out[3] = b.u.x; out[4] = b.u.y; out[5] = b.u.z; out[6] = 0;
}
// thresholded GJK (gjk_cyl_box_decide): the value it returns for threshold thr on the core distance, fp64 and fp32
double geom_gjk_decide_f64(const double* c, const double* a, double r, double h, const double* e, double thr) {
  using namespace pg;
  GjkIn<double> g; g.c = v3t<double>(c[0], c[1], c[2]); g.a = v3t<double>(a[0], a[1], a[2]); g.b = v3t<double>(0.0, 0.0, 0.0);
  g.e = v3t<double>(e[0], e[1], e[2]); g.r = r; g.h = h;
  return gjk_cyl_box_decide<double>(g, thr);
}
double geom_gjk_decide_f32(const double* c, const double* a, double r, double h, const double* e, double thr) {
  using namespace pg;
  GjkIn<float> g; g.c = v3((float)c[0], (float)c[1], (float)c[2]); g.a = v3((float)a[0], (float)a[1], (float)a[2]); g.b = v3(0.f, 0.f, 0.f);
  g.e = v3((float)e[0], (float)e[1], (float)e[2]); g.r = (float)r; g.h = (float)h;
  return (double)gjk_cyl_box_decide<float>(g, (float)thr);
}
// segment vs box: out [0] s, [1] t
void geom_seg_box_f64(const double* c, const double* a, double h, const double* e, double* out) {
  using namespace pg;
  SegBoxT<double> s = seg_box<double>(v3t<double>(c[0], c[1], c[2]), v3t<double>(a[0], a[1], a[2]), h, v3t<double>(e[0], e[1], e[2]));
  out[0] = s.s; out[1] = s.t;
}
}
