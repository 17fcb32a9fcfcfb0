// pg_packed.cuh -- two environments per thread on Blackwell's packed fp32 pipe.
//
// sm_100 adds FFMA2 / FADD2 / FMUL2 (PTX fma.rn.f32x2 ...): one instruction, one issue slot, two independent fp32
// lanes held in an aligned 64-bit register pair.  The step kernel is issue-bound (ncu: issue-active 59 %, fma pipe
// 42 %, profiles/r01_k_step_summary.md), so the dynamics are instantiated a second time on P2 = (environment A,
// environment B): every FFMA/FADD/FMUL of the scalar code becomes one packed instruction that advances both
// environments.  Measured on B200 (tools/ffma2_probe.cu): FFMA2 sustains 57 TFLOP/s of the 67.7 TFLOP/s FP32 peak with
// HALF the instruction issues; FMNMX co-issues almost for free.  ptxas folds negation, broadcast immediates and a
// broadcast scalar register into the FFMA2 operands, so mixed scalar * packed costs nothing extra.
//
// The scalar code relies on the compiler contracting a * b + c; P2 does the same explicitly: a * b yields a deferred
// product (P2Prod) that fuses with the following + / - into one FFMA2.
#pragma once
#include <cuda_runtime.h>

namespace pg {

struct P2 {
  float2 v;
  P2() = default;
  __device__ __forceinline__ P2(float s) { v.x = s; v.y = s; }             // broadcast
  __device__ __forceinline__ P2(float a, float b) { v.x = a; v.y = b; }
  __device__ __forceinline__ explicit P2(float2 f) : v(f) {}
};
struct P2Prod {  // a * b, not yet evaluated
  P2 a, b;
  __device__ __forceinline__ operator P2() const { return P2(__fmul2_rn(a.v, b.v)); }
};

__device__ __forceinline__ P2 operator-(P2 a) { return P2(-a.v.x, -a.v.y); }
__device__ __forceinline__ P2Prod operator-(P2Prod p) { P2Prod r; r.a = -p.a; r.b = p.b; return r; }

__device__ __forceinline__ P2Prod operator*(P2 a, P2 b) { P2Prod r; r.a = a; r.b = b; return r; }
__device__ __forceinline__ P2Prod operator*(float a, P2 b) { P2Prod r; r.a = P2(a); r.b = b; return r; }
__device__ __forceinline__ P2Prod operator*(P2 a, float b) { P2Prod r; r.a = a; r.b = P2(b); return r; }
__device__ __forceinline__ P2Prod operator*(P2Prod p, P2 c) { P2Prod r; r.a = (P2)p; r.b = c; return r; }
__device__ __forceinline__ P2Prod operator*(P2 c, P2Prod p) { P2Prod r; r.a = c; r.b = (P2)p; return r; }
__device__ __forceinline__ P2Prod operator*(P2Prod p, float c) { P2Prod r; r.a = (P2)p; r.b = P2(c); return r; }
__device__ __forceinline__ P2Prod operator*(float c, P2Prod p) { P2Prod r; r.a = P2(c); r.b = (P2)p; return r; }
__device__ __forceinline__ P2Prod operator*(P2Prod p, P2Prod q) { P2Prod r; r.a = (P2)p; r.b = (P2)q; return r; }

__device__ __forceinline__ P2 fma2(P2 a, P2 b, P2 c) { return P2(__ffma2_rn(a.v, b.v, c.v)); }

__device__ __forceinline__ P2 operator+(P2 a, P2 b) { return P2(__fadd2_rn(a.v, b.v)); }
__device__ __forceinline__ P2 operator-(P2 a, P2 b) { return P2(__fadd2_rn(a.v, (-b).v)); }
__device__ __forceinline__ P2 operator+(P2 a, float b) { return a + P2(b); }
__device__ __forceinline__ P2 operator+(float a, P2 b) { return P2(a) + b; }
__device__ __forceinline__ P2 operator-(P2 a, float b) { return a + P2(-b); }
__device__ __forceinline__ P2 operator-(float a, P2 b) { return P2(a) - b; }

__device__ __forceinline__ P2 operator+(P2Prod p, P2 c) { return fma2(p.a, p.b, c); }
__device__ __forceinline__ P2 operator+(P2 c, P2Prod p) { return fma2(p.a, p.b, c); }
__device__ __forceinline__ P2 operator-(P2Prod p, P2 c) { return fma2(p.a, p.b, -c); }
__device__ __forceinline__ P2 operator-(P2 c, P2Prod p) { return fma2(-p.a, p.b, c); }
__device__ __forceinline__ P2 operator+(P2Prod p, float c) { return fma2(p.a, p.b, P2(c)); }
__device__ __forceinline__ P2 operator+(float c, P2Prod p) { return fma2(p.a, p.b, P2(c)); }
__device__ __forceinline__ P2 operator-(P2Prod p, float c) { return fma2(p.a, p.b, P2(-c)); }
__device__ __forceinline__ P2 operator-(float c, P2Prod p) { return fma2(-p.a, p.b, P2(c)); }
__device__ __forceinline__ P2 operator+(P2Prod p, P2Prod q) { return fma2(p.a, p.b, (P2)q); }
__device__ __forceinline__ P2 operator-(P2Prod p, P2Prod q) { return fma2(-q.a, q.b, (P2)p); }

__device__ __forceinline__ P2& operator+=(P2& a, P2 b) { a = a + b; return a; }
__device__ __forceinline__ P2& operator-=(P2& a, P2 b) { a = a - b; return a; }
__device__ __forceinline__ P2& operator+=(P2& a, P2Prod p) { a = a + p; return a; }
__device__ __forceinline__ P2& operator-=(P2& a, P2Prod p) { a = a - p; return a; }
__device__ __forceinline__ P2& operator+=(P2& a, float b) { a = a + b; return a; }

// scalar-generic helpers: the dynamics templates call these for both T = float and T = P2
__device__ __forceinline__ float vfma(float a, float b, float c) { return fmaf(a, b, c); }
__device__ __forceinline__ P2 vfma(P2 a, P2 b, P2 c) { return fma2(a, b, c); }
__device__ __forceinline__ float vmin(float a, float b) { return fminf(a, b); }
__device__ __forceinline__ P2 vmin(P2 a, P2 b) { return P2(fminf(a.v.x, b.v.x), fminf(a.v.y, b.v.y)); }
__device__ __forceinline__ float vmax(float a, float b) { return fmaxf(a, b); }
__device__ __forceinline__ P2 vmax(P2 a, P2 b) { return P2(fmaxf(a.v.x, b.v.x), fmaxf(a.v.y, b.v.y)); }
__device__ __forceinline__ float vabs(float a) { return fabsf(a); }
__device__ __forceinline__ P2 vabs(P2 a) { return P2(fabsf(a.v.x), fabsf(a.v.y)); }
__device__ __forceinline__ float vclamp(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
__device__ __forceinline__ P2 vclamp(P2 x, float lo, float hi) { return P2(fminf(fmaxf(x.v.x, lo), hi), fminf(fmaxf(x.v.y, lo), hi)); }
__device__ __forceinline__ float vsqrt(float a) { return sqrtf(a); }
__device__ __forceinline__ P2 vsqrt(P2 a) { return P2(sqrtf(a.v.x), sqrtf(a.v.y)); }
__device__ __forceinline__ float vrsqrt(float a) { return rsqrtf(a); }
__device__ __forceinline__ P2 vrsqrt(P2 a) { return P2(rsqrtf(a.v.x), rsqrtf(a.v.y)); }
// 1 / x for x > eps, else 0 (btMultiBodyConstraint's jacDiagABInv guard)
__device__ __forceinline__ float vinv_guard(float x) { return x > 1.1920929e-07f ? 1.0f / x : 0.f; }
__device__ __forceinline__ P2 vinv_guard(P2 x) { return P2(vinv_guard(x.v.x), vinv_guard(x.v.y)); }

}  // namespace pg
