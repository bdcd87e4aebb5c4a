// f32x2.cuh -- two Float32 cells per thread on Blackwell's packed FP32 instructions (FADD2 / FMUL2, PTX
// add/sub/mul.rn.f32x2 on a 64-bit register pair; sm_100+).
//
// Why (tools/fp32x2_probe.cu, measured on B200): a scalar FADD/FMUL issues at 0.97 warp-instructions per clock per
// scheduler, the packed form at 0.50 -- the FP32 lane throughput is the same, but a packed instruction takes ONE
// issue slot for two cells.  The stage and closure kernels are bound by issue slots (ncu: 81-84 % issue-slot
// utilisation at 38 % DRAM), with about a third of their instructions FADD/FMUL (-fmad=false keeps them unfused);
// with two cells per thread the FP instructions halve and the address arithmetic, loop control and barriers are
// shared by the two cells.
//
// Contraction: ptxas (12.9) fuses mul.rn.f32x2 + add/sub.rn.f32x2 into FFMA2 even with -fmad=false and explicit .rn
// (it does not for scalar mul.rn/add.rn).  f2t<true> ("exact") therefore multiplies with two scalar mul.rn.f32 on the
// unpacked lanes and only adds / subtracts packed: every lane performs the scalar kernels' IEEE operations, results
// are BIT-IDENTICAL to them.  f2t<false> ("fused") multiplies packed and lets ptxas contract: fewer instructions,
// products feeding a sum are not rounded (never less accurate), results agree with the scalar kernels to round-off.
#pragma once

namespace insb200 {

template <bool EXACT>
struct f2t {
  typedef f2t f2;
  unsigned long long v;
  __device__ __forceinline__ f2t() {}
  __device__ __forceinline__ explicit f2t(float s) { asm("mov.b64 %0, {%1, %1};" : "=l"(v) : "f"(s)); }
  __device__ __forceinline__ f2t(float lo, float hi) { asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(lo), "f"(hi)); }
  __device__ __forceinline__ float lo() const {
    float a, b;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
    return a;
  }
  __device__ __forceinline__ float hi() const {
    float a, b;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
    return b;
  }
};

template <bool E>
__device__ __forceinline__ f2t<E> operator+(f2t<E> a, f2t<E> b) {
  f2t<E> r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}
template <bool E>
__device__ __forceinline__ f2t<E> operator-(f2t<E> a, f2t<E> b) {
  f2t<E> r;
  asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}
template <bool E>
__device__ __forceinline__ f2t<E> operator*(f2t<E> a, f2t<E> b) {
  f2t<E> r;
  if (E)
    asm("{ .reg .f32 x0, x1, y0, y1;\n"
        "  mov.b64 {x0, x1}, %1; mov.b64 {y0, y1}, %2;\n"
        "  mul.rn.f32 x0, x0, y0; mul.rn.f32 x1, x1, y1;\n"
        "  mov.b64 %0, {x0, x1}; }"
        : "=l"(r.v)
        : "l"(a.v), "l"(b.v));
  else
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
  return r;
}
template <bool E>
__device__ __forceinline__ f2t<E> &operator+=(f2t<E> &a, f2t<E> b) { a = a + b; return a; }
template <bool E>
__device__ __forceinline__ f2t<E> &operator-=(f2t<E> &a, f2t<E> b) { a = a - b; return a; }
template <bool E>
__device__ __forceinline__ f2t<E> sqrt(f2t<E> a) { return f2t<E>(sqrtf(a.lo()), sqrtf(a.hi())); }

}  // namespace insb200
