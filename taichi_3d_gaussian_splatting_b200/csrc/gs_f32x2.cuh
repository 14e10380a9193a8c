// Packed binary32 pairs for sm_100a: add / sub / mul / fma on two floats with ONE instruction
// (SASS FADD2 / FMUL2 / FFMA2).  Each half is an ordinary IEEE round-to-nearest operation, so a
// packed kernel is bit-identical to its scalar form; the pair occupies the FMA pipe for two cycles
// but only one issue slot (tools/microbench/packed_fp32.cu: 1.94 packed against 3.75 scalar
// instructions per clock per SM), which is what an issue-bound kernel needs.  pack(x, x) of one
// register is folded by ptxas into a broadcast operand (`R4.F32`), pack / unpack of distinct
// registers are register-pair aliases.
#pragma once
#include <cuda_runtime.h>

struct f32x2 {
    unsigned long long v;
};

__device__ __forceinline__ f32x2 pack2(float lo, float hi)
{
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r.v) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ f32x2 splat2(float x) { return pack2(x, x); }
__device__ __forceinline__ void unpack2(f32x2 a, float &lo, float &hi)
{
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a.v));
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
    return r;
}
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
    return r;
}
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b)
{
    f32x2 r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r.v) : "l"(a.v), "l"(b.v));
    return r;
}
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r.v) : "l"(a.v), "l"(b.v), "l"(c.v));
    return r;
}
