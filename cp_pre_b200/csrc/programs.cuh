// Residual programs: the per-cell algebra of each PDE residual, written once and evaluated by
// either kernel through a context type C that provides
//     c.template D<SLOT, FIELD, I>()   derivative slot SLOT applied to field FIELD at point I
//     c.template V<FIELD, I>()         the field value at the output cell
// (I is the compile-time index of the point inside the thread's register block; the generic
// kernel ignores it).  Field / slot / scalar orders are the ones documented in cppre.h.
//
// The algebra is written once over a value type T = decltype(c.D<..>()): `float` (one grid point,
// generic kernel) or `f2` (two neighbouring grid points in one 64-bit register pair, TMA star
// kernel).  On sm_100a the f2 operations are single FFMA2 / FMUL2 / FADD2 instructions
// (__ffma2_rn & co., crt/sm_100_rt.h) whose scalar operands (coefficients, program scalars) are
// broadcast from one uniform register and whose negations are operand modifiers -- half the
// issue slots of the scalar form, same IEEE results per component.
//
// The formulas restate the reference's residual expressions; see the file:line on each.
#pragma once

#include "common.cuh"

namespace cppre {


// ---- value-type algebra -------------------------------------------------------------------
struct f2 {
    float2 v;
};
__device__ __forceinline__ f2 mk2(float x, float y) { return f2{make_float2(x, y)}; }
__device__ __forceinline__ float2 bc2(float s) { return make_float2(s, s); }

// the scalar forms use the _rn intrinsics so that the compiler cannot contract a product into a following sum:
// the packed forms below never are, and the generic kernel must return what the star kernels return, bit for bit
__device__ __forceinline__ float pmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float padd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float psub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float pfma(float a, float b, float c) { return fmaf(a, b, c); }    // a*b + c
__device__ __forceinline__ float pfnma(float a, float b, float c) { return fmaf(-a, b, c); }  // c - a*b
__device__ __forceinline__ float pinv(float a) { return 1.0f / a; }   // IEEE division (no fast-math)

__device__ __forceinline__ f2 pmul(f2 a, f2 b) { return f2{__fmul2_rn(a.v, b.v)}; }
__device__ __forceinline__ f2 pmul(f2 a, float b) { return f2{__fmul2_rn(a.v, bc2(b))}; }
__device__ __forceinline__ f2 pmul(float a, f2 b) { return f2{__fmul2_rn(b.v, bc2(a))}; }
__device__ __forceinline__ f2 padd(f2 a, f2 b) { return f2{__fadd2_rn(a.v, b.v)}; }
__device__ __forceinline__ f2 psub(f2 a, f2 b) { return f2{__fadd2_rn(a.v, make_float2(-b.v.x, -b.v.y))}; }
__device__ __forceinline__ f2 psub(f2 a, float b) { return f2{__fadd2_rn(a.v, bc2(-b))}; }
__device__ __forceinline__ f2 pfma(f2 a, f2 b, f2 c) { return f2{__ffma2_rn(a.v, b.v, c.v)}; }
__device__ __forceinline__ f2 pfma(f2 a, float b, f2 c) { return f2{__ffma2_rn(a.v, bc2(b), c.v)}; }
__device__ __forceinline__ f2 pfma(float a, f2 b, f2 c) { return f2{__ffma2_rn(b.v, bc2(a), c.v)}; }
__device__ __forceinline__ f2 pfnma(f2 a, f2 b, f2 c) { return f2{__ffma2_rn(make_float2(-a.v.x, -a.v.y), b.v, c.v)}; }
__device__ __forceinline__ f2 pfnma(f2 a, float b, f2 c) { return f2{__ffma2_rn(a.v, bc2(-b), c.v)}; }
__device__ __forceinline__ f2 pinv(f2 a) { return mk2(1.0f / a.v.x, 1.0f / a.v.y); }

// Normaliser-decode fold (CPPRE_FLAG_DECODE): wraps a kernel's context so the program sees the field
// a_f * x + b_f.  A derivative slot is linear, so on cells whose whole stencil lies inside the field
// D(a x + b) = a D(x) + b * sum(taps): one more FFMA(2) per slot evaluation with broadcast operands, and the
// raw field factors are decoded at the output cell.  No pass over the tensor, no decoded copy.
__device__ __forceinline__ float paffine(float x, float a, float b) { return fmaf(x, a, b); }
__device__ __forceinline__ f2 paffine(f2 x, float a, float b) { return f2{__ffma2_rn(x.v, bc2(a), bc2(b))}; }

template <class C>
struct DecodeCtx {
    const C& c;
    const DecodeDev& d;
    template <int SLOT, int FIELD, int I>
    __device__ __forceinline__ auto D() const {
        return paffine(c.template D<SLOT, FIELD, I>(), d.a[FIELD], d.bs[SLOT][FIELD]);
    }
    template <int FIELD, int I>
    __device__ __forceinline__ auto V() const {
        return paffine(c.template V<FIELD, I>(), d.a[FIELD], d.b[FIELD]);
    }
};

template <int KIND>
struct Prog;

// r = K (x) u                                      [ref: Utils/ConvOps_2d.py:135-150]
template <>
struct Prog<CPPRE_PROG_LINEAR> {
    static constexpr int NF = 1, NSLOT = 1;
    template <int I, class C>
    __device__ __forceinline__ static auto eval(const C& c, const float*) {
        return c.template D<0, 0, I>();
    }
};

// dx*D_t(u) + dt*u*D_x(u) - nu*D_xx(u)*(2*dt/dx)   [ref: Joint/Burgers_Residuals_CP.py:182-187]
// slots {D_t, D_x, D_xx}; scalars {dx, dt, nu*2*dt/dx}
template <>
struct Prog<CPPRE_PROG_BURGERS> {
    static constexpr int NF = 1, NSLOT = 3;
    template <int I, class C>
    __device__ __forceinline__ static auto eval(const C& c, const float* s) {
        const auto u = c.template V<0, I>();
        auto r = pmul(c.template D<0, 0, I>(), s[0]);
        r = pfma(pmul(u, s[1]), c.template D<1, 0, I>(), r);
        r = pfnma(c.template D<2, 0, I>(), s[2], r);
        return r;
    }
};

// D_x(u) + (dx/dy)*D_y(v)                          [ref: Joint/NS_Residuals_CP.py:222-228]
// slots {D_x, D_y}; scalars {dx/dy}
template <>
struct Prog<CPPRE_PROG_NS_CONTINUITY> {
    static constexpr int NF = 2, NSLOT = 2;
    template <int I, class C>
    __device__ __forceinline__ static auto eval(const C& c, const float* s) {
        return pfma(c.template D<1, 1, I>(), s[0], c.template D<0, 0, I>());
    }
};

// res_x + res_y of the momentum equations           [ref: Joint/NS_Residuals_CP.py:231-240]
// fields {u,v,p}; slots {D_t, D_x, D_y, D_xx_yy}; scalars {dx*dy, dt*dy, dt*dx, nu*dt}
template <>
struct Prog<CPPRE_PROG_NS_MOMENTUM> {
    static constexpr int NF = 3, NSLOT = 4;
    template <int I, class C>
    __device__ __forceinline__ static auto eval(const C& c, const float* s) {
        const auto u = c.template V<0, I>(), v = c.template V<1, I>();
        const float a = s[0], b = s[1], cc = s[2], d = s[3];
        auto rx = pmul(c.template D<0, 0, I>(), a);
        rx = pfma(pmul(u, c.template D<1, 0, I>()), b, rx);
        rx = pfma(pmul(v, c.template D<2, 0, I>()), cc, rx);
        rx = pfnma(c.template D<3, 0, I>(), d, rx);
        rx = pfma(c.template D<1, 2, I>(), b, rx);
        auto ry = pmul(c.template D<0, 1, I>(), a);
        ry = pfma(pmul(u, c.template D<1, 1, I>()), cc, ry);
        ry = pfma(pmul(v, c.template D<2, 1, I>()), b, ry);
        ry = pfnma(c.template D<3, 1, I>(), d, ry);
        ry = pfma(c.template D<2, 2, I>(), cc, ry);
        return padd(rx, ry);
    }
};

// MHD: slots {D_t, D_x, D_y}                        [ref: Joint/MHD_Residuals_CP.py:225-278]
// continuity: fields {rho,u,v}
template <>
struct Prog<CPPRE_PROG_MHD_CONTINUITY> {
    static constexpr int NF = 3, NSLOT = 3;
    template <int I, class C>
    __device__ __forceinline__ static auto eval(const C& c, const float*) {
        const auto rho = c.template V<0, I>(), u = c.template V<1, I>(), v = c.template V<2, I>();
        auto r = c.template D<0, 0, I>();
        r = pfma(u, c.template D<1, 0, I>(), r);
        r = pfma(rho, c.template D<1, 1, I>(), r);
        r = pfma(v, c.template D<2, 0, I>(), r);
        r = pfma(rho, c.template D<2, 2, I>(), r);
        return r;
    }
};

// momentum: fields {rho,u,v,p,Bx,By}                [ref: Joint/MHD_Residuals_CP.py:234-243]
template <>
struct Prog<CPPRE_PROG_MHD_MOMENTUM> {
    static constexpr int NF = 6, NSLOT = 3;
    template <int I, class C>
    __device__ __forceinline__ static auto eval(const C& c, const float*) {
        const auto rho = c.template V<0, I>(), u = c.template V<1, I>(), v = c.template V<2, I>();
        const auto Bx = c.template V<4, I>(), By = c.template V<5, I>();
        const auto ir = pinv(rho);    // rho is a divisor in the reference
        const auto bxr = pmul(Bx, ir), byr = pmul(By, ir);
        const auto dxBx = c.template D<1, 4, I>(), dyBy = c.template D<2, 5, I>();
        const auto dyBx = c.template D<2, 4, I>(), dxBy = c.template D<1, 5, I>();
        auto rx = c.template D<0, 1, I>();
        rx = pfma(u, c.template D<1, 1, I>(), rx);
        rx = pfma(ir, c.template D<1, 3, I>(), rx);
        rx = pfnma(pmul(bxr, 2.0f), dxBx, rx);
        rx = pfma(v, c.template D<2, 1, I>(), rx);
        rx = pfnma(byr, dyBx, rx);
        rx = pfnma(bxr, dyBy, rx);
        auto ry = c.template D<0, 2, I>();
        ry = pfma(u, c.template D<1, 2, I>(), ry);
        ry = pfma(ir, c.template D<2, 3, I>(), ry);
        ry = pfnma(pmul(byr, 2.0f), dyBy, ry);
        ry = pfma(v, c.template D<2, 2, I>(), ry);
        ry = pfnma(byr, dxBx, ry);
        ry = pfnma(bxr, dxBy, ry);
        return padd(rx, ry);
    }
};

// energy: fields {rho,u,v,p,Bx,By}; scalars {gamma}  [ref: Joint/MHD_Residuals_CP.py:246-256]
template <>
struct Prog<CPPRE_PROG_MHD_ENERGY> {
    static constexpr int NF = 6, NSLOT = 3;
    template <int I, class C>
    __device__ __forceinline__ static auto eval(const C& c, const float* s) {
        const float g = s[0];
        const auto u = c.template V<1, I>(), v = c.template V<2, I>(), p = c.template V<3, I>();
        const auto Bx = c.template V<4, I>(), By = c.template V<5, I>();
        const auto bx2 = pmul(Bx, Bx), by2 = pmul(By, By);
        const auto gp = pmul(pfnma(padd(bx2, by2), 0.5f, p), g);       // g * (p - 0.5*(bx2+by2))
        auto r = c.template D<0, 0, I>();
        r = pfma(u, c.template D<1, 3, I>(), r);
        r = pfma(v, c.template D<2, 3, I>(), r);
        const auto ub = pmul(pfma(v, By, pmul(u, Bx)), g - 2.0f);      // (g-2) * (u*Bx + v*By)
        r = pfma(ub, padd(c.template D<1, 4, I>(), c.template D<2, 5, I>()), r);
        r = pfma(padd(gp, by2), c.template D<1, 1, I>(), r);
        r = pfma(padd(gp, bx2), c.template D<2, 2, I>(), r);
        r = pfnma(pmul(Bx, By), padd(c.template D<2, 1, I>(), c.template D<1, 2, I>()), r);
        return r;
    }
};

// induction: fields {u,v,Bx,By}                      [ref: Joint/MHD_Residuals_CP.py:259-268]
template <>
struct Prog<CPPRE_PROG_MHD_INDUCTION> {
    static constexpr int NF = 4, NSLOT = 3;
    template <int I, class C>
    __device__ __forceinline__ static auto eval(const C& c, const float*) {
        const auto u = c.template V<0, I>(), v = c.template V<1, I>();
        const auto Bx = c.template V<2, I>(), By = c.template V<3, I>();
        auto rx = c.template D<0, 2, I>();
        rx = pfnma(By, c.template D<2, 0, I>(), rx);
        rx = pfma(Bx, c.template D<2, 1, I>(), rx);
        rx = pfnma(v, c.template D<2, 2, I>(), rx);
        rx = pfma(u, c.template D<2, 3, I>(), rx);
        auto ry = c.template D<0, 3, I>();
        ry = pfma(By, c.template D<1, 0, I>(), ry);
        ry = pfnma(Bx, c.template D<1, 1, I>(), ry);
        ry = pfnma(v, c.template D<1, 2, I>(), ry);
        ry = pfma(u, c.template D<1, 3, I>(), ry);
        return padd(rx, ry);
    }
};

// Gauss law: fields {Bx,By}                          [ref: Joint/MHD_Residuals_CP.py:271-278]
template <>
struct Prog<CPPRE_PROG_MHD_GAUSS> {
    static constexpr int NF = 2, NSLOT = 3;
    template <int I, class C>
    __device__ __forceinline__ static auto eval(const C& c, const float*) {
        return padd(c.template D<1, 0, I>(), c.template D<2, 1, I>());
    }
};

// JOREK reduced MHD: fields {rho, phi, T, R}; slots {D_t, D_R, D_Z, D_RR, D_ZZ}
//                                                    [ref: Marginal/JOREK_residuals_CP.py:211-242]
// continuity; scalars {D}:
//   D_t(rho) - R*(D_R(rho)*D_Z(phi) - D_R(phi)*D_Z(rho)) - 2*rho*D_Z(phi) - D*(D_RR(rho) + (1/R)*D_R(rho) + D_ZZ(rho))
template <>
struct Prog<CPPRE_PROG_JOREK_CONTINUITY> {
    static constexpr int NF = 4, NSLOT = 5;
    template <int I, class C>
    __device__ __forceinline__ static auto eval(const C& c, const float* s) {
        const auto rho = c.template V<0, I>(), R = c.template V<3, I>();
        const auto dRrho = c.template D<1, 0, I>(), dZrho = c.template D<2, 0, I>();
        const auto dRphi = c.template D<1, 1, I>(), dZphi = c.template D<2, 1, I>();
        auto r = c.template D<0, 0, I>();
        r = pfnma(R, pfnma(dRphi, dZrho, pmul(dRrho, dZphi)), r);
        r = pfnma(pmul(rho, 2.0f), dZphi, r);
        const auto lap = padd(padd(c.template D<3, 0, I>(), pmul(pinv(R), dRrho)), c.template D<4, 0, I>());
        return pfnma(lap, s[0], r);
    }
};

// temperature; scalars {K, gamma}:
//   T*D_t(rho) + rho*D_t(T) - rho*R*(D_R(T)*D_Z(phi) - D_R(phi)*D_Z(T)) + T*R*(D_R(rho)*D_Z(phi) - D_R(phi)*D_Z(rho))
//   + 2*gamma*rho*T*D_Z(phi) + K*(D_RR(T) + (1/R)*D_R(T) + D_ZZ(T))
template <>
struct Prog<CPPRE_PROG_JOREK_TEMPERATURE> {
    static constexpr int NF = 4, NSLOT = 5;
    template <int I, class C>
    __device__ __forceinline__ static auto eval(const C& c, const float* s) {
        const auto rho = c.template V<0, I>(), T = c.template V<2, I>(), R = c.template V<3, I>();
        const auto dRrho = c.template D<1, 0, I>(), dZrho = c.template D<2, 0, I>();
        const auto dRphi = c.template D<1, 1, I>(), dZphi = c.template D<2, 1, I>();
        const auto dRT = c.template D<1, 2, I>(), dZT = c.template D<2, 2, I>();
        auto r = pmul(T, c.template D<0, 0, I>());
        r = pfma(rho, c.template D<0, 2, I>(), r);
        r = pfnma(pmul(rho, R), pfnma(dRphi, dZT, pmul(dRT, dZphi)), r);
        r = pfma(pmul(T, R), pfnma(dRphi, dZrho, pmul(dRrho, dZphi)), r);
        r = pfma(pmul(pmul(rho, 2.0f * s[1]), T), dZphi, r);
        const auto lap = padd(padd(c.template D<3, 2, I>(), pmul(pinv(R), dRT)), c.template D<4, 2, I>());
        return pfma(lap, s[0], r);
    }
};

}  // namespace cppre
