// Residual programs: the per-cell algebra of each PDE residual, written once and evaluated by
// either kernel through a context type C that provides
//     c.template D<SLOT, FIELD, I>()   derivative slot SLOT applied to field FIELD at point I
//     c.template V<FIELD, I>()         the field value at the output cell
// (I is the compile-time index of the point inside the thread's register block; the generic
// kernel ignores it).  Field / slot / scalar orders are the ones documented in cppre.h.
//
// The formulas restate the reference's residual expressions; see the file:line on each.
#pragma once

#include "common.cuh"

namespace cppre {

template <int KIND>
struct Prog;

// r = K (x) u                                      [ref: Utils/ConvOps_2d.py:135-150]
template <>
struct Prog<CPPRE_PROG_LINEAR> {
    static constexpr int NF = 1, NSLOT = 1;
    template <int I, class C>
    __device__ __forceinline__ static float eval(const C& c, const float*) {
        return c.template D<0, 0, I>();
    }
};

// dx*D_t(u) + dt*u*D_x(u) - nu*D_xx(u)*(2*dt/dx)   [ref: Joint/Burgers_Residuals_CP.py:182-187]
// slots {D_t, D_x, D_xx}; scalars {dx, dt, nu*2*dt/dx}
template <>
struct Prog<CPPRE_PROG_BURGERS> {
    static constexpr int NF = 1, NSLOT = 3;
    template <int I, class C>
    __device__ __forceinline__ static float eval(const C& c, const float* s) {
        const float u = c.template V<0, I>();
        float r = s[0] * c.template D<0, 0, I>();
        r += (s[1] * u) * c.template D<1, 0, I>();
        r -= s[2] * c.template D<2, 0, I>();
        return r;
    }
};

// D_x(u) + (dx/dy)*D_y(v)                          [ref: Joint/NS_Residuals_CP.py:222-228]
// slots {D_x, D_y}; scalars {dx/dy}
template <>
struct Prog<CPPRE_PROG_NS_CONTINUITY> {
    static constexpr int NF = 2, NSLOT = 2;
    template <int I, class C>
    __device__ __forceinline__ static float eval(const C& c, const float* s) {
        return c.template D<0, 0, I>() + s[0] * c.template D<1, 1, I>();
    }
};

// res_x + res_y of the momentum equations           [ref: Joint/NS_Residuals_CP.py:231-240]
// fields {u,v,p}; slots {D_t, D_x, D_y, D_xx_yy}; scalars {dx*dy, dt*dy, dt*dx, nu*dt}
template <>
struct Prog<CPPRE_PROG_NS_MOMENTUM> {
    static constexpr int NF = 3, NSLOT = 4;
    template <int I, class C>
    __device__ __forceinline__ static float eval(const C& c, const float* s) {
        const float u = c.template V<0, I>(), v = c.template V<1, I>();
        const float a = s[0], b = s[1], cc = s[2], d = s[3];
        float rx = c.template D<0, 0, I>() * a;
        rx += (u * c.template D<1, 0, I>()) * b;
        rx += (v * c.template D<2, 0, I>()) * cc;
        rx -= c.template D<3, 0, I>() * d;
        rx += c.template D<1, 2, I>() * b;
        float ry = c.template D<0, 1, I>() * a;
        ry += (u * c.template D<1, 1, I>()) * cc;
        ry += (v * c.template D<2, 1, I>()) * b;
        ry -= c.template D<3, 1, I>() * d;
        ry += c.template D<2, 2, I>() * cc;
        return rx + ry;
    }
};

// MHD: slots {D_t, D_x, D_y}                        [ref: Joint/MHD_Residuals_CP.py:225-278]
// continuity: fields {rho,u,v}
template <>
struct Prog<CPPRE_PROG_MHD_CONTINUITY> {
    static constexpr int NF = 3, NSLOT = 3;
    template <int I, class C>
    __device__ __forceinline__ static float eval(const C& c, const float*) {
        const float rho = c.template V<0, I>(), u = c.template V<1, I>(), v = c.template V<2, I>();
        float r = c.template D<0, 0, I>();
        r += u * c.template D<1, 0, I>();
        r += rho * c.template D<1, 1, I>();
        r += v * c.template D<2, 0, I>();
        r += rho * c.template D<2, 2, I>();
        return r;
    }
};

// momentum: fields {rho,u,v,p,Bx,By}                [ref: Joint/MHD_Residuals_CP.py:234-243]
template <>
struct Prog<CPPRE_PROG_MHD_MOMENTUM> {
    static constexpr int NF = 6, NSLOT = 3;
    template <int I, class C>
    __device__ __forceinline__ static float eval(const C& c, const float*) {
        const float rho = c.template V<0, I>(), u = c.template V<1, I>(), v = c.template V<2, I>();
        const float Bx = c.template V<4, I>(), By = c.template V<5, I>();
        const float ir = 1.0f / rho;  // IEEE division (no fast-math): rho is a divisor in the reference
        const float bxr = Bx * ir, byr = By * ir;
        const float dxBx = c.template D<1, 4, I>(), dyBy = c.template D<2, 5, I>();
        const float dyBx = c.template D<2, 4, I>(), dxBy = c.template D<1, 5, I>();
        float rx = c.template D<0, 1, I>();
        rx += u * c.template D<1, 1, I>();
        rx += ir * c.template D<1, 3, I>();
        rx -= 2.0f * bxr * dxBx;
        rx += v * c.template D<2, 1, I>();
        rx -= byr * dyBx;
        rx -= bxr * dyBy;
        float ry = c.template D<0, 2, I>();
        ry += u * c.template D<1, 2, I>();
        ry += ir * c.template D<2, 3, I>();
        ry -= 2.0f * byr * dyBy;
        ry += v * c.template D<2, 2, I>();
        ry -= byr * dxBx;
        ry -= bxr * dxBy;
        return rx + ry;
    }
};

// energy: fields {rho,u,v,p,Bx,By}; scalars {gamma}  [ref: Joint/MHD_Residuals_CP.py:246-256]
template <>
struct Prog<CPPRE_PROG_MHD_ENERGY> {
    static constexpr int NF = 6, NSLOT = 3;
    template <int I, class C>
    __device__ __forceinline__ static float eval(const C& c, const float* s) {
        const float g = s[0];
        const float u = c.template V<1, I>(), v = c.template V<2, I>(), p = c.template V<3, I>();
        const float Bx = c.template V<4, I>(), By = c.template V<5, I>();
        const float bx2 = Bx * Bx, by2 = By * By;
        const float gp = g * (p - 0.5f * (bx2 + by2));
        float r = c.template D<0, 0, I>();
        r += u * c.template D<1, 3, I>();
        r += v * c.template D<2, 3, I>();
        r += ((g - 2.0f) * (u * Bx + v * By)) * (c.template D<1, 4, I>() + c.template D<2, 5, I>());
        r += (gp + by2) * c.template D<1, 1, I>();
        r += (gp + bx2) * c.template D<2, 2, I>();
        r -= (Bx * By) * (c.template D<2, 1, I>() + c.template D<1, 2, I>());
        return r;
    }
};

// induction: fields {u,v,Bx,By}                      [ref: Joint/MHD_Residuals_CP.py:259-268]
template <>
struct Prog<CPPRE_PROG_MHD_INDUCTION> {
    static constexpr int NF = 4, NSLOT = 3;
    template <int I, class C>
    __device__ __forceinline__ static float eval(const C& c, const float*) {
        const float u = c.template V<0, I>(), v = c.template V<1, I>();
        const float Bx = c.template V<2, I>(), By = c.template V<3, I>();
        float rx = c.template D<0, 2, I>();
        rx -= By * c.template D<2, 0, I>();
        rx += Bx * c.template D<2, 1, I>();
        rx -= v * c.template D<2, 2, I>();
        rx += u * c.template D<2, 3, I>();
        float ry = c.template D<0, 3, I>();
        ry += By * c.template D<1, 0, I>();
        ry -= Bx * c.template D<1, 1, I>();
        ry -= v * c.template D<1, 2, I>();
        ry += u * c.template D<1, 3, I>();
        return rx + ry;
    }
};

// Gauss law: fields {Bx,By}                          [ref: Joint/MHD_Residuals_CP.py:271-278]
template <>
struct Prog<CPPRE_PROG_MHD_GAUSS> {
    static constexpr int NF = 2, NSLOT = 3;
    template <int I, class C>
    __device__ __forceinline__ static float eval(const C& c, const float*) {
        return c.template D<1, 0, I>() + c.template D<2, 1, I>();
    }
};

}  // namespace cppre
