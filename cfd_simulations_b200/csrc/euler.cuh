// euler.cuh — per-cell / per-face device math of the compressible-Euler step (sm_100a).
// Follows python/shockwave_analysis/v1_shock.py: compute_fluxes :84-95, hllc_solver :147-209, roe_solver :97-145,
// update_state :211-223, compute_dt :263-275.  Scalar arithmetic is float64 with float32 rounding exactly where the
// reference materialises float32 arrays (physical fluxes, star states, U* - U, the returned flux, the stored state).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

namespace b200cfd { namespace eu {

struct Phys {
    double gamma, gm1, eps, mv;
    float  epsf, gm1f;          // compute_dt's NumPy f32 operands (v1_shock.py:265,269)
};

// Everything a face needs to know about one of its two cells; computed ONCE per cell per step (the reference
// recomputes it for each of the cell's four faces — same inputs, same results).
struct Cell {
    float4 U;                   // conserved state as stored
    double rho, u, v, E, p, a;  // clamped primitives (v1_shock.py:150-158,168-169)
    float4 Fp, Gp;              // compute_fluxes() results, float32 (v1_shock.py:92-93)
};

__device__ __forceinline__ double clipd(double x, double lo, double hi) { return fmax(lo, fmin(x, hi)); }

// Reciprocal and square root for the default (not exact_division) mode: the hardware seed (MUFU.RCP64H / RSQ64H, about 20
// bits) and two Newton steps, 5 / 9 instructions with no slow path, against ~30 each (plus a call for the special cases)
// for the IEEE-rounded operations.  Accurate to 1-2 ulp of a double; every consumer rounds to float32 afterwards.
// Denominators here are positive and far from the subnormal range (the reference adds eps to each of them).
__device__ __forceinline__ double rcp_fast(double b) {
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(b));
    double e = fma(-b, r, 1.0);
    r = fma(r, e, r);
    e = fma(-b, r, 1.0);
    return fma(r, e, r);
}
__device__ __forceinline__ double sqrt_fast(double x) {      // x > 0
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double hx = 0.5 * x;
    y = y * fma(-(hx * y), y, 1.5);
    y = y * fma(-(hx * y), y, 1.5);
    return x * y;
}
// x rounded to 24 significant bits (what a conversion to float32 and back gives for numbers in float32's normal range) with
// three fp64-pipe operations (Veltkamp splitting) instead of two conversions: the conversion pipe (XU, a quarter of the
// fp64 rate) is this kernel's busiest.
__device__ __forceinline__ double rnd24(double x) {
    const double t = __dmul_rn(x, 536870913.0);          // 2^29 + 1
    return __dsub_rn(t, __dsub_rn(t, x));
}
template <bool EXACT> __device__ __forceinline__ double dv(double a, double b) { return EXACT ? a / b : a * rcp_fast(b); }
// fmax / fmin / clip.  The library forms quiet and drop NaNs the IEEE way, which costs 7-8 instructions each on sm_100a
// (DSETP.MAX, two selects, a NaN fix-up and the moves around them) and there are about thirty per cell; the default mode
// uses a compare and a select (3 instructions).  Same result for ordinary numbers and when the FIRST argument is NaN (the
// second is returned); a NaN in the second argument is returned instead of dropped — a state with NaNs is reported
// through the nan flag either way.
template <bool EXACT> __device__ __forceinline__ double vmax(double a, double b) { return EXACT ? fmax(a, b) : (a > b ? a : b); }
template <bool EXACT> __device__ __forceinline__ double vmin(double a, double b) { return EXACT ? fmin(a, b) : (a < b ? a : b); }
template <bool EXACT> __device__ __forceinline__ double vclip(double x, double lo, double hi) { return vmax<EXACT>(vmin<EXACT>(x, hi), lo); }

// EXACT: IEEE divisions where the reference divides (bit-faithful); otherwise one reciprocal per cell.
template <bool EXACT>
__device__ __forceinline__ Cell make_cell(const float4 q, const Phys& P) {
    Cell c;
    c.U = q;
    c.rho = vmax<EXACT>((double)q.x, P.eps);
    double pr;
    if (EXACT) {
        c.u = vclip<EXACT>((double)q.y / c.rho, -P.mv, P.mv);
        c.v = vclip<EXACT>((double)q.z / c.rho, -P.mv, P.mv);
        c.E = vclip<EXACT>((double)q.w / c.rho, P.eps, P.mv);
        c.p = vmax<EXACT>(P.gm1 * c.rho * (c.E - 0.5 * (c.u * c.u + c.v * c.v)), P.eps);
        pr = c.p / c.rho;
        c.a = sqrt(vmax<EXACT>(P.gamma * c.p / c.rho, P.eps));
    } else {
        const double inv = rcp_fast(c.rho);
        c.u = vclip<EXACT>((double)q.y * inv, -P.mv, P.mv);
        c.v = vclip<EXACT>((double)q.z * inv, -P.mv, P.mv);
        c.E = vclip<EXACT>((double)q.w * inv, P.eps, P.mv);
        c.p = vmax<EXACT>(P.gm1 * c.rho * (c.E - 0.5 * (c.u * c.u + c.v * c.v)), P.eps);
        pr = c.p * inv;
        c.a = sqrt_fast(vmax<EXACT>(P.gamma * pr, P.eps));
    }
    const double H = c.E + pr, ru = c.rho * c.u, rv = c.rho * c.v;
    c.Fp = make_float4((float)ru, (float)(ru * c.u + c.p), (float)(ru * c.v), (float)(ru * H));
    c.Gp = make_float4((float)rv, (float)(rv * c.u), (float)(rv * c.v + c.p), (float)(rv * H));
    return c;
}

__device__ __forceinline__ float4 f4zero() { return make_float4(0.f, 0.f, 0.f, 0.f); }

// hllc_solver (v1_shock.py:147-209). XN: normal (1,0), else (0,1).
template <bool XN, bool EXACT>
__device__ __forceinline__ float4 hllc_face(const Cell& L, const Cell& R, const Phys& P) {
    if (L.rho <= P.eps || R.rho <= P.eps) return f4zero();                         // :152-153,160-161
    const double qL = XN ? L.u : L.v, qR = XN ? R.u : R.v;
    const double SL = vmin<EXACT>(qL - L.a, qR - R.a), SR = vmax<EXACT>(qL + L.a, qR + R.a);      // :172-173
    const float4 fL = XN ? L.Fp : L.Gp, fR = XN ? R.Fp : R.Gp;
    if (SL >= 0) return fL;                                                         // :186-187
    if (SR <= 0) return fR;                                                         // :188-189
    const double mL = L.rho * (SL - qL), mR = R.rho * (SR - qR);
    const double SM = dv<EXACT>(R.rho * qR * (SR - qR) - L.rho * qL * (SL - qL) + L.p - R.p, mR - mL + P.eps);   // :174
    const double pstar = L.rho * (qL - SL) * (qL - SM) + L.p;                       // :191
    // one star state, chosen by the sign of S_M (:199-209); selecting operands keeps the warp converged
    const bool left = SM >= 0;
    const double S = left ? SL : SR, m = left ? mL : mR, q = left ? qL : qR, p = left ? L.p : R.p;
    const double E = left ? L.E : R.E, vt = left ? L.v : R.v;
    const float4 f = left ? fL : fR, U = left ? L.U : R.U;
    const double c = dv<EXACT>(m, S - SM + P.eps);
    const float s0 = (float)c;
    const float s1 = (float)(c * (XN ? SM : 0.0 * SM));
    const float s2 = (float)(c * (XN ? vt : SM));
    const float s3 = (float)(c * (E + dv<EXACT>(pstar * SM - p * q, m + P.eps)));
    return make_float4((float)((double)f.x + S * (double)(s0 - U.x)), (float)((double)f.y + S * (double)(s1 - U.y)),
                       (float)((double)f.z + S * (double)(s2 - U.z)), (float)((double)f.w + S * (double)(s3 - U.w)));
}

// roe_solver (v1_shock.py:97-145): Rusanov-like flux with the sum of three entropy-fixed Roe wave speeds; always
// built from F = compute_fluxes()[0], also for the y-normal (:139-140).
template <bool XN>
__device__ __forceinline__ float4 roe_face(const Cell& L, const Cell& R, const Phys& P) {
    if (L.rho <= P.eps || R.rho <= P.eps) return f4zero();
    const double hL = ((double)L.U.w + L.p) / L.rho, hR = ((double)R.U.w + R.p) / R.rho;
    const double sL = sqrt(L.rho), sR = sqrt(R.rho), den = sL + sR + P.eps;
    const double ur = (L.u * sL + R.u * sR) / den, vr = (L.v * sL + R.v * sR) / den, hr = (hL * sL + hR * sR) / den;
    const double ar = sqrt(fmax(P.gm1 * (hr - 0.5 * (ur * ur + vr * vr)), P.eps));
    const double vn = XN ? ur : vr;
    const double er = 0.05 * ar;
    const double lam = 0.5 * (fmax(er, fabs(vn)) + fmax(er, fabs(vn + ar)) + fmax(er, fabs(vn - ar)));
    float4 o;
    o.x = (float)((double)(0.5f * (L.Fp.x + R.Fp.x)) - lam * (double)(R.U.x - L.U.x));
    o.y = (float)((double)(0.5f * (L.Fp.y + R.Fp.y)) - lam * (double)(R.U.y - L.U.y));
    o.z = (float)((double)(0.5f * (L.Fp.z + R.Fp.z)) - lam * (double)(R.U.z - L.U.z));
    o.w = (float)((double)(0.5f * (L.Fp.w + R.Fp.w)) - lam * (double)(R.U.w - L.U.w));
    return o;
}

template <int SOLVER, bool XN, bool EXACT = true>
__device__ __forceinline__ float4 face_flux(const Cell& L, const Cell& R, const Phys& P) {
    return SOLVER == 0 ? hllc_face<XN, EXACT>(L, R, P) : roe_face<XN>(L, R, P);
}

// update_state (v1_shock.py:215-221) for one interior cell: west/east x-faces, south/north y-faces.
template <bool EXACT = true>
__device__ __forceinline__ float4 update_cell(const float4 U, const float4 Fw, const float4 Fe, const float4 Gs, const float4 Gn,
                                              double ax, double ay, const Phys& P) {
    double r = (double)U.x - ax * (double)(Fe.x - Fw.x) - ay * (double)(Gn.x - Gs.x);
    double mx = (double)U.y - ax * (double)(Fe.y - Fw.y) - ay * (double)(Gn.y - Gs.y);
    double my = (double)U.z - ax * (double)(Fe.z - Fw.z) - ay * (double)(Gn.z - Gs.z);
    double e = (double)U.w - ax * (double)(Fe.w - Fw.w) - ay * (double)(Gn.w - Gs.w);
    r = vmax<EXACT>(r, P.eps);
    mx = vclip<EXACT>(mx, -P.mv * r, P.mv * r);
    my = vclip<EXACT>(my, -P.mv * r, P.mv * r);
    e = vmax<EXACT>(e, P.eps * r);
    return make_float4((float)r, (float)mx, (float)my, (float)e);
}

// compute_dt's per-cell wave speeds (v1_shock.py:265-272): f32 NumPy divisions, then f64.
// EXACT = false: one fast float64 reciprocal serves all four divisions.  The three float32 quotients keep their float32
// rounding (they are the reference's np.float32 divisions; dropping it moved dt by up to 2e-6 in the parity runs): a
// float64 quotient rounded to float32 is the float32 quotient, except where the 1-2 ulp of rcp_fast straddle a float32
// rounding boundary (about one quotient in 2^28).
template <bool EXACT = true>
__device__ __forceinline__ void cell_speeds(const float4 q, const Phys& P, double& sx, double& sy) {
    const float rho = fmaxf(q.x, P.epsf);
    double u, v, E, a;
    if (EXACT) {
        u = clipd((double)__fdiv_rn(q.y, rho), -P.mv, P.mv);
        v = clipd((double)__fdiv_rn(q.z, rho), -P.mv, P.mv);
        E = clipd((double)__fdiv_rn(q.w, rho), P.eps, P.mv);
        const double p = fmax((double)__fmul_rn(P.gm1f, rho) * (E - 0.5 * (u * u + v * v)), P.eps);
        a = sqrt(clipd(P.gamma * p / (double)rho, P.eps, P.mv));
    } else {
        const double inv = rcp_fast((double)rho);
        u = vclip<false>(rnd24((double)q.y * inv), -P.mv, P.mv);
        v = vclip<false>(rnd24((double)q.z * inv), -P.mv, P.mv);
        E = vclip<false>(rnd24((double)q.w * inv), P.eps, P.mv);
        const double p = vmax<false>((double)__fmul_rn(P.gm1f, rho) * (E - 0.5 * (u * u + v * v)), P.eps);
        a = sqrt_fast(vclip<false>(P.gamma * p * inv, P.eps, P.mv));
    }
    sx = fabs(u) + a;
    sy = fabs(v) + a;
}

// ---- warp exchange of a Cell with the lane holding the next column ----
__device__ __forceinline__ float4 shfl_down4(const float4 a) {
    return make_float4(__shfl_down_sync(0xffffffffu, a.x, 1), __shfl_down_sync(0xffffffffu, a.y, 1),
                       __shfl_down_sync(0xffffffffu, a.z, 1), __shfl_down_sync(0xffffffffu, a.w, 1));
}
__device__ __forceinline__ float4 shfl_up4(const float4 a) {
    return make_float4(__shfl_up_sync(0xffffffffu, a.x, 1), __shfl_up_sync(0xffffffffu, a.y, 1),
                       __shfl_up_sync(0xffffffffu, a.z, 1), __shfl_up_sync(0xffffffffu, a.w, 1));
}
template <int SOLVER>
__device__ __forceinline__ Cell shfl_down_cell(const Cell& c) {
    Cell n;
    n.U = shfl_down4(c.U);
    n.rho = __shfl_down_sync(0xffffffffu, c.rho, 1);
    n.u = __shfl_down_sync(0xffffffffu, c.u, 1);
    n.v = __shfl_down_sync(0xffffffffu, c.v, 1);
    n.E = __shfl_down_sync(0xffffffffu, c.E, 1);
    n.p = __shfl_down_sync(0xffffffffu, c.p, 1);
    n.a = __shfl_down_sync(0xffffffffu, c.a, 1);
    if (SOLVER == 0) { n.Gp = shfl_down4(c.Gp); n.Fp = n.Gp; }   // HLLC y-faces read only G
    else             { n.Fp = shfl_down4(c.Fp); n.Gp = n.Fp; }   // the reference's Roe flux reads only F
    return n;
}

}}  // namespace b200cfd::eu
