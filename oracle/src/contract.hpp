// oracle/src/contract.hpp -- TEST INFRASTRUCTURE ONLY (see oracle.hpp header).
// The fp32 arithmetic contract shared by the equaliser and the header path: every operation is an explicit IEEE-754
// single operation, so the CUDA kernels (__fmaf_rn/__fmul_rn/__frcp_rn) can match it bit for bit.
#pragma once
#include "oracle.hpp"
#include <cmath>

namespace orc {

// ---- arithmetic contract (bit-exact target for the CUDA kernel) -------------------------
static inline cf32 c_div(cf32 y, cf32 h)
{
    const float yr = y.real(), yi = y.imag(), hr = h.real(), hi = h.imag();
    const float den = fmaf(hr, hr, hi * hi);
    const float r = 1.0f / den;
    const float nr = fmaf(yr, hr, yi * hi);
    const float ni = fmaf(yi, hr, -(yr * hi));
    return cf32(nr * r, ni * r);
}
static inline cf32 c_mul(cf32 a, cf32 b)
{
    const float ar = a.real(), ai = a.imag(), br = b.real(), bi = b.imag();
    return cf32(fmaf(ar, br, -(ai * bi)), fmaf(ar, bi, ai * br));
}
static inline cf32 c_axpy(float alpha, cf32 h, cf32 u)   // alpha*h + u
{
    return cf32(fmaf(alpha, h.real(), u.real()), fmaf(alpha, h.imag(), u.imag()));
}
static inline cf32 expj_contract(double phase)
{
    const float p32 = (float)phase;                 // gr_expj takes a float
    return cf32((float)std::cos((double)p32), (float)std::sin((double)p32));
}

} // namespace orc
