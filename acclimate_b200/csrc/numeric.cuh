// numeric.cuh -- the reference's fixed-point value algebra, restated for device code.
//
// Follows reference include/types.h:270-665 operation by operation.  With the reference's default
// options (BASED_ON_INT=ON, BANKERS_ROUNDING=ON) FlowQuantity/Quantity (3 digits), Price (6 digits)
// and Time (0 digits) are int64 counts of a quantum, FlowValue/Value are plain f64, and every typed
// operator re-quantises its result with  lrint(f / precision)  -- a DIVISION by a constant that is not
// an exact power of ten (types.h:313-319,328).  Bit-exact parity therefore needs: these exact
// constants, IEEE division, round-half-even, and no FMA contraction (the library is compiled with
// --fmad=false; g++ -O2 on x86-64 emits no FMA either).
//
// Conventions used throughout the engine:
//   q  : int64 count of 1e-3 quanta (FlowQuantity / Quantity)
//   p  : int64 count of 1e-6 quanta (Price); the reference's Price::quiet_NaN() is 0 (types.h:370-374)
//   v  : f64 value (FlowValue / Value)
#pragma once

#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define ACC_HD __host__ __device__ __forceinline__
#else
#define ACC_HD inline
#endif

namespace acc {

typedef long long i64;

// 0.1*(0.1*(0.1*1)) and 0.1*(0.1*(0.1*(0.1*(0.1*(0.1*1))))) evaluated in f64 (types.h:313-315)
#define ACC_PREC3 0x1.0624dd2f1a9fdp-10
#define ACC_PREC6 0x1.0c6f7a0b5ed8fp-20

ACC_HD i64 iround(double x) {  // types.h:288-294, std::lrint under FE_TONEAREST
#if defined(__CUDA_ARCH__)
    return __double2ll_rn(x);
#else
    return llrint(x);
#endif
}

ACC_HD double qf(i64 q) { return (double)q * ACC_PREC3; }   // to_float(FlowQuantity), types.h:336-342
ACC_HD double pf(i64 p) { return (double)p * ACC_PREC6; }   // to_float(Price)
// lrint(f / precision) without the division on the common path.  f * (1/precision rounded) differs from the
// correctly rounded quotient by less than 2^-50 relative (precision is within 2 ulp of 10^-digits, plus the two
// roundings), so both round to the same integer unless a half-integer lies within that distance; only then (and for
// huge or non-finite arguments, where the test fails as well) the division is carried out.  tests/test_numeric.py
// checks the equivalence exhaustively around the rounding boundaries.
ACC_HD i64 quantise(double f, double prec, double inv) {
    const double q1 = f * inv;
    const double r = rint(q1);
    const double margin = 0.5 - fabs(q1 - r);
    if (margin > fabs(q1) * 0x1p-48) return (i64)r;
    return iround(f / prec);
}
ACC_HD i64 mkq(double f) { return quantise(f, ACC_PREC3, 1e3); }  // FlowQuantity(f), types.h:316-319,345
ACC_HD i64 mkp(double f) { return quantise(f, ACC_PREC6, 1e6); }  // Price(f)
ACC_HD i64 mkt(double f) { return iround(f / 1.0); }        // Time(f): 0 digits, precision 1

// std::max / std::min with their exact tie and signed-zero behaviour
ACC_HD double stdmax(double a, double b) { return (a < b) ? b : a; }
ACC_HD double stdmin(double a, double b) { return (b < a) ? b : a; }

struct Flow {  // PricedQuantity<FlowQuantity, FlowValue> and <Quantity, Value> (same arithmetic)
    i64 q;
    double v;
};

ACC_HD Flow mkflow(i64 q, double v) { Flow f; f.q = q; f.v = v; return f; }
// Flow(quantity, price): value = quantity * price, types.h:452,504
ACC_HD double q_times_p(i64 q, i64 p) { return qf(q) * pf(p); }
ACC_HD Flow flow_qp(i64 q, i64 p) { return mkflow(q, q_times_p(q, p)); }
// Flow(quantity): priced at 1.0, types.h:501-502
ACC_HD Flow flow_q(i64 q) { return mkflow(q, qf(q) * pf(1000000)); }
// copy construction / assignment zeroes the value of an empty flow, types.h:509-517,571-581
ACC_HD Flow fcopy(Flow f) { if (f.q <= 0) f.v = 0.0; return f; }
// get_price(): round(Price(value / quantity)), 0 ("NaN") for an empty flow, types.h:520-526
ACC_HD i64 flow_price(Flow f) { return f.q <= 0 ? 0 : mkp(f.v / qf(f.q)); }
// get_price_float(), types.h:527-534
ACC_HD double flow_price_float(Flow f) { return f.q <= 0 ? (double)NAN : f.v / qf(f.q); }
// set_price(), types.h:535-543
ACC_HD Flow set_price(Flow f, i64 p) { f.v = f.q <= 0 ? 0.0 : q_times_p(f.q, p); return f; }
ACC_HD Flow fadd(Flow a, Flow b) { return mkflow(a.q + b.q, a.v + b.v); }  // types.h:552-554
ACC_HD Flow fsub(Flow a, Flow b) { return mkflow(a.q - b.q, a.v - b.v); }  // types.h:555-557
// operator-= snaps to zero when nothing is left, types.h:589-599
ACC_HD Flow fsub_assign(Flow a, Flow b) {
    a.q -= b.q; a.v -= b.v;
    if (a.q <= 0) { a.q = 0; a.v = 0.0; }
    return a;
}
// Flow * Ratio, types.h:558 (quantity re-quantised, value scaled)
ACC_HD Flow fmul(Flow a, double r) { return mkflow(mkq(qf(a.q) * r), a.v * r); }
// FlowQuantity * Ratio / Ratio * FlowQuantity, types.h:391,393
ACC_HD i64 qmul(i64 q, double r) { return mkq(qf(q) * r); }
ACC_HD i64 qmul_l(double r, i64 q) { return mkq(r * qf(q)); }
ACC_HD i64 qdiv(i64 q, double r) { return mkq(qf(q) / r); }
ACC_HD i64 pmul(i64 p, double r) { return mkp(pf(p) * r); }
ACC_HD i64 pdiv(i64 p, double r) { return mkp(pf(p) / r); }
// T / T -> Ratio of the raw counts, types.h:394
ACC_HD double ratio(i64 a, i64 b) { return (double)a / (double)b; }
// Flow * Time -> Stock and Stock / Time -> Flow, types.h:643-651 (time given as to_float(Time))
ACC_HD Flow flow_times_time(Flow f, double t) { return mkflow(mkq(qf(f.q) * t), f.v * t); }
ACC_HD Flow stock_over_time(Flow s, double t) { return mkflow(mkq(qf(s.q) / t), s.v / t); }

}  // namespace acc
