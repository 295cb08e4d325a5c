"""Fixed-point quantisation of csrc/numeric.cuh against the reference's definition lrint(f / precision)
(reference include/types.h:288-294,313-319).  The header is plain C++ outside nvcc, so the check compiles it with
g++ and runs it on the host: random magnitudes, exact multiples, and the neighbourhood of every kind of rounding
boundary (k + 1/2 quanta, +- a few ulp)."""
import os
import subprocess
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SRC = textwrap.dedent(r"""
    #include <cstdio>
    #include <cstdint>
    #include <cmath>
    #include <random>
    #include "numeric.cuh"
    using namespace acc;
    static long long bad = 0, n = 0;
    static void check(double f) {
        ++n;
        if (mkq(f) != llrint(f / ACC_PREC3)) { if (bad < 5) printf("mkq %a\n", f); ++bad; }
        if (mkp(f) != llrint(f / ACC_PREC6)) { if (bad < 5) printf("mkp %a\n", f); ++bad; }
    }
    static void around(double f) {
        double lo = f, hi = f;
        check(f);
        for (int k = 0; k < 6; ++k) { lo = std::nextafter(lo, -INFINITY); hi = std::nextafter(hi, INFINITY); check(lo); check(hi); }
    }
    int main() {
        std::mt19937_64 rng(12345);
        std::uniform_real_distribution<double> u(0.0, 1.0);
        for (int i = 0; i < 4000000; ++i) {
            const double mag = std::pow(10.0, -7.0 + 20.0 * u(rng));
            const double f = (u(rng) - 0.3) * mag;
            check(f);
        }
        for (long long k = -2000; k < 400000; k += (k < 3000 ? 1 : 997)) {
            around(((double)k + 0.5) * ACC_PREC3); around(((double)k + 0.5) * 1e-3); around(((double)k + 0.5) / 1e3);
            around(((double)k + 0.5) * ACC_PREC6); around(((double)k + 0.5) * 1e-6); around(((double)k + 0.5) / 1e6);
            around((double)k * ACC_PREC3); around((double)k * ACC_PREC6);
        }
        for (int e = 30; e < 70; ++e) { around(std::ldexp(1.0, e)); around(-std::ldexp(1.5, e)); around((std::ldexp(1.0, e) + 0.5) * 1e-3); }
        around(0.0); around(-0.0); around(5e-4); around(5e-7); around(1.5e-3); around(2.5e-3); around(2.5e-6);
        printf("%lld checked, %lld bad\n", n, bad);
        return bad != 0;
    }
""")


def test_quantise_equals_division(tmp_path):
    src = tmp_path / "q.cpp"
    src.write_text(SRC)
    exe = tmp_path / "q"
    subprocess.run(["/usr/bin/g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "acclimate_b200", "csrc"), str(src), "-o", str(exe)], check=True)
    r = subprocess.run([str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
