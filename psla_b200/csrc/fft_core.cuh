// In-register radix-2 DIT FFTs with compile-time twiddles, in two-wide (packed complex) fp32 arithmetic.
//
// Everything here is `__host__ __device__` so that the exact per-thread code the sm_100a kernel
// runs can also be compiled by g++ and stepped "one simulated lane at a time" by the CPU test
// harness (tests/emu/): there is no GPU in the build container, so index maps, twiddles and the
// two-pass decomposition are verified on the host before a B200 is ever touched.
//
// All loop indices are compile-time (static_for), which guarantees that the complex arrays live in
// registers and that every twiddle is an immediate operand (FFMA2 imm-form).
#pragma once

#include <cmath>
#include <type_traits>

#if defined(__CUDACC__)
#define PSLA_HD __host__ __device__ __forceinline__
#define PSLA_HDC __host__ __device__ constexpr
#else
#define PSLA_HD inline __attribute__((always_inline))
#define PSLA_HDC constexpr
#endif

namespace psla {

template <int B, int E, class F>
PSLA_HD void static_for(F&& f) {
    if constexpr (B < E) {
        f(std::integral_constant<int, B>{});
        static_for<B + 1, E>(f);
    }
}

// ---- compile-time trigonometry (double Taylor series after octant reduction) ----
constexpr double kPi = 3.14159265358979323846264338327950288;

constexpr double ct_sin_small(double x) {  // |x| <= pi/4
    double x2 = x * x, term = x, sum = x;
    for (int i = 1; i < 14; ++i) {
        term *= -x2 / double((2 * i) * (2 * i + 1));
        sum += term;
    }
    return sum;
}
constexpr double ct_cos_small(double x) {  // |x| <= pi/4
    double x2 = x * x, term = 1.0, sum = 1.0;
    for (int i = 1; i < 14; ++i) {
        term *= -x2 / double((2 * i - 1) * (2 * i));
        sum += term;
    }
    return sum;
}
// cos(2*pi*k/n), sin(2*pi*k/n) for integers, exact symmetries first so that 0, +-1 and
// +-sqrt(1/2) come out as exactly those doubles.
constexpr double ct_cos2pi(int k, int n) {
    k %= n;
    if (k < 0) k += n;
    if (2 * k > n) k = n - k;                      // cos is even around pi
    if (4 * k > n) return -ct_cos2pi(n - 2 * k, 2 * n);   // cos(pi - a) = -cos(a); a = 2pi(n-2k)/(2n)
    if (8 * k > n) return ct_sin_small(2.0 * kPi * double(n - 4 * k) / double(4 * n));  // cos(a)=sin(pi/2-a)
    return ct_cos_small(2.0 * kPi * double(k) / double(n));
}
constexpr double ct_sin2pi(int k, int n) {
    // sin(a) = cos(a - pi/2) = cos(2pi (4k - n) / (4n))
    return ct_cos2pi(4 * k - n, 4 * n);
}

template <int BITS>
PSLA_HDC int brev(int v) {
    int r = 0;
    for (int i = 0; i < BITS; ++i) r |= ((v >> i) & 1) << (BITS - 1 - i);
    return r;
}

// W_N^K = exp(-2*pi*i*K/N) as floats rounded from double
template <int K, int N>
struct Twiddle {
    static constexpr float re = static_cast<float>(ct_cos2pi(K, N));
    static constexpr float im = static_cast<float>(-ct_sin2pi(K, N));
};

// ---------------------------------------------------------------------------------------------
// Packed complex arithmetic.  A complex value is one `cf` = (re, im) in an aligned register pair.  sm_100a has
// two-wide fp32 instructions (FADD2 / FMUL2 / FFMA2: one issue slot, both halves) whose operands can be a register pair,
// the same pair with its halves swapped and / or one half negated, or one scalar broadcast to both halves -- which
// is exactly complex add, multiply by +-i, and multiply by a real.  Every helper below is ONE such instruction on the
// device; on the host (tests/emu) it is the same arithmetic on two floats with the same roundings (fmaf).
// ---------------------------------------------------------------------------------------------
#if defined(__CUDA_ARCH__)
using cf = float2;
PSLA_HD cf mk(float x, float y) { return make_float2(x, y); }
PSLA_HD cf cadd(cf a, cf b) { return __fadd2_rn(a, b); }
PSLA_HD cf csub(cf a, cf b) { return __fadd2_rn(a, make_float2(-b.x, -b.y)); }
PSLA_HD cf cfma(cf a, cf b, cf c) { return __ffma2_rn(a, b, c); }  // halves independently
PSLA_HD cf cmul2(cf a, cf b) { return __fmul2_rn(a, b); }         // halves independently
#else
struct alignas(8) cf {
    float x, y;
};
PSLA_HD cf mk(float x, float y) { return cf{x, y}; }
PSLA_HD cf cadd(cf a, cf b) { return cf{a.x + b.x, a.y + b.y}; }
PSLA_HD cf csub(cf a, cf b) { return cf{a.x - b.x, a.y - b.y}; }
PSLA_HD cf cfma(cf a, cf b, cf c) { return cf{fmaf(a.x, b.x, c.x), fmaf(a.y, b.y, c.y)}; }
PSLA_HD cf cmul2(cf a, cf b) { return cf{a.x * b.x, a.y * b.y}; }
#endif
PSLA_HD cf bc(float s) { return mk(s, s); }                                // scalar broadcast operand
PSLA_HD cf mul_i(cf b) { return mk(-b.y, b.x); }                           //  i * b   (operand swizzle, no instruction)
PSLA_HD cf mul_mi(cf b) { return mk(b.y, -b.x); }                          // -i * b
PSLA_HD cf conj_(cf b) { return mk(b.x, -b.y); }
PSLA_HD cf cscale_add(cf p, float m, cf a) { return cfma(p, bc(m), a); }   // a + m * p
// a * (wr + i wi)  and  a * (wr - i wi): FMUL2 + FFMA2.  Rounding order as in the scalar expressions
//   (a_r*wr - a_i*wi, a_r*wi + a_i*wr)  /  (a_r*wr + a_i*wi, a_i*wr - a_r*wi)  with the second product rounded first.
PSLA_HD cf cmul(cf a, float wr, float wi) {
    const cf t = cmul2(bc(a.y), mk(wi, wr));  // (a_i*wi, a_i*wr)
    return cfma(bc(a.x), mk(wr, wi), mk(-t.x, t.y));
}
PSLA_HD cf cmul_conj(cf a, float wr, float wi) { return cfma(a, bc(wr), cmul2(mul_mi(a), bc(wi))); }

// ---------------------------------------------------------------------------------------------
// Decimation-in-time FFT with FMA-only butterflies (Linzer-Feig): for a twiddle W = c + i*s
//   |c| >= |s|:  W*b = c * (b + t * i*b),   t = s/c      (1 FFMA2)   then  a +- c*(...)   (2 FFMA2)
//   |c| <  |s|:  W*b = s * (t*b + i*b),     t = c/s
// i.e. 3 two-wide instructions per non-trivial butterfly, 2 for W in {1, -i}.  Natural-order input, bit-reversed
// output: X[brev(r)] = z[r] -- the DIT network runs on the bit-reversed view y[i] = x[brev(i)], so "in place" it leaves
// X[k] at index brev(k).
// NZ: inputs with index >= NZ are known to be zero (their first-stage butterflies become copies).
// ---------------------------------------------------------------------------------------------
template <int J, int L>
PSLA_HD void dit_butterfly(cf& a, cf& b) {
    constexpr int j = ((J % L) + L) % L;
    if constexpr (j == 0) {
        const cf x = csub(a, b);
        a = cadd(a, b);
        b = x;
    } else if constexpr (4 * j == L) {  // W = -i:  W*b = (b_i, -b_r)
        const cf x = cadd(a, mul_i(b));  // a - W*b
        a = cadd(a, mul_mi(b));
        b = x;
    } else {
        constexpr double c = ct_cos2pi(j, L), sn = -ct_sin2pi(j, L);
        constexpr bool big_c = (c < 0 ? -c : c) >= (sn < 0 ? -sn : sn);
        cf p;
        float m;
        if constexpr (8 * j == L) {  // c = -s = sqrt(1/2):  p = (b_r + b_i, b_i - b_r)
            p = cadd(b, mul_mi(b));
            m = static_cast<float>(c);
        } else if constexpr (8 * j == 3 * L) {  // c = s = -sqrt(1/2):  p = (b_r - b_i, b_i + b_r)
            p = cadd(b, mul_i(b));
            m = static_cast<float>(c);
        } else if constexpr (big_c) {  // p = (b_r - t b_i, b_i + t b_r)
            constexpr float t = static_cast<float>(sn / c);
            p = cfma(mul_i(b), bc(t), b);
            m = static_cast<float>(c);
        } else {  // p = (t b_r - b_i, t b_i + b_r)
            constexpr float t = static_cast<float>(c / sn);
            p = cfma(b, bc(t), mul_i(b));
            m = static_cast<float>(sn);
        }
        b = cscale_add(p, -m, a);
        a = cscale_add(p, m, a);
    }
}

template <int N, int BITS, int NZ = N, int LEN = 2>
PSLA_HD void fft_dit(cf (&z)[N]) {
    if constexpr (LEN <= N) {
        constexpr int HALF = LEN / 2;
        static_for<0, N / LEN>([&](auto blk_) {
            constexpr int base = decltype(blk_)::value * LEN;
            static_for<0, HALF>([&](auto j_) {
                constexpr int j = decltype(j_)::value;
                constexpr int p = brev<BITS>(base + j), q = brev<BITS>(base + j + HALF);  // registers holding y[base+j], y[..+HALF]
                if constexpr (LEN == 2 && q >= NZ) {  // x[q] == 0: both outputs equal x[p]
                    z[q] = z[p];
                } else {
                    dit_butterfly<j, LEN>(z[p], z[q]);
                }
            });
        });
        fft_dit<N, BITS, NZ, 2 * LEN>(z);
    }
}

}  // namespace psla
