// In-register radix-2 DIT FFTs with compile-time twiddles.
//
// Everything here is `__host__ __device__` so that the exact per-thread code the sm_100a kernel
// runs can also be compiled by g++ and stepped "one simulated lane at a time" by the CPU test
// harness (tests/emu/): there is no GPU in the build container, so index maps, twiddles and the
// two-pass decomposition are verified on the host before a B200 is ever touched.
//
// All loop indices are compile-time (static_for), which guarantees that the complex arrays live in
// registers and that every twiddle is an immediate operand (FFMA imm-form).
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
// Decimation-in-time FFT with FMA-only butterflies (Linzer-Feig): for a twiddle W = c + i*s
//   |c| >= |s|:  W*b = c * (b_r - t*b_i,  b_i + t*b_r),  t = s/c      (2 FMA)   then  a +- c*(...)   (4 FMA)
//   |c| <  |s|:  W*b = s * (t*b_r - b_i,  t*b_i + b_r),  t = c/s
// i.e. 6 instructions per non-trivial butterfly instead of the 8 of the multiply-after-subtract DIF form, and 4 for
// W in {1, -i}.  Natural-order input, bit-reversed output: X[brev(r)] = (re[r], im[r]) -- the DIT network runs
// on the bit-reversed view y[i] = x[brev(i)], so "in place" it leaves X[k] at index brev(k).
// NZ: inputs with index >= NZ are known to be zero (their first-stage butterflies become copies).
// ---------------------------------------------------------------------------------------------
template <int J, int L>
PSLA_HD void dit_butterfly(float& ar, float& ai, float& br, float& bi) {
    constexpr int j = ((J % L) + L) % L;
    if constexpr (j == 0) {
        const float xr = ar - br, xi = ai - bi;
        ar += br;
        ai += bi;
        br = xr;
        bi = xi;
    } else if constexpr (4 * j == L) {  // W = -i:  W*b = (b_i, -b_r)
        const float xr = ar - bi, xi = ai + br;
        ar += bi;
        ai -= br;
        br = xr;
        bi = xi;
    } else {
        constexpr double c = ct_cos2pi(j, L), sn = -ct_sin2pi(j, L);
        constexpr bool big_c = (c < 0 ? -c : c) >= (sn < 0 ? -sn : sn);
        float pr, pi, m;
        if constexpr (8 * j == L) {  // c = -s = sqrt(1/2)
            pr = br + bi;
            pi = bi - br;
            m = static_cast<float>(c);
        } else if constexpr (8 * j == 3 * L) {  // c = s = -sqrt(1/2)
            pr = br - bi;
            pi = bi + br;
            m = static_cast<float>(c);
        } else if constexpr (big_c) {
            constexpr float t = static_cast<float>(sn / c);
            pr = fmaf(-t, bi, br);
            pi = fmaf(t, br, bi);
            m = static_cast<float>(c);
        } else {
            constexpr float t = static_cast<float>(c / sn);
            pr = fmaf(t, br, -bi);
            pi = fmaf(t, bi, br);
            m = static_cast<float>(sn);
        }
        br = fmaf(-m, pr, ar);
        bi = fmaf(-m, pi, ai);
        ar = fmaf(m, pr, ar);
        ai = fmaf(m, pi, ai);
    }
}

template <int N, int BITS, int NZ = N, int LEN = 2>
PSLA_HD void fft_dit(float (&re)[N], float (&im)[N]) {
    if constexpr (LEN <= N) {
        constexpr int HALF = LEN / 2;
        static_for<0, N / LEN>([&](auto blk_) {
            constexpr int base = decltype(blk_)::value * LEN;
            static_for<0, HALF>([&](auto j_) {
                constexpr int j = decltype(j_)::value;
                constexpr int p = brev<BITS>(base + j), q = brev<BITS>(base + j + HALF);  // registers holding y[base+j], y[..+HALF]
                if constexpr (LEN == 2 && q >= NZ) {  // x[q] == 0: both outputs equal x[p]
                    re[q] = re[p];
                    im[q] = im[p];
                } else {
                    dit_butterfly<j, LEN>(re[p], im[p], re[q], im[q]);
                }
            });
        });
        fft_dit<N, BITS, NZ, 2 * LEN>(re, im);
    }
}

}  // namespace psla
