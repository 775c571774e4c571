// In-register radix-2 DIF FFTs with compile-time twiddles.
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

// (ar + i ai) *= W_N^K with the trivial cases folded away
template <int K, int N>
PSLA_HD void mul_twiddle(float& ar, float& ai) {
    constexpr int k = ((K % N) + N) % N;
    if constexpr (k == 0) {
    } else if constexpr (4 * k == N) {  // -i
        float t = ar;
        ar = ai;
        ai = -t;
    } else if constexpr (2 * k == N) {  // -1
        ar = -ar;
        ai = -ai;
    } else if constexpr (4 * k == 3 * N) {  // +i
        float t = ar;
        ar = -ai;
        ai = t;
    } else if constexpr (8 * k == N) {  // (1 - i)/sqrt2
        constexpr float h = Twiddle<1, 8>::re;
        float t = (ar + ai) * h;
        ai = (ai - ar) * h;
        ar = t;
    } else if constexpr (8 * k == 3 * N) {  // (-1 - i)/sqrt2
        constexpr float h = Twiddle<1, 8>::re;
        float t = (ai - ar) * h;
        ai = -(ar + ai) * h;
        ar = t;
    } else {
        constexpr float wr = Twiddle<k, N>::re, wi = Twiddle<k, N>::im;
        float t = ar * wr - ai * wi;
        ai = ar * wi + ai * wr;
        ar = t;
    }
}

// In-place decimation-in-frequency FFT of length N (power of two) on register arrays.
// Input natural order; output bit-reversed: X[brev<log2 N>(r)] = (re[r], im[r]).
template <int N, int LEN = N>
PSLA_HD void fft_dif(float (&re)[N], float (&im)[N]) {
    if constexpr (LEN >= 2) {
        constexpr int HALF = LEN / 2;
        static_for<0, N / LEN>([&](auto blk_) {
            constexpr int base = decltype(blk_)::value * LEN;
            static_for<0, HALF>([&](auto j_) {
                constexpr int j = decltype(j_)::value;
                constexpr int p = base + j, q = base + j + HALF;
                float ur = re[p], ui = im[p], vr = re[q], vi = im[q];
                re[p] = ur + vr;
                im[p] = ui + vi;
                float dr = ur - vr, di = ui - vi;
                mul_twiddle<j, LEN>(dr, di);
                re[q] = dr;
                im[q] = di;
            });
        });
        fft_dif<N, HALF>(re, im);
    }
}

}  // namespace psla
