// In-register FFT building blocks (host+device), used by the framed real-FFT kernels.
//
// Replaces the pocketfft rfft that librosa.stft reaches from
// /root/reference/src/identify_tracks.py:243,389,654 and src/identify_morepork.py:24.
// Everything here is compile-time indexed so that arrays live in registers:
//   * twiddles are constexpr (double Taylor series after exact octant reduction, rounded once to f32)
//   * fft_dif<N>() is a decimation-in-frequency mixed-radix FFT (radix 4, then 2, 3, 5) on N
//     contiguous complex values of type float or double (deduced from the arrays); output X[k]
//     lands at position out_pos(N, k) (mixed-radix digit reversal).
#pragma once
#include <cmath>

#if defined(__CUDACC__)
#define CAC_HD __host__ __device__ __forceinline__
#define CAC_CX __host__ __device__ constexpr
#else
#define CAC_HD inline
#define CAC_CX constexpr
#endif

namespace cac {

constexpr double kPi = 3.14159265358979323846264338327950288;

CAC_CX double cx_sin_taylor(double x) {  // |x| <= pi/4
    double x2 = x * x, term = x, sum = x;
    for (int i = 1; i <= 13; ++i) {
        term *= -x2 / double((2 * i) * (2 * i + 1));
        sum += term;
    }
    return sum;
}
CAC_CX double cx_cos_taylor(double x) {  // |x| <= pi/4
    double x2 = x * x, term = 1.0, sum = 1.0;
    for (int i = 1; i <= 13; ++i) {
        term *= -x2 / double((2 * i - 1) * (2 * i));
        sum += term;
    }
    return sum;
}
// cos / sin of 2*pi*k/n for integers, reduced exactly to the first octant.
CAC_CX double cx_cos2pi(long long k, long long n) {
    k %= n;
    if (k < 0) k += n;
    long long K = 8 * k, N = 8 * n;  // N divisible by 8
    if (K > N / 2) K = N - K;
    bool neg = false;
    if (K > N / 4) { K = N / 2 - K; neg = true; }
    double v = (K > N / 8) ? cx_sin_taylor(2.0 * kPi * double(N / 4 - K) / double(N))
                           : cx_cos_taylor(2.0 * kPi * double(K) / double(N));
    return neg ? -v : v;
}
CAC_CX double cx_sin2pi(long long k, long long n) {
    k %= n;
    if (k < 0) k += n;
    long long K = 8 * k, N = 8 * n;
    bool neg = false;
    if (K > N / 2) { K = N - K; neg = true; }
    if (K > N / 4) K = N / 2 - K;
    double v = (K > N / 8) ? cx_cos_taylor(2.0 * kPi * double(N / 4 - K) / double(N))
                           : cx_sin_taylor(2.0 * kPi * double(K) / double(N));
    return neg ? -v : v;
}

// Radix fft_dif uses to split a length-n transform: 4 while divisible, then 2, 3, 5.
CAC_CX int fft_radix(int n) {
    return n % 4 == 0 ? 4 : (n % 2 == 0 ? 2 : (n % 3 == 0 ? 3 : (n % 5 == 0 ? 5 : n)));
}
// Position of X[k] after fft_dif<N> (digit reversal in the radix sequence above).
CAC_CX int out_pos(int n, int k) {
    return n == 1 ? 0 : (k % fft_radix(n)) * (n / fft_radix(n)) + out_pos(n / fft_radix(n), k / fft_radix(n));
}

CAC_HD float fma_t(float a, float b, float c) { return fmaf(a, b, c); }
CAC_HD double fma_t(double a, double b, double c) { return fma(a, b, c); }

// v *= exp(-2*pi*i*E/N) with trivial cases folded at compile time.
template <int E_, int N, class T>
CAC_HD void mul_twiddle(T& re, T& im) {
    constexpr int E = ((E_ % N) + N) % N;
    if constexpr (E == 0) {
    } else if constexpr (4 * E == N) {          // -i
        T t = re; re = im; im = -t;
    } else if constexpr (2 * E == N) {          // -1
        re = -re; im = -im;
    } else if constexpr (4 * E == 3 * N) {      // +i
        T t = re; re = -im; im = t;
    } else if constexpr (8 * E == N) {          // (1 - i)/sqrt2
        constexpr T c = (T)cx_cos2pi(1, 8);
        T t = re; re = (t + im) * c; im = (im - t) * c;
    } else if constexpr (8 * E == 3 * N) {      // (-1 - i)/sqrt2
        constexpr T c = (T)cx_cos2pi(1, 8);
        T t = re; re = (im - t) * c; im = -(t + im) * c;
    } else {
        constexpr T c = (T)cx_cos2pi(E, N);
        constexpr T s = (T)cx_sin2pi(E, N);
        T t = re;
        re = fma_t(im, s, t * c);      // (re + i im)(c - i s)
        im = fma_t(-t, s, im * c);
    }
}

template <class A>
struct ElemOf;
template <class T, int M>
struct ElemOf<T[M]> { using type = T; };

// One radix-R butterfly of a length-N DIF split (M = N/R): inputs at B+J+r*M, output r gets
// the twiddle W_N^(J*r) and goes back to the same slot.
template <int N, int B, int J>
struct RStep {
    template <class A>
    static CAC_HD void run(A& re, A& im) {
        using T = typename ElemOf<A>::type;
        constexpr int R = fft_radix(N);
        constexpr int Q = N / R;
        if constexpr (R == 4) {
            T a0r = re[B + J], a0i = im[B + J];
            T a1r = re[B + J + Q], a1i = im[B + J + Q];
            T a2r = re[B + J + 2 * Q], a2i = im[B + J + 2 * Q];
            T a3r = re[B + J + 3 * Q], a3i = im[B + J + 3 * Q];
            T t0r = a0r + a2r, t0i = a0i + a2i;
            T t1r = a0r - a2r, t1i = a0i - a2i;
            T t2r = a1r + a3r, t2i = a1i + a3i;
            T t3r = a1r - a3r, t3i = a1i - a3i;
            T y0r = t0r + t2r, y0i = t0i + t2i;
            T y2r = t0r - t2r, y2i = t0i - t2i;
            T y1r = t1r + t3i, y1i = t1i - t3r;   // t1 - i*t3
            T y3r = t1r - t3i, y3i = t1i + t3r;   // t1 + i*t3
            mul_twiddle<J, N>(y1r, y1i);
            mul_twiddle<2 * J, N>(y2r, y2i);
            mul_twiddle<3 * J, N>(y3r, y3i);
            re[B + J] = y0r;          im[B + J] = y0i;
            re[B + J + Q] = y1r;      im[B + J + Q] = y1i;
            re[B + J + 2 * Q] = y2r;  im[B + J + 2 * Q] = y2i;
            re[B + J + 3 * Q] = y3r;  im[B + J + 3 * Q] = y3i;
        } else if constexpr (R == 2) {
            T ar = re[B + J], ai = im[B + J], br = re[B + J + Q], bi = im[B + J + Q];
            T y1r = ar - br, y1i = ai - bi;
            mul_twiddle<J, N>(y1r, y1i);
            re[B + J] = ar + br;      im[B + J] = ai + bi;
            re[B + J + Q] = y1r;      im[B + J + Q] = y1i;
        } else if constexpr (R == 3) {
            constexpr T s3 = (T)cx_sin2pi(1, 3);
            T a0r = re[B + J], a0i = im[B + J];
            T a1r = re[B + J + Q], a1i = im[B + J + Q];
            T a2r = re[B + J + 2 * Q], a2i = im[B + J + 2 * Q];
            T t1r = a1r + a2r, t1i = a1i + a2i;
            T t2r = fma_t((T)-0.5, t1r, a0r), t2i = fma_t((T)-0.5, t1i, a0i);
            T t3r = (a1r - a2r) * s3, t3i = (a1i - a2i) * s3;
            T y1r = t2r + t3i, y1i = t2i - t3r;   // t2 - i*t3
            T y2r = t2r - t3i, y2i = t2i + t3r;   // t2 + i*t3
            mul_twiddle<J, N>(y1r, y1i);
            mul_twiddle<2 * J, N>(y2r, y2i);
            re[B + J] = a0r + t1r;    im[B + J] = a0i + t1i;
            re[B + J + Q] = y1r;      im[B + J + Q] = y1i;
            re[B + J + 2 * Q] = y2r;  im[B + J + 2 * Q] = y2i;
        } else {
            static_assert(R == 5, "fft_dif supports lengths with prime factors 2, 3 and 5 only");
            constexpr T c1 = (T)cx_cos2pi(1, 5), c2 = (T)cx_cos2pi(2, 5);
            constexpr T s1 = (T)cx_sin2pi(1, 5), s2 = (T)cx_sin2pi(2, 5);
            T a0r = re[B + J], a0i = im[B + J];
            T a1r = re[B + J + Q], a1i = im[B + J + Q];
            T a2r = re[B + J + 2 * Q], a2i = im[B + J + 2 * Q];
            T a3r = re[B + J + 3 * Q], a3i = im[B + J + 3 * Q];
            T a4r = re[B + J + 4 * Q], a4i = im[B + J + 4 * Q];
            T t1r = a1r + a4r, t1i = a1i + a4i;
            T t2r = a2r + a3r, t2i = a2i + a3i;
            T t3r = a1r - a4r, t3i = a1i - a4i;
            T t4r = a2r - a3r, t4i = a2i - a3i;
            T m1r = fma_t(c2, t2r, fma_t(c1, t1r, a0r)), m1i = fma_t(c2, t2i, fma_t(c1, t1i, a0i));
            T m2r = fma_t(c1, t2r, fma_t(c2, t1r, a0r)), m2i = fma_t(c1, t2i, fma_t(c2, t1i, a0i));
            T n1r = fma_t(s2, t4r, s1 * t3r), n1i = fma_t(s2, t4i, s1 * t3i);
            T n2r = fma_t(-s1, t4r, s2 * t3r), n2i = fma_t(-s1, t4i, s2 * t3i);
            T y1r = m1r + n1i, y1i = m1i - n1r;   // m1 - i*n1
            T y4r = m1r - n1i, y4i = m1i + n1r;   // m1 + i*n1
            T y2r = m2r + n2i, y2i = m2i - n2r;   // m2 - i*n2
            T y3r = m2r - n2i, y3i = m2i + n2r;   // m2 + i*n2
            mul_twiddle<J, N>(y1r, y1i);
            mul_twiddle<2 * J, N>(y2r, y2i);
            mul_twiddle<3 * J, N>(y3r, y3i);
            mul_twiddle<4 * J, N>(y4r, y4i);
            re[B + J] = a0r + t1r + t2r;  im[B + J] = a0i + t1i + t2i;
            re[B + J + Q] = y1r;      im[B + J + Q] = y1i;
            re[B + J + 2 * Q] = y2r;  im[B + J + 2 * Q] = y2i;
            re[B + J + 3 * Q] = y3r;  im[B + J + 3 * Q] = y3i;
            re[B + J + 4 * Q] = y4r;  im[B + J + 4 * Q] = y4i;
        }
        if constexpr (J + 1 < Q) RStep<N, B, J + 1>::run(re, im);
    }
};

template <int N, int B, int R, int I>
struct FftSub;

template <int N, int B>
struct FftDif {
    template <class A>
    static CAC_HD void run(A& re, A& im) {
        if constexpr (N >= 2) {
            RStep<N, B, 0>::run(re, im);
            if constexpr (N / fft_radix(N) >= 2) FftSub<N, B, fft_radix(N), 0>::run(re, im);
        }
    }
};
template <int N, int B, int R, int I>
struct FftSub {   // the R sub-transforms of length N/R
    template <class A>
    static CAC_HD void run(A& re, A& im) {
        FftDif<N / R, B + I * (N / R)>::run(re, im);
        if constexpr (I + 1 < R) FftSub<N, B, R, I + 1>::run(re, im);
    }
};

// Forward DFT of re/im[0..N) in place; X[k] ends at index out_pos(N, k).
template <int N, class A>
CAC_HD void fft_dif(A& re, A& im) {
    FftDif<N, 0>::run(re, im);
}

}  // namespace cac
