// In-register FFT building blocks (host+device), used by the framed real-FFT kernels.
//
// Replaces the pocketfft rfft that librosa.stft reaches from
// /root/reference/src/identify_tracks.py:243,389,654 and src/identify_morepork.py:24.
// Everything here is compile-time indexed so that arrays live in registers:
//   * twiddles are constexpr (double Taylor series after exact octant reduction, rounded once to f32)
//   * fft_dif<N>() is a decimation-in-frequency radix-4 (radix-2 tail) FFT on N contiguous
//     complex values; output X[k] lands at position out_pos(N, k) (base-4 digit reversal).
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

// Position of X[k] after fft_dif<N> (radix-4 splits while n >= 4, then radix-2).
CAC_CX int out_pos(int n, int k) {
    return n == 1 ? 0 : (n == 2 ? k : (k % 4) * (n / 4) + out_pos(n / 4, k / 4));
}

// v *= exp(-2*pi*i*E/N) with trivial cases folded at compile time.
template <int E_, int N>
CAC_HD void mul_twiddle(float& re, float& im) {
    constexpr int E = ((E_ % N) + N) % N;
    if constexpr (E == 0) {
    } else if constexpr (4 * E == N) {          // -i
        float t = re; re = im; im = -t;
    } else if constexpr (2 * E == N) {          // -1
        re = -re; im = -im;
    } else if constexpr (4 * E == 3 * N) {      // +i
        float t = re; re = -im; im = t;
    } else if constexpr (8 * E == N) {          // (1 - i)/sqrt2
        constexpr float c = (float)cx_cos2pi(1, 8);
        float t = re; re = (t + im) * c; im = (im - t) * c;
    } else if constexpr (8 * E == 3 * N) {      // (-1 - i)/sqrt2
        constexpr float c = (float)cx_cos2pi(1, 8);
        float t = re; re = (im - t) * c; im = -(t + im) * c;
    } else {
        constexpr float c = (float)cx_cos2pi(E, N);
        constexpr float s = (float)cx_sin2pi(E, N);
        float t = re;
        re = fmaf(im, s, t * c);       // (re + i im)(c - i s)
        im = fmaf(-t, s, im * c);
    }
}

template <int N, int B, int J>
struct R4Step {
    template <class A>
    static CAC_HD void run(A& re, A& im) {
        constexpr int Q = N / 4;
        float a0r = re[B + J], a0i = im[B + J];
        float a1r = re[B + J + Q], a1i = im[B + J + Q];
        float a2r = re[B + J + 2 * Q], a2i = im[B + J + 2 * Q];
        float a3r = re[B + J + 3 * Q], a3i = im[B + J + 3 * Q];
        float t0r = a0r + a2r, t0i = a0i + a2i;
        float t1r = a0r - a2r, t1i = a0i - a2i;
        float t2r = a1r + a3r, t2i = a1i + a3i;
        float t3r = a1r - a3r, t3i = a1i - a3i;
        float y0r = t0r + t2r, y0i = t0i + t2i;
        float y2r = t0r - t2r, y2i = t0i - t2i;
        float y1r = t1r + t3i, y1i = t1i - t3r;   // t1 - i*t3
        float y3r = t1r - t3i, y3i = t1i + t3r;   // t1 + i*t3
        mul_twiddle<J, N>(y1r, y1i);
        mul_twiddle<2 * J, N>(y2r, y2i);
        mul_twiddle<3 * J, N>(y3r, y3i);
        re[B + J] = y0r;          im[B + J] = y0i;
        re[B + J + Q] = y1r;      im[B + J + Q] = y1i;
        re[B + J + 2 * Q] = y2r;  im[B + J + 2 * Q] = y2i;
        re[B + J + 3 * Q] = y3r;  im[B + J + 3 * Q] = y3i;
        if constexpr (J + 1 < Q) R4Step<N, B, J + 1>::run(re, im);
    }
};

template <int N, int B>
struct FftDif {
    template <class A>
    static CAC_HD void run(A& re, A& im) {
        if constexpr (N == 2) {
            float ar = re[B], ai = im[B], br = re[B + 1], bi = im[B + 1];
            re[B] = ar + br;      im[B] = ai + bi;
            re[B + 1] = ar - br;  im[B + 1] = ai - bi;
        } else if constexpr (N >= 4) {
            R4Step<N, B, 0>::run(re, im);
            FftDif<N / 4, B>::run(re, im);
            FftDif<N / 4, B + N / 4>::run(re, im);
            FftDif<N / 4, B + N / 2>::run(re, im);
            FftDif<N / 4, B + 3 * N / 4>::run(re, im);
        }
    }
};

// Forward DFT of re/im[0..N) in place; X[k] ends at index out_pos(N, k).
template <int N, class A>
CAC_HD void fft_dif(A& re, A& im) {
    FftDif<N, 0>::run(re, im);
}

}  // namespace cac
