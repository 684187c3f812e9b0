// In-register FFT building blocks on PACKED complex values (float2 = (re, im) in an aligned register pair) for
// sm_100a's packed FP32 instructions (FADD2 / FMUL2 / FFMA2).  One packed instruction does the real and the imaginary
// part of a complex addition; a twiddle multiplication is FMUL2 + FFMA2 (the second with the halves swapped and one
// sign flipped - both are free operand modifiers of the packed instructions, as is the broadcast of a scalar
// constant), so a butterfly needs half the issue slots of fft_regs.cuh's scalar form at the same FP32-pipe time.
// The arithmetic is the scalar form's operation for operation as written in the source; the compiled scalar form
// differs in the last bit here and there because nvcc contracts some of its separate multiplies and additions into
// FMAs (-fmad), which the explicit _rn packed intrinsics never do - both stay far inside the parity tolerances.
//
// Replaces the pocketfft rfft that librosa.stft reaches from /root/reference/src/identify_tracks.py:243, 654.
// Lengths: powers of two (radix 4, then 2); output X[k] lands at out_pos(N, k) like fft_dif.
#pragma once
#include "fft_regs.cuh"

namespace cac {

__device__ __forceinline__ float2 c_add(float2 a, float2 b) { return __fadd2_rn(a, b); }
__device__ __forceinline__ float2 c_sub(float2 a, float2 b) { return __fadd2_rn(a, make_float2(-b.x, -b.y)); }
// a - i b = (a.re + b.im, a.im - b.re);  a + i b = (a.re - b.im, a.im + b.re)
__device__ __forceinline__ float2 c_sub_ib(float2 a, float2 b) { return __fadd2_rn(a, make_float2(b.y, -b.x)); }
__device__ __forceinline__ float2 c_add_ib(float2 a, float2 b) { return __fadd2_rn(a, make_float2(-b.y, b.x)); }
// a * (c - i s): re = fma(im, s, re c), im = fma(-re, s, im c)   (mul_twiddle's general case)
__device__ __forceinline__ float2 c_mul_tw(float2 a, float c, float s) {
    const float2 t = __fmul2_rn(a, make_float2(c, c));
    return __ffma2_rn(make_float2(a.y, a.x), make_float2(s, -s), t);
}
// a * (tr + i ti): re = fma(-im, ti, re tr), im = fma(re, ti, im tr)   (the pass-2 twiddle form of fft4096.cuh)
__device__ __forceinline__ float2 c_mul(float2 a, float tr, float ti) {
    const float2 t = __fmul2_rn(a, make_float2(tr, tr));
    return __ffma2_rn(make_float2(a.y, a.x), make_float2(-ti, ti), t);
}

// v *= exp(-2 pi i E / N), trivial cases folded at compile time (same case split as mul_twiddle)
template <int E_, int N>
__device__ __forceinline__ float2 c_twiddle(float2 v) {
    constexpr int E = ((E_ % N) + N) % N;
    if constexpr (E == 0) {
        return v;
    } else if constexpr (4 * E == N) {          // -i
        return make_float2(v.y, -v.x);
    } else if constexpr (2 * E == N) {          // -1
        return make_float2(-v.x, -v.y);
    } else if constexpr (4 * E == 3 * N) {      // +i
        return make_float2(-v.y, v.x);
    } else if constexpr (8 * E == N) {          // (1 - i)/sqrt2: ((re + im) c, (im - re) c)
        constexpr float c = (float)cx_cos2pi(1, 8);
        return __fmul2_rn(__fadd2_rn(v, make_float2(v.y, -v.x)), make_float2(c, c));
    } else if constexpr (8 * E == 3 * N) {      // (-1 - i)/sqrt2: ((im - re) c, -(re + im) c)
        constexpr float c = (float)cx_cos2pi(1, 8);
        return __fmul2_rn(__fadd2_rn(make_float2(v.y, -v.x), make_float2(-v.x, -v.y)), make_float2(c, c));
    } else {
        constexpr float c = (float)cx_cos2pi(E, N);
        constexpr float s = (float)cx_sin2pi(E, N);
        return c_mul_tw(v, c, s);
    }
}

// One radix-R butterfly of a length-N DIF split (Q = N / R): inputs at B + J + r Q, output r gets W_N^(J r)
template <int N, int B, int J>
struct CStep {
    template <int M>
    static __device__ __forceinline__ void run(float2 (&z)[M]) {
        constexpr int R = fft_radix(N);
        constexpr int Q = N / R;
        static_assert(R == 4 || R == 2, "fft_dif_c: powers of two only");
        if constexpr (R == 4) {
            const float2 a0 = z[B + J], a1 = z[B + J + Q], a2 = z[B + J + 2 * Q], a3 = z[B + J + 3 * Q];
            const float2 t0 = c_add(a0, a2), t1 = c_sub(a0, a2);
            const float2 t2 = c_add(a1, a3), t3 = c_sub(a1, a3);
            z[B + J] = c_add(t0, t2);
            z[B + J + Q] = c_twiddle<J, N>(c_sub_ib(t1, t3));
            z[B + J + 2 * Q] = c_twiddle<2 * J, N>(c_sub(t0, t2));
            z[B + J + 3 * Q] = c_twiddle<3 * J, N>(c_add_ib(t1, t3));
        } else {
            const float2 a = z[B + J], b = z[B + J + Q];
            z[B + J] = c_add(a, b);
            z[B + J + Q] = c_twiddle<J, N>(c_sub(a, b));
        }
        if constexpr (J + 1 < Q) CStep<N, B, J + 1>::run(z);
    }
};

template <int N, int B, int R, int I>
struct CFftSub;

template <int N, int B>
struct CFftDif {
    template <int M>
    static __device__ __forceinline__ void run(float2 (&z)[M]) {
        if constexpr (N >= 2) {
            CStep<N, B, 0>::run(z);
            if constexpr (N / fft_radix(N) >= 2) CFftSub<N, B, fft_radix(N), 0>::run(z);
        }
    }
};
template <int N, int B, int R, int I>
struct CFftSub {
    template <int M>
    static __device__ __forceinline__ void run(float2 (&z)[M]) {
        CFftDif<N / R, B + I * (N / R)>::run(z);
        if constexpr (I + 1 < R) CFftSub<N, B, R, I + 1>::run(z);
    }
};

// Forward DFT of z[0..N) in place; X[k] ends at index out_pos(N, k)
template <int N>
__device__ __forceinline__ void fft_dif_c(float2 (&z)[N]) {
    CFftDif<N, 0>::run(z);
}

}  // namespace cac
