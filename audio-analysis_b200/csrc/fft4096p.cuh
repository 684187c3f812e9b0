// fft4096.cuh's warp-per-frame 4096-point real FFT on sm_100a's PACKED FP32 instructions (FADD2 / FMUL2 / FFMA2):
// every complex value is a float2 in an aligned register pair, a complex addition is one instruction, a twiddle
// multiplication two (fft_pairs.cuh).  Same algorithm and operation order as the scalar form (results agree to the
// last bit or two: the scalar build gets a few extra FMA contractions from nvcc) at roughly half the FP32 issue
// slots; the FP32 pipe time is unchanged
// (a packed instruction occupies the pipe for two cycles), so this helps exactly where the kernel is bound by
// issue slots rather than by the pipe (profiles/r2j_summary.md: issue active 69 %, FMA pipe 52 %).
// The default; -DCAC_FFT_PACKED=0 selects the scalar form (stft_mel.cu, sn_stft.cu).
#pragma once
#include "fft4096.cuh"
#include "fft_pairs.cuh"

namespace cac {

// z[I] = (x[2n] w[2n], x[2n+1] w[2n+1]), n = lane + 32 I; the window by angle addition as in WinLoad
template <int I>
struct WinLoadP {
    static __device__ __forceinline__ void run(float2 (&z)[64], const float* pe, const float* po, float2 hcb, float2 hsb) {
        constexpr float ca = (float)cx_cos2pi(I, 64);
        constexpr float sa = (float)cx_sin2pi(I, 64);
        const float2 w = __ffma2_rn(make_float2(-ca, -ca), hcb, __ffma2_rn(make_float2(sa, sa), hsb, make_float2(0.5f, 0.5f)));
        z[I] = __fmul2_rn(make_float2(pe[32 * I], po[32 * I]), w);
        if constexpr (I + 1 < 64) WinLoadP<I + 1>::run(z, pe, po, hcb, hsb);
    }
};

template <int K1, int BASE>
struct ExWriteP {
    static __device__ __forceinline__ void run(const float2 (&v)[64], float2* ex_lane) {
        ex_lane[K1 * kExStride] = v[out_pos(64, BASE + K1)];
        if constexpr (K1 + 1 < 32) ExWriteP<K1 + 1, BASE>::run(v, ex_lane);
    }
};
template <int N2>
struct ExReadP {
    static __device__ __forceinline__ void run(float2 (&v)[32], const float2* ex_row) {
        v[N2] = ex_row[N2];
        if constexpr (N2 + 1 < 32) ExReadP<N2 + 1>::run(v, ex_row);
    }
};

// pass-2 twiddles, one half at a time (see Pass2TwHalf): z *= A[n2 >> 3] * B[n2 & 7] (* W64^n2 for the second half)
template <int N2, int H>
struct Pass2TwHalfP {
    static __device__ __forceinline__ void run(float2 (&z)[32], const Twid& tw) {
        if constexpr (N2 > 0) {
            constexpr int A = N2 >> 3, B = N2 & 7;
            float2 x = z[N2];
            if constexpr (H == 1) x = c_twiddle<N2, 64>(x);
            if constexpr (A == 0) {
                z[N2] = c_mul(x, tw.br[B], tw.bi[B]);
            } else if constexpr (B == 0) {
                z[N2] = c_mul(x, tw.ar[A], tw.ai[A]);
            } else {
                const float2 t = c_mul(make_float2(tw.ar[A], tw.ai[A]), tw.br[B], tw.bi[B]);
                z[N2] = c_mul(x, t.x, t.y);
            }
        }
        if constexpr (N2 + 1 < 32) Pass2TwHalfP<N2 + 1, H>::run(z, tw);
    }
};

// real-FFT untangle + power of bins k = lane + 32 H + 64 K2 (see Untangle)
template <int K2, int H, int K2N, int POWER>
struct UntangleP {
    static __device__ __forceinline__ void run(const float2 (&z0)[32], const float2 (&z1)[32], int lane, int partner,
                                               float bc, float bs, float* pbuf_lane) {
        constexpr int pown = out_pos(32, K2);
        constexpr int psend = out_pos(32, 31 - K2);
        constexpr int psend0 = out_pos(32, (32 - K2) & 31);
        const float2 a = H ? z1[pown] : z0[pown];
        float2 snd;
        if constexpr (H == 0) snd = (lane == 0) ? z0[psend0] : z1[psend];
        else                  snd = (lane == 0) ? z1[psend] : z0[psend];
        float2 b;
        b.x = __shfl_sync(0xffffffffu, snd.x, partner);
        b.y = -__shfl_sync(0xffffffffu, snd.y, partner);      // conj
        const float2 sum = c_add(a, b);
        const float2 d = c_twiddle<K2, 64>(c_sub(a, b));      // W4096^(64 K2)
        const float2 t = c_mul_tw(d, bc, bs);                 // * W4096^(lane + 32 H) = (bc, -bs)
        const float2 x = c_sub_ib(sum, t);                    // 2 X[k] = (sum.re + t.im, sum.im - t.re)
        float p = fmaf(x.x, x.x, x.y * x.y);
        if constexpr (POWER == 1) p = sqrtf(p);
        pbuf_lane[32 * H + 64 * K2] = p;
        if constexpr (K2 + 1 < K2N) UntangleP<K2 + 1, H, K2N, POWER>::run(z0, z1, lane, partner, bc, bs, pbuf_lane);
    }
};

// The frame in two halves, so that a caller can put other work (a barrier) between them:
// fft_frame_p_load: windowed samples -> radix-64 pass, state = the 64 complex values of the lane;
// fft_frame_p_finish: exchange, twiddles, two radix-32 passes, untangle -> power values in pbuf.
__device__ __forceinline__ void fft_frame_p_load(const float* __restrict__ se, const float* __restrict__ so, int u0,
                                                 int lane, const float* __restrict__ lt, float2 (&a)[64]) {
    const float* pe;
    const float* po;
    if (u0 & 1) { pe = so + (u0 >> 1) + lane; po = se + ((u0 + 1) >> 1) + lane; }
    else        { pe = se + (u0 >> 1) + lane; po = so + (u0 >> 1) + lane; }
    WinLoadP<0>::run(a, pe, po, make_float2(lt[0 * 32], lt[2 * 32]), make_float2(lt[1 * 32], lt[3 * 32]));
    fft_dif_c<64>(a);
}

template <int POWER, int K2N, bool NYQ = false>
__device__ __forceinline__ void fft_frame_p_finish(const float2 (&a)[64], float* __restrict__ ex, float* __restrict__ pbuf,
                                                   int lane, const float* __restrict__ lt) {
    float2 z0[32], z1[32];
    float2* ex2 = reinterpret_cast<float2*>(ex);
    __syncwarp();
    ExWriteP<0, 0>::run(a, ex2 + lane);
    __syncwarp();
    ExReadP<0>::run(z0, ex2 + lane * kExStride);
    __syncwarp();
    ExWriteP<0, 32>::run(a, ex2 + lane);
    __syncwarp();
    ExReadP<0>::run(z1, ex2 + lane * kExStride);
    Twid tw;
#pragma unroll
    for (int j = 0; j < 4; ++j) { tw.ar[j] = lt[(8 + j) * 32]; tw.ai[j] = lt[(12 + j) * 32]; }
#pragma unroll
    for (int j = 0; j < 8; ++j) { tw.br[j] = lt[(16 + j) * 32]; tw.bi[j] = lt[(24 + j) * 32]; }
    Pass2TwHalfP<0, 0>::run(z0, tw);
    __syncwarp();   // pbuf aliases ex: nobody may still be reading the exchange rows when the power values land
    Pass2TwHalfP<0, 1>::run(z1, tw);
    fft_dif_c<32>(z0);
    fft_dif_c<32>(z1);

    const int partner = (32 - lane) & 31;
    UntangleP<0, 0, K2N, POWER>::run(z0, z1, lane, partner, lt[4 * 32], lt[5 * 32], pbuf + lane);
    UntangleP<0, 1, K2N, POWER>::run(z0, z1, lane, partner, lt[6 * 32], lt[7 * 32], pbuf + lane);
    if constexpr (NYQ) {
        if (lane == 0) {
            const float x = 2.0f * (z0[0].x - z0[0].y);
            pbuf[2048] = POWER == 1 ? fabsf(x) : x * x;
        }
    }
}

// Same contract as fft_frame (fft4096.cuh)
template <int POWER, int K2N, bool NYQ = false>
__device__ __forceinline__ void fft_frame_p(const float* __restrict__ se, const float* __restrict__ so, int u0,
                                            float* __restrict__ ex, float* __restrict__ pbuf, int lane,
                                            const float* __restrict__ lt) {
    float2 a[64];
    fft_frame_p_load(se, so, u0, lane, lt, a);
    fft_frame_p_finish<POWER, K2N, NYQ>(a, ex, pbuf, lane, lt);
}

}  // namespace cac
