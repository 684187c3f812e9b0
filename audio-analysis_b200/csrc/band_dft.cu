// Framed DFT of a narrow band of bins for ANY frame length (even or odd): the path behind
// identify_morepork._load_sample (src/identify_morepork.py:19-27) at native sample rates whose
// n_fft = int(sr / 10) is not one of the mixed-radix lengths of stft4800.cu (44.1 kHz -> 4410 = 2 3^2 5 7^2,
// 22.05 kHz -> 2205, 8 kHz -> 800, ...).  Only 60 of the n_fft / 2 + 1 bins are kept (:25), so the band is
// evaluated directly: X[k] = sum_n w[n] x[n] exp(-2 pi i k n / N), 60 N complex multiply-adds per frame.
// librosa.stft semantics: periodic Hann, centre = True with n_fft // 2 zeros on both sides.
//
// One CTA per frame (persistent loop): the windowed frame and the N-entry twiddle table live in shared
// memory; one warp per bin, lane l sums the samples n = l, l + 32, ... with the table index (k n) mod N
// advanced by (32 k) mod N per step (four independent accumulator pairs), then a shuffle reduction - every
// lane adds N / 32 terms, so the rounding error stays at the level of the FFT kernels.
#include "cac_common.cuh"
#include "spec.cuh"

namespace cac {

constexpr int kDftThreads = 256;

__global__ void __launch_bounds__(kDftThreads) band_dft_kernel(const SpecArgs a, const float2* __restrict__ tw_g) {
    extern __shared__ __align__(16) unsigned char dsm[];
    const int N = a.n_fft;
    float2* tw = reinterpret_cast<float2*>(dsm);                 // exp(-2 pi i j / N), j = 0..N-1
    float* xw = reinterpret_cast<float*>(dsm + (size_t)N * sizeof(float2));
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int j = tid; j < N; j += kDftThreads) tw[j] = tw_g[j];
    float vmax = 0.f;

    for (long long g = blockIdx.x; g < a.n_frames_total; g += gridDim.x) {
        int lo = 0, hi = a.n_items - 1;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (a.frame_base[mid] <= g) lo = mid; else hi = mid - 1;
        }
        const long long fbase = a.frame_base[lo];
        const long long f = g - fbase;
        const long long n = a.item_len[lo];
        const long long t0 = f * a.hop - N / 2;                      // centre = True
        const float* p = a.pcm + a.item_off[lo];
        __syncthreads();                                             // table loaded / previous frame's readers done
        for (int j = tid; j < N; j += kDftThreads) {
            const long long t = t0 + j;
            const float v = (t >= 0 && t < n) ? __ldg(p + t) : 0.f;
            xw[j] = v * fmaf(-0.5f, tw[j].x, 0.5f);                  // periodic Hann: 0.5 - 0.5 cos(2 pi j / N)
        }
        __syncthreads();
        for (int b = warp; b < a.nbins; b += kDftThreads / 32) {
            const int k = a.kmin + b;
            const int step = (int)((32LL * k) % N);
            int idx = (int)(((long long)k * lane) % N);
            float r0 = 0.f, i0 = 0.f, r1 = 0.f, i1 = 0.f, r2 = 0.f, i2 = 0.f, r3 = 0.f, i3 = 0.f;
            int j = lane;
            for (; j + 96 < N; j += 128) {
                const float2 t0v = tw[idx]; idx += step; idx -= idx >= N ? N : 0;
                const float2 t1v = tw[idx]; idx += step; idx -= idx >= N ? N : 0;
                const float2 t2v = tw[idx]; idx += step; idx -= idx >= N ? N : 0;
                const float2 t3v = tw[idx]; idx += step; idx -= idx >= N ? N : 0;
                const float x0 = xw[j], x1 = xw[j + 32], x2 = xw[j + 64], x3 = xw[j + 96];
                r0 = fmaf(x0, t0v.x, r0); i0 = fmaf(x0, t0v.y, i0);
                r1 = fmaf(x1, t1v.x, r1); i1 = fmaf(x1, t1v.y, i1);
                r2 = fmaf(x2, t2v.x, r2); i2 = fmaf(x2, t2v.y, i2);
                r3 = fmaf(x3, t3v.x, r3); i3 = fmaf(x3, t3v.y, i3);
            }
            for (; j < N; j += 32) {
                const float2 t = tw[idx]; idx += step; idx -= idx >= N ? N : 0;
                const float x = xw[j];
                r0 = fmaf(x, t.x, r0); i0 = fmaf(x, t.y, i0);
            }
            float re = (r0 + r1) + (r2 + r3), im = (i0 + i1) + (i2 + i3);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                re += __shfl_xor_sync(0xffffffffu, re, o);
                im += __shfl_xor_sync(0xffffffffu, im, o);
            }
            if (lane == 0) {
                float pw = fmaf(re, re, im * im);
                if (a.power == 1) pw = sqrtf(pw);
                vmax = fmaxf(vmax, pw);
                if (a.bin_major) a.out[(size_t)fbase * a.nbins + (size_t)b * (a.frame_base[lo + 1] - fbase) + f] = pw;
                else a.out[(size_t)g * a.nbins + b] = pw;
            }
        }
    }
    if (a.gmax && lane == 0) atomicMax(reinterpret_cast<unsigned*>(a.gmax), __float_as_uint(vmax));
}

// exp(-2 pi i j / N) in long double, rounded once to float
void band_dft_host_table(int n_fft, float2* tw) {
    const long double tau = 6.283185307179586476925286766559005768L;
    for (int j = 0; j < n_fft; ++j) {
        const long double ang = tau * (long double)j / (long double)n_fft;
        tw[j] = make_float2((float)cosl(ang), (float)-sinl(ang));
    }
}

size_t band_dft_smem_bytes(int n_fft) { return (size_t)n_fft * (sizeof(float2) + sizeof(float)); }

cudaError_t launch_band_dft(const SpecArgs& a, const float2* tw, cudaStream_t st) {
    const size_t smem = band_dft_smem_bytes(a.n_fft);
    cudaError_t e = cudaFuncSetAttribute(band_dft_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int per_sm = smem > 110 * 1024 ? 1 : (smem > 72 * 1024 ? 2 : 3);
    long long grid = a.n_frames_total;
    if (grid > (long long)sms * per_sm) grid = (long long)sms * per_sm;
    if (grid < 1) return cudaSuccess;
    band_dft_kernel<<<(unsigned)grid, kDftThreads, smem, st>>>(a, tw);
    return cudaGetLastError();
}

}  // namespace cac
