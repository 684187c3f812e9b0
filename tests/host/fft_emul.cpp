// Host emulation of the warp-per-frame 4096-point real FFT used by the CUDA kernels:
// same fft_regs.cuh code, lanes emulated by loops, smem exchange by arrays, shuffles by
// indexing.  Prints max error vs a long-double DFT.  Built and run by tests/test_host_fft.py.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <complex>
#include "../../audio-analysis_b200/csrc/fft_regs.cuh"
using namespace cac;

template <int N> static double test_small() {
    float re[N], im[N];
    std::vector<std::complex<double>> x(N);
    for (int i = 0; i < N; i++) { re[i] = (float)drand48() - 0.5f; im[i] = (float)drand48() - 0.5f; x[i] = {re[i], im[i]}; }
    fft_dif<N>(re, im);
    double err = 0;
    for (int k = 0; k < N; k++) {
        std::complex<double> s = 0;
        for (int n = 0; n < N; n++) s += x[n] * std::polar(1.0, -2.0 * M_PI * (double)((long)n * k % N) / N);
        int p = out_pos(N, k);
        err = std::max(err, std::abs(s - std::complex<double>(re[p], im[p])));
    }
    return err;
}

int main() {
    printf("small64 %.3e\nsmall32 %.3e\nsmall16 %.3e\nsmall8 %.3e\n", test_small<64>(), test_small<32>(), test_small<16>(), test_small<8>());
    const int NF = 4096, NH = 2048;
    std::vector<float> x(NF);
    for (auto& v : x) v = (float)drand48() * 2 - 1;
    // ---- pass 1: lane l holds z[32*n1 + l], n1 = 0..63, with Hann window via angle addition
    static float Ex_re[64][33], Ex_im[64][33];
    for (int l = 0; l < 32; l++) {
        float re[64], im[64];
        double b0 = 2 * M_PI * (2 * l) / 4096.0, b1 = 2 * M_PI * (2 * l + 1) / 4096.0;
        float hcb0 = 0.5f * (float)cos(b0), hsb0 = 0.5f * (float)sin(b0), hcb1 = 0.5f * (float)cos(b1), hsb1 = 0.5f * (float)sin(b1);
        for (int n1 = 0; n1 < 64; n1++) {
            int n = 32 * n1 + l;
            float ca = (float)cos(2 * M_PI * n1 / 64.0), sa = (float)sin(2 * M_PI * n1 / 64.0);
            float w0 = fmaf(-ca, hcb0, fmaf(sa, hsb0, 0.5f));
            float w1 = fmaf(-ca, hcb1, fmaf(sa, hsb1, 0.5f));
            re[n1] = x[2 * n] * w0; im[n1] = x[2 * n + 1] * w1;
        }
        fft_dif<64>(re, im);
        for (int k1 = 0; k1 < 64; k1++) { Ex_re[k1][l] = re[out_pos(64, k1)]; Ex_im[k1][l] = im[out_pos(64, k1)]; }
    }
    // ---- pass 2: lane l handles k1 = l + 32h; twiddle W2048^(k1*n2) = Tw_l[n2] * W64^(h*n2)
    static float Zr[32][2][32], Zi[32][2][32];  // [lane][h][k2]
    for (int l = 0; l < 32; l++) {
        float twr[32], twi[32];
        for (int n2 = 0; n2 < 32; n2++) { twr[n2] = (float)cos(2 * M_PI * l * n2 / 2048.0); twi[n2] = (float)-sin(2 * M_PI * l * n2 / 2048.0); }
        for (int h = 0; h < 2; h++) {
            float re[32], im[32];
            for (int n2 = 0; n2 < 32; n2++) {
                float ar = Ex_re[l + 32 * h][n2], ai = Ex_im[l + 32 * h][n2];
                if (h) { // * W64^n2
                    float c = (float)cos(2 * M_PI * n2 / 64.0), s = (float)sin(2 * M_PI * n2 / 64.0);
                    float t = ar; ar = fmaf(ai, s, t * c); ai = fmaf(-t, s, ai * c);
                }
                re[n2] = ar * twr[n2] - ai * twi[n2];
                im[n2] = ar * twi[n2] + ai * twr[n2];
            }
            fft_dif<32>(re, im);
            for (int k2 = 0; k2 < 32; k2++) { Zr[l][h][k2] = re[out_pos(32, k2)]; Zi[l][h][k2] = im[out_pos(32, k2)]; }
        }
    }
    // ---- untangle: X[k], k = l + 32h + 64 k2, partner Z[2048-k] from lane (32-l)&31
    double err = 0, ref_max = 0;
    std::vector<std::complex<double>> X(NH + 1);
    for (int k = 0; k <= NH; k++) {
        std::complex<long double> s = 0;
        for (int n = 0; n < NF; n++) {
            long double w = 0.5L - 0.5L * cosl(2 * M_PIl * n / NF);
            s += (long double)x[n] * w * std::polar(1.0L, -2.0L * M_PIl * (long double)((long)n * k % NF) / NF);
        }
        X[k] = {(double)s.real(), (double)s.imag()};
        ref_max = std::max(ref_max, std::abs(X[k]));
    }
    for (int l = 0; l < 32; l++) for (int h = 0; h < 2; h++) for (int k2 = 0; k2 < 16; k2++) {
        int k = l + 32 * h + 64 * k2;
        int pl = (32 - l) & 31, ph, pk2;
        if (l == 0) { ph = h; pk2 = h ? 31 - k2 : (32 - k2) & 31; } else { ph = 1 - h; pk2 = 31 - k2; }
        float ar = Zr[l][h][k2], ai = Zi[l][h][k2], br = Zr[pl][ph][pk2], bi = -Zi[pl][ph][pk2];
        float sr = ar + br, si = ai + bi, dr = ar - br, di = ai - bi;
        float c = (float)cos(2 * M_PI * k / 4096.0), s = (float)sin(2 * M_PI * k / 4096.0);
        float tr = fmaf(di, s, dr * c), ti = fmaf(-dr, s, di * c);
        float xr = 0.5f * (sr + ti), xi = 0.5f * (si - tr);
        err = std::max(err, std::abs(X[k] - std::complex<double>(xr, xi)));
    }
    printf("rfft4096 max_abs_err %.3e ref_max %.3e rel %.3e\n", err, ref_max, err / ref_max);
    return 0;
}
