// Host-side check of gn_fft2.cuh: runs the CTA-level 2048-point FFT (three passes, two exchanges, two transforms in the
// packed lanes) for the 128 "threads" sequentially on the CPU and compares (a) both complex spectra and (b) the four
// packed real power spectra with a direct O(N^2) DFT in double.  Exit code 0 = max relative error below 2e-6 / 4e-6.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../cvml_goalnet_b200/csrc/gn_fft2.cuh"

int main() {
    using namespace gnfft2;
    const double PI = 3.14159265358979323846;
    std::vector<float> xa_re(kN), xa_im(kN), xb_re(kN), xb_im(kN);
    srand(11);
    auto rnd = []() { return (float)rand() / RAND_MAX - 0.5f; };
    for (int i = 0; i < kN; ++i) { xa_re[i] = rnd(); xa_im[i] = rnd(); xb_re[i] = rnd(); xb_im[i] = rnd(); }
    std::vector<float2> tw1(kTw1), tw2(kTw2);
    std::vector<float4> buf0(kN), buf1(kN), spec(kN);
    for (int k1 = 1; k1 < 16; ++k1)
        for (int t = 0; t < 128; ++t) {
            const double a = -2 * PI * (double)(t * k1) / 2048.0;
            tw1[(k1 - 1) * 128 + t] = make_float2((float)cos(a), (float)sin(a));
        }
    for (int k2 = 1; k2 < 16; ++k2)
        for (int n3 = 0; n3 < 8; ++n3) {
            const double a = -2 * PI * (double)(n3 * k2) / 128.0;
            tw2[(k2 - 1) * 8 + n3] = make_float2((float)cos(a), (float)sin(a));
        }
    for (int t = 0; t < kThreads; ++t) {
        C2 z[16];
        for (int n1 = 0; n1 < 16; ++n1) {
            const int n = t + 128 * n1;
            z[n1].re = mk(xa_re[n], xb_re[n]);
            z[n1].im = mk(xa_im[n], xb_im[n]);
        }
        pass1(t, z, tw1.data(), buf0.data());
    }
    for (int t = 0; t < kThreads; ++t) pass2(t, tw2.data(), buf0.data(), buf1.data());
    for (int u = 0; u < kThreads; ++u) pass3(u, buf1.data(), spec.data());

    double max_err = 0, max_ref = 0, perr = 0, pref = 0;
    for (int k = 0; k < kN; ++k) {
        double ra = 0, ia = 0, rb = 0, ib = 0, p[4][2] = {{0}};
        for (int n = 0; n < kN; ++n) {
            const double a = -2 * PI * (double)((long long)n * k % kN) / kN;
            const double c = cos(a), s = sin(a);
            ra += xa_re[n] * c - xa_im[n] * s; ia += xa_re[n] * s + xa_im[n] * c;
            rb += xb_re[n] * c - xb_im[n] * s; ib += xb_re[n] * s + xb_im[n] * c;
            p[0][0] += xa_re[n] * c; p[0][1] += xa_re[n] * s;      // real sequence 1 of A
            p[1][0] += xb_re[n] * c; p[1][1] += xb_re[n] * s;      // real sequence 1 of B
            p[2][0] += xa_im[n] * c; p[2][1] += xa_im[n] * s;      // real sequence 2 of A
            p[3][0] += xb_im[n] * c; p[3][1] += xb_im[n] * s;      // real sequence 2 of B
        }
        const float4 z = spec[spec_index(k)];
        max_err = fmax(max_err, fmax(fmax(fabs(ra - z.x), fabs(ia - z.z)), fmax(fabs(rb - z.y), fabs(ib - z.w))));
        max_ref = fmax(max_ref, fmax(fmax(fabs(ra), fabs(ia)), fmax(fabs(rb), fabs(ib))));
        if (k <= kN / 2) {
            const float4 pw = split_power(spec.data(), k);
            const float got[4] = {pw.x, pw.y, pw.z, pw.w};
            for (int i = 0; i < 4; ++i) {
                const double ref = p[i][0] * p[i][0] + p[i][1] * p[i][1];
                perr = fmax(perr, fabs(ref - got[i]));
                pref = fmax(pref, ref);
            }
        }
    }
    printf("spectrum: max_abs_err %.3e max_ref %.3e rel %.3e\n", max_err, max_ref, max_err / max_ref);
    printf("power   : max_abs_err %.3e max_ref %.3e rel %.3e\n", perr, pref, perr / pref);
    return (max_err / max_ref < 2e-6 && perr / pref < 4e-6) ? 0 : 1;
}
