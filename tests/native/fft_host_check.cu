// Host-side check of gn_fft.cuh: runs the warp-level 2048-point FFT (stage 1, transpose tile, stage 2) for the
// 32 "lanes" sequentially on the CPU and compares (a) the complex spectrum and (b) the two packed real power
// spectra with a direct O(N^2) DFT in double.  Exit code 0 = max relative error below 2e-6.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../cvml_goalnet_b200/csrc/gn_fft.cuh"

int main() {
    using namespace gnfft;
    const double PI = 3.14159265358979323846;
    std::vector<float2> x(kN), tw(kTwLane), tile(kTile);
    srand(7);
    for (int i = 0; i < kN; ++i) x[i] = make_float2((float)rand() / RAND_MAX - 0.5f, (float)rand() / RAND_MAX - 0.5f);
    for (int k1 = 0; k1 < 64; ++k1)
        for (int n2 = 0; n2 < 32; ++n2) {
            const double a = -2 * PI * (double)(k1 * n2) / kN;
            tw[k1 * 32 + n2] = make_float2((float)cos(a), (float)sin(a));
        }
    for (int lane = 0; lane < 32; ++lane) {
        float2 z[64];
        for (int n1 = 0; n1 < 64; ++n1) z[n1] = x[32 * n1 + lane];
        warp_fft_stage1(lane, z, tw.data(), tile.data());
    }
    std::vector<float2> X(kN);
    std::vector<float> pw1(kN / 2 + 1, -1.f), pw2(kN / 2 + 1, -1.f);
    for (int lane = 0; lane < 32; ++lane) {
        float2 a[32], b[32];
        warp_fft_stage2(lane, tile.data(), a, b);
        for (int k2 = 0; k2 < 32; ++k2) {
            X[stage2_row_a(lane) + 64 * k2] = a[inv32(k2)];
            X[stage2_row_b(lane) + 64 * k2] = b[inv32(k2)];
        }
        warp_fft_power(lane, a, b, pw1.data(), pw2.data());
    }
    double max_err = 0, max_ref = 0, perr = 0, pref = 0;
    for (int k = 0; k < kN; ++k) {
        double re = 0, im = 0, r1 = 0, i1 = 0, r2 = 0, i2 = 0;
        for (int n = 0; n < kN; ++n) {
            const double a = -2 * PI * (double)((long long)n * k % kN) / kN;
            const double c = cos(a), s = sin(a);
            re += x[n].x * c - x[n].y * s;
            im += x[n].x * s + x[n].y * c;
            r1 += x[n].x * c; i1 += x[n].x * s;
            r2 += x[n].y * c; i2 += x[n].y * s;
        }
        max_err = fmax(max_err, fmax(fabs(re - X[k].x), fabs(im - X[k].y)));
        max_ref = fmax(max_ref, fmax(fabs(re), fabs(im)));
        if (k <= kN / 2) {
            perr = fmax(perr, fmax(fabs(r1 * r1 + i1 * i1 - pw1[k]), fabs(r2 * r2 + i2 * i2 - pw2[k])));
            pref = fmax(pref, fmax(r1 * r1 + i1 * i1, r2 * r2 + i2 * i2));
        }
    }
    printf("spectrum: max_abs_err %.3e max_ref %.3e rel %.3e\n", max_err, max_ref, max_err / max_ref);
    printf("power   : max_abs_err %.3e max_ref %.3e rel %.3e\n", perr, pref, perr / pref);
    return (max_err / max_ref < 2e-6 && perr / pref < 4e-6) ? 0 : 1;
}
