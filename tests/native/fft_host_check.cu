// Host-side check of gn_fft.cuh: runs the four Stockham passes for all 256 "threads" sequentially on the
// CPU and compares with a direct O(N^2) DFT in double.  Exit code 0 = max relative error below 2e-6.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../../cvml_goalnet_b200/csrc/gn_fft.cuh"

int main() {
    using namespace gnfft;
    std::vector<float2> x(kN), tw(kTwCount);
    srand(7);
    for (int i = 0; i < kN; ++i) x[i] = make_float2((float)rand() / RAND_MAX - 0.5f, (float)rand() / RAND_MAX - 0.5f);
    const double PI = 3.14159265358979323846;
    for (int k = 0; k < 256; ++k) tw[kTw2048 + k] = make_float2((float)cos(-2 * PI * k / 2048), (float)sin(-2 * PI * k / 2048));
    for (int k = 0; k < 32; ++k) tw[kTw256 + k] = make_float2((float)cos(-2 * PI * k / 256), (float)sin(-2 * PI * k / 256));
    for (int k = 0; k < 4; ++k) tw[kTw32 + k] = make_float2((float)cos(-2 * PI * k / 32), (float)sin(-2 * PI * k / 32));
    std::vector<float> are(kBufA), aim(kBufA), bre(kBufB), bim(kBufB);
    for (int tid = 0; tid < kThreads; ++tid) {
        float2 z[8];
        for (int s = 0; s < 8; ++s) z[s] = x[tid + 256 * s];
        pass1(tid, z, are.data(), aim.data());
    }
    for (int tid = 0; tid < kThreads; ++tid) pass8<4, 1, 2>(tid, are.data(), aim.data(), bre.data(), bim.data(), tw.data() + kTw32);
    for (int tid = 0; tid < kThreads; ++tid) pass8<32, 2, 0>(tid, bre.data(), bim.data(), are.data(), aim.data(), tw.data() + kTw256);
    for (int tid = 0; tid < kThreads; ++tid) pass8<256, 0, 0>(tid, are.data(), aim.data(), bre.data(), bim.data(), tw.data() + kTw2048);
    double max_err = 0, max_ref = 0;
    for (int k = 0; k < kN; k += 1) {
        double re = 0, im = 0;
        for (int n = 0; n < kN; ++n) {
            const double a = -2 * PI * (double)((long long)n * k % kN) / kN;
            re += x[n].x * cos(a) - x[n].y * sin(a);
            im += x[n].x * sin(a) + x[n].y * cos(a);
        }
        max_err = fmax(max_err, fmax(fabs(re - bre[k]), fabs(im - bim[k])));
        max_ref = fmax(max_ref, fmax(fabs(re), fabs(im)));
    }
    printf("max_abs_err %.3e max_ref %.3e rel %.3e\n", max_err, max_ref, max_err / max_ref);
    return max_err / max_ref < 2e-6 ? 0 : 1;
}
