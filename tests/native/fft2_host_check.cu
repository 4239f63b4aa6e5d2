// Host-side check of gn_fft2.cuh: runs the CTA-level 2048-point FFT (three passes, two exchanges, two transforms in the
// packed lanes, the last pass fused with the split into the four real columns' power spectra) for the 128 "threads"
// sequentially on the CPU and compares the four power spectra with a direct O(N^2) DFT in double; checks that every
// bin 0..1024 is written exactly once, and that the shared-memory accesses of the last pass put the 8 lanes of each
// quarter-warp on 8 different 16-byte slots.  Exit code 0 = max relative error below 4e-6, full coverage, no conflict.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
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
    std::vector<float4> buf0(kN), buf1(kN);
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
    // pass 3: all loads, (barrier), then the power spectra over the same buffer
    std::vector<C2> va(kThreads * 8), vb(kThreads * 8);
    auto R = [](std::vector<C2>& v, int t) -> C2(&)[8] { return *reinterpret_cast<C2(*)[8]>(&v[t * 8]); };
    for (int u = 0; u < kThreads; ++u) pass3_load(u, buf1.data(), R(va, u), R(vb, u));
    const float kUnset = -12345.f;
    for (auto& e : buf1) e = make_float4(kUnset, kUnset, kUnset, kUnset);
    std::vector<int> writes(kN, 0);
    for (int u = 0; u < kThreads; ++u) {
        std::vector<float4> before(buf1);
        pass3_power(u, R(va, u), R(vb, u), buf1.data());
        for (int i = 0; i < kN; ++i)
            if (memcmp(&before[i], &buf1[i], sizeof(float4)) != 0) ++writes[i];
    }
    int coverage_errors = 0;
    for (int k = 0; k <= kN / 2; ++k) coverage_errors += writes[k] != 1;
    int total_writes = 0;
    for (int i = 0; i < kN; ++i) total_writes += writes[i];
    coverage_errors += total_writes != kN / 2 + 1;

    double perr = 0, pref = 0;
    for (int k = 0; k <= kN / 2; ++k) {
        double p[4][2] = {{0}};
        for (int n = 0; n < kN; ++n) {
            const double a = -2 * PI * (double)((long long)n * k % kN) / kN;
            const double c = cos(a), s = sin(a);
            p[0][0] += xa_re[n] * c; p[0][1] += xa_re[n] * s;      // real sequence 1 of A
            p[1][0] += xb_re[n] * c; p[1][1] += xb_re[n] * s;      // real sequence 1 of B
            p[2][0] += xa_im[n] * c; p[2][1] += xa_im[n] * s;      // real sequence 2 of A
            p[3][0] += xb_im[n] * c; p[3][1] += xb_im[n] * s;      // real sequence 2 of B
        }
        const float4 pw = buf1[k];
        const float got[4] = {pw.x, pw.y, pw.z, pw.w};
        for (int i = 0; i < 4; ++i) {
            const double ref = p[i][0] * p[i][0] + p[i][1] * p[i][1];
            perr = fmax(perr, fabs(ref - got[i]));
            pref = fmax(pref, ref);
        }
    }
    // bank conflicts of pass 3: loads (16 instructions: 8 of unit a, 8 of unit b) and the regular threads' stores
    int conflicts = 0;
    auto check = [&](const char* name, int n_instr, auto index_of) {
        for (int i = 0; i < n_instr; ++i)
            for (int q = 0; q < kThreads / 8; ++q) {
                unsigned seen = 0;
                bool skip = false;
                for (int l = 0; l < 8; ++l) {
                    const int idx = index_of(q * 8 + l, i);
                    if (idx < 0) { skip = true; break; }
                    seen |= 1u << (idx & 7);
                }
                if (!skip && seen != 0xffu) {
                    if (conflicts < 10) printf("bank conflict: %s instr %d quarter-warp %d mask %02x\n", name, i, q, seen);
                    ++conflicts;
                }
            }
    };
    check("pass3 load", 16, [](int u, int i) {
        int pa, pb;
        pass3_units(u, pa, pb);
        const int p = (i < 8) ? pa : pb, n3 = i & 7;
        return p * 8 + (n3 ^ ((p >> 4) & 7));
    });
    check("pass3 store", 8, [](int u, int i) {
        if ((u >> 3) == (kSelfMirrorThread >> 3)) return -1;   // the quarter-warp with the self-mirrored units
        int pa, pb;
        pass3_units(u, pa, pb);
        const int ka = (pa >> 4) + 16 * (pa & 15), kb = (pb >> 4) + 16 * (pb & 15);
        return (i < 4) ? ka + 256 * i : kb + 256 * (7 - i);
    });
    printf("power   : max_abs_err %.3e max_ref %.3e rel %.3e\n", perr, pref, perr / pref);
    printf("coverage errors %d, bank conflicts %d\n", coverage_errors, conflicts);
    return (perr / pref < 4e-6 && coverage_errors == 0 && conflicts == 0) ? 0 : 1;
}
