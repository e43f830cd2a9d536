// Host harness for tests/test_mel_fft_host.py: runs the butterflies of csrc/mel_fft.cuh on the CPU
// exactly as the mel kernel composes them (25 x rdft16 -> twiddle -> 9 x dft25 -> power[201]).
#include "mel_fft.cuh"
#include <cmath>
using namespace q3a::fft;
extern "C" void fft400_power(const float* x, float* power) {
    static cpx z[9][25];
    for (int n1 = 0; n1 < 25; ++n1) {
        float in[16];
        for (int n2 = 0; n2 < 16; ++n2) in[n2] = x[n1 + 25 * n2];
        cpx y[9];
        rdft16(in, y);
        for (int k2 = 0; k2 < 9; ++k2) {
            double ang = -2.0 * M_PI * (double)(n1 * k2) / 400.0;
            z[k2][n1] = cmul(y[k2], mk((float)std::cos(ang), (float)std::sin(ang)));
        }
    }
    for (int k2 = 0; k2 < 9; ++k2) {
        cpx a[25];
        for (int i = 0; i < 25; ++i) a[i] = z[k2][i];
        dft25(a);
        for (int k1 = 0; k1 < 25; ++k1) {
            int k = k2 + 16 * k1;
            float p = a[k1].re * a[k1].re + a[k1].im * a[k1].im;
            if (k <= 200) power[k] = p;
            else if (k2 != 0 && k2 != 8) power[400 - k] = p;
        }
    }
}
