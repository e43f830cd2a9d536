// Register-resident butterflies of the 400-point real FFT used by the log-mel frontend.
//
// N = 400 = 25 x 16 (Cooley-Tukey, n = n1 + 25*n2, k = k2 + 16*k1):
//   stage 1  Y[n1][k2] = sum_n2 x[n1 + 25 n2] W16^(n2 k2)        25 real 16-point DFTs (k2 = 0..8 kept:
//                                                                 the input is real, Y[16-k2] = conj Y[k2])
//   twiddle  Z[n1][k2] = Y[n1][k2] * W400^(n1 k2)
//   stage 2  X[k2 + 16 k1] = sum_n1 Z[n1][k2] W25^(n1 k1)         9 complex 25-point DFTs (5 x 5)
// The 9 x 25 computed bins cover every k in [0,200] either directly or as the mirror 400-k, and the
// power spectrum the reference needs (mel_spectrogram.cpp:583-592) is symmetric.
//
// The functions are __host__ __device__ so that tests/test_mel_fft_host.py can compile this header
// with g++ and check the arithmetic against numpy without a GPU.
#pragma once
#ifdef __CUDACC__
#define Q3A_HD __host__ __device__ __forceinline__
#else
#define Q3A_HD inline
#endif

namespace q3a { namespace fft {

struct cpx { float re, im; };
Q3A_HD cpx mk(float re, float im) { cpx c; c.re = re; c.im = im; return c; }
Q3A_HD cpx operator+(cpx a, cpx b) { return mk(a.re + b.re, a.im + b.im); }
Q3A_HD cpx operator-(cpx a, cpx b) { return mk(a.re - b.re, a.im - b.im); }
Q3A_HD cpx cmul(cpx a, cpx b) { return mk(a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re); }
Q3A_HD cpx conj(cpx a) { return mk(a.re, -a.im); }
Q3A_HD cpx mul_neg_i(cpx a) { return mk(a.im, -a.re); }   // a * (-i)

// real 8-point DFT, bins 0..4 (bins 5..7 are conjugates of 3..1)
Q3A_HD void rdft8(const float (&a)[8], cpx (&r)[5]) {
    const float c = 0.70710678118654752440f;
    float f0 = a[0] + a[4], f1 = a[0] - a[4], f2 = a[2] + a[6], f3 = a[2] - a[6];
    float g0 = a[1] + a[5], g1 = a[1] - a[5], g2 = a[3] + a[7], g3 = a[3] - a[7];
    float F0 = f0 + f2, F2 = f0 - f2, G0 = g0 + g2, G2 = g0 - g2;
    r[0] = mk(F0 + G0, 0.0f);
    r[1] = mk(f1 + c * (g1 - g3), -f3 - c * (g1 + g3));
    r[2] = mk(F2, -G2);
    r[3] = mk(f1 + c * (g3 - g1), f3 - c * (g1 + g3));
    r[4] = mk(F0 - G0, 0.0f);
}

// real 16-point DFT, bins 0..8
Q3A_HD void rdft16(const float (&x)[16], cpx (&y)[9]) {
    float e[8], o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { e[j] = x[2 * j]; o[j] = x[2 * j + 1]; }
    cpx E[5], O[5];
    rdft8(e, E);
    rdft8(o, O);
    // W16^k = (cos(2 pi k/16), -sin(2 pi k/16))
    const float c1 = 0.92387953251128675613f, s1 = 0.38268343236508977173f;   // k = 1
    const float c2 = 0.70710678118654752440f;                                  // k = 2
    y[0] = mk(E[0].re + O[0].re, 0.0f);
    y[1] = E[1] + cmul(mk(c1, -s1), O[1]);
    y[2] = E[2] + cmul(mk(c2, -c2), O[2]);
    y[3] = E[3] + cmul(mk(s1, -c1), O[3]);
    y[4] = E[4] + mul_neg_i(O[4]);
    y[5] = conj(E[3]) + cmul(mk(-s1, -c1), conj(O[3]));
    y[6] = conj(E[2]) + cmul(mk(-c2, -c2), conj(O[2]));
    y[7] = conj(E[1]) + cmul(mk(-c1, -s1), conj(O[1]));
    y[8] = mk(E[0].re - O[0].re, 0.0f);
}

// complex 5-point DFT (forward, W5 = exp(-2 pi i/5))
Q3A_HD void dft5(cpx a0, cpx a1, cpx a2, cpx a3, cpx a4, cpx& x0, cpx& x1, cpx& x2, cpx& x3, cpx& x4) {
    const float c1 = 0.30901699437494742410f, c2 = -0.80901699437494742410f;
    const float s1 = 0.95105651629515357212f, s2 = 0.58778525229247312917f;
    cpx t1 = a1 + a4, t2 = a2 + a3, t3 = a1 - a4, t4 = a2 - a3;
    x0 = mk(a0.re + t1.re + t2.re, a0.im + t1.im + t2.im);
    cpx m1 = mk(a0.re + c1 * t1.re + c2 * t2.re, a0.im + c1 * t1.im + c2 * t2.im);
    cpx m2 = mk(a0.re + c2 * t1.re + c1 * t2.re, a0.im + c2 * t1.im + c1 * t2.im);
    cpx u1 = mk(s1 * t3.re + s2 * t4.re, s1 * t3.im + s2 * t4.im);
    cpx u2 = mk(s2 * t3.re - s1 * t4.re, s2 * t3.im - s1 * t4.im);
    // x1 = m1 - i u1, x4 = m1 + i u1, x2 = m2 - i u2, x3 = m2 + i u2
    x1 = mk(m1.re + u1.im, m1.im - u1.re);
    x4 = mk(m1.re - u1.im, m1.im + u1.re);
    x2 = mk(m2.re + u2.im, m2.im - u2.re);
    x3 = mk(m2.re - u2.im, m2.im + u2.re);
}

// W25^j = (cos(2 pi j/25), -sin(2 pi j/25)) for j = q*r, q,r in 1..4 (j in {1,2,3,4,6,8,9,12,16})
Q3A_HD cpx w25(int j) {
    switch (j) {
        case 1:  return mk(0.96858316112863108f, -0.24868988716485479f);
        case 2:  return mk(0.87630668004386358f, -0.48175367410171532f);
        case 3:  return mk(0.72896862742141155f, -0.68454710592868873f);
        case 4:  return mk(0.53582679497899666f, -0.84432792550201508f);
        case 6:  return mk(0.06279051952931337f, -0.99802672842827156f);
        case 8:  return mk(-0.42577929156507272f, -0.90482705246601958f);
        case 9:  return mk(-0.63742398974868975f, -0.77051324277578925f);
        case 12: return mk(-0.99211470131447788f, -0.12533323356430426f);
        case 16: return mk(-0.63742398974868975f, 0.77051324277578925f);
        default: return mk(1.0f, 0.0f);
    }
}

// complex 25-point DFT, in place: a[n1] -> a[k1]
Q3A_HD void dft25(cpx (&a)[25]) {
    cpx b[5][5];   // b[q][r]
#pragma unroll
    for (int q = 0; q < 5; ++q)
        dft5(a[q], a[5 + q], a[10 + q], a[15 + q], a[20 + q], b[q][0], b[q][1], b[q][2], b[q][3], b[q][4]);
#pragma unroll
    for (int q = 1; q < 5; ++q)
#pragma unroll
        for (int r = 1; r < 5; ++r) b[q][r] = cmul(b[q][r], w25(q * r));
#pragma unroll
    for (int r = 0; r < 5; ++r)
        dft5(b[0][r], b[1][r], b[2][r], b[3][r], b[4][r], a[r], a[r + 5], a[r + 10], a[r + 15], a[r + 20]);
}

}}  // namespace q3a::fft
