// Host check of helm_b200/csrc/fft_regs.cuh (compiled by tests/test_dft_forms_cpu.py against tests/tools/stub): the
// decimation-in-time FMA-form transform (dft_regs_dit), its twist-carrying variant (dft_regs_dit_twisted) and the
// original DIF form (dft_regs) against a direct O(R^2) DFT in long double, R = 16 and 32, both signs, on integer
// inputs of digit size.  Prints "ok <max abs error>"; fails above 1e-9 (the kernels need < 0.5 after 9 x 512 terms).
#include <cstdio>
#include <cstdlib>
#include "fft_regs.cuh"
using namespace hb;

static double worst = 0.0;
static void upd(double a, long double b) { const double e = std::fabs((double)((long double)a - b)); if (e > worst) worst = e; }

template <int R, int SIGN, int TWA> static void run(unsigned seed) {
    double2 in[R], a[R], b[R], c[R];
    for (int i = 0; i < R; ++i) {
        seed = seed * 1664525u + 1013904223u; const int p = (int)((seed >> 8) % 65) - 32;
        seed = seed * 1664525u + 1013904223u; const int q = (int)((seed >> 8) % 65) - 32;
        in[i] = make_double2(p, q); a[i] = in[i]; b[i] = in[i]; c[i] = in[i];
    }
    dft_regs<R, SIGN>(a);
    dft_regs_dit<R, SIGN>(b);
    const long double PI = 3.141592653589793238462643383279502884L;
    for (int k = 0; k < R; ++k) {
        long double re = 0, im = 0;
        for (int j = 0; j < R; ++j) {
            const long double ang = SIGN * 2 * PI * j * k / R, cs = cosl(ang), sn = sinl(ang);
            re += in[j].x * cs - in[j].y * sn; im += in[j].x * sn + in[j].y * cs;
        }
        upd(a[brev<R>(k)].x, re); upd(a[brev<R>(k)].y, im);
        upd(b[brev<R>(k)].x, re); upd(b[brev<R>(k)].y, im);
    }
    if (SIGN > 0) {       // the forward transforms carry the negacyclic twist e^{2 pi i j TWA / 256}
        dft_regs_dit_twisted<R, 1, TWA>(c);
        for (int k = 0; k < R; ++k) {
            long double re = 0, im = 0;
            for (int j = 0; j < R; ++j) {
                const long double ang = 2 * PI * j * k / R + 2 * PI * j * TWA / 256, cs = cosl(ang), sn = sinl(ang);
                re += in[j].x * cs - in[j].y * sn; im += in[j].x * sn + in[j].y * cs;
            }
            upd(c[brev<R>(k)].x, re); upd(c[brev<R>(k)].y, im);
        }
    }
}

int main() {
    for (unsigned s = 1; s <= 200; ++s) { run<16, 1, 4>(s); run<16, -1, 4>(s); run<32, 1, 2>(s); run<32, -1, 2>(s); }
    printf("ok %.3g\n", worst);
    return worst < 1e-9 ? 0 : 1;
}
