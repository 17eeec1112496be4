#include <math.h>
#include <stdio.h>
#include <stdint.h>
#include <string.h>
#include <stdlib.h>
static inline double trig(double a, int add) {  // add=0: sin, add=1: cos
  const double MAGIC = 6755399441055744.0;
  double q = fma(a, 0x1.45f306dc9c883p-1, MAGIC);
  int64_t qi; memcpy(&qi, &q, 8);
  int n = (int)(uint32_t)qi + add;
  double j = q - MAGIC;
  double r = fma(-j, 0x1.921fb54442d18p+0, a);
  r = fma(-j, 0x1.1a62633145c07p-54, r);
  r = fma(-j, -0x1.f1976b7ed8fbcp-110, r);
  double x2 = r * r;
  int odd = n & 1;
  const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03, S3 = -1.98412698298579493134e-04,
               S4 = 2.75573137070700676789e-06, S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
  const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03, C3 = 2.48015872894767294178e-05,
               C4 = -2.75573143513906633035e-07, C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
  double z = odd ? C6 : S6;
  z = fma(z, x2, odd ? C5 : S5);
  z = fma(z, x2, odd ? C4 : S4);
  z = fma(z, x2, odd ? C3 : S3);
  z = fma(z, x2, odd ? C2 : S2);
  z = fma(z, x2, odd ? C1 : S1);
  double w = fma(z, x2, odd ? -0.5 : 0.0);
  double res = fma(w, odd ? x2 : r, odd ? 1.0 : r);
  return (n & 2) ? -res : res;
}
static inline double div_c(double a, double c, double rc) { double q0 = a * rc; double r = fma(-q0, c, a); return fma(r, rc, q0); }
static double ulpdiff(double a, double b) { int64_t x, y; memcpy(&x,&a,8); memcpy(&y,&b,8); if (x<0) x = INT64_MIN - x; if (y<0) y = INT64_MIN - y; return (double)llabs(x-y); }
int main() {
  srand48(1);
  long N = 20000000, mism_s = 0, mism_c = 0; double maxs = 0, maxc = 0;
  for (long i = 0; i < N; ++i) {
    double a = (drand48() * 2 - 1) * (i % 3 == 0 ? 1.6 : (i % 3 == 1 ? 10.0 : 1000.0));
    double s = trig(a, 0), c = trig(a, 1), S = sin(a), C = cos(a);
    double ds = ulpdiff(s, S), dc = ulpdiff(c, C);
    if (ds > 0) mism_s++; if (dc > 0) mism_c++;
    if (ds > maxs) maxs = ds; if (dc > maxc) maxc = dc;
  }
  printf("sin mismatch %.4f%% max ulp %g; cos mismatch %.4f%% max ulp %g\n", 100.0*mism_s/N, maxs, 100.0*mism_c/N, maxc);
  // division by eps
  const double eps = 1.e-3, rc = 1.0/eps; long bad = 0;
  for (long i = 0; i < N; ++i) {
    double a = (drand48()*2-1) * pow(10.0, -8 + 10*drand48());
    if (div_c(a, eps, rc) != a/eps) bad++;
    double b = a/eps; if (div_c(b, eps, rc) != b/eps) bad++;
  }
  printf("div mismatches %ld of %ld\n", bad, 2*N);
  return 0;
}
